"""GPU parity at BASELINE.json's FULL sizes (c2: 4096 envs x 3000 steps, c5:
65 536 envs, 100-step fused launches) through properties that do not need the
oracle to run 10^9 vehicle-steps: invariance to sharding and to fusing, envs
with equal ring length being bit-identical replicas (noise 0), vehicle order
on the ring, conservation of the step counters — plus the 3000-step fp32
aggregates of north_star (mean speed, stop-and-go wave amplitude, reward within
1 % of the fp64 oracle, noise 0) on the hot kernel."""
import numpy as np
import pytest

from oracle import sim as OS
from tests import scenarios as S

pytestmark = pytest.mark.gpu
KW = dict(env_kind=OS.ENV_WAVE_ATTENUATION_PO, length_lo=220., length_hi=270., **S.C2)


def ring_order_ok(pos, circ):
    """No overtaking: going round the ring from vehicle 0, positions increase
    (mod C) and the gaps sum to exactly one lap."""
    gaps = (np.roll(pos, -1, axis=0) - pos) % circ[None, :]
    return np.abs(gaps.sum(0) - circ) < 1e-2 * np.maximum(circ, 1)


def test_c5_full_size_properties():
    import torch
    E, K = 65536, 100
    eng = S.make_engine(S.c2_groups(), E, 'f32', seed=1234, **KW)
    eng.reset()
    gen = torch.Generator(device='cuda'); gen.manual_seed(7)
    act = torch.rand(3 * K, 1, E, device='cuda', generator=gen) * 2 - 1
    rew = torch.empty(3 * K, E, device='cuda')
    done = torch.empty(3 * K, E, device='cuda', dtype=torch.uint8)
    eng.stats(reset=True)
    for c in range(3):
        eng.step(act[c * K:(c + 1) * K], k=K, reward_seq=rew[c * K:(c + 1) * K],
                 done_seq=done[c * K:(c + 1) * K])
    st = eng.stats()
    d = done.cpu().numpy().astype(bool)
    # counters: every env advanced until (and including) the step it finished on
    alive_steps = (~d).sum(0) + d.any(0)
    assert st['env_steps'] == alive_steps.sum()
    assert st['vehicle_steps'] == 22 * st['env_steps']
    assert st['crashes'] == d[-1].sum() == eng.crash.cpu().numpy().sum()
    assert (np.diff(d.astype(np.int8), axis=0) >= 0).all()          # done is absorbing
    r = rew.cpu().numpy()
    assert (r[d & np.roll(d, 1, axis=0)] == 0).all()                 # frozen envs earn nothing
    pos, circ = eng.pos.cpu().numpy(), eng.length.cpu().numpy()
    live = ~d[-1]
    assert ring_order_ok(pos, circ)[live].all()
    assert (eng.time_counter.cpu().numpy() == 750 + alive_steps).all()
    # sharding: env 41 234 of the big batch == env 0 of an engine based at that global id
    for e0 in (0, 41234, 65535):
        small = S.make_engine(S.c2_groups(), 3 if e0 < 65534 else 1, 'f32', seed=1234,
                              env_id_base=e0, **KW)
        small.reset()
        n = small.n_envs
        for c in range(3):
            small.step(act[c * K:(c + 1) * K, :, e0:e0 + n].contiguous(), k=K)
        assert torch.equal(small.pos, eng.pos[:, e0:e0 + n]), e0
        assert torch.equal(small.vel, eng.vel[:, e0:e0 + n]), e0


def test_c2_full_size_replicas_and_fusing():
    """4096 envs x 3000 steps, noise 0, the SAME action sequence in every env: envs
    that drew the same ring length are bit-identical replicas, whatever warp they
    sit in; one fused 3000-step launch == 30 launches of 100."""
    import torch
    E, K = 4096, 3000
    a = S.make_engine(S.c2_groups(), E, 'f32', seed=99, **KW)
    b = S.make_engine(S.c2_groups(), E, 'f32', seed=99, **KW)
    a.reset(); b.reset()
    gen = torch.Generator(device='cuda'); gen.manual_seed(3)
    seq = torch.rand(K, 1, 1, device='cuda', generator=gen) * 2 - 1
    act = seq.expand(K, 1, E).contiguous()
    rew_a = torch.empty(K, E, device='cuda')
    rew_b = torch.empty(K, E, device='cuda')
    a.step(act, k=K, reward_seq=rew_a)
    for c in range(0, K, 100):
        b.step(act[c:c + 100], k=100, reward_seq=rew_b[c:c + 100])
    assert torch.equal(rew_a, rew_b) and torch.equal(a.pos, b.pos) and torch.equal(a.vel, b.vel)
    L = np.round(a.length.cpu().numpy() - 0.4).astype(int)
    assert L.min() >= 220 and L.max() <= 270 and len(np.unique(L)) == 51
    pos, vel, ret = a.pos.cpu().numpy(), a.vel.cpu().numpy(), rew_a.sum(0).cpu().numpy()
    for length in np.unique(L):
        idx = np.nonzero(L == length)[0]
        assert (pos[:, idx] == pos[:, idx[:1]]).all(), length
        assert (vel[:, idx] == vel[:, idx[:1]]).all(), length
        assert (ret[idx] == ret[idx[0]]).all(), length
    assert (a.time_counter.cpu().numpy() <= 3750).all()


def test_c2_3000_step_fp32_aggregates_vs_oracle():
    """north_star: 3000-step aggregate statistics within 1 % (noise 0)."""
    import torch
    K, E, length = 3000, 8, 250
    eng = S.make_engine(S.c2_groups(), E, 'f32', **KW)
    eng.reset(lengths=[length] * E)
    ora = S.make_oracle(S.c2_groups(), length, env_kind=OS.ENV_WAVE_ATTENUATION_PO, bunching=50.0,
                        warmup_steps=750, horizon=3000, act_lo=-1, act_hi=1, obs_max_length=270.0)
    rng = np.random.default_rng(11)
    # a braking-biased bounded random walk keeps the RL car from ramming its leader for 3000 steps
    acts = np.clip(0.3 * np.sin(np.arange(K) / 40.0) + rng.normal(0, 0.3, K) - 0.05, -1, 1)
    v_eng, r_eng, v_ora, r_ora = [], [], [], []
    rew = torch.empty(10, E, device='cuda')
    for c in range(0, K, 10):
        a = torch.tensor(np.repeat(acts[c:c + 10, None, None], E, 2), device='cuda',
                         dtype=torch.float32)
        eng.step(a, k=10, reward_seq=rew)
        v_eng.append(eng.vel.cpu().numpy()[:, 3].copy()); r_eng.append(rew[:, 3].sum().item())
        for k in range(10):
            _, r, dn = ora.step([acts[c + k]])
            r_ora.append(r)
        v_ora.append(np.array(ora.vel))
        assert not ora.crash, c
    assert dn and eng.done.cpu().numpy().all()            # horizon reached on the last step ...
    assert not eng.crash.cpu().numpy().any()              # ... without a collision
    v_eng, v_ora = np.array(v_eng), np.array(v_ora)
    m_e, m_o = v_eng[-100:].mean(), v_ora[-100:].mean()
    amp_e = (v_eng[-100:].max(1) - v_eng[-100:].min(1)).mean()
    amp_o = (v_ora[-100:].max(1) - v_ora[-100:].min(1)).mean()
    assert abs(m_e - m_o) <= 0.01 * abs(m_o), (m_e, m_o)
    assert abs(amp_e - amp_o) <= 0.01 * max(abs(amp_o), 0.05), (amp_e, amp_o)
    assert abs(sum(r_eng) - sum(r_ora)) <= 0.01 * abs(sum(r_ora)), (sum(r_eng), sum(r_ora))
