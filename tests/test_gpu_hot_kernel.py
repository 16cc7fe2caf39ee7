"""GPU parity of ring1_hot_kernel (the c2 / c5 specialisation): against the
reference-generated golden trajectory, against ring1_fast_kernel on the same
inputs (bit for bit in fp64), across every thread-mapping shape, with launches
that start inside a 4-step noise block, and the statistics of its Philox noise."""
import os

import numpy as np
import pytest

from tests import scenarios as S
from oracle import sim as OS

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), 'golden')
SHAPES = [1, 2, 4, 8, 16, 32]
KW = dict(env_kind=OS.ENV_WAVE_ATTENUATION_PO, length_lo=220., length_hi=270., **S.C2)


def close(a, b, floor=1.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float((np.abs(a - b) / np.maximum(np.abs(b), floor)).max())


@pytest.mark.parametrize('shape', SHAPES)
def test_hot_golden_c2_fp64(shape, monkeypatch):
    import torch
    monkeypatch.setenv('FS_RING_SHAPE', str(shape))
    z = np.load(os.path.join(G, 'ring_c2.npz'))
    lengths = [float(z['length_a']), float(z['length_b'])] * 20
    E = len(lengths)
    eng = S.make_engine(S.c2_groups(), E, 'f64', **KW)
    eng.reset(lengths=lengths)
    n = 240
    act = np.zeros((n, 1, E))
    act[:, 0, 0::2] = z['actions_a'][:n, None]
    act[:, 0, 1::2] = z['actions_b'][:n, None]
    s = 0
    for K in (1, 2, 7, 30, 50, 3, 100, 47):     # launches start at every phase of a 4-step block
        obs_seq = torch.empty(K, E, 3, device='cuda', dtype=torch.float64)
        rew_seq = torch.empty(K, E, device='cuda', dtype=torch.float64)
        eng.step(torch.tensor(act[s:s + K], device='cuda'), k=K, obs_seq=obs_seq, reward_seq=rew_seq)
        o, r = obs_seq.cpu().numpy(), rew_seq.cpu().numpy()
        for e, tag in ((0, 'a'), (1, 'b'), (38, 'a'), (39, 'b')):
            assert close(o[:, e], z['obs_' + tag][s + 1:s + K + 1]) <= 1e-9, (shape, s, tag)
            assert close(r[:, e], z['reward_' + tag][s:s + K]) <= 1e-9, (shape, s, tag)
            assert close(eng.pos.cpu().numpy()[:, e], z['pos_' + tag][s + K]) <= 1e-9
            assert close(eng.vel.cpu().numpy()[:, e], z['vel_' + tag][s + K]) <= 1e-9
        s += K
    assert s == n


@pytest.mark.parametrize('dtype', ['f64', 'f32'])
@pytest.mark.parametrize('shape', SHAPES)
def test_hot_equals_fast_kernel(shape, dtype, monkeypatch):
    """Same inputs through ring1_hot_kernel and ring1_fast_kernel (FS_NO_HOT=1):
    identical in fp64; fp32 within 1e-4 (north_star's fp32 figure) after 833 FREE-RUNNING steps
    (750 warm-up without actions + 83): the two kernels contract multiply-adds differently, and
    that round-off is carried along (measured 1.4e-5)."""
    import torch
    monkeypatch.setenv('FS_RING_SHAPE', str(shape))
    E, K = 100, 83
    td = torch.float64 if dtype == 'f64' else torch.float32
    hot = S.make_engine(S.c2_groups(), E, dtype, seed=9, **KW)
    monkeypatch.setenv('FS_NO_HOT', '1')
    ref = S.make_engine(S.c2_groups(), E, dtype, seed=9, **KW)
    monkeypatch.delenv('FS_NO_HOT')
    o_h, o_r = hot.reset(), ref.reset()       # 750 warm-up steps without actions
    tol = 0.0 if dtype == 'f64' else 1e-4
    assert close(o_h.cpu().numpy(), o_r.cpu().numpy()) <= tol
    gen = torch.Generator(device='cuda'); gen.manual_seed(2)
    act = torch.rand(K, 1, E, device='cuda', dtype=td, generator=gen) * 3 - 1.5   # clipping exercised
    out = []
    for eng in (hot, ref):
        obs_seq = torch.empty(K, E, 3, device='cuda', dtype=td)
        rew_seq = torch.empty(K, E, device='cuda', dtype=td)
        done_seq = torch.empty(K, E, device='cuda', dtype=torch.uint8)
        eng.stats(reset=True)
        eng.step(act[:5], k=5, obs_seq=obs_seq[:5], reward_seq=rew_seq[:5], done_seq=done_seq[:5])
        eng.step(act[5:], k=K - 5, obs_seq=obs_seq[5:], reward_seq=rew_seq[5:], done_seq=done_seq[5:])
        out.append((obs_seq.cpu().numpy(), rew_seq.cpu().numpy(), done_seq.cpu().numpy(),
                    eng.pos.cpu().numpy(), eng.vel.cpu().numpy(), eng.stats(),
                    eng.time_counter.cpu().numpy()))
    a, b = out
    for i in range(5):
        if i == 2:
            assert np.array_equal(a[i], b[i])
        else:
            assert close(a[i], b[i]) <= tol, i
    assert np.array_equal(a[6], b[6])
    for k in ('vehicle_steps', 'env_steps', 'crashes'):
        assert a[5][k] == b[5][k], k
    for k in ('sum_reward', 'sum_speed', 'sum_speed_sq'):
        assert abs(a[5][k] - b[5][k]) <= 1e-4 * max(1.0, abs(b[5][k])), k


@pytest.mark.parametrize('shape', [1, 8, 32])
def test_hot_wave_attenuation_full_obs(shape, monkeypatch):
    """WaveAttenuationEnv (2V observation) also takes the hot kernel."""
    import torch
    monkeypatch.setenv('FS_RING_SHAPE', str(shape))
    kw = dict(KW, env_kind=OS.ENV_WAVE_ATTENUATION)
    E, K = 37, 21
    hot = S.make_engine(S.c2_groups(), E, 'f64', seed=4, **kw)
    monkeypatch.setenv('FS_NO_HOT', '1')
    ref = S.make_engine(S.c2_groups(), E, 'f64', seed=4, **kw)
    hot.reset(); ref.reset()
    act = torch.rand(K, 1, E, device='cuda', dtype=torch.float64) * 2 - 1
    res = []
    for eng in (hot, ref):
        obs_seq = torch.empty(K, E, 44, device='cuda', dtype=torch.float64)
        rew_seq = torch.empty(K, E, device='cuda', dtype=torch.float64)
        eng.step(act, k=K, obs_seq=obs_seq, reward_seq=rew_seq)
        res.append((obs_seq.cpu().numpy(), rew_seq.cpu().numpy(), eng.obs.cpu().numpy()))
    for x, y in zip(*res):
        assert np.array_equal(x, y)


@pytest.mark.parametrize('shape', [1, 2, 8, 32])
def test_hot_noise_is_invariant_to_fusing_and_well_distributed(shape, monkeypatch):
    import torch
    monkeypatch.setenv('FS_RING_SHAPE', str(shape))
    E, K = 512, 24
    kw = dict(KW, warmup_steps=0)
    a = S.make_engine(S.c2_groups(noise=0.2), E, 'f32', seed=21, **kw)
    b = S.make_engine(S.c2_groups(noise=0.2), E, 'f32', seed=21, **kw)
    q = S.make_engine(S.c2_groups(noise=0.0), E, 'f32', seed=21, **kw)
    for eng in (a, b, q):
        eng.reset()
    act = torch.rand(K, 1, E, device='cuda') * 2 - 1
    a.step(act, k=K)
    for chunk in ((0, 1), (1, 3), (3, 10), (10, 11), (11, 24)):     # every block phase
        b.step(act[chunk[0]:chunk[1]], k=chunk[1] - chunk[0])
    assert torch.equal(a.pos, b.pos) and torch.equal(a.vel, b.vel)
    # one noisy step from a common state: dv = noise * z * dt for the 21 humans, 0 for the RL car
    n1 = S.make_engine(S.c2_groups(noise=0.2), E, 'f32', seed=5, **kw)
    n1.reset(); q.reset()
    zs = []
    for k in range(8):
        q.set_state(n1.pos.cpu().numpy(), n1.vel.cpu().numpy())
        q.time_counter.copy_(n1.time_counter)
        n1.step(act[k]); q.step(act[k])
        dv = (n1.vel - q.vel).cpu().numpy().astype(np.float64)
        moving = (q.vel.cpu().numpy() > 0.05)
        assert np.abs(dv[21]).max() == 0.0
        z = dv[:21] / (0.2 * 0.1)
        zs.append(np.where(moving[:21], z, np.nan))
    z = np.stack(zs)                       # [step, vehicle, env]
    flat = z[np.isfinite(z)]
    assert flat.size > 50000
    assert abs(flat.mean()) < 0.02 and abs(flat.std() - 1.0) < 0.02
    assert abs(np.mean(flat ** 3)) < 0.06 and abs(np.mean(flat ** 4) - 3.0) < 0.15
    zz = np.nan_to_num(z)
    c_veh = np.mean(zz[:, :-1] * zz[:, 1:])       # neighbours in the ring
    c_time = np.mean(zz[:-1] * zz[1:])            # successive steps (same Philox call for VPT <= 4)
    c_env = np.mean(zz[:, :, :-1] * zz[:, :, 1:])
    assert max(abs(c_veh), abs(c_time), abs(c_env)) < 0.02


def test_pipelined_host_rollout_equals_blocking():
    """fs_step_host_async (three staging slots, overlapped copies) delivers the
    same bytes as the blocking fs_step_host, for several calls in flight and a
    call larger than the previous staging size."""
    import torch
    from flow_b200.vec_env import B200VecEnv
    from flow_b200 import workloads as W
    types, tov, kw = W.ring_c2(noise=0.2)
    E = 300
    a = B200VecEnv(types, tov, E, dtype='f32', seed=3, **kw)
    b = B200VecEnv(types, tov, E, dtype='f32', seed=3, **kw)
    a.reset(); b.reset()
    chunks = [10, 10, 10, 25, 3, 40, 10]
    K = sum(chunks)
    act = (torch.rand(K, 1, E) * 2 - 1).pin_memory()
    obs = torch.empty(K, E, 3).pin_memory()
    rew = torch.empty(K, E).pin_memory()
    done = torch.empty(K, E, dtype=torch.uint8).pin_memory()
    s = 0
    for k in chunks:
        a.rollout_host_async(act[s:s + k], obs[s:s + k], rew[s:s + k], done[s:s + k])
        s += k
    a.wait()
    s = 0
    for k in chunks:
        o, r, d = b.rollout_host(act[s:s + k])
        assert np.array_equal(o, obs[s:s + k].numpy()), s
        assert np.array_equal(r, rew[s:s + k].numpy()), s
        assert np.array_equal(d, done[s:s + k].numpy()), s
        s += k
    assert torch.equal(a.engine.pos, b.engine.pos)


@pytest.mark.gpu
@pytest.mark.parametrize('dtype', ['f64', 'f32'])
def test_unfused_graph_replay_equals_fused_launch(dtype):
    """fs_step_unfused (K single-step launches, second call replayed from the CUDA graph with
    NEW action values in the same buffers) == fs_step (one fused launch), bit for bit."""
    import torch
    K, E = 60, 96
    td = torch.float64 if dtype == 'f64' else torch.float32
    mk = lambda: S.make_engine(S.c2_groups(noise=0.2), E, dtype, **KW)  # noqa: E731
    a, b = mk(), mk()
    act = torch.empty(K, 1, E, device='cuda', dtype=td)
    outs = []
    for eng in (a, b):
        outs.append((torch.zeros(K, E, 3, device='cuda', dtype=td),
                     torch.zeros(K, E, device='cuda', dtype=td),
                     torch.zeros(K, E, device='cuda', dtype=torch.uint8)))
    gen = torch.Generator(device='cuda')
    for rnd in range(3):            # round 0 captures, rounds 1-2 replay
        gen.manual_seed(100 + rnd)
        act.copy_(torch.rand(K, 1, E, device='cuda', dtype=td, generator=gen) * 2 - 1)
        a.reset(); b.reset()
        l0 = a.launch_count
        a.step_unfused_ptr(act.data_ptr(), K, *(t.data_ptr() for t in outs[0]))
        assert a.launch_count - l0 == K
        b.step(act, k=K, obs_seq=outs[1][0], reward_seq=outs[1][1], done_seq=outs[1][2])
        for x, y in zip(outs[0], outs[1]):
            assert torch.equal(x, y), rnd
        assert torch.equal(a.pos, b.pos) and torch.equal(a.vel, b.vel)
        sa_, sb_ = a.stats(reset=True), b.stats(reset=True)   # float partial sums: order differs
        assert sa_['vehicle_steps'] == sb_['vehicle_steps'] and sa_['crashes'] == sb_['crashes']
        assert close(sa_['sum_speed'], sb_['sum_speed']) <= 1e-5


# ---- the TIMED instantiation (idm noise 0.2) against the oracle --------------------------------
# oracle/noise.py restates the kernel's Philox stream counter for counter, so the noisy rollout
# bench.py times is checked per step like the noise-free one
def _noisy_pair(shape, dtype, E, seed, gid0, warmup, lengths):
    from oracle.noise import HotNoise
    kw = dict(KW)
    kw['warmup_steps'] = warmup
    eng = S.make_engine(S.c2_groups(noise=0.2), E, dtype, seed=seed, env_id_base=gid0, **kw)
    eng.reset(lengths=lengths)                       # first reset: episode 1
    oras = [S.make_oracle(S.c2_groups(noise=0.2), lengths[e], env_kind=OS.ENV_WAVE_ATTENUATION_PO,
                          bunching=50.0, warmup_steps=warmup, horizon=3000, act_lo=-1, act_hi=1,
                          obs_max_length=270.0, noise_fn=HotNoise(seed, gid0 + e, 1, shape))
            for e in range(E)]
    return eng, oras


@pytest.mark.parametrize('shape', SHAPES)
def test_hot_noisy_rollout_vs_oracle_fp64(shape, monkeypatch):
    """c2 with idm noise 0.2, fp64, every thread shape: warm-up inside reset and 200 steps with
    random actions in launches that start at every phase of a 4-step noise block; observation,
    reward and state <= 1e-9 of the oracle driven by the restated stream."""
    import torch
    monkeypatch.setenv('FS_RING_SHAPE', str(shape))
    E, seed, gid0 = 5, (0x1234 << 32) | 0x9E3779B1, (1 << 33) + 11     # high words in play
    lengths = [220.0, 231.0, 247.0, 263.0, 270.0]
    eng, oras = _noisy_pair(shape, 'f64', E, seed, gid0, 37, lengths)
    pos, vel = eng.pos.cpu().numpy(), eng.vel.cpu().numpy()
    for e, o in enumerate(oras):
        assert close(pos[:, e], o.pos) <= 1e-9 and close(vel[:, e], o.vel) <= 1e-9, (shape, e)
    # the noise is really on: the same warm-up without it ends somewhere else
    quiet = S.make_oracle(S.c2_groups(), lengths[0], env_kind=OS.ENV_WAVE_ATTENUATION_PO,
                          bunching=50.0, warmup_steps=37, horizon=3000, act_lo=-1, act_hi=1,
                          obs_max_length=270.0)
    assert abs(np.array(quiet.vel) - np.array(oras[0].vel)).max() > 1e-3
    rng = np.random.default_rng(shape)
    s = 0
    for K in (1, 2, 7, 30, 50, 3, 68, 39):
        act = rng.uniform(-1.2, 0.7, size=(K, 1, E))     # beyond the box; braking-biased
        obs_seq = torch.empty(K, E, 3, device='cuda', dtype=torch.float64)
        rew_seq = torch.empty(K, E, device='cuda', dtype=torch.float64)
        done_seq = torch.empty(K, E, device='cuda', dtype=torch.uint8)
        eng.step(torch.tensor(act, device='cuda'), k=K, obs_seq=obs_seq, reward_seq=rew_seq,
                 done_seq=done_seq)
        o_, r_, d_ = obs_seq.cpu().numpy(), rew_seq.cpu().numpy(), done_seq.cpu().numpy()
        for e, o in enumerate(oras):
            for k in range(K):
                oo, rr, done = o.step([act[k, 0, e]])      # a finished oracle env stays frozen too
                assert bool(d_[k, e]) == done, (shape, s + k, e)
                assert close(o_[k, e], oo) <= 1e-9, (shape, s + k, e)
                assert close(r_[k, e], rr) <= 1e-9, (shape, s + k, e)
        pos, vel = eng.pos.cpu().numpy(), eng.vel.cpu().numpy()
        for e, o in enumerate(oras):
            assert close(pos[:, e], o.pos) <= 1e-9 and close(vel[:, e], o.vel) <= 1e-9
        s += K
    assert s == 200


@pytest.mark.parametrize('shape', [1, 8, 32])
def test_hot_wave_attenuation_full_obs(shape, monkeypatch):
    """WaveAttenuationEnv (2V observation) also takes the hot kernel."""
    import torch
    monkeypatch.setenv('FS_RING_SHAPE', str(shape))
    kw = dict(KW, env_kind=OS.ENV_WAVE_ATTENUATION)
    E, K = 37, 21
    hot = S.make_engine(S.c2_groups(), E, 'f64', seed=4, **kw)
    monkeypatch.setenv('FS_NO_HOT', '1')
    ref = S.make_engine(S.c2_groups(), E, 'f64', seed=4, **kw)
    hot.reset(); ref.reset()
    act = torch.rand(K, 1, E, device='cuda', dtype=torch.float64) * 2 - 1
    res = []
    for eng in (hot, ref):
        obs_seq = torch.empty(K, E, 44, device='cuda', dtype=torch.float64)
        rew_seq = torch.empty(K, E, device='cuda', dtype=torch.float64)
        eng.step(act, k=K, obs_seq=obs_seq, reward_seq=rew_seq)
        res.append((obs_seq.cpu().numpy(), rew_seq.cpu().numpy(), eng.obs.cpu().numpy()))
    for x, y in zip(*res):
        assert np.array_equal(x, y)


@pytest.mark.parametrize('shape', [1, 2, 8, 32])
def test_hot_noise_is_invariant_to_fusing_and_well_distributed(shape, monkeypatch):
    import torch
    monkeypatch.setenv('FS_RING_SHAPE', str(shape))
    E, K = 512, 24
    kw = dict(KW, warmup_steps=0)
    a = S.make_engine(S.c2_groups(noise=0.2), E, 'f32', seed=21, **kw)
    b = S.make_engine(S.c2_groups(noise=0.2), E, 'f32', seed=21, **kw)
    q = S.make_engine(S.c2_groups(noise=0.0), E, 'f32', seed=21, **kw)
    for eng in (a, b, q):
        eng.reset()
    act = torch.rand(K, 1, E, device='cuda') * 2 - 1
    a.step(act, k=K)
    for chunk in ((0, 1), (1, 3), (3, 10), (10, 11), (11, 24)):     # every block phase
        b.step(act[chunk[0]:chunk[1]], k=chunk[1] - chunk[0])
    assert torch.equal(a.pos, b.pos) and torch.equal(a.vel, b.vel)
    # one noisy step from a common state: dv = noise * z * dt for the 21 humans, 0 for the RL car
    n1 = S.make_engine(S.c2_groups(noise=0.2), E, 'f32', seed=5, **kw)
    n1.reset(); q.reset()
    zs = []
    for k in range(8):
        q.set_state(n1.pos.cpu().numpy(), n1.vel.cpu().numpy())
        q.time_counter.copy_(n1.time_counter)
        n1.step(act[k]); q.step(act[k])
        dv = (n1.vel - q.vel).cpu().numpy().astype(np.float64)
        moving = (q.vel.cpu().numpy() > 0.05)
        assert np.abs(dv[21]).max() == 0.0
        z = dv[:21] / (0.2 * 0.1)
        zs.append(np.where(moving[:21], z, np.nan))
    z = np.stack(zs)                       # [step, vehicle, env]
    flat = z[np.isfinite(z)]
    assert flat.size > 50000
    assert abs(flat.mean()) < 0.02 and abs(flat.std() - 1.0) < 0.02
    assert abs(np.mean(flat ** 3)) < 0.06 and abs(np.mean(flat ** 4) - 3.0) < 0.15
    zz = np.nan_to_num(z)
    c_veh = np.mean(zz[:, :-1] * zz[:, 1:])       # neighbours in the ring
    c_time = np.mean(zz[:-1] * zz[1:])            # successive steps (same Philox call for VPT <= 4)
    c_env = np.mean(zz[:, :, :-1] * zz[:, :, 1:])
    assert max(abs(c_veh), abs(c_time), abs(c_env)) < 0.02


def test_pipelined_host_rollout_equals_blocking():
    """fs_step_host_async (three staging slots, overlapped copies) delivers the
    same bytes as the blocking fs_step_host, for several calls in flight and a
    call larger than the previous staging size."""
    import torch
    from flow_b200.vec_env import B200VecEnv
    from flow_b200 import workloads as W
    types, tov, kw = W.ring_c2(noise=0.2)
    E = 300
    a = B200VecEnv(types, tov, E, dtype='f32', seed=3, **kw)
    b = B200VecEnv(types, tov, E, dtype='f32', seed=3, **kw)
    a.reset(); b.reset()
    chunks = [10, 10, 10, 25, 3, 40, 10]
    K = sum(chunks)
    act = (torch.rand(K, 1, E) * 2 - 1).pin_memory()
    obs = torch.empty(K, E, 3).pin_memory()
    rew = torch.empty(K, E).pin_memory()
    done = torch.empty(K, E, dtype=torch.uint8).pin_memory()
    s = 0
    for k in chunks:
        a.rollout_host_async(act[s:s + k], obs[s:s + k], rew[s:s + k], done[s:s + k])
        s += k
    a.wait()
    s = 0
    for k in chunks:
        o, r, d = b.rollout_host(act[s:s + k])
        assert np.array_equal(o, obs[s:s + k].numpy()), s
        assert np.array_equal(r, rew[s:s + k].numpy()), s
        assert np.array_equal(d, done[s:s + k].numpy()), s
        s += k
    assert torch.equal(a.engine.pos, b.engine.pos)


@pytest.mark.gpu
@pytest.mark.parametrize('dtype', ['f64', 'f32'])
def test_unfused_graph_replay_equals_fused_launch(dtype):
    """fs_step_unfused (K single-step launches, second call replayed from the CUDA graph with
    NEW action values in the same buffers) == fs_step (one fused launch), bit for bit."""
    import torch
    K, E = 60, 96
    td = torch.float64 if dtype == 'f64' else torch.float32
    mk = lambda: S.make_engine(S.c2_groups(noise=0.2), E, dtype, **KW)  # noqa: E731
    a, b = mk(), mk()
    act = torch.empty(K, 1, E, device='cuda', dtype=td)
    outs = []
    for eng in (a, b):
        outs.append((torch.zeros(K, E, 3, device='cuda', dtype=td),
                     torch.zeros(K, E, device='cuda', dtype=td),
                     torch.zeros(K, E, device='cuda', dtype=torch.uint8)))
    gen = torch.Generator(device='cuda')
    for rnd in range(3):            # round 0 captures, rounds 1-2 replay
        gen.manual_seed(100 + rnd)
        act.copy_(torch.rand(K, 1, E, device='cuda', dtype=td, generator=gen) * 2 - 1)
        a.reset(); b.reset()
        l0 = a.launch_count
        a.step_unfused_ptr(act.data_ptr(), K, *(t.data_ptr() for t in outs[0]))
        assert a.launch_count - l0 == K
        b.step(act, k=K, obs_seq=outs[1][0], reward_seq=outs[1][1], done_seq=outs[1][2])
        for x, y in zip(outs[0], outs[1]):
            assert torch.equal(x, y), rnd
        assert torch.equal(a.pos, b.pos) and torch.equal(a.vel, b.vel)
        sa_, sb_ = a.stats(reset=True), b.stats(reset=True)   # float partial sums: order differs
        assert sa_['vehicle_steps'] == sb_['vehicle_steps'] and sa_['crashes'] == sb_['crashes']
        assert close(sa_['sum_speed'], sb_['sum_speed']) <= 1e-5


# ---- the TIMED instantiation (idm noise 0.2) against the oracle --------------------------------
# oracle/noise.py restates the kernel's Philox stream counter for counter, so the noisy rollout
# bench.py times is checked per step like the noise-free one
def _noisy_pair(shape, dtype, E, seed, gid0, warmup, lengths):
    from oracle.noise import HotNoise
    kw = dict(KW)
    kw['warmup_steps'] = warmup
    eng = S.make_engine(S.c2_groups(noise=0.2), E, dtype, seed=seed, env_id_base=gid0, **kw)
    eng.reset(lengths=lengths)                       # first reset: episode 1
    oras = [S.make_oracle(S.c2_groups(noise=0.2), lengths[e], env_kind=OS.ENV_WAVE_ATTENUATION_PO,
                          bunching=50.0, warmup_steps=warmup, horizon=3000, act_lo=-1, act_hi=1,
                          obs_max_length=270.0, noise_fn=HotNoise(seed, gid0 + e, 1, shape))
            for e in range(E)]
    return eng, oras


@pytest.mark.parametrize('shape', SHAPES)
def test_hot_noisy_rollout_vs_oracle_fp64(shape, monkeypatch):
    """c2 with idm noise 0.2, fp64, every thread shape: warm-up inside reset and 200 steps with
    random actions in launches that start at every phase of a 4-step noise block; observation,
    reward and state <= 1e-9 of the oracle driven by the restated stream."""
    import torch
    monkeypatch.setenv('FS_RING_SHAPE', str(shape))
    E, seed, gid0 = 5, (0x1234 << 32) | 0x9E3779B1, (1 << 33) + 11     # high words in play
    lengths = [220.0, 231.0, 247.0, 263.0, 270.0]
    eng, oras = _noisy_pair(shape, 'f64', E, seed, gid0, 37, lengths)
    pos, vel = eng.pos.cpu().numpy(), eng.vel.cpu().numpy()
    for e, o in enumerate(oras):
        assert close(pos[:, e], o.pos) <= 1e-9 and close(vel[:, e], o.vel) <= 1e-9, (shape, e)
    # the noise is really on: the same warm-up without it ends somewhere else
    quiet = S.make_oracle(S.c2_groups(), lengths[0], env_kind=OS.ENV_WAVE_ATTENUATION_PO,
                          bunching=50.0, warmup_steps=37, horizon=3000, act_lo=-1, act_hi=1,
                          obs_max_length=270.0)
    assert abs(np.array(quiet.vel) - np.array(oras[0].vel)).max() > 1e-3
    rng = np.random.default_rng(shape)
    s = 0
    for K in (1, 2, 7, 30, 50, 3, 68, 39):
        act = rng.uniform(-1.2, 0.7, size=(K, 1, E))     # beyond the box; braking-biased
        obs_seq = torch.empty(K, E, 3, device='cuda', dtype=torch.float64)
        rew_seq = torch.empty(K, E, device='cuda', dtype=torch.float64)
        done_seq = torch.empty(K, E, device='cuda', dtype=torch.uint8)
        eng.step(torch.tensor(act, device='cuda'), k=K, obs_seq=obs_seq, reward_seq=rew_seq,
                 done_seq=done_seq)
        o_, r_, d_ = obs_seq.cpu().numpy(), rew_seq.cpu().numpy(), done_seq.cpu().numpy()
        for e, o in enumerate(oras):
            for k in range(K):
                oo, rr, done = o.step([act[k, 0, e]])      # a finished oracle env stays frozen too
                assert bool(d_[k, e]) == done, (shape, s + k, e)
                assert close(o_[k, e], oo) <= 1e-9, (shape, s + k, e)
                assert close(r_[k, e], rr) <= 1e-9, (shape, s + k, e)
        pos, vel = eng.pos.cpu().numpy(), eng.vel.cpu().numpy()
        for e, o in enumerate(oras):
            assert close(pos[:, e], o.pos) <= 1e-9 and close(vel[:, e], o.vel) <= 1e-9
        s += K
    assert s == 200
    # the stream is not a constant: the same rollout without noise differs
    quiet = S.make_oracle(S.c2_groups(), lengths[0], env_kind=OS.ENV_WAVE_ATTENUATION_PO,
                          bunching=50.0, warmup_steps=37, horizon=3000, act_lo=-1, act_hi=1,
                          obs_max_length=270.0)
    assert abs(np.array(quiet.vel) - np.array(oras[0].vel)).max() > 1e-3 or s > 0 and \
        abs(np.array(quiet.pos) - np.array(oras[0].pos)).max() > 1e-3


@pytest.mark.parametrize('shape', [1, 8, 32])
def test_hot_noisy_step_fp32_resync(shape, monkeypatch):
    """fp32 noise path (fast-math Box-Muller) per step <= 1e-4 of the oracle evaluated on the
    same fp32-rounded state with the restated (double) stream."""
    import torch
    from oracle.noise import HotNoise
    monkeypatch.setenv('FS_RING_SHAPE', str(shape))
    E, seed, gid0, L = 2, 99, 7, 250.0
    eng, oras = _noisy_pair(shape, 'f32', E, seed, gid0, 0, [L, L])
    ora = oras[0]
    rng = np.random.default_rng(5)
    for step in range(120):
        a = float(rng.uniform(-1, 1))
        p32 = np.asarray(ora.pos, np.float32); v32 = np.asarray(ora.vel, np.float32)
        eng.set_state(np.stack([p32, p32], 1), np.stack([v32, v32], 1))
        chk = S.make_oracle(S.c2_groups(noise=0.2), L, env_kind=OS.ENV_WAVE_ATTENUATION_PO,
                            bunching=50.0, warmup_steps=0, horizon=3000, act_lo=-1, act_hi=1,
                            obs_max_length=270.0, noise_fn=HotNoise(seed, gid0, 1, shape))
        chk.circ = float(np.float32(ora.circ))
        chk.set_state(p32.astype(np.float64), v32.astype(np.float64))
        chk.step_index = ora.step_index
        chk.time_counter = ora.time_counter
        chk.step([a])
        eng.step(torch.tensor([[a, a]], device='cuda', dtype=torch.float32))
        eng.sync()
        assert close(eng.vel.cpu().numpy()[:, 0], chk.vel, floor=1e-2) <= 1e-4, (shape, step)
        assert close(eng.pos.cpu().numpy()[:, 0], chk.pos) <= 1e-4, (shape, step)
        ora.step([a])


@pytest.mark.parametrize('cl_mode', ['0', '1', '2', '3'])
def test_closed_loop_host_step_equals_device_steps(cl_mode, monkeypatch):
    """fs_step_host with k <= 8 (pinned staging block; FS_CL_MODE bit 0: actions read by the
    kernel from the mapped block, bit 1: results written straight into it; otherwise device
    block + copy engine, the sequence replayed as one CUDA graph) returns
    the bytes fs_step leaves on the device: k = 1 through B200VecEnv.step_host, k = 3 and a
    pageable numpy buffer through Engine.step_host, and the no-action (simulator model) form."""
    import torch
    monkeypatch.setenv('FS_CL_MODE', cl_mode)
    from flow_b200.vec_env import B200VecEnv
    from flow_b200 import workloads as W
    types, tov, kw = W.ring_c2(noise=0.2)
    kw['warmup_steps'] = 10
    E, n = 257, 40
    a = B200VecEnv(types, tov, E, dtype='f32', seed=3, **kw)
    b = B200VecEnv(types, tov, E, dtype='f32', seed=3, **kw)
    a.reset(); b.reset()
    rng = np.random.default_rng(0)
    acts = rng.uniform(-1, 1, size=(n, 1, E)).astype(np.float32)
    for s in range(n):
        obs, rew, done, _ = a.step_host(acts[s] if s % 7 else None)
        o2, r2, d2, _ = b.step(torch.tensor(acts[s], device='cuda') if s % 7 else None)
        assert np.array_equal(obs, o2.cpu().numpy()), s
        assert np.array_equal(rew, r2.cpu().numpy()) and np.array_equal(done, d2.cpu().numpy()), s
    # k = 3, caller-owned pageable buffers (copied through the staging block)
    o3, r3, d3 = a.engine.step_host(acts[:3], k=3)
    obs_seq = torch.empty(3, E, 3, device='cuda'); rew_seq = torch.empty(3, E, device='cuda')
    b.engine.step(torch.tensor(acts[:3], device='cuda'), k=3, obs_seq=obs_seq, reward_seq=rew_seq)
    assert np.array_equal(o3, obs_seq.cpu().numpy()) and np.array_equal(r3, rew_seq.cpu().numpy())
    # the graph replays count their kernel launches
    l0 = a.engine.launch_count
    a.step_host(acts[0])
    assert a.engine.launch_count == l0 + 1


def test_unfused_graph_is_dropped_when_the_inflows_change():
    """fs_step_unfused replays a CUDA graph whose nodes hold the inflow rows by value;
    fs_set_inflows (and fs_set_types / fs_set_stream) must drop it."""
    import ctypes as C
    import torch
    from flow_b200 import _lib
    from flow_b200 import workloads as W
    from flow_b200.vec_env import B200VecEnv
    types, tov, kw = W.highway_c4(noise=0.0)
    E, K = 8, 60
    outs = []
    for variant in ('changed', 'fresh'):
        env = B200VecEnv(types, tov, E, dtype='f64', seed=1, **kw)
        eng = env.engine
        rows = (_lib.FsInflow * 1)()
        rows[0].vtype, rows[0].vehs_per_hour, rows[0].depart_speed, rows[0].begin = 0, 7200.0, 15.0, 1.0
        obs = torch.empty(K, E, env.obs_dim, device='cuda', dtype=torch.float64)
        rew = torch.empty(K, E, device='cuda', dtype=torch.float64)
        dn = torch.empty(K, E, device='cuda', dtype=torch.uint8)
        if variant == 'changed':
            env.reset()
            eng.step_unfused_ptr(0, K, obs.data_ptr(), rew.data_ptr(), dn.data_ptr())   # captured with 2000 veh/h
            eng.step_unfused_ptr(0, K, obs.data_ptr(), rew.data_ptr(), dn.data_ptr())   # replayed
        eng._check(eng.lib.fs_set_inflows(eng.h, rows, 1))
        env.reset()
        eng.step_unfused_ptr(0, K, obs.data_ptr(), rew.data_ptr(), dn.data_ptr())       # same buffers
        eng.sync()
        outs.append((eng.counters.cpu().numpy().copy(), obs.cpu().numpy().copy()))
    assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])
    assert outs[0][0][1].min() > 5      # 7200 veh/h: a departure every 5 steps
