"""Seeded differential fuzz of the two junction networks: random merge networks
(edge lengths, zone, flows on both arms, RL penetration, controlled rows, initial vehicles in any
order, controller mixes with fail-safes and speed modes, dt, sims_per_step) and random figure
eights (radius, zone, vehicle counts across the slot-row boundaries, RL vehicles anywhere in the
order, shuffled initial order) stepped by the CUDA engine in fp64 and by the oracles
(oracle/merge.py, oracle/figure8.py): states, ids, leader views, observations, rewards and done
flags agree every Env.step, collisions included."""
import numpy as np
import pytest

from tests import scenarios as S
from tests.test_gpu_merge import check_state as check_merge, close

pytestmark = pytest.mark.gpu

HUMANS = ['IDM', 'OVM', 'LINEAR_OVM', 'CFM', 'GIPPS', 'FOLLOWER_STOPPER', 'SIM']


def draw_types(rng, n_kinds):
    kinds = [str(k) for k in rng.choice(HUMANS, size=n_kinds, replace=False)]
    if 'IDM' not in kinds and rng.random() < 0.6:
        kinds[0] = 'IDM'
    out = []
    for k in kinds:
        fs = str(rng.choice(['instantaneous', 'safe_velocity'])) \
            if k in ('IDM', 'OVM', 'LINEAR_OVM', 'CFM', 'GIPPS') and rng.random() < 0.35 else None
        out.append(S.vt(k, 0, fail_safe=fs, speed_mode=int(rng.choice([0, 1, 1, 7, 25, 31])),
                        max_decel=float(rng.choice([1.5, 4.5]))))
    return out


def draw_merge(seed):
    rng = np.random.default_rng(7000 + seed)
    groups = draw_types(rng, int(rng.integers(1, 4)))
    rl_t = len(groups)
    groups.append(S.vt('RL', 0, speed_mode=int(rng.choice([0, 1, 25]))))
    pre, ramp, post = float(rng.uniform(80, 400)), float(rng.uniform(30, 150)), float(rng.uniform(40, 150))
    flows = []
    for _ in range(int(rng.integers(1, 5))):
        arm = int(rng.random() < 0.35)
        t = rl_t if rng.random() < 0.3 else int(rng.integers(0, rl_t))
        flows.append((t, arm, float(rng.uniform(200, 1500)), float(rng.uniform(3, 15)),
                      float(rng.choice([0.2, 1.0, 3.7]))))
    initial = []
    for _ in range(int(rng.integers(0, 7))):
        arm = int(rng.random() < 0.3)
        t = rl_t if rng.random() < 0.25 else int(rng.integers(0, rl_t))
        initial.append((t, arm, float(rng.uniform(0, (ramp if arm else pre) + 100.0))))
    # keep the initial vehicles apart (a start with overlapping vehicles is a crashed start)
    initial.sort(key=lambda r: (r[1], r[2]))
    spaced, last = [], {0: -1e9, 1: -1e9}
    for t, arm, p in initial:
        p = max(p, last[arm] + 9.0)
        last[arm] = p
        spaced.append((t, arm, p))
    order = rng.permutation(len(spaced))
    initial = [spaced[i] for i in order]
    return dict(groups=groups, pre=pre, ramp=ramp, post=post, inflows=flows, initial=initial,
                num_rl=int(rng.choice([1, 2, 3, 5, 8])), dt=float(rng.choice([0.1, 0.2, 0.4])),
                sims_per_step=int(rng.choice([1, 2, 5])), capacity=int(rng.choice([64, 128, 256])),
                zone=float(rng.uniform(20, ramp + 22.5)) if rng.random() < 0.5 else None,
                horizon=int(rng.integers(25, 70)), target_velocity=float(rng.uniform(5, 25)),
                evaluate=bool(rng.random() < 0.15), act=float(rng.choice([0.5, 1.5, 3.0])))


@pytest.mark.parametrize('seed', range(24))
def test_fuzz_merge_vs_oracle_fp64(seed):
    import torch
    c = draw_merge(seed)
    act = c.pop('act')
    groups = c.pop('groups')
    eng = S.merge_engine(2, 'f64', groups=groups, **c)
    sim = S.merge_sim(groups=groups, **c)
    rng = np.random.default_rng(seed)
    A = c['num_rl']
    for s in range(c['horizon'] + 2):
        a = rng.uniform(-act, act, A)
        if rng.random() < 0.1:
            a[:] = np.nan if False else a            # (rows without a vehicle are ignored anyway)
        obs, rew, done = eng.step(torch.tensor(np.repeat(a[:, None], 2, 1), device='cuda'))
        o_obs, o_rew, o_done = sim.step(a)
        tag = (seed, s)
        if not sim.done or s < c['horizon'] + 1:
            check_merge(eng, 1, sim, 1e-9, tag)
        assert close(obs[0].cpu().numpy(), o_obs) <= 1e-9, tag
        assert close(rew[1].item(), o_rew) <= 1e-9, tag
        assert bool(done[0].item()) == o_done and bool(eng.crash[1].item()) == sim.crash, tag


def draw_fig8(seed):
    rng = np.random.default_rng(9000 + seed)
    V = int(rng.choice([1, 2, 5, 14, 22, 31, 32, 33, 40, 64, 65]))
    radius = float(rng.uniform(max(20.0, V * 1.6), max(40.0, V * 2.5)))
    types = draw_types(rng, int(rng.integers(1, 4)))
    n_rl = int(rng.integers(0, min(V, 3) + 1))
    counts = rng.multinomial(V - n_rl, np.ones(len(types)) / len(types))
    groups = [dict(t, count=int(n)) for t, n in zip(types, counts) if n > 0]
    for _ in range(n_rl):
        groups.insert(int(rng.integers(0, len(groups) + 1)), S.vt('RL', 1, speed_mode=int(rng.choice([0, 1, 25]))))
    return dict(groups=groups, V=V, radius=radius, dt=float(rng.choice([0.1, 0.2, 0.5])),
                sims_per_step=int(rng.choice([1, 1, 3])), horizon=int(rng.integers(40, 120)),
                zone=float(rng.uniform(8.0, radius)) if rng.random() < 0.5 else None,
                target_velocity=float(rng.uniform(5, 25)), evaluate=bool(rng.random() < 0.15),
                clip_actions=bool(rng.random() < 0.7), shuffle=bool(rng.random() < 0.3),
                v0=float(rng.choice([0.0, 4.0, 9.0])), act=float(rng.choice([0.5, 2.0, 4.0])))


@pytest.mark.parametrize('seed', range(30))
def test_fuzz_figure8_vs_oracle_fp64(seed):
    import torch
    from tests.test_gpu_figure8 import check_state
    c = draw_fig8(seed)
    V, act, groups, shuffle, v0 = c.pop('V'), c.pop('act'), c.pop('groups'), c.pop('shuffle'), c.pop('v0')
    eng = S.figure8_engine(2, 'f64', groups=groups, **c)
    eng.reset()
    L = float(eng.length[0])
    rng = np.random.default_rng(seed)
    # evenly spread (the reference's own placement stacks vehicles near the first junction pass),
    # jittered, optionally in a shuffled index order (all-pairs leader search), moving at v0
    pos = (np.arange(V) * (L / V) + rng.uniform(0, 0.3 * L / V, V) + rng.uniform(0, L)) % L
    if shuffle:
        pos = pos[rng.permutation(V)]
    vel = np.full(V, v0)
    eng.set_state(pos=np.repeat(pos[:, None], 2, 1), vel=np.repeat(vel[:, None], 2, 1))
    sim = S.figure8_sim(groups, pos.tolist(), **c)
    sim.set_state(vel=vel.tolist())
    n_rl = len(sim.rl_idx)
    for s in range(c['horizon'] + 2):
        a = rng.uniform(-act, act, max(n_rl, 1))
        t = torch.tensor(np.repeat(a[:n_rl, None], 2, 1), device='cuda') if n_rl else None
        obs, rew, done = eng.step(t)
        o_obs, o_rew, o_done = sim.step(a[:n_rl] if n_rl else None)
        tag = (seed, s)
        check_state(eng, 1, sim, 1e-9, tag)
        assert close(obs[0].cpu().numpy(), o_obs) <= 1e-9 and close(rew[1].item(), o_rew) <= 1e-9, tag
        assert bool(done[0].item()) == o_done and bool(eng.crash[1].item()) == sim.crash, tag


@pytest.mark.parametrize('net', ['merge', 'figure8'])
def test_fp32_aggregates_within_one_percent_of_fp64(net):
    """Free-running fp32 against fp64 over a whole episode (no re-synchronisation): mean speed
    and mean reward of 64 envs agree within 1 % (BASELINE.json's aggregate bar)."""
    import torch
    E = 64
    rng = np.random.default_rng(3)
    out = {}
    for dtype in ('f64', 'f32'):
        td = torch.float64 if dtype == 'f64' else torch.float32
        if net == 'merge':
            eng = S.merge_engine(E, dtype, initial=[], warmup_steps=40, horizon=260)
            K, A = 260, 3
        else:
            eng = S.figure8_engine(E, dtype, horizon=1500)
            eng.reset()
            K, A = 1500, 1
        acts = torch.tensor(np.random.default_rng(17).uniform(-0.5, 1.0, (K, A, E)), device='cuda', dtype=td)
        eng.stats(reset=True)
        eng.step(acts, k=K)
        st = eng.stats()
        out[dtype] = (st['sum_speed'] / st['vehicle_steps'], st['sum_reward'] / st['env_steps'],
                      st['crashes'], st['vehicle_steps'])
    (v64, r64, c64, n64), (v32, r32, c32, n32) = out['f64'], out['f32']
    assert c64 == c32 == 0
    assert abs(v32 - v64) <= 0.01 * v64 and abs(r32 - r64) <= 0.01 * r64, out
    assert abs(n32 - n64) <= 0.01 * n64
