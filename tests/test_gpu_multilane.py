"""GPU parity: multi-lane ring (c3) vs the reference-generated golden
trajectory and the oracle; lanes / leaders / followers bit-exact."""
import os

import numpy as np
import pytest

from tests import scenarios as S
from tests.test_oracle_golden import c3_groups
from oracle import sim as OS

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), 'golden')


def close(a, b, floor=1.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float((np.abs(a - b) / np.maximum(np.abs(b), floor)).max())


def mk(E, dtype, **kw):
    return S.make_engine(c3_groups(), E, dtype, length=230, env_kind=OS.ENV_LANE_CHANGE_ACCEL,
                         n_lanes=3, act_lo=-3.0, act_hi=3.0, **kw)


def test_c3_golden_fp64_stepwise():
    import torch
    z = np.load(os.path.join(G, 'ring_c3.npz'))
    E = 5
    eng = mk(E, 'f64', record_views=1)
    eng.reset()
    assert close(eng.pos.cpu().numpy()[:, 4], z['pos'][0]) <= 1e-12
    assert (eng.lane.cpu().numpy()[:, 0] == z['lane'][0]).all()
    o = S.make_oracle(c3_groups(), 230, env_kind=OS.ENV_LANE_CHANGE_ACCEL, n_lanes=3,
                      act_lo=-3, act_hi=3)
    for s, a in enumerate(z['actions']):
        if np.isnan(a).any():
            obs, rew, done = eng.step(None)
            o.step(None)
        else:
            act = torch.tensor(np.repeat(a[:, None], E, 1), device='cuda')
            obs, rew, done = eng.step(act)
            o.step(list(a))
        for e in (0, 3):
            assert close(eng.pos.cpu().numpy()[:, e], z['pos'][s + 1]) <= 1e-9, s
            assert close(eng.vel.cpu().numpy()[:, e], z['vel'][s + 1]) <= 1e-9, s
            assert (eng.lane.cpu().numpy()[:, e] == z['lane'][s + 1]).all(), s
            assert close(obs.cpu().numpy()[e], z['obs'][s]) <= 1e-9, s
            assert close(rew.cpu().numpy()[e], z['reward'][s]) <= 1e-9, s
        assert (eng.leader.cpu().numpy()[:, 1] == np.array(o.leader)).all(), s
        assert (eng.follower.cpu().numpy()[:, 1] == np.array(o.follower)).all(), s
        assert close(eng.headway.cpu().numpy()[:, 2], o.headway) <= 1e-9, s
    assert bool(done.cpu().numpy()[0]) == bool(z['crashed'])


def test_c3_fused_equals_stepwise_and_host():
    import torch
    E, K = 70, 120
    for dt in ('f32', 'f64'):
        td = torch.float32 if dt == 'f32' else torch.float64
        a, b = mk(E, dt), mk(E, dt)
        a.reset(); b.reset()
        gen = torch.Generator(device='cuda'); gen.manual_seed(1)
        act = torch.empty(K, 6, E, device='cuda', dtype=td)
        act[:, 0::2] = torch.rand(K, 3, E, device='cuda', dtype=td, generator=gen) * 1.6 - 0.8
        act[:, 1::2] = torch.randint(-1, 2, (K, 3, E), device='cuda', generator=gen).to(td)
        obs_seq = torch.empty(K, E, 198, device='cuda', dtype=td)
        rew_seq = torch.empty(K, E, device='cuda', dtype=td)
        a.step(act, k=K, obs_seq=obs_seq, reward_seq=rew_seq)
        for k in range(K):
            obs, rew, _ = b.step(act[k])
            assert torch.equal(obs, obs_seq[k]) and torch.equal(rew, rew_seq[k]), (dt, k)
        assert torch.equal(a.lane, b.lane) and torch.equal(a.pos, b.pos)


def test_c3_fp32_aggregates_vs_oracle():
    import torch
    K = 150
    eng = mk(4, 'f32')
    eng.reset()
    o = S.make_oracle(c3_groups(), 230, env_kind=OS.ENV_LANE_CHANGE_ACCEL, n_lanes=3,
                      act_lo=-3, act_hi=3)
    rng = np.random.default_rng(3)
    acts = np.zeros((K, 6))
    acts[:, 0::2] = rng.uniform(-0.5, 0.5, (K, 3))
    acts[:, 1::2] = rng.integers(-1, 2, (K, 3))
    rew_seq = torch.empty(K, 4, device='cuda')
    eng.step(torch.tensor(np.repeat(acts[:, :, None], 4, 2), device='cuda', dtype=torch.float32),
             k=K, reward_seq=rew_seq)
    r_o = 0.0
    for k in range(K):
        _, r, d = o.step(list(acts[k]))
        r_o += r
    assert (eng.lane.cpu().numpy()[:, 0] == np.array(o.lane)).all()
    assert abs(eng.vel.cpu().numpy()[:, 0].mean() - np.mean(o.vel)) <= 0.01 * np.mean(o.vel)
    assert abs(float(rew_seq[:, 0].sum()) - r_o) <= 0.01 * abs(r_o)


# ---- LaneChangeAccelPOEnv (per-lane leaders / followers of the RL vehicles) ----------
from tests.test_oracle_golden import PO_CASES


def mk_po(tag, E, dtype, **kw):
    groups, dist = PO_CASES[tag]
    kw.setdefault('n_lanes', 3)
    return S.make_engine([S.vt(c, n) for c, n in groups], E, dtype, length=230,
                         env_kind=OS.ENV_LANE_CHANGE_ACCEL_PO, act_lo=-3.0, act_hi=3.0,
                         lanes_distribution=dist, **kw)


@pytest.mark.parametrize('tag', ['a', 'b'])
def test_lane_change_po_golden_fp64(tag):
    """Observation of the reference's LaneChangeAccelPOEnv.get_state fed by its own
    _multi_lane_headways (tests/golden/ring_c3_po.npz): fused launches with obs_seq and
    single steps, 1e-9; the 1000-m empty-lane defaults and the zero block exactly."""
    import torch
    z = np.load(os.path.join(G, 'ring_c3_po.npz'))
    acts, ref = z['actions_' + tag], z['obs_' + tag]
    n_rl = acts.shape[1] // 2
    E = 3
    eng = mk_po(tag, E, 'f64')
    eng.reset()
    assert eng.obs_dim == ref.shape[1] == 4 * n_rl * 3 + 1
    warm = int(np.isnan(acts).any(1).sum())
    for s in range(warm):
        obs, rew, done = eng.step(None)
        assert close(obs.cpu().numpy()[1], ref[s]) <= 1e-9, s
    s, n = warm, len(acts)
    for K in (1, 5, 60, 2, 100, 112):
        K = min(K, n - s)
        act = torch.tensor(np.repeat(acts[s:s + K, :, None], E, 2), device='cuda')
        obs_seq = torch.empty(K, E, eng.obs_dim, device='cuda', dtype=torch.float64)
        rew_seq = torch.empty(K, E, device='cuda', dtype=torch.float64)
        eng.step(act, k=K, obs_seq=obs_seq, reward_seq=rew_seq)
        o = obs_seq.cpu().numpy()
        for e in (0, 2):
            assert close(o[:, e], ref[s:s + K]) <= 1e-9, (s, K)
            assert np.array_equal(o[:, e] == 1000, ref[s:s + K] == 1000)
            assert np.array_equal(o[:, e, 12:-1], ref[s:s + K, 12:-1])       # zero blocks
            assert close(rew_seq.cpu().numpy()[:, e], z['reward_' + tag][s:s + K]) <= 1e-9
        assert close(eng.obs.cpu().numpy()[0], ref[s + K - 1]) <= 1e-9
        assert (eng.lane.cpu().numpy()[:, 0] == z['lane_' + tag][s + K]).all()
        s += K
    assert s == n


@pytest.mark.parametrize('corrected', [0, 1])
@pytest.mark.parametrize('lanes,groups', [(1, [('IDM', 5), ('RL', 1)]),
                                          (4, [('IDM', 9), ('RL', 3)]),
                                          (3, [('RL', 2)])])
def test_lane_change_po_vs_oracle(lanes, groups, corrected):
    """Other shapes against the oracle: one lane, 4 lanes with 3 RL vehicles, RL vehicles
    alone on the ring (own leader one lap away, empty lanes); reference behaviour and the
    corrected observation (lc_corrected=1: every block, normalised)."""
    import torch
    E, K = 4, 80
    g = [S.vt(c, n) for c, n in groups]
    n_rl = sum(n for c, n in groups if c == 'RL')
    kw = dict(env_kind=OS.ENV_LANE_CHANGE_ACCEL_PO, n_lanes=lanes, act_lo=-3, act_hi=3,
              lanes_distribution=float('inf'))
    eng = S.make_engine(g, E, 'f64', length=120, lc_corrected=corrected, **kw)
    o = S.make_oracle(g, 120, **kw)
    o.lc_bug_compat = not corrected
    obs0 = eng.reset()
    assert eng.obs_dim == 4 * n_rl * lanes + (n_rl if corrected else 1)
    assert close(obs0.cpu().numpy()[0], o.get_state()) <= 1e-9
    rng = np.random.default_rng(7 + lanes)
    acts = np.zeros((K, 2 * n_rl))
    acts[:, 0::2] = rng.uniform(-0.6, 1.0, (K, n_rl))
    acts[:, 1::2] = rng.integers(-1, 2, (K, n_rl))
    obs_seq = torch.empty(K, E, eng.obs_dim, device='cuda', dtype=torch.float64)
    rew_seq = torch.empty(K, E, device='cuda', dtype=torch.float64)
    done_seq = torch.empty(K, E, device='cuda', dtype=torch.uint8)
    eng.step(torch.tensor(np.repeat(acts[:, :, None], E, 2), device='cuda'), k=K,
             obs_seq=obs_seq, reward_seq=rew_seq, done_seq=done_seq)
    og, rg, dg = obs_seq.cpu().numpy(), rew_seq.cpu().numpy(), done_seq.cpu().numpy()
    for k in range(K):
        ob, r, d = o.step(list(acts[k]))
        assert close(og[k, 3], ob) <= 1e-9, k
        assert close(rg[k, 3], r) <= 1e-9 and bool(dg[k, 3]) == d, k
        if d:
            break
    assert k > 20


def test_lane_change_po_fp32_and_batch():
    """fp32 against fp64 on the same actions (short horizon, before rounding can flip a
    lane-change decision), envs of a batch independent of each other."""
    import torch
    E, K = 64, 40
    a32, a64 = mk_po('a', E, 'f32'), mk_po('a', E, 'f64')
    a32.reset(); a64.reset()
    gen = torch.Generator(device='cuda'); gen.manual_seed(5)
    act = torch.empty(K, 2, E, device='cuda', dtype=torch.float64)
    act[:, 0] = torch.rand(K, E, device='cuda', dtype=torch.float64, generator=gen) - 0.5
    act[:, 1] = torch.randint(-1, 2, (K, E), device='cuda', generator=gen).to(torch.float64)
    o32 = torch.empty(K, E, 13, device='cuda')
    o64 = torch.empty(K, E, 13, device='cuda', dtype=torch.float64)
    a32.step(act.float(), k=K, obs_seq=o32)
    a64.step(act, k=K, obs_seq=o64)
    same = (a32.lane == a64.lane).all(0)
    assert same.float().mean() > 0.9
    d = (o32.double() - o64).abs()[:, same]
    assert float(d.max()) < 5e-3
    solo = mk_po('a', 1, 'f64')
    solo.reset()
    o1 = torch.empty(K, 1, 13, device='cuda', dtype=torch.float64)
    solo.step(act[:, :, 17:18].contiguous(), k=K, obs_seq=o1)
    assert torch.equal(o1[:, 0], o64[:, 17])


def test_multilane_known_answers_of_the_reference_test():
    """TestMultiLaneData.test_no_junctions_ring (tests/fast_tests/test_vehicles.py:205-253)
    through the CUDA observation: headways [27.857142857142858, -5, -5], tailways 28.577143."""
    eng = S.make_engine([S.vt('RL', 21)], 2, 'f64', length=230, env_kind=OS.ENV_LANE_CHANGE_ACCEL_PO,
                        n_lanes=3, junction_len=0.18, lanes_distribution=float('inf'))
    obs = eng.reset().cpu().numpy()[1]
    assert obs.shape == (4 * 21 * 3 + 1,)
    np.testing.assert_allclose(obs[0:3], [27.85714285714286, -5, -5], rtol=1e-13)
    np.testing.assert_array_almost_equal(obs[3:6], [28.577143] * 3)
    assert (obs[6:] == 0).all()                     # everyone at rest, other blocks untouched


def test_uniform_type_instantiation_equals_the_table_driven_one(monkeypatch):
    """kGeneric = false (one IDM-law human type + one RL type, parameters as kernel-argument
    constants; the default for c3) against FS_ML_GENERIC=1 (per-vehicle type table), with noise
    and lane changes, on c3 and on a sparser ring where the RL vehicles find gaps often:
    identical lanes, positions, observations and rewards in fp64."""
    import torch
    from flow_b200 import workloads as W
    from flow_b200.vec_env import B200VecEnv
    types, _, kw = W.ring_c3(noise=0.2)
    E, K = 33, 300
    n_changed = 0
    for tov in ([0] * 63 + [1] * 3, [0] * 9 + [1] + [0] * 10 + [1, 1] + [0] * 8):
        n_rl = sum(tov)
        gen = torch.Generator(device='cuda'); gen.manual_seed(3)
        act = torch.rand(K, 2 * n_rl, E, device='cuda', dtype=torch.float64, generator=gen) * 0.6 - 0.4
        act[:, 1::2] = torch.round(torch.rand(K, n_rl, E, device='cuda', dtype=torch.float64,
                                              generator=gen) * 3 - 1.5)
        out = []
        for generic in ('0', '1'):
            monkeypatch.setenv('FS_ML_GENERIC', generic)
            env = B200VecEnv(types, tov, E, dtype='f64', seed=5, **kw)
            env.reset()
            lane0 = env.engine.lane.cpu().numpy().copy()
            rew = torch.empty(K, E, device='cuda', dtype=torch.float64)
            obs = torch.empty(K, E, env.obs_dim, device='cuda', dtype=torch.float64)
            env.rollout(act, obs, rew)
            out.append((obs.cpu().numpy(), rew.cpu().numpy(), env.engine.lane.cpu().numpy(),
                        env.engine.pos.cpu().numpy()))
        n_changed += int((out[0][2] != lane0).sum())
        assert np.array_equal(out[0][2], out[1][2])
        assert close(out[0][3], out[1][3]) <= 1e-12 and close(out[0][0], out[1][0]) <= 1e-12
        assert close(out[0][1], out[1][1]) <= 1e-12
    assert n_changed > 0      # the rebuild-on-lane-change path ran
