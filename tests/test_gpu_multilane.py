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
