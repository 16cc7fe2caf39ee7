"""GPU parity: every thread-mapping shape of the ring kernel and the generic
controller path (OVM, LinearOVM, BCM, CFM, LAC, FollowerStopper,
NonLocalFollowerStopper, PISaturation, both fail-safes) against the golden
trajectories produced by the reference's own controller code, and the
reference's known-answer vectors through the C-ABI."""
import json
import os

import numpy as np
import pytest

from tests import scenarios as S
from tests.test_oracle_golden import ZOO, zoo_groups, KAT, kat_groups, ring_for_headways
from oracle import sim as OS

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), 'golden')


def close(a, b, floor=1.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float((np.abs(a - b) / np.maximum(np.abs(b), floor)).max())


@pytest.mark.parametrize('shape', [1, 2, 4, 8, 16, 32])
def test_c2_golden_every_shape_fp64(shape, monkeypatch):
    """ring_c2.npz (reference-generated) through each (VPT, G) instantiation."""
    import torch
    monkeypatch.setenv('FS_RING_SHAPE', str(shape))
    z = np.load(os.path.join(G, 'ring_c2.npz'))
    lengths = [float(z['length_a']), float(z['length_b'])] * 20   # 40 envs: partial warps
    eng = S.make_engine(S.c2_groups(), len(lengths), 'f64', env_kind=OS.ENV_WAVE_ATTENUATION_PO,
                        length_lo=220., length_hi=270., record_views=1, **S.C2)
    obs = eng.reset(lengths=lengths).cpu().numpy()
    for e, tag in ((0, 'a'), (1, 'b'), (38, 'a'), (39, 'b')):
        assert close(eng.pos.cpu().numpy()[:, e], z['pos_' + tag][0]) <= 1e-9
        assert close(eng.vel.cpu().numpy()[:, e], z['vel_' + tag][0]) <= 1e-9
        assert close(obs[e], z['obs_' + tag][0]) <= 1e-9
    n = 200
    act = np.zeros((n, 1, len(lengths)))
    act[:, 0, 0::2] = z['actions_a'][:n, None]
    act[:, 0, 1::2] = z['actions_b'][:n, None]
    K = 50
    obs_seq = torch.empty(K, len(lengths), 3, device='cuda', dtype=torch.float64)
    rew_seq = torch.empty(K, len(lengths), device='cuda', dtype=torch.float64)
    for s in range(0, n, K):
        eng.step(torch.tensor(act[s:s + K], device='cuda'), k=K, obs_seq=obs_seq, reward_seq=rew_seq)
        o, r = obs_seq.cpu().numpy(), rew_seq.cpu().numpy()
        for e, tag in ((0, 'a'), (1, 'b'), (38, 'a'), (39, 'b')):
            assert close(o[:, e], z['obs_' + tag][s + 1:s + K + 1]) <= 1e-9, (shape, s, tag)
            assert close(r[:, e], z['reward_' + tag][s:s + K]) <= 1e-9, (shape, s, tag)
            assert close(eng.pos.cpu().numpy()[:, e], z['pos_' + tag][s + K]) <= 1e-9
    assert (eng.leader.cpu().numpy()[:, 0] == (np.arange(22) + 1) % 22).all()
    assert (eng.follower.cpu().numpy()[:, 0] == (np.arange(22) - 1) % 22).all()


@pytest.mark.parametrize('tag,ctrl,fs', ZOO, ids=[z[0] for z in ZOO])
def test_zoo_golden_fp64(tag, ctrl, fs):
    z = np.load(os.path.join(G, 'ring_zoo.npz'))
    eng = S.make_engine(zoo_groups(ctrl, fs), 5, 'f64', length=240, bunching=30.0, record_views=1)
    eng.reset()
    n = len(z['reward_' + tag])
    tol = 1e-9
    for s in range(n):
        _, rew, _ = eng.step()
        assert close(eng.vel.cpu().numpy()[:, 3], z['vel_' + tag][s + 1]) <= tol, (tag, s)
        assert close(eng.accel.cpu().numpy()[:, 4], z['accel_' + tag][s]) <= tol, (tag, s)
        assert close(rew.cpu().numpy()[0], z['reward_' + tag][s]) <= tol, (tag, s)
        if (s + 1) % 10 == 0:
            assert close(eng.pos.cpu().numpy()[:, 0], z['pos_' + tag][(s + 1) // 10]) <= tol


@pytest.mark.parametrize('tag,ctrl,fs', ZOO, ids=[z[0] for z in ZOO])
def test_zoo_fused_fp32_aggregates(tag, ctrl, fs):
    """fp32, 300 fused steps: mean speed and summed reward within 1 % of the reference."""
    import torch
    z = np.load(os.path.join(G, 'ring_zoo.npz'))
    eng = S.make_engine(zoo_groups(ctrl, fs), 33, 'f32', length=240, bunching=30.0)
    eng.reset()
    n = len(z['reward_' + tag])
    rew_seq = torch.empty(n, 33, device='cuda')
    eng.step(None, k=n, reward_seq=rew_seq)
    v = eng.vel.cpu().numpy()[:, 32]
    assert abs(v.mean() - z['vel_' + tag][n].mean()) <= 0.01 * max(z['vel_' + tag][n].mean(), 0.1)
    r = float(rew_seq[:, 32].sum())
    assert abs(r - z['reward_' + tag].sum()) <= 0.01 * max(abs(z['reward_' + tag].sum()), 0.1)


@pytest.mark.parametrize('case', [c for c in KAT['controllers'] if min(c['headways']) > 0],
                         ids=lambda c: c['name'])
def test_reference_known_answers_through_abi(case):
    """tests/fast_tests/test_controllers.py vectors: inject the state with
    fs_set_state, take one step, read back the applied acceleration."""
    x, circ = ring_for_headways(case['headways'])
    for dt, tol in (('f64', 1e-9), ('f32', 1e-4)):
        eng = S.make_engine(kat_groups(case), 2, dt, length=230, record_views=1)
        eng.reset()
        sp = case['speeds'] or [0.0] * 5
        eng.set_state(np.stack([x, x], 1), np.stack([sp, sp], 1), lengths=[circ - 0.4] * 2)
        assert close(eng.headway.cpu().numpy()[:, 0], case['headways']) <= tol
        eng.step()
        got = eng.accel.cpu().numpy()[:, 1]
        assert close(got, case['reference_value']) <= tol, (case['name'], dt, got)
