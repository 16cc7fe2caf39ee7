"""SUMO speed modes (SumoCarFollowingParams.speed_mode, flow/core/params.py:12-18, default
'right_of_way' = 25): the restated clamp on speeds Flow commands. CPU: the oracle against the
golden produced with the reference's own controllers (tests/golden/ring_c2_speed_mode.npz) and
the behaviour VERDICT r01 asks for (the RL vehicle of singleagent_ring.py cannot ram its leader
under random actions). GPU: the engine's kernels against the same golden and the oracle."""
import os

import numpy as np
import pytest

from tests import scenarios as S
from oracle import sim as OS
from oracle import controllers as OC

G = os.path.join(os.path.dirname(__file__), 'golden')
KW = dict(env_kind=OS.ENV_WAVE_ATTENUATION_PO, bunching=50.0, horizon=3000, act_lo=-1, act_hi=1,
          obs_max_length=270.0)


def close(a, b, floor=1.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float((np.abs(a - b) / np.maximum(np.abs(b), floor)).max())


def test_speed_mode_bits_on_the_oracle():
    vt = OC.VType(OC.RL, max_accel=2.6, max_decel=4.5, sim_min_gap=2.5, speed_mode=0)
    k = type('K', (), dict(vel=[5.0, 5.0], leader=[1, 0], headway=[3.0, 200.0]))()
    free = OC.sim_model_accel(k, 1, vt, 30.0)
    tight = OC.sim_model_accel(k, 0, vt, 30.0)
    assert tight < -5 < 0 < free <= 2.6
    for mode, a, i, want in ((0, 9.0, 0, 9.0), (1, 9.0, 0, tight), (1, -9.0, 0, min(-9.0, tight)),
                             (2, 9.0, 1, 2.6), (4, -9.0, 1, -4.5), (7, 9.0, 1, min(free, 2.6)),
                             (25, 9.0, 1, free), (25, -9.0, 1, -9.0), (31, -9.0, 1, -4.5)):
        vt.speed_mode = mode
        assert OC.speed_mode_accel(k, i, vt, a, 30.0) == want, (mode, a, i)


def test_oracle_matches_the_reference_driven_golden():
    z = np.load(os.path.join(G, 'ring_c2_speed_mode.npz'))
    for tag in 'ab':
        mode = int(z['mode_' + tag])
        o = S.make_oracle(S.c2_groups(speed_mode=mode), float(z['length_' + tag]), warmup_steps=100, **KW)
        assert np.array_equal(np.array(o.pos), z['pos_' + tag][0])
        for s, a in enumerate(z['actions_' + tag]):
            obs, r, done = o.step([a])
            assert not done
            assert np.array_equal(np.array(o.pos), z['pos_' + tag][s + 1]), (tag, s)
            assert np.array_equal(np.array(o.vel), z['vel_' + tag][s + 1]), (tag, s)
            assert np.array_equal(obs, z['obs_' + tag][s + 1]) and r == z['reward_' + tag][s]
    # the clamp did something: the same actions without it end in a collision
    o = S.make_oracle(S.c2_groups(), float(z['length_a']), warmup_steps=100, **KW)
    assert any(o.step([a])[2] for a in z['actions_a'])


def test_rl_vehicle_of_singleagent_ring_cannot_ram_its_leader():
    """mode 25 on every type (the reference's default, what singleagent_ring.py leaves in place),
    idm noise 0.2, U(-1, 1) actions for a whole 3000-step horizon: no negative headway."""
    rng = np.random.default_rng(3)
    o = S.make_oracle(S.c2_groups(noise=0.2, speed_mode=25), 230.0, warmup_steps=0,
                      noise_fn=lambda s, i: rng.normal(), **KW)
    low = 1e9
    for a in rng.uniform(-1, 1, 3000):
        _, _, done = o.step([a])
        low = min(low, min(o.headway))
    assert o.time_counter == 3000 and not o.crash and low >= 0


def test_flow_api_hands_the_mode_to_the_engine():
    from flow_b200 import flow_api as F
    assert F.SumoCarFollowingParams().speed_mode == 25
    assert F.SumoCarFollowingParams(speed_mode='aggressive').speed_mode == 0
    veh = F.VehicleParams()
    veh.add('human', acceleration_controller=(F.IDMController, {}), num_vehicles=2,
            car_following_params=F.SumoCarFollowingParams(speed_mode='no_collide'))
    veh.add('rl', acceleration_controller=(F.RLController, {}), num_vehicles=1)
    rows = [F._vtype_row(r, False) for r in veh.initial]
    assert [r.speed_mode for r in rows] == [7, 25]


# ------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize('shape', [1, 2, 4, 8, 16, 32])
def test_gpu_hot_kernel_speed_mode_golden_fp64(shape, monkeypatch):
    import torch
    monkeypatch.setenv('FS_RING_SHAPE', str(shape))
    z = np.load(os.path.join(G, 'ring_c2_speed_mode.npz'))
    for tag in 'ab':
        mode, L = int(z['mode_' + tag]), float(z['length_' + tag])
        E = 3
        eng = S.make_engine(S.c2_groups(speed_mode=mode), E, 'f64', length_lo=220., length_hi=270.,
                            env_kind=OS.ENV_WAVE_ATTENUATION_PO, **dict(S.C2, warmup_steps=100))
        eng.reset(lengths=[L] * E)
        assert close(eng.pos.cpu().numpy()[:, 1], z['pos_' + tag][0]) <= 1e-9
        s = 0
        for K in (1, 6, 93, 250, 350):
            act = np.repeat(z['actions_' + tag][s:s + K, None, None], E, 2)
            obs_seq = torch.empty(K, E, 3, device='cuda', dtype=torch.float64)
            rew_seq = torch.empty(K, E, device='cuda', dtype=torch.float64)
            eng.step(torch.tensor(act, device='cuda'), k=K, obs_seq=obs_seq, reward_seq=rew_seq)
            assert close(obs_seq.cpu().numpy()[:, 2], z['obs_' + tag][s + 1:s + K + 1]) <= 1e-9, (shape, tag, s)
            assert close(rew_seq.cpu().numpy()[:, 0], z['reward_' + tag][s:s + K]) <= 1e-9
            s += K
            assert close(eng.pos.cpu().numpy()[:, 0], z['pos_' + tag][s]) <= 1e-9
            assert close(eng.vel.cpu().numpy()[:, 0], z['vel_' + tag][s]) <= 1e-9
        assert s == 700 and int(eng.crash.sum()) == 0


@pytest.mark.gpu
@pytest.mark.parametrize('kernel', ['hot', 'fast', 'generic'])
def test_gpu_ring_kernels_speed_mode_vs_oracle(kernel, monkeypatch):
    """The three single-lane ring kernels (hot: c2 pattern; fast: FS_NO_HOT=1; generic: an OVM
    vehicle in the table) with mixed modes per type against the oracle, fp64."""
    import torch
    if kernel == 'fast':
        monkeypatch.setenv('FS_NO_HOT', '1')
    groups = [S.vt('IDM', 21, sim_min_gap=0.0, speed_mode=25), S.vt('RL', 1, speed_mode=7)]
    if kernel == 'generic':
        groups = [S.vt('IDM', 10, speed_mode=31), S.vt('OVM', 1, speed_mode=1),
                  S.vt('IDM', 10, speed_mode=0), S.vt('RL', 1, speed_mode=25)]
    E, L, K = 2, 240.0, 400
    eng = S.make_engine(groups, E, 'f64', length=L, env_kind=OS.ENV_WAVE_ATTENUATION_PO,
                        **dict(S.C2, warmup_steps=20))
    eng.reset()
    o = S.make_oracle(groups, L, warmup_steps=20, **KW)
    rng = np.random.default_rng(1)
    act = rng.uniform(-1.0, 1.6, size=(K, 1, E))
    act[:, :, 1] = act[:, :, 0]
    obs_seq = torch.empty(K, E, 3, device='cuda', dtype=torch.float64)
    rew_seq = torch.empty(K, E, device='cuda', dtype=torch.float64)
    eng.step(torch.tensor(act, device='cuda'), k=K, obs_seq=obs_seq, reward_seq=rew_seq)
    og, rg = obs_seq.cpu().numpy(), rew_seq.cpu().numpy()
    for k in range(K):
        ob, r, d = o.step([act[k, 0, 0]])
        assert not d
        assert close(og[k, 1], ob) <= 1e-9 and close(rg[k, 1], r) <= 1e-9, (kernel, k)
    assert close(eng.pos.cpu().numpy()[:, 0], o.pos) <= 1e-9


@pytest.mark.gpu
def test_gpu_c2_default_mode_keeps_every_ring_alive_fp32():
    """256 rings, fp32, noise 0.2, mode 25, U(-1, 1) actions, 3000 fused steps: no collision
    (without the mode about 28 % of the rings of bench.py's c2 workload collided)."""
    import torch
    from flow_b200 import workloads as W
    from flow_b200.vec_env import B200VecEnv
    types, tov, kw = W.ring_c2(noise=0.2)
    assert all(t.speed_mode == 25 for t in types)
    E, K = 256, 3000
    env = B200VecEnv(types, tov, E, dtype='f32', seed=9, **kw)
    env.reset()
    act = torch.rand(K, 1, E, device='cuda') * 2 - 1
    done = torch.empty(K, E, device='cuda', dtype=torch.uint8)
    env.rollout(act, done_seq=done)
    assert int(env.engine.crash.sum()) == 0
    assert int(done[:-1].sum()) == 0 and int(done[-1].sum()) == E      # horizon only
