"""GPU parity of the multi-agent ring front:
MultiAgentWaveAttenuationPOEnv — per-agent observation rows, shared reward —
against the trajectory the reference's own get_state / compute_reward produced
(tests/golden/ring_ma.npz), through the C-ABI engine (IDM-ring kernel, every
thread shape; generic-controller kernel vs the oracle) and through the
dict-keyed env of flow_b200.flow_api."""
import os

import numpy as np
import pytest

from oracle import sim as OS
from tests import scenarios as S
from tests.test_oracle_golden import MA, ma_groups

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), 'golden')


def close(a, b, floor=1.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float((np.abs(a - b) / np.maximum(np.abs(b), floor)).max())


@pytest.mark.parametrize('shape', [1, 2, 4, 8, 16, 32])
def test_ma_golden_fp64_every_shape(shape, monkeypatch):
    import torch
    monkeypatch.setenv('FS_RING_SHAPE', str(shape))
    z = np.load(os.path.join(G, 'ring_ma.npz'))
    E = 37
    eng = S.make_engine(ma_groups(), E, 'f64', length=float(z['length']),
                        env_kind=OS.ENV_MA_WAVE_ATTENUATION_PO, **MA)
    assert (eng.obs_dim, eng.act_dim) == (6, 2)
    eng.reset()
    assert close(eng.pos.cpu().numpy()[:, 36], z['pos'][0]) <= 1e-12
    acts = z['actions']
    s = 0
    while s < len(acts):
        warm = np.isnan(acts[s]).any()
        K = 1
        while s + K < len(acts) and np.isnan(acts[s + K]).any() == warm and K < 64:
            K += 1
        obs_seq = torch.empty(K, E, 6, device='cuda', dtype=torch.float64)
        rew_seq = torch.empty(K, E, device='cuda', dtype=torch.float64)
        a = None if warm else torch.tensor(np.repeat(acts[s:s + K, :, None], E, 2), device='cuda')
        eng.step(a, k=K, obs_seq=obs_seq, reward_seq=rew_seq)
        o, r = obs_seq.cpu().numpy(), rew_seq.cpu().numpy()
        for e in (0, 17, 36):
            assert close(o[:, e], z['obs'][s:s + K]) <= 1e-9, (shape, s)
            assert close(r[:, e], z['reward'][s:s + K]) <= 1e-9, (shape, s)
        s += K
        assert close(eng.pos.cpu().numpy()[:, 5], z['pos'][s]) <= 1e-9
        assert close(eng.vel.cpu().numpy()[:, 5], z['vel'][s]) <= 1e-9
    assert not eng.done.cpu().numpy().any()


def test_ma_generic_controller_kernel_vs_oracle():
    """Humans on OVM + safe_velocity: the generic ring kernel writes the same rows."""
    import torch
    groups = [S.vt('RL', 1), S.vt('OVM', 8, fail_safe='safe_velocity'), S.vt('RL', 1),
              S.vt('IDM', 7), S.vt('RL', 1), S.vt('OVM', 4)]
    kw = dict(MA, bunching=20.0)
    eng = S.make_engine(groups, 9, 'f64', length=250, env_kind=OS.ENV_MA_WAVE_ATTENUATION_PO, **kw)
    okw = dict(kw); okw.pop('init_min_gap')
    ora = S.make_oracle(groups, 250, env_kind=OS.ENV_MA_WAVE_ATTENUATION_PO, **okw)
    assert (eng.obs_dim, eng.act_dim) == (9, 3)
    o0 = eng.reset().cpu().numpy()
    assert close(o0[3], ora.get_state()) <= 1e-12
    rng = np.random.default_rng(0)
    for s in range(120):
        a = rng.uniform(-0.3, 0.3, 3)
        obs, rew, done = eng.step(torch.tensor(np.repeat(a[:, None], 9, 1), device='cuda'))
        o_obs, o_rew, o_done = ora.step(list(a))
        assert close(obs.cpu().numpy()[8], o_obs) <= 1e-9, s
        assert close(rew.cpu().numpy()[0], o_rew) <= 1e-9, s
        assert bool(done.cpu().numpy()[4]) == o_done


def test_ma_env_dict_api_matches_golden():
    """make_create_env(multiagent_ring.py-style flow_params) -> dict-keyed gym API."""
    from flow_b200 import flow_api as F
    z = np.load(os.path.join(G, 'ring_ma.npz'))
    v = F.VehicleParams()
    for i in range(2):
        # goldens carry no SUMO speed-mode clamp: 'aggressive'
        v.add(veh_id='rl_%d' % i, acceleration_controller=(F.RLController, {}),
              car_following_params=F.SumoCarFollowingParams(speed_mode='aggressive'),
              routing_controller=(F.ContinuousRouter, {}), num_vehicles=1)
        v.add(veh_id='human_%d' % i, acceleration_controller=(F.IDMController, {}),
              car_following_params=F.SumoCarFollowingParams(min_gap=0, speed_mode='aggressive'),
              routing_controller=(F.ContinuousRouter, {}), num_vehicles=10)
    L = float(z['length'])
    fp = dict(exp_tag='multiagent_ring', env_name=F.MultiAgentWaveAttenuationPOEnv,
              network=F.RingNetwork, simulator='traci', sim=F.SumoParams(sim_step=0.1),
              env=F.EnvParams(horizon=3000, warmup_steps=0, clip_actions=False, additional_params={
                  'max_accel': 1, 'max_decel': 1, 'ring_length': [220, 270]}),
              net=F.NetParams(additional_params={'length': 260, 'lanes': 1, 'speed_limit': 30,
                                                 'resolution': 40}),
              veh=v, initial=F.InitialConfig(bunching=50, min_gap=0))
    env = F.make_create_env(fp)[0]()
    assert env.rl_ids == ['rl_0_0', 'rl_1_0']
    assert env.observation_space.shape == (3,) and env.action_space.shape == (1,)
    env.reset()
    env.engine.reset(lengths=[L])
    for s, a in enumerate(z['actions'][:500]):
        acts = None if np.isnan(a).any() else {'rl_1_0': [a[1]], 'rl_0_0': [a[0]]}
        obs, rew, done, info = env.step(acts)
        assert set(obs) == {'rl_0_0', 'rl_1_0'} and set(done) == {'rl_0_0', 'rl_1_0', '__all__'}
        np.testing.assert_allclose(np.concatenate([obs['rl_0_0'], obs['rl_1_0']]), z['obs'][s],
                                   rtol=1e-9, atol=1e-12)
        if acts is None:
            assert rew == 0
        else:
            assert rew['rl_0_0'] == rew['rl_1_0']
            assert abs(rew['rl_0_0'] - z['reward'][s]) <= 1e-9 * max(1.0, abs(z['reward'][s]))
        assert not done['__all__']
    with pytest.raises(KeyError):
        env.step({'rl_0_0': [0.0]})
    env.terminate()

    # batched: tensors per agent, env 0 identical to the single env
    envb = F.make_create_env(fp, n_envs=16, dtype='f64', seed=2)[0]()
    ob = envb.reset()
    assert ob['rl_1_0'].shape == (16, 3)
    import torch
    o, r, d, _ = envb.step({'rl_0_0': torch.zeros(16), 'rl_1_0': torch.full((16,), 0.5)})
    assert r['rl_0_0'].shape == (16,) and d['__all__'].shape == (16,)
    envb.terminate()


# ---- MultiAgentAccelPOEnv (flow/envs/multiagent/ring/accel.py:87-226) --------------------
def test_multiagent_accel_po_golden_fp64():
    """6 numbers per RL vehicle incl. follower speed / headway and the reference's unwrapped
    leader distance, every thread shape of the generic ring kernel; reward shared."""
    import torch
    from tests.test_oracle_golden import ma_accel_groups, MA_ACCEL
    z = np.load(os.path.join(G, 'ring_ma_accel.npz'))
    E = 5
    eng = S.make_engine(ma_accel_groups(), E, 'f64', length=180, env_kind=OS.ENV_MA_ACCEL_PO,
                        **MA_ACCEL)
    obs = eng.reset()
    assert eng.obs_dim == 18 and eng.act_dim == 3
    assert close(obs.cpu().numpy()[4], z['obs'][0]) <= 1e-12
    acts = z['actions']
    warm = int(np.isnan(acts).any(1).sum())
    for s in range(warm):
        obs, rew, done = eng.step(None)
        assert close(obs.cpu().numpy()[0], z['obs'][s + 1]) <= 1e-9, s
    s, n = warm, len(acts)
    for K in (1, 7, 100, n):
        K = min(K, n - s)
        obs_seq = torch.empty(K, E, 18, device='cuda', dtype=torch.float64)
        rew_seq = torch.empty(K, E, device='cuda', dtype=torch.float64)
        eng.step(torch.tensor(np.repeat(acts[s:s + K, :, None], E, 2), device='cuda'), k=K,
                 obs_seq=obs_seq, reward_seq=rew_seq)
        assert close(obs_seq.cpu().numpy()[:, 2], z['obs'][s + 1:s + K + 1]) <= 1e-9, s
        assert close(rew_seq.cpu().numpy()[:, 2], z['reward'][s:s + K]) <= 1e-9, s
        s += K
    assert s == n


def test_multiagent_accel_po_small_rings_vs_oracle():
    """V = 1 (nobody ahead or behind: vmax / 0 / L defaults), V = 2 and V = 40 (other thread
    mappings of the generic kernel), fp32 within 1e-4."""
    import torch
    for groups, L in (([S.vt('RL', 1)], 60), ([S.vt('RL', 2)], 60),
                      ([S.vt('IDM', 19), S.vt('RL', 1), S.vt('IDM', 19), S.vt('RL', 1)], 420)):
        kw = dict(env_kind=OS.ENV_MA_ACCEL_PO, act_lo=-1.0, act_hi=1.0, target_velocity=20.0)
        o = S.make_oracle(groups, L, **kw)
        A = len(o.rl_idx)
        rng = np.random.default_rng(A)
        acts = rng.uniform(-0.2, 0.6, (60, A))
        ref = [o.step(list(a)) for a in acts]
        for dt, tol in (('f64', 1e-9), ('f32', 1e-4)):
            eng = S.make_engine(groups, 3, dt, length=L, **kw)
            eng.reset()
            td = torch.float64 if dt == 'f64' else torch.float32
            obs_seq = torch.empty(60, 3, 6 * A, device='cuda', dtype=td)
            rew_seq = torch.empty(60, 3, device='cuda', dtype=td)
            eng.step(torch.tensor(np.repeat(acts[:, :, None], 3, 2), device='cuda', dtype=td), k=60,
                     obs_seq=obs_seq, reward_seq=rew_seq)
            assert close(obs_seq.cpu().numpy()[:, 1], np.array([r[0] for r in ref]), floor=0.05) <= tol
            assert close(rew_seq.cpu().numpy()[:, 1], np.array([r[1] for r in ref]), floor=0.05) <= tol


def test_multiagent_accel_po_dict_api_matches_golden():
    from flow_b200 import flow_api as F
    z = np.load(os.path.join(G, 'ring_ma_accel.npz'))
    v = F.VehicleParams()
    for name, ctrl, n in (('rl_a', F.RLController, 1), ('ha', F.IDMController, 6),
                          ('rl_b', F.RLController, 1), ('hb', F.IDMController, 7),
                          ('rl_c', F.RLController, 1)):
        v.add(veh_id=name, acceleration_controller=(ctrl, {}),
              car_following_params=F.SumoCarFollowingParams(speed_mode='aggressive'),
              routing_controller=(F.ContinuousRouter, {}), num_vehicles=n)
    fp = dict(exp_tag='ma_accel', env_name=F.MultiAgentAccelPOEnv, network=F.RingNetwork,
              simulator='traci', sim=F.SumoParams(sim_step=0.1),
              env=F.EnvParams(horizon=3000, additional_params={
                  'max_accel': 1, 'max_decel': 1, 'target_velocity': 20}),
              net=F.NetParams(additional_params={'length': 180, 'lanes': 1, 'speed_limit': 30,
                                                 'resolution': 40}),
              veh=v, initial=F.InitialConfig())
    env = F.make_create_env(fp)[0]()
    ids = ['rl_a_0', 'rl_b_0', 'rl_c_0']
    assert env.rl_ids == ids and env.observation_space.shape == (6,)
    obs = env.reset()
    np.testing.assert_allclose(np.concatenate([obs[i] for i in ids]), z['obs'][0], atol=1e-12)
    for s, a in enumerate(z['actions'][:300]):
        acts = None if np.isnan(a).any() else {i: [a[q]] for q, i in enumerate(ids)}
        obs, rew, done, info = env.step(acts)
        np.testing.assert_allclose(np.concatenate([obs[i] for i in ids]), z['obs'][s + 1],
                                   rtol=1e-9, atol=1e-12)
        assert set(rew) == set(ids)                       # a dict even during the warm-up
        assert abs(rew['rl_b_0'] - z['reward'][s]) <= 1e-9
        assert not done['__all__']
    env.terminate()
    with pytest.raises(KeyError):            # multiagent/ring/accel.py:125-128
        bad = dict(fp, env=F.EnvParams(horizon=10, additional_params={'max_accel': 1, 'max_decel': 1}))
        F.make_create_env(bad)[0]()
