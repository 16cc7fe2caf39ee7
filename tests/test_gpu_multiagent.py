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
        v.add(veh_id='rl_%d' % i, acceleration_controller=(F.RLController, {}),
              routing_controller=(F.ContinuousRouter, {}), num_vehicles=1)
        v.add(veh_id='human_%d' % i, acceleration_controller=(F.IDMController, {}),
              car_following_params=F.SumoCarFollowingParams(min_gap=0),
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
