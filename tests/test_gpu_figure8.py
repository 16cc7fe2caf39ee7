"""GPU parity of the figure-eight network under AccelEnv through the C-ABI engine:
the device placement against the reference's gen_even_start_pos (start_pos_figure8.json), the
rollout against the golden recorded with the reference's own controllers / AccelEnv methods
(tests/golden/figure8.npz), and seeded rollouts against oracle/figure8.py."""
import json
import os

import numpy as np
import pytest

from tests import scenarios as S

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), 'golden')
CASES = json.load(open(os.path.join(G, 'start_pos_figure8.json')))


def close(a, b, floor=1.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float((np.abs(a - b) / np.maximum(np.abs(b), floor)).max()) if a.size else 0.0


@pytest.mark.parametrize('c', CASES, ids=['r%d-n%d' % (c['radius'], c['n']) for c in CASES])
def test_device_placement_is_the_reference_s(c):
    groups = [S.vt('IDM', c['n'])]
    kw = c['kw']
    eng = S.figure8_engine(3, 'f64', groups=groups, radius=float(c['radius']), x0=kw.get('x0', 0.0),
                           min_gap=kw.get('min_gap', 0.0), bunching=kw.get('bunching', 0.0))
    eng.reset()
    assert eng.pos[:, 2].cpu().numpy().tolist() == c['x']           # bit for bit
    assert float(eng.length[1]) == pytest.approx(S.figure8_sim(groups, c['x'], radius=c['radius']).g.L, abs=0)


@pytest.mark.parametrize('K', [1, 9])
def test_figure8_golden_fp64(K):
    import torch
    z = np.load(os.path.join(G, 'figure8.npz'))
    E, V = 4, 14
    eng = S.figure8_engine(E, 'f64')
    assert (eng.obs_dim, eng.act_dim) == (2 * V, 1)
    eng.reset()
    assert eng.pos[:, 3].cpu().numpy().tolist() == z['pos'][0].tolist()
    acts, n = z['actions'], len(z['actions'])
    s = 0
    while s < n:
        k = min(K, n - s)
        a = torch.tensor(np.repeat(acts[s:s + k, :, None], E, 2), device='cuda')
        obs_seq = torch.empty(k, E, 2 * V, device='cuda', dtype=torch.float64)
        rew_seq = torch.empty(k, E, device='cuda', dtype=torch.float64)
        done_seq = torch.empty(k, E, device='cuda', dtype=torch.uint8)
        eng.step(a, k=k, obs_seq=obs_seq, reward_seq=rew_seq, done_seq=done_seq)
        o, r, d = obs_seq.cpu().numpy(), rew_seq.cpu().numpy(), done_seq.cpu().numpy()
        for e in (0, E - 1):
            assert close(o[:, e], z['obs'][s:s + k]) <= 1e-9, s
            assert close(r[:, e], z['reward'][s:s + k]) <= 1e-9, s
            assert (d[:, e] != 0).tolist() == z['done'][s:s + k].tolist(), s
        s += k
        assert close(eng.pos[:, 1].cpu().numpy(), z['pos'][s], floor=10.0) <= 1e-9, s
    assert close(eng.vel[:, 2].cpu().numpy(), z['vel_final']) <= 1e-9
    assert eng.done.cpu().numpy().all()


def check_state(eng, e, sim, tol, tag):
    assert close(eng.pos[:, e].cpu().numpy(), sim.pos, floor=10.0) <= tol, tag
    assert close(eng.vel[:, e].cpu().numpy(), sim.vel) <= tol, tag
    assert eng.leader[:, e].cpu().numpy().tolist() == sim.leader, tag
    assert eng.follower[:, e].cpu().numpy().tolist() == sim.follower, tag
    assert close(eng.headway[:, e].cpu().numpy(), sim.headway) <= tol * 1e3, tag


@pytest.mark.parametrize('dtype,tol', [('f64', 1e-9), ('f32', 1e-4)])
def test_figure8_vs_oracle_per_env_actions(dtype, tol):
    """Two RL vehicles, OVM + fail-safe humans among the IDM ones, unclipped and clipped commands;
    fp32 is re-synchronised from the oracle every step (single-step error)."""
    import torch
    groups = [S.vt('IDM', 5, speed_mode=1, max_decel=1.5), S.vt('RL', 1, speed_mode=1),
              S.vt('OVM', 4, fail_safe='safe_velocity', speed_mode=25), S.vt('IDM', 6, speed_mode=7),
              S.vt('RL', 1, speed_mode=25)]
    E, steps = 3, 250
    eng = S.figure8_engine(E, dtype, groups=groups, radius=40.0, zone=25.0)
    eng.reset()
    # the placement puts vehicles exactly ON edge starts, where fp32 and fp64 disagree about the
    # edge (internal or not): start a few centimetres further on
    pos0 = eng.pos.double().cpu().numpy() + 0.037
    eng.set_state(pos=pos0)
    pos0 = eng.pos.double().cpu().numpy()
    sims = [S.figure8_sim(groups, pos0[:, e].tolist(), radius=40.0, zone=25.0) for e in range(E)]
    rng = np.random.default_rng(11)
    tdt = torch.float64 if dtype == 'f64' else torch.float32
    n_cross = 0
    for s in range(steps):
        a = rng.uniform(-2.0, 4.0, (2, E))
        obs, rew, done = eng.step(torch.tensor(a, device='cuda', dtype=tdt))
        obs, rew, done = obs.cpu().numpy(), rew.cpu().numpy(), done.cpu().numpy()
        for e, sim in enumerate(sims):
            a_e = a[:, e].astype(np.float32).astype(np.float64) if dtype == 'f32' else a[:, e]
            o_obs, o_rew, o_done = sim.step(a_e)
            tag = (dtype, s, e)
            check_state(eng, e, sim, tol, tag)
            assert close(obs[e], o_obs) <= tol and close(rew[e], o_rew) <= tol, tag
            assert bool(done[e]) == o_done, tag
            n_cross += sum(sim.cross)
        if dtype == 'f32':
            eng.set_state(pos=np.array([sim.pos for sim in sims]).T, vel=np.array([sim.vel for sim in sims]).T)
        if all(sim.done for sim in sims):
            break
    assert n_cross > 50


def test_figure8_collision_sizes_and_sims_per_step():
    """Speed mode 0 and full throttle: somebody is hit (on the loop or in the box) -> done, reward 0,
    frozen; 70 vehicles (three slot rows) and sims_per_step 3 against the oracle."""
    import torch
    groups = S.figure8_groups(n_human=6, n_rl=2, speed_mode=0)
    eng = S.figure8_engine(2, 'f64', groups=groups)
    eng.reset()
    sim = S.figure8_sim(groups, eng.pos[:, 0].cpu().numpy().tolist())
    hit = None
    for s in range(400):
        a = np.array([3.0, 2.0])
        obs, rew, done = eng.step(torch.tensor(np.repeat(a[:, None], 2, 1), device='cuda'))
        o_obs, o_rew, o_done = sim.step(a)
        assert bool(done[0].item()) == o_done and close(rew[1].item(), o_rew) <= 1e-9, s
        assert close(obs[0].cpu().numpy(), o_obs) <= 1e-9, s
        if o_done:
            hit = s
            break
    assert hit is not None and sim.crash and bool(eng.crash[1].item()) and float(eng.reward[0]) == 0.0
    before = eng.pos[:, 0].cpu().numpy().copy()
    eng.step(torch.zeros(2, 2, device='cuda', dtype=torch.float64))
    np.testing.assert_array_equal(eng.pos[:, 0].cpu().numpy(), before)
    groups = [S.vt('IDM', 68, speed_mode=1), S.vt('RL', 2, speed_mode=1)]
    eng = S.figure8_engine(2, 'f64', groups=groups, radius=60.0, sims_per_step=3, horizon=50)
    eng.reset()
    # (the reference's placement stacks every vehicle due between 7.04 m and the first junction
    # pass on the start of 'top' - test_device_placement...; 70 vehicles would start crashed)
    L = float(eng.length[0])
    spread = np.repeat((np.arange(70) * (L / 70) + 1.0)[:, None], 2, 1)
    eng.set_state(pos=spread)
    sim = S.figure8_sim(groups, eng.pos[:, 1].cpu().numpy().tolist(), radius=60.0, sims_per_step=3, horizon=50)
    rng = np.random.default_rng(5)
    for s in range(50):
        a = rng.uniform(-1, 1, 2)
        obs, rew, done = eng.step(torch.tensor(np.repeat(a[:, None], 2, 1), device='cuda'))
        o_obs, o_rew, o_done = sim.step(a)
        check_state(eng, 0, sim, 1e-9, s)
        assert close(obs[1].cpu().numpy(), o_obs) <= 1e-9 and close(rew[0].item(), o_rew) <= 1e-9
        assert bool(done[0].item()) == o_done and not sim.crash, s
    assert o_done and bool(eng.done.all()) and int(eng.time_counter[0]) == 150


def test_figure8_noise_stats_and_validation():
    import torch
    from flow_b200 import engine as EN
    groups = S.figure8_groups(noise=0.2)
    eng = S.figure8_engine(8, 'f32', groups=groups, seed=5)
    eng.reset()
    eng.stats(reset=True)
    eng.step(torch.zeros(100, 1, 8, device='cuda'), k=100)
    st = eng.stats()
    assert st['env_steps'] == 800 and st['vehicle_steps'] == 800 * 14 and st['crashes'] == 0
    p = eng.pos.cpu().numpy()
    assert not np.array_equal(p[:, 0], p[:, 1])
    with pytest.raises(EN.FlowStepError):
        S.figure8_engine(1, groups=[S.vt('BCM', 5)])
    with pytest.raises(EN.FlowStepError):
        S.figure8_engine(1, radius=30.0, zone=31.0)
    with pytest.raises(EN.FlowStepError):          # 90 vehicles do not fit on a radius-10 loop
        S.figure8_engine(1, groups=[S.vt('IDM', 90)], radius=10.0).reset()


def figure8_flow_params(F, horizon=600, noise=0.0):
    """examples/exp_configs/rl/singleagent/singleagent_figure_eight.py against module F."""
    vehicles = F.VehicleParams()
    vehicles.add(veh_id='human', acceleration_controller=(F.IDMController, {'noise': noise}),
                 routing_controller=(F.ContinuousRouter, {}),
                 car_following_params=F.SumoCarFollowingParams(speed_mode='obey_safe_speed', decel=1.5),
                 num_vehicles=13)
    vehicles.add(veh_id='rl', acceleration_controller=(F.RLController, {}),
                 routing_controller=(F.ContinuousRouter, {}),
                 car_following_params=F.SumoCarFollowingParams(speed_mode='obey_safe_speed', decel=1.5),
                 num_vehicles=1)
    return dict(
        exp_tag='singleagent_figure_eight', env_name=F.AccelEnv, network=F.FigureEightNetwork,
        simulator='traci', sim=F.SumoParams(sim_step=0.1, render=False),
        env=F.EnvParams(horizon=horizon, additional_params={
            'target_velocity': 20, 'max_accel': 3, 'max_decel': 3, 'sort_vehicles': False}),
        net=F.NetParams(additional_params=dict(F.FIGURE_EIGHT_NET_PARAMS)), veh=vehicles,
        initial=F.InitialConfig())


def test_figure8_env_through_the_flow_api():
    """make_create_env on singleagent_figure_eight.py-style flow_params against the golden the
    reference's own classes produced for the same scenario."""
    import torch
    from flow_b200 import flow_api as F
    z = np.load(os.path.join(G, 'figure8.npz'))
    env = F.make_create_env(figure8_flow_params(F))[0]()
    assert env.action_space.shape == (1,) and env.observation_space.shape == (28,)
    obs = env.reset()
    kv = env.k.vehicle
    assert [[kv.get_edge(v), kv.get_position(v)] for v in kv.get_ids()] == \
        [[e, pytest.approx(p, abs=1e-9)] for e, p in CASES[0]['pos']]
    assert env.k.network.length() == pytest.approx(416.2633388230813, abs=1e-9)
    for k in range(len(z['actions'])):
        obs, rew, done, _ = env.step(z['actions'][k])
        assert close(obs, z['obs'][k]) <= 1e-9 and abs(rew - z['reward'][k]) <= 1e-9, k
        assert done == bool(z['done'][k]), k
    assert kv.get_x_by_id('rl_0') == pytest.approx(z['obs'][-1][27] * env.k.network.length(), abs=1e-9)
    assert kv.get_edge('human_3') in env.k.network.get_edge_list() + env.k.network.get_junction_list()
    env.terminate()
    # a batch, noisy humans as in the shipped config; RL vehicle left to the simulator model
    env = F.make_create_env(figure8_flow_params(F, horizon=300, noise=0.2), n_envs=256, dtype='f32')[0]()
    env.reset()
    for s in range(300):
        obs, rew, done, _ = env.step(None)
    assert tuple(obs.shape) == (256, 28) and bool(done.all())
    assert float(env.engine.crash.float().mean()) < 0.05 and float(rew.mean()) > 0.05
    env.terminate()
    bad = figure8_flow_params(F)
    bad['net'].additional_params['lanes'] = 2
    with pytest.raises(NotImplementedError):
        F.make_create_env(bad)[0]()
    bad = figure8_flow_params(F)
    bad['env_name'] = F.WaveAttenuationPOEnv
    bad['env'].additional_params['ring_length'] = [220, 270]
    with pytest.raises(NotImplementedError):
        F.make_create_env(bad)[0]()
