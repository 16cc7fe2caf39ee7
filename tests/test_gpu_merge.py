"""GPU parity of the merge network under MergePOEnv through the C-ABI engine:
against the golden rollout recorded with the reference's own MergePOEnv methods
(tests/golden/merge.npz) and against oracle/merge.py on seeded actions — observations, rewards,
done flags, the controlled-vehicle rows, counters and every slot's position / arm / id."""
import os

import numpy as np
import pytest

from tests import scenarios as S

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), 'golden', 'merge.npz')


def close(a, b, floor=1.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float((np.abs(a - b) / np.maximum(np.abs(b), floor)).max()) if a.size else 0.0


def engine_state(eng, e):
    n = int(eng.counters[0, e].item())
    return dict(n=n, departed=int(eng.counters[1, e].item()), arrived=int(eng.counters[2, e].item()),
                pos=eng.pos[:n, e].cpu().numpy(), vel=eng.vel[:n, e].cpu().numpy(),
                arm=eng.lane[:n, e].cpu().numpy(), ids=eng.veh_id[e, :n].cpu().numpy(),
                ctrl=eng.agent_id[e].cpu().numpy())


def check_state(eng, e, sim, tol, tag):
    st = engine_state(eng, e)
    assert [st['n'], st['departed'], st['arrived']] == [sim.num_vehicles, sim.departed, sim.arrived], tag
    assert list(st['ids']) == list(sim.ids), tag
    assert list(st['arm']) == list(sim.arm), tag
    assert list(st['ctrl']) == list(sim.rl_veh) + [-1] * (sim.num_rl - len(sim.rl_veh)), tag
    assert close(st['pos'], sim.pos) <= tol, tag
    assert close(st['vel'], sim.vel) <= tol, tag
    n = st['n']
    assert eng.leader[:n, e].cpu().numpy().tolist() == sim.leader, tag
    assert eng.follower[:n, e].cpu().numpy().tolist() == sim.follower, tag
    assert close(eng.headway[:n, e].cpu().numpy(), sim.headway) <= tol * 1e3, tag


@pytest.mark.parametrize('K', [1, 7])
def test_merge_golden_fp64(K):
    """300 Env.steps of 5 simulation steps, stepwise and in fused launches of 7."""
    import torch
    z = np.load(GOLD)
    E = 5
    eng = S.merge_engine(E, 'f64')
    assert (eng.obs_dim, eng.act_dim) == (15, 3)
    assert not eng.obs.cpu().numpy().any()            # no controlled vehicle yet
    acts, n = z['actions'], len(z['actions'])
    s = 0
    while s < n:
        k = min(K, n - s)
        a = torch.tensor(np.repeat(acts[s:s + k, :, None], E, 2), device='cuda')
        obs_seq = torch.empty(k, E, 15, device='cuda', dtype=torch.float64)
        rew_seq = torch.empty(k, E, device='cuda', dtype=torch.float64)
        done_seq = torch.empty(k, E, device='cuda', dtype=torch.uint8)
        eng.step(a, k=k, obs_seq=obs_seq, reward_seq=rew_seq, done_seq=done_seq)
        o, r, d = obs_seq.cpu().numpy(), rew_seq.cpu().numpy(), done_seq.cpu().numpy()
        for e in (0, E - 1):
            assert close(o[:, e], z['obs'][s:s + k]) <= 1e-9, s
            assert close(r[:, e], z['reward'][s:s + k]) <= 1e-9, s
            assert (d[:, e] != 0).tolist() == z['done'][s:s + k].tolist(), s
        s += k
        assert eng.agent_id[2].cpu().numpy().tolist() == z['ctrl'][s - 1].tolist(), s
        assert eng.counters[:, 3].cpu().numpy().tolist() == z['counts'][s - 1].tolist(), s
    st = engine_state(eng, 1)
    assert st['ids'].tolist() == z['ids_final'].tolist()
    assert st['arm'].tolist() == z['arm_final'].tolist()
    assert close(st['pos'], z['pos_final']) <= 1e-9
    assert eng.done.cpu().numpy().all()               # the horizon ended the episode


@pytest.mark.parametrize('dtype,tol', [('f64', 1e-9), ('f32', 1e-4)])
def test_merge_vs_oracle_per_env_actions(dtype, tol):
    """Every env gets its own action stream; fp32 is compared step by step from the oracle's
    state (fs_set_state resynchronises the positions so one step's error is what is measured)."""
    import torch
    E, steps = 3, 120
    eng = S.merge_engine(E, dtype)
    sims = [S.merge_sim() for _ in range(E)]
    rng = np.random.default_rng(7)
    tdt = torch.float64 if dtype == 'f64' else torch.float32
    for s in range(steps):
        a = rng.uniform(-1.2, 1.5, (3, E))
        obs, rew, done = eng.step(torch.tensor(a, device='cuda', dtype=tdt))
        obs, rew, done = obs.cpu().numpy(), rew.cpu().numpy(), done.cpu().numpy()
        for e, sim in enumerate(sims):
            a_e = a[:, e].astype(np.float32).astype(np.float64) if dtype == 'f32' else a[:, e]
            o_obs, o_rew, o_done = sim.step(a_e)
            tag = (dtype, s, e)
            check_state(eng, e, sim, tol, tag)
            assert close(obs[e], o_obs) <= tol, tag
            assert close(rew[e], o_rew) <= tol, tag
            assert bool(done[e]) == o_done, tag
        if dtype == 'f32':      # carry the oracle's state: single-step error, not drift
            n = max(sim.num_vehicles for sim in sims)
            pos = eng.pos.cpu().numpy().copy(); vel = eng.vel.cpu().numpy().copy()
            for e, sim in enumerate(sims):
                pos[:sim.num_vehicles, e] = sim.pos; vel[:sim.num_vehicles, e] = sim.vel
            eng.set_state(pos=pos, vel=vel)
    assert max(sim.arrived for sim in sims) > 10


def test_merge_capacity_classes_and_empty_network():
    """Slot classes 64 / 128 / 256 run the same rollout; a network that starts empty fills up
    from its inflows (no initial vehicle, nobody to observe at first)."""
    import torch
    z = np.load(GOLD)
    for cap in (64, 256):
        eng = S.merge_engine(2, 'f64', capacity=cap)
        a = torch.tensor(np.repeat(z['actions'][:60, :, None], 2, 2), device='cuda')
        obs_seq = torch.empty(60, 2, 15, device='cuda', dtype=torch.float64)
        eng.step(a, k=60, obs_seq=obs_seq)
        assert close(obs_seq.cpu().numpy()[:, 1], z['obs'][:60]) <= 1e-9, cap
    # 4-warp blocks x 256 slots in fp64: more than 48 KB of dynamic shared memory per block
    big = S.merge_engine(1500, 'f64', capacity=256)
    a = torch.tensor(np.repeat(z['actions'][:20, :, None], 1500, 2), device='cuda')
    obs_seq = torch.empty(20, 1500, 15, device='cuda', dtype=torch.float64)
    big.step(a, k=20, obs_seq=obs_seq)
    assert close(obs_seq.cpu().numpy()[:, 1499], z['obs'][:20]) <= 1e-9
    eng = S.merge_engine(2, 'f64', initial=[])
    sim = S.merge_sim(initial=[])
    for s in range(40):
        a = np.full(3, 0.5)
        obs, rew, done = eng.step(torch.tensor(np.repeat(a[:, None], 2, 1), device='cuda'))
        o_obs, o_rew, _ = sim.step(a)
        check_state(eng, 1, sim, 1e-9, s)
        assert close(obs.cpu().numpy()[0], o_obs) <= 1e-9 and close(rew.cpu().numpy()[1], o_rew) <= 1e-9
    assert sim.num_vehicles > 5


def test_merge_collision_ends_the_env_and_reset_restarts_it():
    """Full-throttle commands under speed mode 0 run into the leader: crash -> done, reward 0,
    the env freezes; a masked reset starts it again while the other env runs on."""
    import torch
    groups = S.merge_groups(speed_mode=0)
    E = 2
    over = dict(inflows=[(0, 0, 1800.0, 10.0, 1.0), (1, 0, 900.0, 10.0, 1.0)], horizon=400)
    eng = S.merge_engine(E, 'f64', groups=groups, **over)
    sim = S.merge_sim(groups=groups, **over)
    a = np.full(3, 1.5)
    crashed_at = None
    for s in range(200):
        act = np.repeat(a[:, None], E, 1)
        act[:, 1] = 0.0                       # env 1 coasts
        obs, rew, done = eng.step(torch.tensor(act, device='cuda'))
        o_obs, o_rew, o_done = sim.step(a)
        assert bool(done[0].item()) == o_done, s
        assert close(rew[0].item(), o_rew) <= 1e-9, s
        if o_done:
            crashed_at = s
            break
    assert crashed_at is not None and sim.crash and bool(eng.crash[0].item())
    assert not bool(eng.done[1].item())
    before = engine_state(eng, 0)
    eng.step(torch.zeros(3, E, device='cuda', dtype=torch.float64))
    after = engine_state(eng, 0)
    assert before['pos'].tolist() == after['pos'].tolist()       # frozen
    assert eng.reward[0].item() == 0.0
    t1 = int(eng.time_counter[1].item())
    mask = np.array([1, 0], np.uint8)
    eng.reset_stage(eng.RESET_PLACE, mask=mask)
    st = engine_state(eng, 0)
    assert st['n'] == 3 and st['departed'] == 3 and st['arrived'] == 0
    assert int(eng.time_counter[1].item()) == t1 and int(eng.time_counter[0].item()) == 0


def test_merge_noise_and_stats():
    """Noisy humans: the engine's stream is the documented Philox stream (oracle/noise.py ring
    tests pin it); here the run is reproducible, differs across envs, and the statistics count
    one sample per Env.step."""
    import torch
    groups = S.merge_groups(noise=0.2)
    eng = S.merge_engine(4, 'f32', groups=groups, seed=3)
    eng.stats(reset=True)
    a = torch.zeros(50, 3, 4, device='cuda')
    eng.step(a, k=50)
    p1 = eng.pos.cpu().numpy().copy()
    st = eng.stats()
    assert st['env_steps'] == 4 * 50
    assert not np.array_equal(p1[:3, 0], p1[:3, 1])
    eng2 = S.merge_engine(4, 'f32', groups=groups, seed=3)
    eng2.step(a, k=50)
    np.testing.assert_array_equal(eng2.pos.cpu().numpy(), p1)


def merge_flow_params(F, n_initial=5, flow_rate=2000.0, rl_pen=0.1, num_rl=5, horizon=120):
    """examples/exp_configs/rl/singleagent/singleagent_merge.py, written against module F
    (flow_b200.flow_api here; the reference's own modules take the same calls)."""
    vehicles = F.VehicleParams()
    vehicles.add(veh_id='human', acceleration_controller=(F.IDMController, {}),
                 car_following_params=F.SumoCarFollowingParams(speed_mode='obey_safe_speed'),
                 num_vehicles=n_initial)
    vehicles.add(veh_id='rl', acceleration_controller=(F.RLController, {}),
                 car_following_params=F.SumoCarFollowingParams(speed_mode='obey_safe_speed'),
                 num_vehicles=0)
    inflow = F.InFlows()
    inflow.add(veh_type='human', edge='inflow_highway', vehs_per_hour=(1 - rl_pen) * flow_rate,
               depart_lane='free', depart_speed=10)
    inflow.add(veh_type='rl', edge='inflow_highway', vehs_per_hour=rl_pen * flow_rate,
               depart_lane='free', depart_speed=10)
    inflow.add(veh_type='human', edge='inflow_merge', vehs_per_hour=100, depart_lane='free',
               depart_speed=7.5)
    net = dict(F.MERGE_NET_PARAMS)
    net.update(merge_lanes=1, highway_lanes=1, pre_merge_length=500)
    return dict(
        exp_tag='stabilizing_open_network_merges', env_name=F.MergePOEnv, network=F.MergeNetwork,
        simulator='traci', sim=F.SumoParams(sim_step=0.2, render=False, restart_instance=True),
        env=F.EnvParams(horizon=horizon, sims_per_step=5, warmup_steps=0, additional_params={
            'max_accel': 1.5, 'max_decel': 1.5, 'target_velocity': 20, 'num_rl': num_rl}),
        net=F.NetParams(inflows=inflow, additional_params=net), veh=vehicles,
        initial=F.InitialConfig())


def test_merge_env_through_the_flow_api():
    """make_create_env on singleagent_merge.py-style flow_params: spaces, the initial placement
    (the reference-written start_pos_merge.json case), the gym step against the oracle, names."""
    import json
    from flow_b200 import flow_api as F
    from flow_b200 import placement as PL
    from oracle import merge as OM
    params = merge_flow_params(F)
    create_env, name = F.make_create_env(params)
    assert name.startswith('MergePOEnv-v')
    env = create_env()
    assert env.action_space.shape == (5,) and env.observation_space.shape == (25,)
    assert float(env.action_space.low[0]) == -1.5 and float(env.action_space.high[0]) == 1.5
    obs = env.reset()
    assert obs.shape == (25,) and not obs.any()
    case = json.load(open(os.path.join(os.path.dirname(GOLD), 'start_pos_merge.json')))[0]
    assert (case['pre'], case['n']) == (500, 5)
    assert [[env.k.vehicle.get_edge(v), env.k.vehicle.get_position(v)]
            for v in env.k.vehicle.get_ids()] == [[e, pytest.approx(p, abs=1e-9)] for e, p in case['pos']]
    assert env.k.vehicle.get_ids() == ['human_%d' % i for i in range(5)]
    assert env.k.network.length() == pytest.approx(200 + 500 + 100 + 100 + 0.2 + 22.5)
    # the same scenario on the oracle
    groups = S.merge_groups()
    initial = [(0,) + PL.merge_arm_position(e, p, 500) for e, p in case['pos']]
    sim = S.merge_sim(groups=groups, initial=initial, num_rl=5, horizon=120,
                      inflows=[(0, 0, 1800.0, 10.0, 1.0), (1, 0, 200.0, 10.0, 1.0),
                               (0, 1, 100.0, 7.5, 1.0)])
    rng = np.random.default_rng(3)
    for s in range(120):
        a = rng.uniform(-1.5, 1.5, 5)
        o, r, d, _ = env.step(a)
        o_obs, o_rew, o_done = sim.step(a)
        assert close(o, o_obs) <= 1e-9, s
        assert abs(r - o_rew) <= 1e-9 and d == o_done, s
        names = ['flow_%d.%d' % ((v >> 24) - 1, v & 0xFFFFFF) if v >= (1 << 24) else 'human_%d' % v
                 for v in sim.rl_veh]
        assert env.rl_veh == names, s
    assert d and len(env.rl_veh) >= 3
    kv = env.k.vehicle
    ids = kv.get_ids()
    assert len(ids) == sim.num_vehicles and kv.get_rl_ids() == sorted(
        n for n in ids if n.startswith('flow_1.'))
    v0 = env.rl_veh[0]
    s0 = sim.ids.index(sim.rl_veh[0])
    assert kv.get_type(v0) == 'rl'
    assert kv.get_x_by_id(v0) == pytest.approx(sim.g.x_abs(sim.arm[s0], sim.pos[s0]), abs=1e-9)
    lead = kv.get_leader(v0)
    assert (lead is None) == (sim.leader[s0] < 0)
    if lead is not None:
        assert kv.get_speed(lead) == pytest.approx(sim.vel[sim.leader[s0]], abs=1e-9)
    assert kv.get_headway(v0) == pytest.approx(sim.headway[s0], abs=1e-9)
    assert kv.get_inflow_rate(20) > 0 and kv.get_outflow_rate(20) >= 0
    env.terminate()


def test_merge_env_batched_and_error_cases():
    import torch
    from flow_b200 import flow_api as F
    params = merge_flow_params(F, horizon=50)
    create_env, _ = F.make_create_env(params, n_envs=64, dtype='f32')
    env = create_env()
    obs = env.reset()
    assert tuple(obs.shape) == (64, 25)
    for s in range(50):
        obs, rew, done, _ = env.step(torch.zeros(64, 5, device='cuda'))
    assert bool(done.all()) and float(rew.min()) >= 0.0
    assert int(env.engine.counters[2].min()) > 0           # vehicles reached the end in every env
    env.terminate()
    bad = merge_flow_params(F)
    bad['net'].additional_params['highway_lanes'] = 2
    with pytest.raises(NotImplementedError):
        F.make_create_env(bad)[0]()
    bad = merge_flow_params(F)
    bad['net'].inflows.add(veh_type='human', edge='center', vehs_per_hour=10, depart_speed=5)
    with pytest.raises(NotImplementedError):
        F.make_create_env(bad)[0]()
    bad = merge_flow_params(F)
    del bad['env'].additional_params['num_rl']
    with pytest.raises(KeyError):
        F.make_create_env(bad)[0]()
    bad = merge_flow_params(F)
    bad['network'] = F.RingNetwork
    bad['net'] = F.NetParams(additional_params=dict(F.RING_NET_PARAMS))
    with pytest.raises(NotImplementedError):
        F.make_create_env(bad)[0]()


def test_experiment_runs_on_the_junction_networks(capsys):
    """flow/core/experiment.py's loop (returns, mean speeds, outflow) over MergePOEnv and over
    AccelEnv on the figure eight, RL vehicles left to the simulator model."""
    from flow_b200 import flow_api as F
    from tests.test_gpu_figure8 import figure8_flow_params
    info = F.Experiment(merge_flow_params(F, horizon=80)).run(1)
    assert len(info['returns']) == 1 and info['returns'][0] > 0 and info['velocities'][0] > 1.0
    assert info['outflows'][0] >= 0
    info = F.Experiment(figure8_flow_params(F, horizon=80)).run(2)
    assert len(info['returns']) == 2 and info['returns'][0] == info['returns'][1] > 0
    assert 'Average, std returns' in capsys.readouterr().out
