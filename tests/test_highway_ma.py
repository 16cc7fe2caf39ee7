"""MultiAgentHighwayPOEnv on the single-lane open highway.

Golden: tests/golden/highway_ma.npz holds, step by step, what the REFERENCE's own
MultiAgentHighwayPOEnv.get_state / compute_reward (flow/envs/multiagent/highway.py:97-172)
return on the restated highway (tests/golden/make_golden.py::highway_ma; SUMO's move,
insertion and arrival are the restatement, "parity unpinned"). CPU: the oracle replays it
bit-exactly. GPU: the CUDA path through the C-ABI against golden and oracle - agent ids
and slot occupancy bit-exact, fp64 1e-9, fp32 1e-4 per step."""
import os

import numpy as np
import pytest

from tests import scenarios as S

G = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'highway_ma.npz'))
N_STEPS = G['actions'].shape[0]
A = G['ids'].shape[1]


def close(a, b, floor=1.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float((np.abs(a - b) / np.maximum(np.abs(b), floor)).max()) if a.size else 0.0


def golden_actions(t):
    """{vehicle id: action} of golden step t, or None (no command)."""
    row = G['actions'][t]
    if np.isnan(row).all():
        return None
    return {int(v): float(a) for v, a in zip(G['ids'][t], row) if v >= 0}


def test_oracle_replays_the_reference_golden():
    sim = S.highway_ma_sim()
    seen = set()
    for t in range(N_STEPS + 1):
        st = sim.ma_state()
        ids = [int(v) for v in G['ids'][t] if v >= 0]
        assert list(st) == ids, t
        for a, vid in enumerate(ids):
            assert np.array_equal(st[vid], G['obs'][t, a]), t
        if t > 0:
            rw = sim.ma_reward(False)
            assert [rw[v] for v in ids] == list(G['reward'][t, :len(ids)]), t
        assert len(sim.pos) == G['n_veh'][t]
        seen |= set(ids)
        if t < N_STEPS:
            sim.step(golden_actions(t))
    # initial RL vehicle (id 1) and vehicles of inflow 0 (ids (1 << 24) | k)
    assert 1 in seen and sum(v >= (1 << 24) for v in seen) >= 10
    assert all((v >> 24) in (0, 1) for v in seen)      # the RL flow is inflow 0


def test_oracle_inflows_are_served_in_definition_order():
    sim = S.highway_ma_sim(inflows=[(1, 3600.0, 10.0, 1.0), (0, 3600.0, 10.0, 1.0)], x0=300.0)
    for _ in range(200):
        sim.step()
    # both due every second; the entry takes one vehicle per 7.5 m of free road
    assert sim.spawned_by_flow[0] >= sim.spawned_by_flow[1] > 0
    assert sorted(sim.pos, reverse=True) == sim.pos
    assert len(set(sim.ids)) == len(sim.ids)


def test_oracle_agent_without_action_keeps_the_simulator_model():
    a, b = S.highway_ma_sim(), S.highway_ma_sim()
    for t in range(60):
        a.step(None)
        b.step({vid: float('nan') for vid in b.ids})
    assert a.pos == b.pos and a.vel == b.vel


# --------------------------------------------------------------------------- GPU
def _engine_actions(eng, acts):
    """{id: accel} -> [A, E] tensor in the engine's agent-row order (NaN = none)."""
    import torch
    ids = eng.agent_id.cpu().numpy()                    # [E, A]
    out = np.full(ids.shape, np.nan)
    for e in range(ids.shape[0]):
        for a, v in enumerate(ids[e]):
            if v >= 0 and acts and int(v) in acts:
                out[e, a] = acts[int(v)]
    td = torch.float64 if eng.dtype == 'f64' else torch.float32
    return torch.as_tensor(out.T.copy(), dtype=td, device='cuda')


@pytest.mark.gpu
@pytest.mark.parametrize('dtype,tol', [('f64', 1e-9), ('f32', 1e-4)])
def test_gpu_ma_highway_vs_reference_golden(dtype, tol):
    eng = S.highway_ma_engine(3, dtype)
    eng.reset()
    assert eng.act_dim == A and eng.obs_dim == 5 * A
    sim = S.highway_ma_sim()
    n_check = N_STEPS if dtype == 'f64' else 250   # fp32 drifts; ids stay exact while it is close
    for t in range(n_check + 1):
        ids = eng.agent_id.cpu().numpy()
        assert (ids == G['ids'][t]).all(), t
        obs = eng.agent_obs.cpu().numpy()
        assert close(obs, np.broadcast_to(G['obs'][t], obs.shape), floor=1e-2 if dtype == 'f64' else 1.0) <= tol, t
        if t > 0:
            rw = eng.agent_reward.cpu().numpy()
            assert close(rw, np.broadcast_to(G['reward'][t], rw.shape)) <= tol, t
            n = len(sim.pos)
            assert (eng.counters.cpu().numpy()[0] == n).all(), t
            assert (eng.veh_id.cpu().numpy()[1, :n] == sim.ids).all(), t
            assert close(eng.pos.cpu().numpy()[:n, 2], sim.pos) <= tol, t
        if t == n_check:
            break
        acts = golden_actions(t)
        eng.step(None if acts is None else _engine_actions(eng, acts))
        sim.step(acts)
    assert not eng.crash.cpu().numpy().any()


@pytest.mark.gpu
def test_gpu_ma_highway_evaluate_and_missing_actions():
    """evaluate: reward = the agent's speed; an agent with a NaN action row keeps the restated
    simulator model (the reference applies nothing to a vehicle missing from rl_actions)."""
    eng = S.highway_ma_engine(2, 'f64', evaluate=True)
    eng.reset()
    sim = S.highway_ma_sim(evaluate=True)
    rng = np.random.default_rng(3)
    for t in range(400):
        acts = {}
        for slot in sim.rl_slots():
            if rng.random() < 0.7:
                acts[sim.ids[slot]] = S.highway_ma_action(sim, slot, rng.normal(0, 0.3))
        eng.step(_engine_actions(eng, acts) if acts else None)
        sim.step(acts or None)
        assert not sim.crash, t
        n = len(sim.pos)
        assert close(eng.pos.cpu().numpy()[:n, 0], sim.pos) <= 1e-9, t
        rw = sim.ma_reward(False)
        ids = [int(v) for v in eng.agent_id.cpu().numpy()[1] if v >= 0]
        assert ids == list(rw), t
        assert close(eng.agent_reward.cpu().numpy()[1, :len(ids)], [rw[v] for v in ids]) <= 1e-9, t
    assert sim.arrived > 10


@pytest.mark.gpu
def test_gpu_ma_highway_fused_launch_equals_single_steps():
    import torch
    a = S.highway_ma_engine(5, 'f64')
    b = S.highway_ma_engine(5, 'f64')
    a.reset(); b.reset()
    K = 300
    acts = torch.full((K, A, 5), float('nan'), dtype=torch.float64, device='cuda')
    acts[:, 0] = 0.4      # the front-most agent gets a command, the others none
    obs_seq = torch.zeros((K, 5, 5 * A), dtype=torch.float64, device='cuda')
    a.step(acts, k=K, obs_seq=obs_seq)
    for s in range(K):
        b.step(acts[s])
        assert torch.equal(obs_seq[s], b.obs), s
    for name in ('pos', 'vel', 'veh_id', 'agent_id', 'agent_reward', 'counters'):
        assert torch.equal(getattr(a, name), getattr(b, name)), name


@pytest.mark.gpu
def test_gpu_rl_vehicles_need_the_multi_agent_env():
    from flow_b200 import engine as EN
    types = S.engine_types(S.highway_ma_groups())
    with pytest.raises(EN.FlowStepError, match='FS_ENV_MA_HIGHWAY_PO'):
        EN.Engine(types, [0, 1, 0], 2, capacity=64, topology=EN.TOPO_HIGHWAY,
                  env_kind=EN.ENV_HIGHWAY, length_lo=600.0, length_hi=600.0)
    with pytest.raises(EN.FlowStepError, match='0..4 flows'):
        S.highway_ma_engine(1, inflows=[(0, 100.0, 10.0, 1.0)] * 5)


@pytest.mark.gpu
def test_gpu_flow_api_multiagent_highway_dicts():
    """The reference-facing class: dict-keyed obs / reward / done with SUMO's vehicle names,
    arrivals reported once with done = True, reward 0, observation None."""
    from flow_b200 import flow_api as F
    veh = F.VehicleParams()
    veh.add('human', acceleration_controller=(F.IDMController, {}),
            car_following_params=F.SumoCarFollowingParams(speed_mode='aggressive'), num_vehicles=0)
    veh.add('rl', acceleration_controller=(F.RLController, {}),
            car_following_params=F.SumoCarFollowingParams(speed_mode='aggressive'), num_vehicles=1)
    flows = F.InFlows()
    flows.add(veh_type='rl', edge='highway_0', vehs_per_hour=360, depart_speed=10, begin=5,
              name='rl_in')
    flows.add(veh_type='human', edge='highway_0', vehs_per_hour=720, depart_speed=12, begin=5)
    net = F.NetParams(inflows=flows, additional_params=dict(
        length=600, lanes=1, speed_limit=30, num_edges=1))
    env_params = F.EnvParams(horizon=1500, additional_params=dict(
        max_accel=1, max_decel=1, target_velocity=25, max_agents=8))
    network = F.HighwayNetwork('hw', veh, net, F.InitialConfig(x0=100))
    env = F.MultiAgentHighwayPOEnv(env_params, F.SumoParams(sim_step=0.2), network)
    obs = env.reset()
    assert list(obs) == ['rl_0'] and obs['rl_0'].shape == (5,)
    assert env.action_space.shape == (1,) and env.observation_space.shape == (5,)
    o, r, d, _ = env.step(None)
    assert r == {} and d == {'rl_0': False, '__all__': False}
    names, arrived = set(), []
    for t in range(600):
        # speed matching on the observed closing speed (30 m/s "leader" when there is none)
        acts = {k: [float(np.clip(2.0 * v[1] * 30, -3.0, 0.5))] for k, v in o.items()
                if v is not None}
        o, r, d, info = env.step(acts)
        names |= set(o)
        for k, v in o.items():
            if v is None:
                arrived.append(k)
                assert d[k] is True and r[k] == 0
            else:
                assert d[k] is False and 0 <= r[k] <= 1 and v.shape == (5,)
        if d['__all__']:
            break
    assert 'rl_0' in arrived and len(arrived) == len(set(arrived)) >= 3
    assert any(n.startswith('rl_in_0.') for n in names)
    env.terminate()


@pytest.mark.gpu
def test_gpu_flow_api_multiagent_highway_batched_tensors():
    """n_envs > 1: tensors [E, A, 5] / [E, A] / [E] + agent ids; every env of the batch equals
    the single-env dict run under the same policy (fp64)."""
    import torch
    from flow_b200 import flow_api as F

    def build(n_envs):
        veh = F.VehicleParams()
        veh.add('human', acceleration_controller=(F.IDMController, {}),
            car_following_params=F.SumoCarFollowingParams(speed_mode='aggressive'), num_vehicles=0)
        veh.add('rl', acceleration_controller=(F.RLController, {}),
            car_following_params=F.SumoCarFollowingParams(speed_mode='aggressive'), num_vehicles=1)
        flows = F.InFlows()
        flows.add(veh_type='rl', edge='highway_0', vehs_per_hour=360, depart_speed=10, begin=5,
                  name='rl_in')
        flows.add(veh_type='human', edge='highway_0', probability=0.15, depart_speed=12, begin=5)
        net = F.NetParams(inflows=flows, additional_params=dict(
            length=600, lanes=1, speed_limit=30, num_edges=1))
        env_params = F.EnvParams(horizon=1500, additional_params=dict(
            max_accel=1, max_decel=1, target_velocity=25, max_agents=8))
        network = F.HighwayNetwork('hw', veh, net, F.InitialConfig(x0=100))
        return F.MultiAgentHighwayPOEnv(env_params, F.SumoParams(sim_step=0.2), network,
                                        n_envs=n_envs, dtype='f64', seed=9)
    one, many = build(1), build(3)
    o1, om = one.reset(), many.reset()
    assert om.shape == (3, 8, 5)
    for t in range(250):
        a = torch.clamp(2.0 * om[:, :, 1] * 30, -3.0, 0.5)
        a1 = {k: [float(np.clip(2.0 * v[1] * 30, -3.0, 0.5))] for k, v in o1.items() if v is not None}
        om, rm, dm, info = many.step(a)
        o1, r1, d1, _ = one.step(a1)
        ids = info['agent_id'][0].cpu().numpy()        # env 0 of the batch = the single env (gid 0)
        names = [one.vehicle_name(v) for v in ids if v >= 0]
        assert names == [k for k, v in o1.items() if v is not None], t
        for row, k in enumerate(names):
            assert np.array_equal(om[0, row].cpu().numpy(), o1[k]), (t, k)
            assert float(rm[0, row]) == r1[k], (t, k)
        assert bool(dm[0]) == d1['__all__']
    # the probabilistic human inflow is keyed by the env id: the envs of the batch differ
    c = many.engine.counters.cpu().numpy()
    assert len(set(c[1])) > 1
    one.terminate(); many.terminate()
