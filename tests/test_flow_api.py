"""The reference-facing Python interface (flow_b200.flow_api): a Flow user's
flow_params dict must compile to the same engine configuration the parity
tests use, raise the reference's exceptions, and — on the GPU — reproduce the
reference-generated golden trajectories through `make_create_env` /
`Env.reset` / `Env.step` / `env.k.vehicle.*` / `Experiment.run`."""
import ctypes
import os

import numpy as np
import pytest

from flow_b200 import engine as EN
from flow_b200 import flow_api as F
from flow_b200 import workloads as W

G = os.path.join(os.path.dirname(__file__), 'golden')
REF = '/root/reference'


def ring_flow_params(api=F, ctrl=None, env=None, noise=0.0, length=260, ring=(220, 270),
                     lanes=1, env_extra=None, n_human=21, n_rl=1, warmup=750, speed_mode=None):
    """examples/exp_configs/rl/singleagent/singleagent_ring.py, spelled with `api`'s classes.
    speed_mode None = the config's (the reference's default 'right_of_way'); the goldens were
    produced without SUMO's clamp, i.e. 'aggressive'."""
    ctrl = ctrl or api.IDMController
    sm = {} if speed_mode is None else {'speed_mode': speed_mode}
    v = api.VehicleParams()
    v.add(veh_id='human', acceleration_controller=(ctrl, {'noise': noise} if noise else {}),
          car_following_params=api.SumoCarFollowingParams(min_gap=0, **sm),
          routing_controller=(api.ContinuousRouter, {}), num_vehicles=n_human)
    if n_rl:
        v.add(veh_id='rl', acceleration_controller=(api.RLController, {}),
              car_following_params=api.SumoCarFollowingParams(**sm),
              routing_controller=(api.ContinuousRouter, {}), num_vehicles=n_rl)
    add = {'max_accel': 1, 'max_decel': 1, 'ring_length': list(ring)}
    add.update(env_extra or {})
    return dict(
        exp_tag='stabilizing_the_ring', env_name=env or api.WaveAttenuationPOEnv,
        network=api.RingNetwork, simulator='traci',
        sim=api.SumoParams(sim_step=0.1, render=False, restart_instance=False),
        env=api.EnvParams(horizon=3000, warmup_steps=warmup, clip_actions=False,
                          additional_params=add),
        net=api.NetParams(additional_params={'length': length, 'lanes': lanes, 'speed_limit': 30,
                                             'resolution': 40}),
        veh=v, initial=api.InitialConfig(bunching=50, min_gap=0), detector_params=None)


def c1_flow_params(api=F, speed_mode=None):
    """examples/exp_configs/non_rl/ring.py: 22 IDM, 230 m, AccelEnv."""
    v = api.VehicleParams()
    sm = {} if speed_mode is None else {'speed_mode': speed_mode}
    v.add(veh_id='idm', acceleration_controller=(api.IDMController, {}),
          car_following_params=api.SumoCarFollowingParams(**sm),
          routing_controller=(api.ContinuousRouter, {}), num_vehicles=22)
    return dict(
        exp_tag='ring', env_name=api.AccelEnv, network=api.RingNetwork, simulator='traci',
        sim=api.SumoParams(render=False, sim_step=0.1),
        env=api.EnvParams(horizon=3000, additional_params={
            'max_accel': 3, 'max_decel': 3, 'target_velocity': 10, 'sort_vehicles': False}),
        net=api.NetParams(additional_params={'length': 230, 'lanes': 1, 'speed_limit': 30,
                                             'resolution': 40}),
        veh=v, initial=api.InitialConfig(bunching=20), detector_params=None)


def spec_of(fp):
    net = F._resolve(fp['network'], F._NET_BY_NAME, 'network')(
        fp['exp_tag'], fp['veh'], fp['net'], fp.get('initial'))
    env = F._resolve(fp['env_name'], F._ENV_BY_NAME, 'environment')
    return F.compile_spec(env, fp['env'], fp['sim'], net)


def struct_bytes(s):
    return ctypes.string_at(ctypes.addressof(s), ctypes.sizeof(s))


def test_c2_flow_params_compile_to_the_bench_workload():
    spec = spec_of(ring_flow_params(noise=0.2))
    types, type_of, kw = W.ring_c2(noise=0.2)
    assert spec.type_of == type_of
    assert [struct_bytes(a) for a in spec.types] == [struct_bytes(b) for b in types]
    for k, v in kw.items():
        assert spec.kw[k] == v, k
    assert spec.veh_ids[0] == 'human_0' and spec.veh_ids[-1] == 'rl_0'


def test_c1_flow_params_compile():
    spec = spec_of(c1_flow_params())
    types, type_of, kw = W.ring_c1()
    assert spec.type_of == type_of
    t = spec.types[0]
    assert t.controller == EN.CTRL_IDM and list(t.p)[:6] == [30, 1, 1, 1.5, 4, 2]
    for k, v in kw.items():
        assert spec.kw[k] == v, k


@pytest.mark.skipif(not os.path.isdir(REF), reason='reference checkout not on this machine')
def test_reference_own_param_classes_are_accepted():
    """flow_params built with the REAL flow.core.params / flow.controllers /
    flow.envs / flow.networks classes compile to the same spec."""
    from tests.golden import ref_shim
    ref_shim.install()
    import types as _t
    import flow.controllers as rc
    import flow.core.params as rp
    import flow.envs as re_
    import flow.networks as rn
    api = _t.SimpleNamespace(
        VehicleParams=rp.VehicleParams, SumoCarFollowingParams=rp.SumoCarFollowingParams,
        SumoParams=rp.SumoParams, EnvParams=rp.EnvParams, NetParams=rp.NetParams,
        InitialConfig=rp.InitialConfig, IDMController=rc.IDMController,
        RLController=rc.RLController, ContinuousRouter=rc.ContinuousRouter,
        WaveAttenuationPOEnv=re_.WaveAttenuationPOEnv, RingNetwork=rn.RingNetwork,
        AccelEnv=re_.AccelEnv)
    for mk in (lambda a: ring_flow_params(a, noise=0.2), c1_flow_params):
        ours, theirs = spec_of(mk(F)), spec_of(mk(api))
        assert ours.kw == theirs.kw and ours.type_of == theirs.type_of
        assert [struct_bytes(a) for a in ours.types] == [struct_bytes(b) for b in theirs.types]
        assert ours.veh_ids == theirs.veh_ids


def test_reference_error_behaviour():
    fp = ring_flow_params()
    del fp['env'].additional_params['ring_length']
    with pytest.raises(KeyError):            # wave_attenuation.py:88-92
        spec_of(fp)
    fp = ring_flow_params()
    del fp['net'].additional_params['lanes']
    with pytest.raises(KeyError):            # ring.py:63-65
        spec_of(fp)
    with pytest.raises(F.FatalFlowError):    # network/base.py:603-605: 60 vehicles on 220 m
        spec_of(ring_flow_params(n_human=60))

    class MyController:
        pass
    with pytest.raises(NotImplementedError):  # user get_accel cannot be compiled: loud, no fallback
        spec_of(ring_flow_params(ctrl=MyController))
    with pytest.raises(TypeError):
        F.IDMController('x', bogus=1)


def test_sumo_autonomous_lane_changes_are_refused_not_dropped():
    fp = lane_change_po_flow_params()
    fp['veh'].initial[0]['lane_change_params'] = F.SumoLaneChangeParams(lane_change_mode='strategic')
    with pytest.raises(NotImplementedError):
        spec_of(fp)
    fp['veh'].initial[0]['lane_change_params'] = F.SumoLaneChangeParams(lane_change_mode='aggressive')
    spec_of(fp)


def test_controller_tables():
    fs = F.FollowerStopper('a', v_des=7)
    assert (fs.delay, fs.fail_safe, fs.engine_params()) == (1.0, 'safe_velocity', [7])
    assert F.PISaturation('a').delay == 1.0
    assert F.GippsController('a', delay=0.5).delay == 0.5
    assert F.OVMController('a', h_go=20).engine_params() == [1, 1, 2, 20, 30]
    cf = F.SumoCarFollowingParams(minGap=1.5, speed_mode='aggressive')
    assert cf.controller_params['minGap'] == 1.5 and cf.speed_mode == 0
    v = F.VehicleParams()
    v.add('h', num_vehicles=3)
    v.add('rl', acceleration_controller=(F.RLController, {}), num_vehicles=2)
    assert (v.num_vehicles, v.num_rl_vehicles, v.num_types) == (5, 2, 2)
    assert v.get_type('rl_1') == 'rl' and v.minGap['h'] == 2.5


def test_lane_change_direction_check_is_the_references():
    fp = ring_flow_params(env=F.LaneChangeAccelEnv, lanes=3, length=230, env_extra={
        'max_accel': 3, 'max_decel': 3, 'lane_change_duration': 5, 'target_velocity': 10,
        'sort_vehicles': False}, n_human=20, n_rl=2, warmup=0)
    spec = spec_of(fp)
    assert spec.types[1].lc_controller == EN.LC_RL and spec.kw['n_lanes'] == 3
    env = F.LaneChangeAccelEnv.__new__(F.LaneChangeAccelEnv)
    env.spec = spec
    with pytest.raises(ValueError):          # vehicle/traci.py:973-975
        env._check_host_actions(np.array([0.5, 2, 0.1, 0]))
    env._check_host_actions(np.array([0.5, -1, 0.1, 1]))


def lane_change_po_flow_params(api=F):
    """3-lane 230 m ring, 21 IDM + 1 RL, LaneChangeAccelPOEnv (tests/golden/ring_c3_po.npz, case a)."""
    v = api.VehicleParams()
    # the golden was produced without SUMO's speed-mode clamp: 'aggressive'
    v.add(veh_id='human', acceleration_controller=(api.IDMController, {}),
          car_following_params=api.SumoCarFollowingParams(speed_mode='aggressive'),
          routing_controller=(api.ContinuousRouter, {}), num_vehicles=21)
    v.add(veh_id='rl', acceleration_controller=(api.RLController, {}),
          car_following_params=api.SumoCarFollowingParams(speed_mode='aggressive'),
          routing_controller=(api.ContinuousRouter, {}), num_vehicles=1)
    return dict(
        exp_tag='lc_po', env_name=api.LaneChangeAccelPOEnv, network=api.RingNetwork,
        simulator='traci', sim=api.SumoParams(sim_step=0.1, render=False),
        env=api.EnvParams(horizon=3000, additional_params={
            'max_accel': 3, 'max_decel': 3, 'lane_change_duration': 5, 'target_velocity': 10,
            'sort_vehicles': False}),
        net=api.NetParams(additional_params={'length': 230, 'lanes': 3, 'speed_limit': 30,
                                             'resolution': 40}),
        veh=v, initial=api.InitialConfig(lanes_distribution=float('inf')), detector_params=None)


def test_lane_change_po_env_compiles():
    spec = spec_of(lane_change_po_flow_params())
    assert spec.kw['env_kind'] == EN.ENV_LANE_CHANGE_ACCEL_PO and spec.kw['n_lanes'] == 3
    assert spec.types[1].lc_controller == EN.LC_RL


def test_evaluate_and_sort_vehicles_compile_to_env_flags():
    fp = c1_flow_params()
    fp['env'].evaluate = True
    fp['env'].additional_params['sort_vehicles'] = True
    kw = spec_of(fp).kw
    assert kw['evaluate'] == 1 and kw['sort_vehicles'] == 1
    fp = ring_flow_params()                      # WaveAttenuationPOEnv reads neither
    fp['env'].evaluate = True
    fp['env'].additional_params['sort_vehicles'] = True
    kw = spec_of(fp).kw
    assert kw['evaluate'] == 0 and kw['sort_vehicles'] == 0


def test_highway_inflow_compiles():
    v = F.VehicleParams()
    v.add('human', acceleration_controller=(F.IDMController, {}), num_vehicles=10)
    inflow = F.InFlows()
    inflow.add(veh_type='human', edge='highway_0', vehs_per_hour=2000, depart_lane='free',
               depart_speed=20)
    net = F.HighwayNetwork('hw', v, F.NetParams(inflows=inflow, additional_params={
        'length': 1000, 'lanes': 1, 'speed_limit': 30, 'num_edges': 1}), F.InitialConfig())
    spec = F.compile_spec(F.AccelEnv, F.EnvParams(horizon=3000, additional_params=dict(
        F.AccelEnv.ADDITIONAL)), F.SumoParams(), net)
    assert spec.kw['topology'] == EN.TOPO_HIGHWAY and spec.kw['env_kind'] == EN.ENV_HIGHWAY
    assert spec.kw['inflow_vehs_per_hour'] == 2000 and spec.kw['inflow_depart_speed'] == 20
    assert spec.capacity % 32 == 0 and spec.capacity >= 160


# ------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_gpu_experiment_c1_matches_reference_golden():
    """Experiment(flow_params).run through the mirrored API == the trajectory the
    REAL reference controllers produced (tests/golden/ring_c1.npz)."""
    z = np.load(os.path.join(G, 'ring_c1.npz'))
    fp = c1_flow_params(speed_mode='aggressive')
    create_env, name = F.make_create_env(fp)
    assert name.startswith('AccelEnv-v')
    env = create_env()
    obs = env.reset()
    assert obs.shape == (44,) and env.observation_space.shape == (44,)
    assert env.action_space.shape == (0,)
    ids = env.k.vehicle.get_ids()
    assert ids[:2] == ['idm_0', 'idm_1'] and len(ids) == 22
    rets = 0.0
    for s in range(300):
        obs, r, done, _ = env.step(None)
        rets += r
        np.testing.assert_allclose(obs, z['obs'][s], rtol=1e-9, atol=1e-12)
        assert abs(r - z['reward'][s]) <= 1e-9 * max(1.0, abs(z['reward'][s]))
        np.testing.assert_allclose(env.k.vehicle.get_speed(ids), z['vel'][s + 1], rtol=1e-9,
                                   atol=1e-12)
        np.testing.assert_allclose(env.k.vehicle.get_x_by_id(ids), z['pos'][s + 1], rtol=1e-9,
                                   atol=1e-12)
    # getters: leader of i is i+1 on a single-lane ring; headway = gap to it
    assert env.k.vehicle.get_leader('idm_3') == 'idm_4'
    assert env.k.vehicle.get_follower('idm_0') == 'idm_21'
    x = np.array(env.k.vehicle.get_x_by_id(ids))
    C = env.k.network.length()
    assert abs(C - 230.4) < 1e-12 and env.k.network.non_internal_length() == pytest.approx(230.0)
    h = env.k.vehicle.get_headway('idm_3')
    assert h == pytest.approx((x[4] - x[3]) % C - 5.0, abs=1e-9)
    assert env.k.vehicle.get_speed('nobody') == -1001
    e, p = env.k.vehicle.get_edge('idm_0'), env.k.vehicle.get_position('idm_0')
    assert e in ('bottom', 'right', 'top', 'left', ':right_0', ':top_0', ':left_0', ':bottom_0')
    assert 0 <= p <= 57.5 + 1e-9
    assert not env.k.simulation.check_collision()
    env.terminate()

    exp = F.Experiment(c1_flow_params(speed_mode='aggressive'))
    exp.env.env_params.horizon = 300
    info = exp.run(1)
    assert info['returns'][0] == pytest.approx(float(z['reward'][:300].sum()), rel=1e-9)


@pytest.mark.gpu
def test_gpu_wave_attenuation_po_env_matches_golden():
    """WaveAttenuationPOEnv through make_create_env vs ring_c2.npz (reference
    controllers + reward + observation; warm-up replaced by a state injection
    because the golden run starts from a stated position vector)."""
    z = np.load(os.path.join(G, 'ring_c2.npz'))
    for tag in ('a', 'b'):
        fp = ring_flow_params(warmup=0, speed_mode='aggressive')
        env = F.make_create_env(fp)[0]()
        env.reset()
        env.engine.reset(lengths=[float(z['length_' + tag])])   # the golden run's ring length
        env.engine.set_state(pos=z['pos_' + tag][0][:, None], vel=z['vel_' + tag][0][:, None])
        for s in range(600):
            a = z['actions_' + tag][s]
            obs, r, done, _ = env.step(None if np.isnan(a) else [a])
            np.testing.assert_allclose(obs, z['obs_' + tag][s + 1], rtol=1e-9, atol=1e-12)
            assert abs(r - z['reward_' + tag][s]) <= 1e-9 * max(1.0, abs(z['reward_' + tag][s]))
        assert env.k.vehicle.get_rl_ids() == ['rl_0']
        assert env.k.vehicle.num_rl_vehicles == 1
        env.terminate()


@pytest.mark.gpu
def test_gpu_lane_change_po_env_matches_golden():
    """LaneChangeAccelPOEnv through the mirrored API == the reference's get_state on its own
    multi-lane bookkeeping; spaces as the reference declares them."""
    z = np.load(os.path.join(G, 'ring_c3_po.npz'))
    env = F.make_create_env(lane_change_po_flow_params())[0]()
    obs = env.reset()
    assert env.observation_space.shape == (13,) and env.action_space.shape == (2,)
    assert obs.shape == (13,) and env.num_lanes == 3
    for s, a in enumerate(z['actions_a'][:150]):
        obs, r, done, _ = env.step(None if np.isnan(a).any() else a)
        np.testing.assert_allclose(obs, z['obs_a'][s], rtol=1e-9, atol=1e-9)
        assert abs(r - z['reward_a'][s]) <= 1e-9
        assert not done
    assert env.k.vehicle.get_lane('rl_0') == int(z['lane_a'][150][21])
    with pytest.raises(ValueError):          # vehicle/traci.py:973-975
        env.step(np.array([0.1, 0.5]))
    env.terminate()
    full = F.make_create_env(lane_change_po_flow_params(), lc_corrected=True)[0]()
    o = full.reset()
    assert o.shape == (13,) and 0 <= o.min() and o.max() <= 1     # one RL: same length, normalised
    full.terminate()


@pytest.mark.gpu
def test_gpu_batched_env_and_action_layout():
    import torch
    fp = ring_flow_params(noise=0.0)
    env = F.make_create_env(fp, n_envs=64, dtype='f64', seed=5)[0]()
    one = F.make_create_env(fp, n_envs=1, dtype='f64', seed=5)[0]()
    obs = env.reset()
    o1 = one.reset()
    assert obs.shape == (64, 3) and obs.is_cuda
    np.testing.assert_allclose(obs[0].cpu().numpy(), o1, rtol=0, atol=0)   # env 0 == the single env
    rng = np.random.default_rng(0)
    for s in range(20):
        a = rng.uniform(-1, 1, size=(64, 1))
        obs, rew, done, _ = env.step(a)
        o1, r1, d1, _ = one.step(a[0])
        np.testing.assert_allclose(obs[0].cpu().numpy(), o1, rtol=0, atol=0)
        assert float(rew[0]) == r1
    with pytest.raises(ValueError):
        env.step(np.zeros(7))
    with pytest.raises(RuntimeError):
        env.k.vehicle.get_headway('human_0')      # views are off on a batched env by default
    env.terminate(); one.terminate()


@pytest.mark.gpu
@pytest.mark.parametrize('n_init', [0, 10])
def test_gpu_highway_env_rates(n_init):
    v = F.VehicleParams()
    v.add('human', acceleration_controller=(F.IDMController, {}), num_vehicles=n_init)
    inflow = F.InFlows()
    inflow.add(veh_type='human', edge='highway_0', vehs_per_hour=2000, depart_speed=20)
    fp = dict(exp_tag='hw', env_name=F.AccelEnv, network=F.HighwayNetwork, simulator='traci',
              sim=F.SumoParams(sim_step=0.1),
              env=F.EnvParams(horizon=3000, additional_params=dict(F.AccelEnv.ADDITIONAL)),
              net=F.NetParams(inflows=inflow, additional_params={
                  'length': 1000, 'lanes': 1, 'speed_limit': 30, 'num_edges': 1}),
              veh=v, initial=F.InitialConfig())
    env = F.make_create_env(fp)[0]()
    env.reset()
    assert env.k.vehicle.num_vehicles == n_init
    for _ in range(1500):
        obs, r, done, _ = env.step(None)
    assert not done
    n = env.k.vehicle.num_vehicles
    assert n > n_init and len(env.k.vehicle.get_speed(env.k.vehicle.get_ids())) == n
    c = env.engine.counters[:, 0].cpu().numpy()
    assert n == c[0] == c[1] - c[2]
    rate_in, rate_out = env.k.vehicle.get_inflow_rate(500), env.k.vehicle.get_outflow_rate(500)
    if n_init == 0:
        # free entry: 2000 veh/h = one departure per 1.8 s (vehicle/traci.py:493-505 formulas)
        assert rate_in == pytest.approx(2000, rel=0.03)
        assert rate_out > 0
    else:
        # ten stationary vehicles queue at the entry: departures are held back (blocked insertion)
        assert 0 < rate_in < 2000
    env.terminate()


@pytest.mark.gpu
def test_gpu_emission_csv_matches_the_reference_sumo_fixture(tmp_path):
    """Experiment.run(convert_to_csv=True) writes `<name>-emission.csv` with the
    reference's columns; the rows reproduce tests/fast_tests/test_files/
    ring_230_emission.csv (real SUMO output) where SUMO prints them: speed to
    0.01, edge ids, displacement to 0.01 for t <= 0.5."""
    import csv
    import json
    from flow_b200 import emission
    fp = c1_flow_params()
    fp['sim'].emission_path = str(tmp_path)
    fp['env'].horizon = 6
    exp = F.Experiment(fp)
    exp.run(1, convert_to_csv=True)
    path = os.path.join(str(tmp_path), 'ring-emission.csv')
    rows = list(csv.DictReader(open(path)))
    assert list(rows[0].keys()) == emission.COLUMNS
    assert len(rows) == 22 * 7 and [r['id'] for r in rows] == sorted(r['id'] for r in rows)
    ours = {(round(float(r['time']), 1), r['id']): r for r in rows}
    ref = json.load(open(os.path.join(G, 'sumo_ring_230.json')))['rows']
    by = {(round(r['time'], 1), r['id']): r for r in ref}
    n = 0
    for (t, vid), r in by.items():
        if t > 0.5 or (t, vid) not in ours:
            continue
        o, o0, r0 = ours[(t, vid)], ours[(0.1, vid)], by[(0.1, vid)]
        assert abs(float(o['speed']) - r['speed']) <= 0.0101, (t, vid)
        assert o['type'] == 'idm' and o['lane_number'] == '0'
        if r['edge'] == r0['edge'] and o['edge_id'] == o0['edge_id']:
            d_ref = r['rel_pos'] - r0['rel_pos']
            d_our = float(o['relative_position']) - float(o0['relative_position'])
            assert abs(d_our - d_ref) <= 0.0301, (t, vid)
        n += 1
    assert n >= 22 * 5
    # the reference's time-space reader needs exactly these columns
    for k in ('time', 'id', 'edge_id', 'relative_position', 'speed', 'lane_number'):
        assert k in rows[0]
