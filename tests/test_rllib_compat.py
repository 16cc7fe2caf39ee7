"""flow_params.json round trip (FlowParamsEncoder / get_flow_params,
flow/utils/rllib.py:22-192) and the RLlib VectorEnv adapter."""
import json
import os

import numpy as np
import pytest

from flow_b200 import flow_api as F
from flow_b200 import rllib_compat as RC
from tests.test_flow_api import c1_flow_params, ring_flow_params, spec_of, struct_bytes
from tests.test_gpu_figure8 import figure8_flow_params
from tests.test_gpu_merge import merge_flow_params

G = os.path.join(os.path.dirname(__file__), 'golden')


def highway_flow_params(api=F):
    v = api.VehicleParams()
    v.add('human', acceleration_controller=(api.IDMController, {}), num_vehicles=10)
    inflow = api.InFlows()
    inflow.add(veh_type='human', edge='highway_0', vehs_per_hour=2000, depart_lane='free',
               depart_speed=20)
    return dict(exp_tag='highway', env_name=api.AccelEnv, network=api.HighwayNetwork,
                simulator='traci', sim=api.SumoParams(sim_step=0.1),
                env=api.EnvParams(horizon=3000, additional_params={
                    'max_accel': 3, 'max_decel': 3, 'target_velocity': 10, 'sort_vehicles': False}),
                net=api.NetParams(inflows=inflow, additional_params={
                    'length': 1000, 'lanes': 1, 'speed_limit': 30, 'num_edges': 1}),
                veh=v, initial=api.InitialConfig(), detector_params=None)


def same_spec(a, b):
    assert a.kw == b.kw and a.type_of == b.type_of and a.veh_ids == b.veh_ids
    assert [struct_bytes(x) for x in a.types] == [struct_bytes(y) for y in b.types]
    assert a.capacity == b.capacity


@pytest.mark.parametrize('mk', [lambda: ring_flow_params(noise=0.2), c1_flow_params,
                                highway_flow_params, lambda: merge_flow_params(F),
                                lambda: figure8_flow_params(F, noise=0.2)],
                         ids=['c2', 'c1', 'highway', 'merge', 'figure8'])
def test_round_trip_through_json(mk, tmp_path):
    fp = mk()
    txt = RC.dump_flow_params(fp, str(tmp_path / 'flow_params.json'))
    back = RC.get_flow_params(str(tmp_path / 'flow_params.json'))
    same_spec(spec_of(fp), spec_of(back))
    assert back['env_name'] is fp['env_name'] and back['network'] is fp['network']
    # the RLlib config form: JSON text under env_config.flow_params
    back2 = RC.get_flow_params({'env_config': {'flow_params': txt}})
    same_spec(spec_of(fp), spec_of(back2))
    # decoding adds the optional keys (tls) like the reference; from there it is a fixed point
    txt2 = RC.dump_flow_params(back)
    assert RC.dump_flow_params(RC.get_flow_params({'env_config': {'flow_params': txt2}})) == txt2


@pytest.mark.parametrize('name,mk', [('c2', lambda: ring_flow_params(noise=0.2)),
                                     ('c1', c1_flow_params), ('highway', highway_flow_params),
                                     ('merge', lambda: merge_flow_params(F)),
                                     ('figure8', lambda: figure8_flow_params(F, noise=0.2))])
def test_reads_files_written_by_the_reference_encoder(name, mk):
    """tests/golden/flow_params_*.json were produced by the reference's own
    FlowParamsEncoder from its own classes (make_flow_params_json.py)."""
    path = os.path.join(G, 'flow_params_%s.json' % name)
    back = RC.get_flow_params(path)
    same_spec(spec_of(mk()), spec_of(back))
    ref = json.load(open(path))
    ours = json.loads(RC.dump_flow_params(mk()))
    # same document up to the module path of the env / network classes
    for k in ('env_name', 'network'):
        assert ref[k].split('.')[-1] == ours[k].split('.')[-1]
        ref.pop(k); ours.pop(k)
    assert ref == ours


@pytest.mark.gpu
def test_vector_env_adapter():
    fp = ring_flow_params(noise=0.0)
    fp['env'].horizon = 30
    fp['env'].warmup_steps = 20
    n = 48
    vec = RC.B200VectorEnv(fp, n, dtype='f64', seed=4)
    one = F.make_create_env(fp, n_envs=1, dtype='f64', seed=4)[0]()
    obs = vec.vector_reset()
    assert len(obs) == n and obs[0].shape == (3,) and vec.get_unwrapped() == []
    np.testing.assert_array_equal(obs[0], one.reset())
    rng = np.random.default_rng(1)
    for s in range(30):
        a = rng.uniform(-1, 1, (n, 1))
        o, r, d, info = vec.vector_step(list(a))
        o1, r1, d1, _ = one.step(a[0])
        np.testing.assert_array_equal(o[0], o1)
        assert r[0] == r1 and bool(d[0]) == d1 and len(info) == n
    assert all(d)                                  # horizon reached everywhere
    # RLlib resets finished sub-envs one by one ...
    L_before = vec.env.engine.length.cpu().numpy().copy()
    o5 = vec.reset_at(5)
    tc = vec.env.engine.time_counter.cpu().numpy()
    assert tc[5] == 20 and tc[4] == 50 and o5.shape == (3,)
    assert not vec.env.engine.done.cpu().numpy()[5]
    # ... or all at once
    idx = vec.reset_done()
    assert len(idx) == n - 1 and (vec.env.engine.time_counter.cpu().numpy() == 20).all()
    assert (vec.env.engine.length.cpu().numpy() != L_before).any()    # fresh ring-length draws
    o, r, d, _ = vec.vector_step(np.zeros((n, 1)))
    assert not any(d)
    vec.close(); one.terminate()
