"""InitialConfig(perturbation / spacing='random' / shuffle) - SURVEY 8a row a27's stochastic
branches. The reference draws them from Python's `random` and numpy's global generator on the
host; flow_b200/placement.py makes the same calls in the same order, so with the same seeds the
placements are the reference's, number for number (tests/golden/start_pos_stochastic.json was
written by the reference's own gen_even_start_pos / gen_random_start_pos / random.shuffle)."""
import json
import os
import random

import numpy as np
import pytest

from flow_b200 import placement as PL
from oracle import sim as OS

G = os.path.join(os.path.dirname(__file__), 'golden')
CASES = json.load(open(os.path.join(G, 'start_pos_stochastic.json')))


def _kw(c):
    return {k: (float('inf') if v == 'inf' else v) for k, v in c['kw'].items()}


@pytest.mark.parametrize('c', CASES, ids=['%s-%d' % ('random' if 'spacing' in c['kw'] else 'perturb',
                                                     c['seed']) for c in CASES])
def test_host_restatement_reproduces_the_reference_streams(c):
    kw = _kw(c)
    random.seed(c['seed']); np.random.seed(c['seed'])
    edges = PL.ring_edges(c['length'])
    if kw.get('spacing') == 'random':
        x, lane = PL.random_positions(edges, c['lanes'], kw.get('lanes_distribution', float('inf')),
                                      kw.get('min_gap', 0), kw.get('bunching', 0), c['n'])
    else:
        even = {k: v for k, v in kw.items() if k != 'perturbation'}
        x, lane = OS.gen_even_start_pos(c['length'], c['n'], c['lanes'], **even)
        x = PL.perturb(x, edges, kw['perturbation'])
    assert list(lane) == c['lane']
    assert x == c['x']                                       # bit for bit
    assert PL.shuffled(['test_%d' % i for i in range(c['n'])]) == c['shuffled_ids']


def test_perturbation_keeps_a_vehicle_on_its_edge_and_random_spacing_has_room_for_everyone():
    edges = PL.ring_edges(230.0)
    xs = PL.perturb([0.5, 57.0, 57.7, 229.9], edges, 1.0, normal=lambda m, s: -3.0)
    assert xs[0] == 0.0 and xs[1] == 54.0 and xs[2] == 57.6 and abs(xs[3] - 226.9) < 1e-12
    xs = PL.perturb([57.0, 229.9], edges, 1.0, normal=lambda m, s: 9.0)
    assert xs == [57.5, 57.5 * 3 + 0.3 + 57.5]               # clipped to the end of its edge
    rng = random.Random(3)
    x, lane = PL.random_positions(edges, 1, 1, 0.0, 0.0, 22, rand=rng.random)
    assert lane == [0] * 22 and all(b - a >= 5.0 for a, b in zip(x, x[1:]))
    with pytest.raises(ValueError):
        PL.random_positions(edges, 1, 1, 0.0, 0.0, 60)


def _ring_params(F, n=22, lanes=1, length=230, **init_kw):
    v = F.VehicleParams()
    v.add(veh_id='test', acceleration_controller=(F.IDMController, {}),
          car_following_params=F.SumoCarFollowingParams(min_gap=0),
          routing_controller=(F.ContinuousRouter, {}), num_vehicles=n)
    return dict(
        exp_tag='ring', env_name=F.AccelEnv, network=F.RingNetwork, simulator='traci',
        sim=F.SumoParams(render=False, sim_step=0.1),
        env=F.EnvParams(horizon=100, additional_params={
            'max_accel': 3, 'max_decel': 3, 'target_velocity': 10, 'sort_vehicles': False}),
        net=F.NetParams(additional_params={'length': length, 'lanes': lanes, 'speed_limit': 30,
                                           'resolution': 40}),
        veh=v, initial=F.InitialConfig(**init_kw), detector_params=None)


def test_compile_rules():
    from flow_b200 import flow_api as F
    from tests.test_flow_api import spec_of, ring_flow_params
    spec_of(_ring_params(F, perturbation=1.0))
    spec_of(_ring_params(F, spacing='random', shuffle=True))
    with pytest.raises(NotImplementedError):
        spec_of(_ring_params(F, spacing='custom'))
    fp = ring_flow_params()                       # two types (human + rl)
    fp['initial'].shuffle = True
    with pytest.raises(NotImplementedError):
        spec_of(fp)


# ------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize('c', [c for c in CASES if c['n'] in (22, 21, 5, 40)],
                         ids=lambda c: 'seed%d' % c['seed'])
def test_gpu_env_starts_where_the_reference_would(c):
    """make_create_env with the stochastic InitialConfig and the same seeds: the engine's state
    after reset (no warm-up) is the reference's placement (fp64, <= 1e-9 on the even part that
    the device generates, exact lanes), the ids are the reference's shuffled ids."""
    from flow_b200 import flow_api as F
    kw = _kw(c)
    fp = _ring_params(F, n=c['n'], lanes=c['lanes'], length=c['length'], **kw)
    random.seed(c['seed']); np.random.seed(c['seed'])
    env = F.make_create_env(fp, dtype='f64')[0]()          # __init__ draws: THE golden draw
    env.reset()                                            # no shuffle: re-uses that draw
    x = env.engine.pos.cpu().numpy()[:, 0]
    assert np.abs(x - np.array(c['x'])).max() <= 1e-9
    assert list(env.engine.lane.cpu().numpy()[:, 0]) == c['lane']
    x1 = x.copy()
    env.reset()
    assert np.array_equal(env.engine.pos.cpu().numpy()[:, 0], x1)      # same draw again
    env.terminate()


@pytest.mark.gpu
def test_gpu_shuffle_reorders_ids_and_keeps_the_positions():
    """The reference's own shuffle test (tests/fast_tests/test_environment_base_class.py:62-76):
    after a reset the ids sit elsewhere, the set of positions is the same."""
    from flow_b200 import flow_api as F
    fp = _ring_params(F, n=5, shuffle=True, x0=5)
    random.seed(1)
    env = F.make_create_env(fp, dtype='f64')[0]()
    env.reset()
    ids0 = env.k.vehicle.get_ids()
    before = sorted(env.k.vehicle.get_x_by_id(ids0))
    pos0 = dict(zip(ids0, env.k.vehicle.get_x_by_id(ids0)))
    orders = set()
    for _ in range(6):
        env.reset()
        ids = env.k.vehicle.get_ids()
        orders.add(tuple(ids))
        assert sorted(ids) == sorted(ids0)
        after = env.k.vehicle.get_x_by_id(ids)
        assert np.allclose(sorted(after), before, atol=1e-9)
    assert len(orders) > 1
    assert any(env.k.vehicle.get_x_by_id([v])[0] != pos0[v] for v in ids0)
    # and the stream is Python's: the same seed gives the same orders
    random.seed(1)
    env2 = F.make_create_env(fp, dtype='f64')[0]()
    env2.reset()
    assert env2.k.vehicle.get_ids() == ids0
    env.terminate(); env2.terminate()


@pytest.mark.gpu
def test_gpu_random_spacing_batched_and_stepping():
    """64 rings with spacing='random': every env has its own draw, all start collision-free,
    and the rollout runs (warm-up included)."""
    from flow_b200 import flow_api as F
    fp = _ring_params(F, n=22, spacing='random')
    fp['env'].warmup_steps = 20
    random.seed(5)
    env = F.make_create_env(fp, n_envs=64, dtype='f32')[0]()
    env.reset()
    x = env.engine.pos.cpu().numpy()
    assert len({tuple(np.round(x[:, e], 3)) for e in range(64)}) == 64
    for _ in range(50):
        env.step(None)
    assert int(env.engine.crash.sum()) == 0 and int(env.engine.time_counter[0]) == 70
    env.terminate()


MERGE_CASES = json.load(open(os.path.join(G, 'start_pos_merge.json')))


@pytest.mark.parametrize('c', MERGE_CASES, ids=['%d-%d-%d-n%d' % (c['pre'], c['merge'], c['post'], c['n'])
                                                for c in MERGE_CASES])
def test_even_placement_on_the_merge_network(c):
    """tests/golden/start_pos_merge.json: the reference's gen_even_start_pos on its own
    MergeNetwork edge tables."""
    real, internal = PL.merge_edges(c['pre'], c['merge'], c['post'])
    kw = c['kw']
    pos, lane = PL.even_positions(real, internal, kw.get('x0', 0), kw.get('min_gap', 0),
                                  kw.get('bunching', 0), c['n'])
    assert [[e, p] for e, p in pos] == c['pos']              # bit for bit
    assert lane == c['lane']
    for e, p in pos:
        arm, d = PL.merge_arm_position(e, p, c['pre'])
        assert arm == (1 if e in ('inflow_merge', 'bottom') else 0) and d >= p


def test_even_placement_refuses_an_overfull_network():
    real, internal = PL.merge_edges(50, 10, 20)
    with pytest.raises(ValueError):
        PL.even_positions(real, internal, 0, 0, 0, 80)
