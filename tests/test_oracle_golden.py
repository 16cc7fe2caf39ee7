"""CPU: the oracle restatement against the committed golden vectors
(tests/golden/*, produced by the reference's own code — make_golden.py) and
the reference's known-answer tests. Bit-exact / 1e-12 where stated."""
import json
import os

import numpy as np
import pytest

from oracle import controllers as OC
from oracle import sim as OS
from tests import scenarios as S

G = os.path.join(os.path.dirname(__file__), 'golden')
KAT = json.load(open(os.path.join(G, 'kat.json')))

PARAM_ORDER = dict(
    CFM=['k_d', 'k_v', 'k_c', 'd_des', 'v_des'], BCM=['k_d', 'k_v', 'k_c', 'd_des', 'v_des'],
    OVM=['alpha', 'beta', 'h_st', 'h_go', 'v_max'], LINEAR_OVM=['v_max', 'adaptation', 'h_st'],
    IDM=['v0', 'T', 'a', 'b', 'delta', 's0'], LAC=['k_1', 'k_2', 'h', 'tau'],
    GIPPS=['v0', 'acc', 'b', 'b_l', 's0', 'tau'], FOLLOWER_STOPPER=['v_des'],
    NONLOCAL_FOLLOWER_STOPPER=[], PI_SATURATION=[])


def kat_groups(case):
    ctrl = case['controller']
    p = [case['params'][k] for k in PARAM_ORDER[ctrl]]
    if ctrl == 'LAC':
        p.append(0.0)
    return [S.vt(ctrl, 5, p=p, max_accel=case['cf']['accel'], max_decel=case['cf']['decel'])]


def ring_for_headways(headways):
    """Loop positions whose derived headways equal the injected ones of the
    reference tests (set_headway): x_{i+1} = x_i + h_i + 5, C = sum(h_i + 5)."""
    x = [0.0]
    for h in headways[:-1]:
        x.append(x[-1] + h + 5.0)
    return x, float(sum(headways) + 5.0 * len(headways))


@pytest.mark.parametrize('case', KAT['controllers'], ids=lambda c: c['name'])
def test_controller_known_answers(case):
    """tests/fast_tests/test_controllers.py golden accelerations."""
    g = kat_groups(case)
    x, circ = ring_for_headways(case['headways'])
    o = S.make_oracle(g, 230)
    if min(case['headways']) <= 0:
        # headway injection with no consistent geometry (zero headway case)
        o.headway = [float(h) for h in case['headways']]
        if case['speeds']:
            o.vel = [float(v) for v in case['speeds']]
    else:
        o.circ = circ
        o.set_state(x, case['speeds'] or [0.0] * 5)
        assert np.allclose(o.headway, case['headways'], atol=1e-12)
    got = [OC.get_action(o, i, o.types[0], o.ctrl_state[i], 0.1, 30.0) for i in range(5)]
    assert np.allclose(got, case['reference_value'], rtol=1e-12, atol=1e-12)
    if case['expected'] is not None:
        np.testing.assert_array_almost_equal(got, case['expected'])


def test_reward_known_answers():
    r = {c['name']: c for c in KAT['rewards']}
    c = r['desired_velocity']
    o = S.make_oracle([S.vt('IDM', c['n'])], 230, target_velocity=c['target'])
    assert o.desired_velocity(fail=True) == 0
    assert o.desired_velocity() == 0
    o.vel = list(c['speeds'])
    assert abs(o.desired_velocity() - c['reference_value']) < 1e-15
    np.testing.assert_almost_equal(o.desired_velocity(), c['expected'])
    for case in r['wave_attenuation_reward']['cases']:
        o = S.make_oracle([S.vt('IDM', 1), S.vt('RL', 1)], 230,
                          env_kind=OS.ENV_WAVE_ATTENUATION_PO, act_lo=-1, act_hi=1)
        o.vel = [float(v) for v in case['speeds']]
        got = o.compute_reward(case['actions'], case['fail'])
        np.testing.assert_almost_equal(got, case['expected'])
        assert abs(got - case['reference_value']) < 1e-15


def test_start_positions_match_reference():
    for c in json.load(open(os.path.join(G, 'start_pos.json'))):
        kw = {k: (float('inf') if v == 'inf' else v) for k, v in c['kw'].items()}
        x, lane = OS.gen_even_start_pos(c['length'], c['n'], c['lanes'], **kw)
        assert lane == c['lane']
        assert np.array_equal(np.array(x), np.array(c['x'])), c


def test_not_enough_space_raises():
    with pytest.raises(ValueError):
        OS.gen_even_start_pos(100, 22, 1, bunching=20)


def test_c1_trajectory_bit_exact():
    z = np.load(os.path.join(G, 'ring_c1.npz'))
    o = S.make_oracle(S.c1_groups(), 230, bunching=20.0)
    assert np.array_equal(np.array(o.pos), z['pos'][0])
    rew = []
    for s in range(3000):
        obs, r, done = o.step(None)
        rew.append(r)
        if s < 300:
            assert np.array_equal(np.array(o.pos), z['pos'][s + 1]), s
            assert np.array_equal(np.array(o.vel), z['vel'][s + 1]), s
            assert np.array_equal(np.array(o.accel), z['accel'][s]), s
            assert np.array_equal(obs, z['obs'][s]), s
    assert done
    assert np.array_equal(np.array(rew), z['reward'])
    assert np.array_equal(np.array(o.pos), z['pos_final'])
    assert np.array_equal(np.array(o.vel), z['vel_final'])


def test_c2_trajectory_bit_exact():
    z = np.load(os.path.join(G, 'ring_c2.npz'))
    for tag in 'ab':
        o = S.make_oracle(S.c2_groups(), float(z['length_' + tag]),
                          env_kind=OS.ENV_WAVE_ATTENUATION_PO, bunching=50.0, warmup_steps=750,
                          horizon=3000, act_lo=-1, act_hi=1, obs_max_length=270.0)
        assert o.time_counter == 750
        assert np.array_equal(np.array(o.pos), z['pos_' + tag][0])
        assert np.array_equal(np.array(o.vel), z['vel_' + tag][0])
        assert np.array_equal(o.get_state(), z['obs_' + tag][0])
        for s, a in enumerate(z['actions_' + tag]):
            obs, r, done = o.step([a])
            assert np.array_equal(np.array(o.pos), z['pos_' + tag][s + 1]), (tag, s)
            assert np.array_equal(np.array(o.vel), z['vel_' + tag][s + 1]), (tag, s)
            assert np.array_equal(obs, z['obs_' + tag][s + 1]), (tag, s)
            assert r == z['reward_' + tag][s], (tag, s)


ZOO = [('ovm', 'OVM', None), ('ovm_inst', 'OVM', 'instantaneous'),
       ('ovm_safe', 'OVM', 'safe_velocity'), ('lovm_inst', 'LINEAR_OVM', 'instantaneous'),
       ('bcm', 'BCM', None), ('cfm_safe', 'CFM', 'safe_velocity'), ('lac', 'LAC', None),
       ('idm_inst', 'IDM', 'instantaneous'), ('idm_safe', 'IDM', 'safe_velocity'),
       ('fs', 'FOLLOWER_STOPPER', None), ('nlfs', 'NONLOCAL_FOLLOWER_STOPPER', None),
       ('pi', 'PI_SATURATION', None), ('gipps', 'GIPPS', 'instantaneous')]


def zoo_groups(ctrl, fail_safe):
    p = [5.0] if ctrl == 'FOLLOWER_STOPPER' else None
    if ctrl == 'GIPPS':     # v0, acc, b, b_l, s0, tau (make_golden.ZOO 'gipps')
        p = [30, 1.5, -3.0, -3.5, 5, 0.5]
    x = S.vt(ctrl, 1, p=p, fail_safe=fail_safe)
    return [S.vt('IDM', 10), x, S.vt('IDM', 10), dict(x)]


@pytest.mark.parametrize('tag,ctrl,fs', ZOO, ids=[z[0] for z in ZOO])
def test_zoo_trajectories(tag, ctrl, fs):
    z = np.load(os.path.join(G, 'ring_zoo.npz'))
    o = S.make_oracle(zoo_groups(ctrl, fs), 240, bunching=30.0)
    n = len(z['reward_' + tag])
    for s in range(n):
        _, r, _ = o.step(None)
        # PISaturation: np.mean (pairwise) vs running sum -> last-bit differences
        tol = 1e-12 if ctrl == 'PI_SATURATION' else 0.0
        assert np.allclose(o.vel, z['vel_' + tag][s + 1], rtol=tol, atol=tol), (tag, s)
        assert np.allclose(o.accel, z['accel_' + tag][s], rtol=tol, atol=tol * 10), (tag, s)
        assert abs(r - z['reward_' + tag][s]) <= tol, (tag, s)
        if (s + 1) % 10 == 0:
            assert np.allclose(o.pos, z['pos_' + tag][(s + 1) // 10], rtol=tol, atol=tol)


def test_sumo_trajectory_fixture():
    """The only SUMO-produced numbers the reference holds for this path
    (tests/fast_tests/test_files/ring_230_emission.csv): speeds to 0.01 and
    displacement since t=0.1 to 0.01 (+ print rounding) for t <= 0.5."""
    rows = json.load(open(os.path.join(G, 'sumo_ring_230.json')))['rows']
    o = S.make_oracle(S.c1_groups(), 230, bunching=20.0)
    ids = ['idm_%d' % i for i in range(22)]
    by = {(round(r['time'], 1), r['id']): r for r in rows}
    x0 = list(o.pos)
    travelled = [0.0] * 22
    for step, t in enumerate([0.2, 0.3, 0.4, 0.5]):
        prev = list(o.pos)
        o.step(None)
        for i, vid in enumerate(ids):
            travelled[i] += (o.pos[i] - prev[i]) % o.circ
            r, r0 = by[(t, vid)], by[(0.1, vid)]
            assert abs(o.vel[i] - r['speed']) <= 0.0101, (t, vid)
            if r['edge'] == r0['edge']:
                assert abs(travelled[i] - (r['rel_pos'] - r0['rel_pos'])) <= 0.0201, (t, vid)
    assert x0[0] == 0.0


def c3_groups():
    return [S.vt('IDM', 63), S.vt('RL', 3)]


def test_c3_lane_change_trajectory_bit_exact():
    """LaneChangeAccelEnv on the 3-lane ring: reference env code (get_state,
    compute_reward, _apply_rl_actions, apply_lane_change, get_last_lc bug)."""
    z = np.load(os.path.join(G, 'ring_c3.npz'))
    o = S.make_oracle(c3_groups(), 230, env_kind=OS.ENV_LANE_CHANGE_ACCEL, n_lanes=3,
                      act_lo=-3, act_hi=3)
    assert np.array_equal(np.array(o.pos), z['pos'][0]) and o.lane == list(z['lane'][0])
    for s, a in enumerate(z['actions']):
        obs, r, done = o.step(None if np.isnan(a).any() else list(a))
        assert np.array_equal(np.array(o.pos), z['pos'][s + 1]), s
        assert np.array_equal(np.array(o.vel), z['vel'][s + 1]), s
        assert o.lane == list(z['lane'][s + 1]), s
        assert np.array_equal(obs, z['obs'][s]) and r == z['reward'][s], s
    assert done == bool(z['crashed'])


PO_CASES = {'a': ([('IDM', 21), ('RL', 1)], float('inf')), 'b': ([('IDM', 4), ('RL', 2)], 2)}


@pytest.mark.parametrize('tag', ['a', 'b'])
def test_c3_lane_change_po_observation(tag):
    """LaneChangeAccelPOEnv.get_state fed by the reference's own _multi_lane_headways
    walking the ring's 4 edges + 4 junctions (tests/golden/ring_c3_po.npz) vs the
    oracle's closed-loop restatement. Trajectory bit-exact; the observation to 1e-12
    relative (the reference adds per-edge offsets, the oracle subtracts loop
    coordinates - same numbers up to the last bits), leaders' speeds exactly."""
    z = np.load(os.path.join(G, 'ring_c3_po.npz'))
    groups, dist = PO_CASES[tag]
    o = S.make_oracle([S.vt(c, n) for c, n in groups], 230, env_kind=OS.ENV_LANE_CHANGE_ACCEL_PO,
                      n_lanes=3, act_lo=-3, act_hi=3, lanes_distribution=dist)
    assert np.array_equal(np.array(o.pos), z['pos_' + tag][0])
    assert o.lane == list(z['lane_' + tag][0])
    n_rl = sum(n for c, n in groups if c == 'RL')
    assert z['obs_' + tag].shape[1] == 4 * n_rl * 3 + 1      # the reference's early return
    for s, a in enumerate(z['actions_' + tag]):
        obs, r, done = o.step(None if np.isnan(a).any() else list(a))
        assert np.array_equal(np.array(o.pos), z['pos_' + tag][s + 1]), s
        assert np.array_equal(np.array(o.vel), z['vel_' + tag][s + 1]), s
        assert o.lane == list(z['lane_' + tag][s + 1]), s
        ref = z['obs_' + tag][s]
        assert obs.shape == ref.shape
        assert np.abs(obs - ref).max() <= 1e-12 * 230, (s, obs, ref)
        assert np.array_equal(obs[6:], ref[6:]), s          # speeds, zero block, ego speed
        assert np.array_equal(obs == 1000, ref == 1000)
        assert r == z['reward_' + tag][s], s
    assert done == bool(z['crashed_' + tag])
    if tag == 'b':
        assert (z['obs_b'] == 1000).any()                     # empty lanes were seen
        assert np.abs(z['obs_b'][:, :6] - (230.4 - 5)).min() < 1e-9   # own leader, one lap away


SORT_GROUPS = [('IDM', 5), ('RL', 1), ('IDM', 4), ('RL', 1), ('IDM', 1)]


def test_sort_vehicles_two_episodes_bit_exact():
    """additional_params['sort_vehicles'] with the reference's own sorted_ids /
    additional_command / _apply_rl_actions (tests/golden/ring_sorted.npz): observation rows
    and the RL action slots follow the absolute-position keys; the keys survive a reset."""
    z = np.load(os.path.join(G, 'ring_sorted.npz'))
    o = None
    swaps = 0
    for ep in (0, 1):
        warm = int(z['warm_%d' % ep])
        kw = dict(env_kind=OS.ENV_ACCEL, act_lo=-3, act_hi=3, sort_vehicles=True, warmup_steps=warm)
        if o is None:
            o = S.make_oracle([S.vt(c, n) for c, n in SORT_GROUPS], 120, **kw)
            # the constructor ran reset + warm-up; redo it stepwise to see every observation
            o.abs_position = o.prev_pos = None
        obs = o.reset(120, run_warmup=False)
        # reset(run_warmup=False) already stored the keys; the reference does that after warm-up
        assert np.array_equal(np.array(o.pos), z['pos_%d' % ep][0])
        acts = z['actions_%d' % ep]
        if ep == 0:
            o.abs_position = o.prev_pos = None
            assert np.array_equal(o.get_state(), z['obs_0'][0])          # id order: no keys yet
        else:
            o.abs_position, o.prev_pos = stale
            assert np.array_equal(o.get_state(), z['obs_1'][0])          # keys of episode 0
        for s, a in enumerate(acts):
            if s == warm:
                o.abs_position, o.prev_pos = list(o.pos), list(o.pos)      # accel.py:178-181
            obs, r, done = o.step(None if np.isnan(a).any() else list(a))
            assert np.array_equal(np.array(o.pos), z['pos_%d' % ep][s + 1]), (ep, s)
            assert np.array_equal(obs, z['obs_%d' % ep][s + 1]), (ep, s)
            assert r == z['reward_%d' % ep][s], (ep, s)
            order = o.sorted_idx()
            swaps += order.index(10) < order.index(5)
        stale = (list(o.abs_position), list(o.prev_pos))
    assert swaps > 20           # the two RL vehicles did trade action slots
    # the packaged reset (warm-up inside) ends where the stepwise one did
    o2 = S.make_oracle([S.vt(c, n) for c, n in SORT_GROUPS], 120, env_kind=OS.ENV_ACCEL, act_lo=-3,
                       act_hi=3, sort_vehicles=True, warmup_steps=30)
    w = 30
    assert np.array_equal(o2.reset(120), z['obs_0'][w])
    assert o2.abs_position == list(z['pos_0'][w])


def test_sort_vehicles_lane_change_env_bit_exact():
    z = np.load(os.path.join(G, 'ring_sorted.npz'))
    o = S.make_oracle([S.vt('IDM', 9), S.vt('RL', 3)], 230, env_kind=OS.ENV_LANE_CHANGE_ACCEL,
                      n_lanes=3, act_lo=-3, act_hi=3, x0=200.0, sort_vehicles=True)
    o.abs_position = o.prev_pos = None
    assert np.array_equal(np.array(o.pos), z['lc_pos'][0])
    for s, a in enumerate(z['lc_actions']):
        if s == 20:
            o.abs_position, o.prev_pos = list(o.pos), list(o.pos)
        obs, r, done = o.step(None if np.isnan(a).any() else list(a))
        assert np.array_equal(np.array(o.pos), z['lc_pos'][s + 1]), s
        assert o.lane == list(z['lc_lane'][s + 1]), s
        assert np.array_equal(obs, z['lc_obs'][s]) and r == z['lc_reward'][s], s
    assert o.sorted_idx() != list(range(12))


def test_evaluate_reward_is_mean_speed():
    z = np.load(os.path.join(G, 'ring_sorted.npz'))
    o = S.make_oracle(S.c1_groups(), 230, bunching=20.0, evaluate=True)
    r = [o.step(None)[1] for _ in range(120)]
    assert np.abs(np.array(r) - z['eval_reward']).max() <= 1e-15 * 30


def test_multilane_known_answers_of_the_reference_test():
    """tests/fast_tests/test_vehicles.py:205-253 (TestMultiLaneData.test_no_junctions_ring):
    3-lane 230 m ring, 21 vehicles 3-abreast; for test_0: lane leaders [test_3, test_1, test_2],
    headways [27.857142857142858, -5, -5], followers [test_18, test_19, test_20], tailways
    28.577143. The tailway = 27.857143 + 0.72 mixes the three 0.1 m junction offsets of
    RingNetwork.specify_edge_starts in front of edge "left" with the 0.42 m netconvert gave
    the junction the walk crosses; a closed loop of 230.72 m (junction 0.18) is the same sum."""
    o = S.make_oracle([S.vt('RL', 21)], 230, env_kind=OS.ENV_LANE_CHANGE_ACCEL_PO, n_lanes=3,
                      junction=0.18, lanes_distribution=float('inf'))
    nb = o.lane_neighbours(0)
    assert [t[2] for t in nb] == [3, 1, 2] and [t[3] for t in nb] == [18, 19, 20]
    assert [t[0] for t in nb] == [27.85714285714286, -5, -5]
    np.testing.assert_array_almost_equal([t[1] for t in nb], [28.577143] * 3)


def ma_accel_groups():
    return [S.vt('RL', 1), S.vt('IDM', 6), S.vt('RL', 1), S.vt('IDM', 7), S.vt('RL', 1)]


MA_ACCEL = dict(act_lo=-1.0, act_hi=1.0, target_velocity=20.0)


def test_multiagent_accel_po_trajectory_bit_exact():
    """MultiAgentAccelPOEnv.get_state / compute_reward of the reference
    (tests/golden/ring_ma_accel.npz) vs the oracle restatement, incl. the unwrapped
    leader distance across x = 0."""
    z = np.load(os.path.join(G, 'ring_ma_accel.npz'))
    o = S.make_oracle(ma_accel_groups(), 180, env_kind=OS.ENV_MA_ACCEL_PO, **MA_ACCEL)
    assert np.array_equal(np.array(o.pos), z['pos'][0])
    assert np.array_equal(o.get_state(), z['obs'][0])
    for s, a in enumerate(z['actions']):
        obs, rew, done = o.step(None if np.isnan(a).any() else list(a))
        assert np.array_equal(np.array(o.pos), z['pos'][s + 1]), s
        assert np.array_equal(np.asarray(obs), z['obs'][s + 1]), s
        assert rew == z['reward'][s] and not done, s


def test_lane_change_direction_validation():
    o = S.make_oracle(c3_groups(), 230, env_kind=OS.ENV_LANE_CHANGE_ACCEL, n_lanes=3,
                      act_lo=-3, act_hi=3)
    for _ in range(40):
        o.step(None)
    with pytest.raises(ValueError):   # vehicle/traci.py:973-975
        o.step([0, 0.5, 0, 0, 0, 0])


def ma_groups():
    """examples/exp_configs/rl/multiagent/multiagent_ring.py vehicle order."""
    return [S.vt('RL', 1), S.vt('IDM', 10, sim_min_gap=0.0), S.vt('RL', 1),
            S.vt('IDM', 10, sim_min_gap=0.0)]


MA = dict(bunching=50.0, init_min_gap=0.0, warmup_steps=0, horizon=3000, act_lo=-1.0, act_hi=1.0,
          obs_max_length=270.0)


def test_multiagent_ring_trajectory_bit_exact():
    """MultiAgentWaveAttenuationPOEnv.get_state / compute_reward of the reference
    (tests/golden/ring_ma.npz) vs the oracle restatement."""
    z = np.load(os.path.join(G, 'ring_ma.npz'))
    kw = dict(MA); kw.pop('init_min_gap')
    o = S.make_oracle(ma_groups(), float(z['length']), env_kind=OS.ENV_MA_WAVE_ATTENUATION_PO,
                      **kw)
    assert np.array_equal(np.array(o.pos), z['pos'][0])
    for s, a in enumerate(z['actions']):
        obs, rew, done = o.step(None if np.isnan(a).any() else list(a))
        assert np.array_equal(np.array(o.pos), z['pos'][s + 1]), s
        assert np.array_equal(np.array(o.vel), z['vel'][s + 1]), s
        assert np.array_equal(np.asarray(obs), z['obs'][s]), s
        assert rew == z['reward'][s], s
        assert not done
