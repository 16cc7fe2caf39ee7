"""Generate tests/golden/*.npz|json by running the REAL reference code
(imported from /root/reference through ref_shim) — run in the build container:

    python -m tests.golden.make_golden

Outputs (committed):
  kat.json          known-answer vectors of the reference's own unit tests
                    (tests/fast_tests/test_controllers.py, test_rewards.py,
                    test_environments.py) re-evaluated with the reference's
                    controller / reward code on the stated inputs;
  ring_c1.npz       22 IDM, 230 m, bunching 20 (examples/exp_configs/non_rl/ring.py)
  ring_c2.npz       21 IDM + 1 RL WaveAttenuationPO, two lengths, warm-up + actions
  ring_zoo.npz      one trajectory per controller / fail-safe combination
  start_pos.json    gen_even_start_pos outputs for several configurations
  sumo_ring_230.json  the SUMO-produced trajectory fixture of the reference
                    (tests/fast_tests/test_files/ring_230_emission.csv), reduced
"""
import csv
import json
import os

import numpy as np

from .ref_harness import RefRing, RefRingLC, RefRingSorted, RefRingLCSorted, ref_ctrl, \
    ref_start_positions, ref_rewards
from . import ref_shim

HERE = os.path.dirname(os.path.abspath(__file__))
IDM, OVM, LOVM, BCM, CFM, LAC, GIPPS, FS, NLFS, PI = (
    ref_ctrl.IDMController, ref_ctrl.OVMController, ref_ctrl.LinearOVM,
    ref_ctrl.BCMController, ref_ctrl.CFMController, ref_ctrl.LACController,
    ref_ctrl.GippsController, ref_ctrl.FollowerStopper,
    ref_ctrl.NonLocalFollowerStopper, ref_ctrl.PISaturation)


def _inject(ring, headways, speeds=None):
    """test hook of the reference tests: set_headway / test_set_speed."""
    ring.headway = list(map(float, headways))
    if speeds is not None:
        ring.vel = list(map(float, speeds))


def kat():
    """Known-answer tests: (controller, kwargs, cf kwargs, headways, speeds,
    expected of the reference test file, line)."""
    H = [5, 10, 15, 20, 25]
    SP = [5, 7.5, 7.5, 8, 7]
    cf = dict(accel=20, decel=5)
    cases = [
        ('CFM', CFM, dict(time_delay=0, k_d=1, k_v=1, k_c=1, d_des=1, v_des=8, noise=0), cf, H, None,
         [12., 17., 22., 27., 32.], 'test_controllers.py:70'),
        ('BCM', BCM, dict(time_delay=0, k_d=1, k_v=1, k_c=1, d_des=1, v_des=8, noise=0), cf, H, None,
         [-12., 13., 13., 13., 13.], 'test_controllers.py:119'),
        ('OVM', OVM, dict(time_delay=0, alpha=1, beta=1, h_st=2, h_go=15, v_max=30, noise=0), cf,
         [0, 10, 5, 5, 5], None, [0., 20.319073, 3.772339, 3.772339, 3.772339],
         'test_controllers.py:174'),
        ('LINEAR_OVM', LOVM, dict(time_delay=0, v_max=30, adaptation=0.65, h_st=5, noise=0), cf,
         [5, 10, 10, 15, 0], None, [0., 12.992308, 12.992308, 25.984615, 0.],
         'test_controllers.py:224'),
        ('IDM', IDM, dict(v0=30, T=1, a=1, b=1.5, delta=4, s0=2, noise=0), cf,
         [10, 20, 30, 40, 50], None, [0.96, 0.99, 0.995556, 0.9975, 0.9984],
         'test_controllers.py:271'),
        ('IDM_zero_headway', IDM, dict(v0=30, T=1, a=1, b=1.5, delta=4, s0=2, noise=0), cf,
         [0, 0, 0, 0, 0], None, None, 'test_controllers.py:275-285'),
        ('FOLLOWER_STOPPER', FS, dict(v_des=7.5), cf, H, SP, [0, 0, 0, -5, 5],
         'test_controllers.py:507'),
        ('NONLOCAL_FOLLOWER_STOPPER', NLFS, dict(v_des=7.5), cf, H, SP,
         [-3.33333333333333, -5.0, -5.0, -10.0, 0.0], 'test_controllers.py:579'),
        ('PI_SATURATION', PI, dict(), cf, H, SP,
         [20., -36.847826, -35.76087, -37.173913, -31.086957], 'test_controllers.py:654'),
        ('LAC', LAC, dict(time_delay=0, k_1=0.3, k_2=0.4, h=1, tau=0.1, noise=0), cf,
         H, None, [0., 1.5, 3., 4.5, 6.], 'test_controllers.py:707'),
        ('GIPPS', GIPPS, dict(v0=30, acc=1.5, b=-1, b_l=-1, s0=2, tau=1, noise=0), cf,
         [2, 4, 6, 8, 10], None, [0., 5.929271, 5.929271, 5.929271, 5.929271],
         'test_controllers.py:763'),
    ]
    out = []
    for name, cls, kw, cfk, hw, sp, expected, cite in cases:
        ring = RefRing([('test', cls, kw, cfk, 5)], 230, 'accel',
                       env_additional={'target_velocity': 8, 'max_accel': 3, 'max_decel': 3,
                                       'sort_vehicles': False})
        _inject(ring, hw, sp)
        got = [float(ring.ctrl[i].get_action(ring)) for i in range(5)]
        if expected is not None:
            np.testing.assert_array_almost_equal(got, expected)
        out.append(dict(name=name, controller=name.split('_zero')[0], params=kw, cf=cfk,
                        headways=hw, speeds=sp, expected=expected, reference_value=got,
                        cite='tests/fast_tests/' + cite))
    return out


def reward_kat():
    out = []
    # test_rewards.py:18-58 (whole-network part)
    ring = RefRing([('test', IDM, {}, {}, 10)], 230, 'accel',
                   env_additional={'target_velocity': float(np.sqrt(10)), 'max_accel': 1,
                                   'max_decel': 1, 'sort_vehicles': False})
    assert ref_rewards.desired_velocity(ring, fail=True) == 0
    assert ref_rewards.desired_velocity(ring, fail=False) == 0
    ring.vel[0] = float(np.sqrt(10))
    got = float(ref_rewards.desired_velocity(ring, fail=False))
    np.testing.assert_almost_equal(got, 1 - np.sqrt(90) / 10)
    out.append(dict(name='desired_velocity', n=10, target=float(np.sqrt(10)),
                    speeds=list(ring.vel), expected=float(1 - np.sqrt(90) / 10),
                    reference_value=got, cite='tests/fast_tests/test_rewards.py:18-41'))
    # test_environments.py:552-613
    wa = []
    for vh, vr, act, fail, exp in [(0, 0, [0], False, 0), (0, 1, [0], False, 0.1),
                                   (1, 1, [0], False, 0.2), (1, 1, [0], True, 0),
                                   (1, 1, None, False, 0), (1, 1, [1], False, -3.8)]:
        ring = RefRing([('human', IDM, {}, {}, 1), ('rl', 'rl', {}, {}, 1)], 230, 'wave_po',
                       env_additional={'max_accel': 1, 'max_decel': 1, 'ring_length': [220, 270]})
        ring.vel = [float(vh), float(vr)]
        got = float(ring.compute_reward(act, fail))
        np.testing.assert_almost_equal(got, exp)
        wa.append(dict(speeds=[vh, vr], actions=act, fail=fail, expected=exp, reference_value=got))
    out.append(dict(name='wave_attenuation_reward', cases=wa,
                    cite='tests/fast_tests/test_environments.py:552-613'))
    return out


def trajectory(ring, n_steps, actions=None, sample_every=1):
    """Run the reference-driven ring; returns dict of arrays."""
    P, Vv, A, O, R, Cr = [], [], [], [], [], []
    P.append(list(ring.pos)); Vv.append(list(ring.vel))
    for s in range(n_steps):
        a = None if actions is None or np.isnan(actions[s]).any() else actions[s]
        obs, rew, crash = ring.step(a)
        if s % sample_every == 0 or s == n_steps - 1:
            P.append(list(ring.pos)); Vv.append(list(ring.vel)); A.append(list(ring.accel))
            O.append(np.asarray(obs, dtype=np.float64))
        R.append(float(rew)); Cr.append(bool(crash))
        if crash:
            break
    return dict(pos=np.array(P), vel=np.array(Vv), accel=np.array(A), obs=np.array(O),
                reward=np.array(R), crash=np.array(Cr))


def ring_c1():
    ring = RefRing([('idm', IDM, {}, {}, 22)], 230, 'accel', init_kw=dict(bunching=20),
                   env_additional={'target_velocity': 10, 'max_accel': 3, 'max_decel': 3,
                                   'sort_vehicles': False})
    t = trajectory(ring, 3000)
    v = t['vel'][1:]
    np.savez_compressed(
        os.path.join(HERE, 'ring_c1.npz'),
        pos=t['pos'][:301], vel=t['vel'][:301], accel=t['accel'][:300], obs=t['obs'][:300],
        reward=t['reward'], pos_final=t['pos'][-1], vel_final=t['vel'][-1],
        mean_speed=v[-1000:].mean(), wave_amp=(v[-1000:].max(1) - v[-1000:].min(1)).mean(),
        sum_reward=t['reward'].sum())


def ring_c2_speed_mode():
    """c2 with the speed modes singleagent_ring.py leaves at the reference's default
    ('right_of_way' = 25 on humans and on the RL vehicle) and, second case, 'no_collide' (7):
    actions well beyond what the gap allows, so the restated clamp binds on the RL vehicle; the
    simulator model behind the clamp is evaluated by the reference's own IDMController."""
    out = {}
    for tag, length, mode, seed in (('a', 260, 'right_of_way', 11), ('b', 223, 'no_collide', 12)):
        ring = RefRing([('human', IDM, {}, {'min_gap': 0, 'speed_mode': mode}, 21),
                        ('rl', 'rl', {}, {'speed_mode': mode}, 1)], length, 'wave_po',
                       init_kw=dict(bunching=50, min_gap=0),
                       env_additional={'max_accel': 1, 'max_decel': 1, 'ring_length': [220, 270]},
                       use_speed_mode=True)
        rng = np.random.default_rng(seed)
        n_w, n_a = 100, 700
        acts = np.full((n_w + n_a, 1), np.nan)
        acts[n_w:, 0] = rng.uniform(-0.6, 1.5, n_a)     # accelerating on average: rams without the clamp
        t = trajectory(ring, n_w + n_a, acts)
        assert not t['crash'].any() and len(t['reward']) == n_w + n_a
        out.update({'length_' + tag: length, 'mode_' + tag: ring.speed_mode[0],
                    'actions_' + tag: acts[n_w:, 0],
                    'pos_' + tag: t['pos'][n_w:], 'vel_' + tag: t['vel'][n_w:],
                    'obs_' + tag: t['obs'][n_w - 1:], 'reward_' + tag: t['reward'][n_w:],
                    'pos0_' + tag: t['pos'][0]})
    np.savez_compressed(os.path.join(HERE, 'ring_c2_speed_mode.npz'), **out)


def ring_c2():
    out = {}
    for tag, length, seed in (('a', 260, 1), ('b', 223, 2)):
        ring = RefRing([('human', IDM, {}, {'min_gap': 0}, 21), ('rl', 'rl', {}, {}, 1)], length,
                       'wave_po', init_kw=dict(bunching=50, min_gap=0),
                       env_additional={'max_accel': 1, 'max_decel': 1, 'ring_length': [220, 270]})
        rng = np.random.default_rng(seed)
        n_w, n_a = 750, 600
        acts = np.full((n_w + n_a, 1), np.nan)
        acts[n_w:, 0] = rng.uniform(-1.5, 1.5, n_a)   # beyond the [-1,1] box: exercises the clip
        t = trajectory(ring, n_w + n_a, acts)
        out.update({'length_' + tag: length, 'actions_' + tag: acts[n_w:, 0],
                    'pos_' + tag: t['pos'][n_w:], 'vel_' + tag: t['vel'][n_w:],
                    'obs_' + tag: t['obs'][n_w - 1:], 'reward_' + tag: t['reward'][n_w:],
                    'pos0_' + tag: t['pos'][0]})
    np.savez_compressed(os.path.join(HERE, 'ring_c2.npz'), **out)


def ring_ma():
    """examples/exp_configs/rl/multiagent/multiagent_ring.py: two RL vehicles, each
    followed by ten IDM humans, MultiAgentWaveAttenuationPOEnv observation/reward.
    Actions come from a noisy gap-keeping rule fed with the previous observation
    (purely random accelerations run the RL car into its leader within ~120 steps);
    a good part of them lies outside the [-1, 1] box, which exercises the clip."""
    groups = [('rl_0', 'rl', {}, {}, 1), ('human_0', IDM, {}, {'min_gap': 0}, 10),
              ('rl_1', 'rl', {}, {}, 1), ('human_1', IDM, {}, {'min_gap': 0}, 10)]
    ring = RefRing(groups, 247, 'ma_wave_po', init_kw=dict(bunching=50, min_gap=0),
                   env_additional={'max_accel': 1, 'max_decel': 1, 'ring_length': [220, 270]})
    rng = np.random.default_rng(5)
    n_w, n_a = 200, 600
    acts = np.full((n_w + n_a, 2), np.nan)
    P, Vv, O, Rw = [list(ring.pos)], [list(ring.vel)], [], []
    obs = ring.get_state()
    for s in range(n_w + n_a):
        a = None
        if s >= n_w:
            o = np.asarray(obs).reshape(2, 3)
            a = 2.0 * o[:, 1] * 15 + 0.08 * (o[:, 2] * 270 - 8.0) + rng.normal(0, 0.6, 2)
            acts[s] = a
        obs, rew, crash = ring.step(a)
        assert not crash, s
        P.append(list(ring.pos)); Vv.append(list(ring.vel)); O.append(np.asarray(obs, float))
        Rw.append(float(rew))
    assert (np.abs(acts[n_w:]) > 1).mean() > 0.1
    np.savez_compressed(os.path.join(HERE, 'ring_ma.npz'), length=247, actions=acts,
                        pos=np.array(P), vel=np.array(Vv), obs=np.array(O), reward=np.array(Rw))


def ring_ma_accel():
    """MultiAgentAccelPOEnv.get_state / compute_reward of the reference on a 180 m ring:
    3 RL vehicles (first, middle and LAST index, so one leader sits across x = 0) among 13
    IDM humans; actions from a noisy gap-keeping rule, some outside the [-1, 1] box."""
    groups = [('rl_a', 'rl', {}, {}, 1), ('ha', IDM, {}, {}, 6), ('rl_b', 'rl', {}, {}, 1),
              ('hb', IDM, {}, {}, 7), ('rl_c', 'rl', {}, {}, 1)]
    ring = RefRing(groups, 180, 'ma_accel_po',
                   env_additional={'max_accel': 1, 'max_decel': 1, 'target_velocity': 20})
    rng = np.random.default_rng(8)
    n_w, n_a = 50, 500
    acts = np.full((n_w + n_a, 3), np.nan)
    P, Vv, O, Rw = [list(ring.pos)], [list(ring.vel)], [np.asarray(ring.get_state(), float)], []
    for s in range(n_w + n_a):
        a = None
        if s >= n_w:
            a = np.array([0.3 * (ring.headway[i] - 5 - 2.5 * ring.vel[i])
                          + 0.8 * (ring.vel[ring.leader[i]] - ring.vel[i]) for i in ring.rl_idx]) \
                + rng.normal(0, 0.4, 3)
            acts[s] = a
        obs, rew, crash = ring.step(a)
        assert not crash, s
        P.append(list(ring.pos)); Vv.append(list(ring.vel)); O.append(np.asarray(obs, float))
        Rw.append(float(rew))
    assert (np.abs(acts[n_w:]) > 1).any()      # the Box clip is exercised
    O = np.array(O)
    assert (O[:, 3::6] < 0).any()          # the unwrapped lead_head went negative across x = 0
    np.savez_compressed(os.path.join(HERE, 'ring_ma_accel.npz'), actions=acts, pos=np.array(P),
                        vel=np.array(Vv), obs=O, reward=np.array(Rw))


ZOO = [
    ('ovm', OVM, {}, 'OVM', None),
    ('ovm_inst', OVM, {'fail_safe': 'instantaneous'}, 'OVM', 'instantaneous'),
    ('ovm_safe', OVM, {'fail_safe': 'safe_velocity'}, 'OVM', 'safe_velocity'),
    ('lovm_inst', LOVM, {'fail_safe': 'instantaneous'}, 'LINEAR_OVM', 'instantaneous'),
    ('bcm', BCM, {}, 'BCM', None),
    ('cfm_safe', CFM, {'fail_safe': 'safe_velocity'}, 'CFM', 'safe_velocity'),
    ('lac', LAC, {}, 'LAC', None),
    ('idm_inst', IDM, {'fail_safe': 'instantaneous'}, 'IDM', 'instantaneous'),
    ('idm_safe', IDM, {'fail_safe': 'safe_velocity'}, 'IDM', 'safe_velocity'),
    ('fs', FS, {'v_des': 5.0}, 'FOLLOWER_STOPPER', 'safe_velocity'),
    ('nlfs', NLFS, {}, 'NONLOCAL_FOLLOWER_STOPPER', 'safe_velocity'),
    ('pi', PI, {}, 'PI_SATURATION', None),
    # Gipps with a SHORT reaction time and hard braking: the square root's argument goes negative
    # on the bunched start (np.sqrt -> NaN, skipped by min) and positive later
    ('gipps', GIPPS, {'tau': 0.5, 'b': -3.0, 'b_l': -3.5, 's0': 5, 'fail_safe': 'instantaneous'},
     'GIPPS', 'instantaneous'),
]


def ring_zoo():
    """20 IDM + 2 vehicles of the controller under test, 240 m, 300 steps
    (pos every 10th step, vel/accel/reward every step)."""
    out = {}
    for tag, cls, kw, _, _ in ZOO:
        ring = RefRing([('idm', IDM, {}, {}, 10), ('x', cls, kw, {}, 1), ('idm2', IDM, {}, {}, 10),
                        ('y', cls, kw, {}, 1)], 240, 'accel', init_kw=dict(bunching=30),
                       env_additional={'target_velocity': 10, 'max_accel': 3, 'max_decel': 3,
                                       'sort_vehicles': False})
        t = trajectory(ring, 300)
        out['pos_' + tag] = t['pos'][::10]; out['vel_' + tag] = t['vel']
        out['accel_' + tag] = t['accel']; out['reward_' + tag] = t['reward']
    np.savez_compressed(os.path.join(HERE, 'ring_zoo.npz'), **out)


def ring_c3():
    """3-lane 230 m ring, 63 IDM + 3 RL (one per lane), LaneChangeAccelEnv:
    reference get_state / compute_reward / _apply_rl_actions / apply_lane_change."""
    ring = RefRingLC([('human', IDM, {}, {}, 63), ('rl', 'rl', {}, {}, 3)], 230, lanes=3,
                     init_kw=dict(lanes_distribution=float('inf')),
                     env_additional={'max_accel': 3, 'max_decel': 3, 'target_velocity': 10})
    rng = np.random.default_rng(11)
    P, Vv, Ln, O, R, A = [list(ring.pos)], [list(ring.vel)], [list(ring.lane)], [], [], []
    for s in range(260):
        a = []
        for q in range(3):
            a += [float(rng.uniform(-0.8, 0.8)), float(rng.integers(-1, 2))]
        act = None if s < 20 else a
        obs, rew, crash = ring.step(act)
        A.append([np.nan] * 6 if act is None else a)
        P.append(list(ring.pos)); Vv.append(list(ring.vel)); Ln.append(list(ring.lane))
        O.append(np.asarray(obs)); R.append(float(rew))
        if crash:
            break
    np.savez_compressed(os.path.join(HERE, 'ring_c3.npz'), pos=np.array(P), vel=np.array(Vv),
                        lane=np.array(Ln, dtype=np.int8), obs=np.array(O), reward=np.array(R),
                        actions=np.array(A), crashed=bool(crash))


def ring_c3_po():
    """LaneChangeAccelPOEnv.get_state driven by the reference's own
    _multi_lane_headways on the 4-edge + 4-junction ring. Case a: 3 lanes, 21 IDM + 1 RL.
    Case b: 3 lanes with lane 2 empty at the start, 4 IDM + 2 RL (empty-lane 1000s, an RL
    vehicle alone in a lane being its own leader, the reference's early return leaving the
    second RL block zero)."""
    out = {}
    for tag, groups, init_kw, seed, steps in (
            ('a', [('human', IDM, {}, {}, 21), ('rl', 'rl', {}, {}, 1)],
             dict(lanes_distribution=float('inf')), 5, 300),
            ('b', [('human', IDM, {}, {}, 4), ('rl', 'rl', {}, {}, 2)],
             dict(lanes_distribution=2), 6, 300)):
        ring = RefRingLC(groups, 230, lanes=3, init_kw=init_kw, env_kind='lc_po',
                         env_additional={'max_accel': 3, 'max_decel': 3, 'target_velocity': 10})
        n_rl = len(ring.rl_idx)
        rng = np.random.default_rng(seed)
        P, Vv, Ln, O, R, A = [list(ring.pos)], [list(ring.vel)], [list(ring.lane)], [], [], []
        for s in range(steps):
            a = []
            for q in range(n_rl):
                a += [float(rng.uniform(-0.8, 0.8)), float(rng.integers(-1, 2))]
            act = None if s < 20 else a
            obs, rew, crash = ring.step(act)
            A.append([np.nan] * (2 * n_rl) if act is None else a)
            P.append(list(ring.pos)); Vv.append(list(ring.vel)); Ln.append(list(ring.lane))
            O.append(np.asarray(obs, dtype=np.float64)); R.append(float(rew))
            if crash:
                break
        out.update({'pos_' + tag: np.array(P), 'vel_' + tag: np.array(Vv),
                    'lane_' + tag: np.array(Ln, dtype=np.int8), 'obs_' + tag: np.array(O),
                    'reward_' + tag: np.array(R), 'actions_' + tag: np.array(A),
                    'crashed_' + tag: bool(crash)})
    np.savez_compressed(os.path.join(HERE, 'ring_c3_po.npz'), **out)


def ring_sorted():
    """additional_params['sort_vehicles'] = True and env_params.evaluate = True.
    Part 1 (AccelEnv, 120 m, 10 IDM + 2 RL apart from each other): the reference's
    sorted_ids / additional_command / _apply_rl_actions over two episodes - the first
    reset observes in id order (no keys yet), the second with the keys the first episode
    left behind. Part 2 (LaneChangeAccelEnv, 3 lanes, 9 IDM + 3 RL): keys frozen at the
    reset positions, ties between side-by-side vehicles. Part 3: c1 with evaluate."""
    out = {}
    groups = [('ha', IDM, {}, {}, 5), ('rla', 'rl', {}, {}, 1), ('hb', IDM, {}, {}, 4),
              ('rlb', 'rl', {}, {}, 1), ('hc', IDM, {}, {}, 1)]
    ring = RefRingSorted(groups, 120, env_additional={'max_accel': 3, 'max_decel': 3,
                                                       'target_velocity': 10})
    rng = np.random.default_rng(4)
    for ep, (warm, steps) in enumerate(((30, 400), (10, 80))):
        if ep:
            ring.reset(120)
        P, Vv, O, R, A = [list(ring.pos)], [list(ring.vel)], [np.asarray(ring.get_state())], [], []
        for s in range(warm + steps):
            if s == warm:
                ring.finish_reset()
            if s < warm:
                act = None
            else:   # noisy gap-keeping rule so that nobody rams its leader; slot order = sorted order
                act = []
                for vid in [v for v in ring.sorted_ids if v in ring.k.vehicle.get_rl_ids()]:
                    i = ring.index[vid]
                    act.append(float(np.clip(0.3 * (ring.headway[i] - 4 - 1.2 * ring.vel[i])
                                             + rng.normal(0, 0.4), -3.5, 3.5)))
            obs, rew, crash = ring.step(act)
            assert not crash
            A.append([np.nan, np.nan] if act is None else act)
            P.append(list(ring.pos)); Vv.append(list(ring.vel))
            O.append(np.asarray(obs)); R.append(float(rew))
        out.update({'pos_%d' % ep: np.array(P), 'vel_%d' % ep: np.array(Vv), 'obs_%d' % ep: np.array(O),
                    'reward_%d' % ep: np.array(R), 'actions_%d' % ep: np.array(A), 'warm_%d' % ep: warm})
    lc = RefRingLCSorted([('human', IDM, {}, {}, 9), ('rl', 'rl', {}, {}, 3)], 230, lanes=3,
                         init_kw=dict(lanes_distribution=float('inf'), x0=200.0),
                         env_additional={'max_accel': 3, 'max_decel': 3, 'target_velocity': 10})
    P, Ln, O, R, A = [list(lc.pos)], [list(lc.lane)], [], [], []
    for s in range(170):
        if s == 20:
            lc.finish_reset()
        a = []
        for q in range(3):
            a += [float(rng.uniform(-0.8, 0.8)), float(rng.integers(-1, 2))]
        act = None if s < 20 else a
        obs, rew, crash = lc.step(act)
        assert not crash
        A.append([np.nan] * 6 if act is None else a)
        P.append(list(lc.pos)); Ln.append(list(lc.lane)); O.append(np.asarray(obs)); R.append(float(rew))
    out.update(lc_pos=np.array(P), lc_lane=np.array(Ln, dtype=np.int8), lc_obs=np.array(O),
               lc_reward=np.array(R), lc_actions=np.array(A))
    ev = RefRing([('idm', IDM, {}, {}, 22)], 230, init_kw=dict(bunching=20),
                 env_additional={'max_accel': 3, 'max_decel': 3, 'target_velocity': 10})
    ev.env_params.evaluate = True
    out['eval_reward'] = np.array([ev.step(None)[1] for _ in range(120)])
    np.savez_compressed(os.path.join(HERE, 'ring_sorted.npz'), **out)


def highway_ma():
    """MultiAgentHighwayPOEnv.get_state / compute_reward of the reference, step by step, on the
    restated open highway with a human and an RL inflow: agents enter and leave, the head agent
    has no leader, the tail agent no follower. 40 steps without actions first."""
    from .ref_harness import RefHighwayMA
    from tests.scenarios import highway_ma_sim, highway_ma_action
    A = 8
    sim = highway_ma_sim()
    ref = RefHighwayMA(sim, 25.0)
    rng = np.random.default_rng(11)
    n_w, n_a = 40, 700
    ids = np.full((n_w + n_a + 1, A), -1, np.int64)
    obs = np.zeros((n_w + n_a + 1, A, 5))
    rew = np.zeros((n_w + n_a + 1, A))
    acts = np.full((n_w + n_a, A), np.nan)
    nveh = np.zeros(n_w + n_a + 1, np.int64)

    def record(t, with_reward):
        st = ref.get_state()
        mine = sim.ma_state()
        assert list(st) == list(mine) and len(st) <= A
        for a, (vid, o) in enumerate(st.items()):
            assert np.array_equal(o, mine[vid])          # oracle restatement == reference
            ids[t, a], obs[t, a] = vid, o
        if with_reward:
            r = ref.compute_reward({}, fail=False)
            mr = sim.ma_reward(False)
            for a, vid in enumerate(st):
                assert r[vid] == mr[vid]
                rew[t, a] = r[vid]
        nveh[t] = len(sim.pos)
        return st

    st = record(0, False)
    seen_head = seen_tail = False
    for s in range(n_w + n_a):
        a = None
        if s >= n_w:
            a = {}
            for row, (vid, slot) in enumerate(zip(st, sim.rl_slots())):
                assert sim.ids[slot] == vid
                acts[s, row] = highway_ma_action(sim, slot, rng.normal(0, 0.5))
                a[vid] = acts[s, row]
        sim.step(a)
        assert not sim.crash, s
        st = record(s + 1, True)
        slots = sim.rl_slots()
        seen_head |= bool(slots) and slots[0] == 0
        seen_tail |= bool(slots) and slots[-1] == len(sim.pos) - 1
    assert seen_head and seen_tail and sim.arrived > 20 and (np.abs(acts[n_w:]) > 1).any()
    assert (ids >= 0).sum(axis=1).max() >= 3
    np.savez_compressed(os.path.join(HERE, 'highway_ma.npz'), ids=ids, obs=obs, reward=rew,
                        actions=acts, n_veh=nveh)


def start_pos():
    cases = [dict(length=230, n=22, lanes=1, kw=dict(bunching=20)),
             dict(length=230, n=22, lanes=1, kw=dict()),
             dict(length=260, n=22, lanes=1, kw=dict(bunching=50, min_gap=0)),
             dict(length=221, n=22, lanes=1, kw=dict(bunching=50, min_gap=0)),
             dict(length=270, n=22, lanes=1, kw=dict(bunching=50, min_gap=0)),
             dict(length=230, n=5, lanes=1, kw=dict(x0=7.5, min_gap=1.0)),
             dict(length=230, n=66, lanes=3, kw=dict(lanes_distribution=float('inf'))),
             dict(length=230, n=21, lanes=3, kw=dict(lanes_distribution=float('inf'))),
             dict(length=230, n=20, lanes=3, kw=dict(lanes_distribution=2, bunching=10))]
    out = []
    for c in cases:
        x, lane = ref_start_positions(c['length'], c['n'], c['lanes'], **c['kw'])
        kw = {k: (v if v != float('inf') else 'inf') for k, v in c['kw'].items()}
        out.append(dict(length=c['length'], n=c['n'], lanes=c['lanes'], kw=kw,
                        x=[float(v) for v in x], lane=[int(v) for v in lane]))
    json.dump(out, open(os.path.join(HERE, 'start_pos.json'), 'w'), indent=1)


def start_pos_stochastic():
    """InitialConfig(perturbation / spacing='random' / shuffle): the reference's generators read
    Python's `random` and numpy's global stream; seeded here, so the host restatement
    (flow_b200/placement.py) can be checked number for number."""
    import random
    cases = [dict(length=230, n=22, lanes=1, seed=7, kw=dict(bunching=20, perturbation=1.0)),
             dict(length=260, n=22, lanes=1, seed=8, kw=dict(perturbation=4.0)),      # clips at edge ends
             dict(length=230, n=21, lanes=3, seed=9, kw=dict(lanes_distribution=float('inf'),
                                                             perturbation=2.0)),
             dict(length=230, n=22, lanes=1, seed=10, kw=dict(spacing='random')),
             dict(length=300, n=10, lanes=1, seed=11, kw=dict(spacing='random', min_gap=1.5)),
             dict(length=230, n=5, lanes=4, seed=12, kw=dict(spacing='random', lanes_distribution=2)),
             dict(length=230, n=40, lanes=3, seed=13, kw=dict(spacing='random', bunching=10))]
    out = []
    for c in cases:
        random.seed(c['seed']); np.random.seed(c['seed'])
        x, lane = ref_start_positions(c['length'], c['n'], c['lanes'], **c['kw'])
        ids = ['test_%d' % i for i in range(c['n'])]
        random.shuffle(ids)                 # flow/envs/base.py:275-277, same stream, drawn after
        kw = {k: (v if v != float('inf') else 'inf') for k, v in c['kw'].items()}
        out.append(dict(length=c['length'], n=c['n'], lanes=c['lanes'], seed=c['seed'], kw=kw,
                        x=[float(v) for v in x], lane=[int(v) for v in lane], shuffled_ids=ids))
    json.dump(out, open(os.path.join(HERE, 'start_pos_stochastic.json'), 'w'), indent=1)


def start_pos_merge():
    """gen_even_start_pos of the reference on its own MergeNetwork tables (five edges of unequal
    length, three junction slots): the initial vehicles of singleagent_merge.py and denser cases."""
    from .ref_harness import ref_merge_start_positions
    cases = [dict(pre=500, merge=100, post=100, n=5, kw={}),            # singleagent_merge.py
             dict(pre=200, merge=100, post=100, n=5, kw={}),            # the network's defaults
             dict(pre=500, merge=100, post=100, n=40, kw=dict(min_gap=1.0, bunching=30)),
             dict(pre=300, merge=60, post=80, n=17, kw=dict(x0=77.0)),
             dict(pre=500, merge=100, post=100, n=1, kw={}),
             dict(pre=50, merge=4, post=100, n=9, kw={})]               # an edge too short to use
    out = []
    for c in cases:
        pos, lane = ref_merge_start_positions(c['pre'], c['merge'], c['post'], c['n'], **c['kw'])
        out.append(dict(c, pos=[[e, p] for e, p in pos], lane=lane))
    json.dump(out, open(os.path.join(HERE, 'start_pos_merge.json'), 'w'), indent=1)


def figure8():
    """AccelEnv on FigureEightNetwork (singleagent_figure_eight.py: 13 IDM + 1 RL, radius 30):
    the reference's gen_even_start_pos on its own network tables, then 600 steps of the
    reference's IDMController.get_action (junction guard included), AccelEnv._apply_rl_actions /
    get_state / compute_reward and get_x on the restated loop (RefFigure8)."""
    import flow.controllers as RC
    from .ref_harness import RefFigure8, ref_figure8_start_positions
    from tests.scenarios import figure8_sim, FIGURE8
    cases = [dict(radius=30, n=14, kw={}), dict(radius=50, n=22, kw=dict(x0=3.0, min_gap=1.0)),
             dict(radius=30, n=40, kw=dict(bunching=20)), dict(radius=20, n=1, kw={})]
    out = []
    for c in cases:
        pos, xs = ref_figure8_start_positions(c['radius'], c['n'], **c['kw'])
        out.append(dict(c, pos=[[e, p] for e, p in pos], x=xs))
    json.dump(out, open(os.path.join(HERE, 'start_pos_figure8.json'), 'w'), indent=1)
    sim = figure8_sim(positions=out[0]['x'])
    cf = dict(speed_mode='obey_safe_speed', decel=1.5)
    ref = RefFigure8(sim, [('human', RC.IDMController, {}, cf, 13), ('rl', 'rl', {}, cf, 1)])
    rng = np.random.default_rng(2)
    A, O, R, D, P = [], [], [], [], [np.array(sim.pos)]
    for k in range(FIGURE8['horizon']):
        act = rng.uniform(-0.6, 1.2, 1)
        obs, rew, done = ref.step(act)
        A.append(act); O.append(obs); R.append(rew); D.append(done); P.append(np.array(sim.pos))
        if done:
            break
    np.savez_compressed(os.path.join(HERE, 'figure8.npz'), actions=np.array(A), obs=np.array(O),
                        reward=np.array(R), done=np.array(D), pos=np.array(P),
                        vel_final=np.array(sim.vel))


def merge():
    """MergePOEnv: the reference's own _apply_rl_actions / additional_command / get_state /
    compute_reward (RefMerge) on the restated merge network, 300 Env.steps of 5 simulation steps;
    controlled vehicles come and go, rows close up, a vehicle that left is observed once."""
    from .ref_harness import RefMerge
    from tests.scenarios import merge_sim, MERGE
    sim = merge_sim()
    ref = RefMerge(sim, 1.5, 1.5)
    rng = np.random.default_rng(1)
    A, O, R, D, CT, N = [], [], [], [], [], []
    for k in range(MERGE['horizon']):
        act = np.clip(rng.uniform(-1.0, 1.5, MERGE['num_rl']), -1.5, 1.5)
        obs, rew, done = ref.step(act)
        A.append(act); O.append(obs); R.append(rew); D.append(done)
        CT.append([int(v) for v in ref.rl_veh] + [-1] * (MERGE['num_rl'] - len(ref.rl_veh)))
        N.append([sim.num_vehicles, sim.departed, sim.arrived])
        if done:
            break
    np.savez_compressed(os.path.join(HERE, 'merge.npz'), actions=np.array(A), obs=np.array(O),
                        reward=np.array(R), done=np.array(D), ctrl=np.array(CT), counts=np.array(N),
                        pos_final=np.array(sim.pos), arm_final=np.array(sim.arm),
                        ids_final=np.array(sim.ids))


def sumo_fixture():
    src = os.path.join(ref_shim.REF, 'tests/fast_tests/test_files/ring_230_emission.csv')
    rows = list(csv.DictReader(open(src)))
    out = [dict(time=float(r['time']), id=r['id'], edge=r['edge_id'],
                rel_pos=float(r['relative_position']), speed=float(r['speed'])) for r in rows]
    json.dump(dict(source='tests/fast_tests/test_files/ring_230_emission.csv (SUMO output)',
                   rows=out), open(os.path.join(HERE, 'sumo_ring_230.json'), 'w'))


def main():
    json.dump(dict(controllers=kat(), rewards=reward_kat()),
              open(os.path.join(HERE, 'kat.json'), 'w'), indent=1)
    ring_c1(); ring_c2(); ring_c2_speed_mode(); ring_c3(); ring_c3_po(); ring_sorted(); ring_zoo(); ring_ma(); ring_ma_accel(); start_pos(); start_pos_stochastic(); start_pos_merge()
    highway_ma()
    merge()
    figure8()
    sumo_fixture()
    print('golden vectors written to', HERE)


if __name__ == '__main__':
    main()
