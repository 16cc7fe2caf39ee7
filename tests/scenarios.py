"""Shared scenario builders: the same vehicle/env description instantiates the
CPU oracle (oracle/sim.py) and the CUDA engine (flow_b200.engine)."""
import numpy as np

from oracle import controllers as OC
from oracle import sim as OS

CTRL = dict(SIM=0, IDM=1, OVM=2, LINEAR_OVM=3, BCM=4, CFM=5, LAC=6, GIPPS=7,
            FOLLOWER_STOPPER=8, NONLOCAL_FOLLOWER_STOPPER=9, PI_SATURATION=10, RL=11)
FAILSAFE = {None: 0, 'instantaneous': 1, 'safe_velocity': 2}

DEFAULT_P = dict(
    IDM=[30, 1, 1, 1.5, 4, 2], OVM=[1, 1, 2, 15, 30], LINEAR_OVM=[30, 0.65, 5],
    BCM=[1, 1, 1, 1, 8], CFM=[1, 1, 1, 1, 8], LAC=[0.3, 0.4, 1, 0.1, 0],
    GIPPS=[30, 1.5, -1, -1, 2, 1], FOLLOWER_STOPPER=[15],
    NONLOCAL_FOLLOWER_STOPPER=[], PI_SATURATION=[], SIM=[], RL=[])


def vt(controller, count, p=None, fail_safe=None, noise=0.0, delay=0.0,
       max_accel=2.6, max_decel=4.5, sim_min_gap=2.5, sim_tau=1.0,
       sim_max_speed=30.0, speed_mode=0):
    if controller in ('FOLLOWER_STOPPER', 'NONLOCAL_FOLLOWER_STOPPER'):
        # velocity_controllers.py:31-33: forced delay=1.0 + safe_velocity
        fail_safe, delay = 'safe_velocity', 1.0
    if controller == 'PI_SATURATION':
        delay = 1.0
    return dict(controller=controller, count=count,
                p=list(DEFAULT_P[controller] if p is None else p),
                fail_safe=fail_safe, noise=noise, delay=delay, max_accel=max_accel,
                max_decel=max_decel, sim_min_gap=sim_min_gap, sim_tau=sim_tau,
                sim_max_speed=sim_max_speed, speed_mode=speed_mode)


def type_table(groups):
    type_of = []
    for t, g in enumerate(groups):
        type_of += [t] * g['count']
    return type_of


def oracle_types(groups):
    return [OC.VType(CTRL[g['controller']], g['p'], FAILSAFE[g['fail_safe']], g['noise'],
                     g['delay'], g['max_accel'], g['max_decel'], 5.0, g['sim_tau'],
                     g['sim_min_gap'], g['sim_max_speed'], speed_mode=g.get('speed_mode', 0))
            for g in groups]


def engine_types(groups):
    from flow_b200 import engine as EN
    return [EN.make_vtype(CTRL[g['controller']], g['p'], g['fail_safe'], g['noise'],
                          g['delay'], g['max_accel'], g['max_decel'], 5.0, g['sim_tau'],
                          g['sim_min_gap'], g['sim_max_speed'], speed_mode=g.get('speed_mode', 0))
            for g in groups]


def make_oracle(groups, length, env_kind=OS.ENV_ACCEL, **kw):
    return OS.RingSim(oracle_types(groups), type_table(groups), length,
                      env_kind=env_kind, **kw)


def make_engine(groups, n_envs, dtype='f64', length=None, **kw):
    from flow_b200 import engine as EN
    if length is not None:
        kw.setdefault('length_lo', float(length))
        kw.setdefault('length_hi', float(length))
    if 'lanes_distribution' in kw:
        kw['lanes_distribution'] = int(min(kw['lanes_distribution'], 1 << 30))
    return EN.Engine(engine_types(groups), type_table(groups), n_envs, dtype=dtype, **kw)


# the BASELINE.json configurations -------------------------------------------
def c1_groups():
    """examples/exp_configs/non_rl/ring.py: 22 IDM (noise set 0 for parity)."""
    return [vt('IDM', 22)]


C1 = dict(length=230, bunching=20.0, target_velocity=10.0, horizon=3000)


def c2_groups(noise=0.0, speed_mode=0):
    """examples/exp_configs/rl/singleagent/singleagent_ring.py: 21 IDM + 1 RL. The config leaves
    speed_mode at the reference's default (25); 0 = the goldens' setting (no SUMO clamp)."""
    return [vt('IDM', 21, noise=noise, sim_min_gap=0.0, speed_mode=speed_mode),
            vt('RL', 1, speed_mode=speed_mode)]


C2 = dict(bunching=50.0, init_min_gap=0.0, warmup_steps=750, horizon=3000,
          act_lo=-1.0, act_hi=1.0, obs_max_length=270.0)


def rel_err(a, b, floor=1e-12):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor))) if a.size else 0.0


# multi-agent open highway (tests/golden/highway_ma.npz) ------------------------
HIGHWAY_MA = dict(capacity=64, length=600.0, dt=0.2, speed_limit=30.0, horizon=1500,
                  target_velocity=25.0, x0=100.0, max_agents=8,
                  # (type, vehs_per_hour, depart_speed, begin): RL flow first
                  inflows=[(1, 360.0, 10.0, 5.0), (0, 720.0, 12.0, 5.0)])


def highway_ma_groups():
    return [vt('IDM', 2), vt('RL', 1)]


def highway_ma_sim(evaluate=False, **over):
    """600 m single-lane open highway, 3 initial vehicles (IDM, RL, IDM), an RL inflow of
    360 veh/h served before an IDM inflow of 720 veh/h, both from t = 5 s (with a saturated
    entry the first flow in definition order wins every free gap)."""
    from oracle import highway as OH
    c = dict(HIGHWAY_MA, **over)
    return OH.HighwaySim(oracle_types(highway_ma_groups()), [0, 1, 0], capacity=c['capacity'],
                         length=c['length'], dt=c['dt'], speed_limit=c['speed_limit'],
                         horizon=c['horizon'], target_velocity=c['target_velocity'], x0=c['x0'],
                         evaluate=evaluate,
                         inflows=[dict(type=t, vehs_per_hour=q, depart_speed=v, begin=b)
                                  for t, q, v, b in c['inflows']])


def highway_ma_engine(n_envs, dtype='f64', evaluate=False, **over):
    from flow_b200 import engine as EN
    c = dict(HIGHWAY_MA, **over)
    return EN.Engine(engine_types(highway_ma_groups()), [0, 1, 0], n_envs, dtype=dtype,
                     capacity=c['capacity'], topology=EN.TOPO_HIGHWAY,
                     env_kind=EN.ENV_MA_HIGHWAY_PO, length_lo=c['length'], length_hi=c['length'],
                     sim_step=c['dt'], speed_limit=c['speed_limit'], horizon=c['horizon'],
                     target_velocity=c['target_velocity'], x0=c['x0'], evaluate=int(evaluate),
                     max_agents=c['max_agents'], inflows=c['inflows'], record_views=1)


def highway_ma_action(sim, slot, z):
    """Noisy gap-keeping rule on the agent's own gap and closing speed (unclipped on purpose;
    the observation holds the LEADER's headway, not the agent's, so it cannot drive safely)."""
    v = sim.vel[slot]
    if slot == 0:
        return float(np.clip(0.5 * (20.0 - v), -3.0, 1.2)) + z
    return float(np.clip(0.1 * (sim.headway[slot] - 5.0 - 1.5 * v) + 0.6 * (sim.vel[slot - 1] - v),
                         -6.0, 1.2)) + z


# merge network under MergePOEnv (tests/golden/merge.npz) -----------------------------------
MERGE = dict(pre=500.0, ramp=100.0, post=100.0, capacity=128, num_rl=3, dt=0.2, sims_per_step=5,
             target_velocity=20.0, horizon=300, speed_limit=30.0,
             initial=[(0, 0, 50.0), (0, 0, 300.0), (0, 1, 30.0)],          # (type, arm, position)
             # (type, arm, vehs_per_hour, depart_speed, begin): singleagent_merge.py's three flows
             inflows=[(0, 0, 1800.0, 10.0, 1.0), (1, 0, 200.0, 10.0, 1.0), (0, 1, 100.0, 7.5, 1.0)])


def merge_groups(noise=0.0, speed_mode=1):
    """singleagent_merge.py: IDM humans and RL vehicles, both speed_mode 'obey_safe_speed'."""
    return [vt('IDM', 0, noise=noise, speed_mode=speed_mode), vt('RL', 0, speed_mode=speed_mode)]


def merge_sim(groups=None, **over):
    from oracle import merge as OM
    c = dict(MERGE, **over)
    return OM.MergeSim(oracle_types(groups or merge_groups()), OM.MergeGeometry(c['pre'], c['ramp'], c['post']),
                       initial=c['initial'],
                       inflows=[dict(type=t, arm=a, vehs_per_hour=q, depart_speed=v, begin=b)
                                for t, a, q, v, b in c['inflows']],
                       dt=c['dt'], speed_limit=c['speed_limit'], capacity=c['capacity'],
                       num_rl=c['num_rl'], target_velocity=c['target_velocity'], horizon=c['horizon'],
                       sims_per_step=c['sims_per_step'], zone=c.get('zone'),
                       evaluate=c.get('evaluate', False), warmup_steps=c.get('warmup_steps', 0))


def merge_engine(n_envs, dtype='f64', groups=None, **over):
    """Engine on the merge network, reset with the scenario's initial vehicles in place."""
    from flow_b200 import engine as EN
    c = dict(MERGE, **over)
    eng = EN.Engine(engine_types(groups or merge_groups()), [t for t, _, _ in c['initial']], n_envs,
                    dtype=dtype, capacity=c['capacity'], topology=EN.TOPO_MERGE,
                    env_kind=EN.ENV_MERGE_PO, merge_pre=c['pre'], merge_ramp=c['ramp'],
                    merge_post=c['post'], merge_zone=c.get('zone') or 0.0, max_agents=c['num_rl'],
                    sim_step=c['dt'], speed_limit=c['speed_limit'], horizon=c['horizon'],
                    sims_per_step=c['sims_per_step'], target_velocity=c['target_velocity'],
                    warmup_steps=c.get('warmup_steps', 0), evaluate=int(c.get('evaluate', False)),
                    record_views=1, seed=over.get('seed', 0),
                    inflows=[(t, q, v, b, 0.0, 0.0, a) for t, a, q, v, b in c['inflows']])
    merge_reset(eng, c['initial'])
    return eng


def merge_reset(eng, initial):
    """fs_reset in its two halves with the initial vehicles' (arm, position) set in between."""
    eng.reset_stage(eng.RESET_PLACE)
    pos = np.zeros((eng.n_veh, eng.n_envs), eng.np_dtype)
    arm = np.zeros((eng.n_veh, eng.n_envs), np.uint8)
    for i, (_, a, p) in enumerate(initial):
        pos[i], arm[i] = p, a
    eng.set_state(pos=pos, lane=arm)
    return eng.reset_stage(eng.RESET_WARMUP)


# figure-eight network under AccelEnv (tests/golden/figure8.npz) -------------------------------
FIGURE8 = dict(radius=30.0, dt=0.1, horizon=600, target_velocity=20.0, speed_limit=30.0,
               act_lo=-3.0, act_hi=3.0)


def figure8_groups(n_human=13, n_rl=1, noise=0.0, speed_mode=1):
    """singleagent_figure_eight.py: IDM humans and one RL vehicle, 'obey_safe_speed', decel 1.5"""
    return [vt('IDM', n_human, noise=noise, speed_mode=speed_mode, max_decel=1.5),
            vt('RL', n_rl, speed_mode=speed_mode, max_decel=1.5)]


def figure8_sim(groups=None, positions=None, **over):
    from oracle import figure8 as F8
    c = dict(FIGURE8, **over)
    groups = groups or figure8_groups()
    type_of = [t for t, g in enumerate(groups) for _ in range(g['count'])]
    return F8.Figure8Sim(oracle_types(groups), type_of, c['radius'], positions, dt=c['dt'],
                         speed_limit=c['speed_limit'], target_velocity=c['target_velocity'],
                         horizon=c['horizon'], warmup_steps=c.get('warmup_steps', 0),
                         sims_per_step=c.get('sims_per_step', 1), zone=c.get('zone'),
                         act_lo=c['act_lo'], act_hi=c['act_hi'],
                         clip_actions=c.get('clip_actions', True), evaluate=c.get('evaluate', False))


def figure8_engine(n_envs, dtype='f64', groups=None, **over):
    from flow_b200 import engine as EN
    c = dict(FIGURE8, **over)
    groups = groups or figure8_groups()
    type_of = [t for t, g in enumerate(groups) for _ in range(g['count'])]
    return EN.Engine(engine_types(groups), type_of, n_envs, dtype=dtype, topology=EN.TOPO_FIGURE8,
                     env_kind=EN.ENV_ACCEL, fig8_radius=c['radius'], fig8_zone=c.get('zone') or 0.0,
                     sim_step=c['dt'], speed_limit=c['speed_limit'], horizon=c['horizon'],
                     warmup_steps=c.get('warmup_steps', 0), sims_per_step=c.get('sims_per_step', 1),
                     target_velocity=c['target_velocity'], act_lo=c['act_lo'], act_hi=c['act_hi'],
                     clip_actions=int(c.get('clip_actions', True)), evaluate=int(c.get('evaluate', False)),
                     x0=c.get('x0', 0.0), init_min_gap=c.get('min_gap', 0.0), bunching=c.get('bunching', 0.0),
                     record_views=1, seed=over.get('seed', 0))
