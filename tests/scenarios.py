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
       sim_max_speed=30.0):
    if controller in ('FOLLOWER_STOPPER', 'NONLOCAL_FOLLOWER_STOPPER'):
        # velocity_controllers.py:31-33: forced delay=1.0 + safe_velocity
        fail_safe, delay = 'safe_velocity', 1.0
    if controller == 'PI_SATURATION':
        delay = 1.0
    return dict(controller=controller, count=count,
                p=list(DEFAULT_P[controller] if p is None else p),
                fail_safe=fail_safe, noise=noise, delay=delay, max_accel=max_accel,
                max_decel=max_decel, sim_min_gap=sim_min_gap, sim_tau=sim_tau,
                sim_max_speed=sim_max_speed)


def type_table(groups):
    type_of = []
    for t, g in enumerate(groups):
        type_of += [t] * g['count']
    return type_of


def oracle_types(groups):
    return [OC.VType(CTRL[g['controller']], g['p'], FAILSAFE[g['fail_safe']], g['noise'],
                     g['delay'], g['max_accel'], g['max_decel'], 5.0, g['sim_tau'],
                     g['sim_min_gap'], g['sim_max_speed']) for g in groups]


def engine_types(groups):
    from flow_b200 import engine as EN
    return [EN.make_vtype(CTRL[g['controller']], g['p'], g['fail_safe'], g['noise'],
                          g['delay'], g['max_accel'], g['max_decel'], 5.0, g['sim_tau'],
                          g['sim_min_gap'], g['sim_max_speed']) for g in groups]


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


def c2_groups(noise=0.0):
    """examples/exp_configs/rl/singleagent/singleagent_ring.py: 21 IDM + 1 RL."""
    return [vt('IDM', 21, noise=noise, sim_min_gap=0.0), vt('RL', 1)]


C2 = dict(bunching=50.0, init_min_gap=0.0, warmup_steps=750, horizon=3000,
          act_lo=-1.0, act_hi=1.0, obs_max_length=270.0)


def rel_err(a, b, floor=1e-12):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor))) if a.size else 0.0
