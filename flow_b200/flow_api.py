"""Mirror of the reference's operator interface for the step path (the gym
seam): the parameter classes, controller classes, networks,
environments, `make_create_env` and `Experiment` a Flow user already writes,
backed by the CUDA engine instead of SUMO/TraCI.

What each piece stands for in the reference (parthjaggi/flow):

  SumoCarFollowingParams / SumoLaneChangeParams / VehicleParams / SumoParams /
  EnvParams / NetParams / InitialConfig / InFlows    flow/core/params.py:199-1221
  *Controller classes (constructor signatures)          flow/controllers/*.py
  RingNetwork / HighwayNetwork                          flow/networks/ring.py, highway.py
  AccelEnv, WaveAttenuationEnv, WaveAttenuationPOEnv    flow/envs/ring/accel.py, wave_attenuation.py
  LaneChangeAccelEnv                                    flow/envs/ring/lane_change_accel.py
  MultiAgentWaveAttenuationPOEnv                        flow/envs/multiagent/ring/wave_attenuation.py:143-312
  MultiAgentAccelPOEnv                                  flow/envs/multiagent/ring/accel.py:87-226
  MultiAgentHighwayPOEnv                                flow/envs/multiagent/highway.py:14-196
  Env.reset / Env.step / clip_actions                   flow/envs/base.py:294-615
  env.k.vehicle.* / env.k.network.*                     flow/core/kernel/vehicle/traci.py, network/traci.py
  make_create_env                                       flow/utils/registry.py:13-134
  Experiment.run                                        flow/core/experiment.py:60-200

Controllers are *parameter carriers* here: user-defined Python `get_accel`
cannot run inside a CUDA kernel, so the built-in classes are recognised BY NAME
(the reference's own classes from `flow.controllers` are accepted as well as
the ones below) and compiled into the engine's per-type table; anything else
raises NotImplementedError — never a silent CPU fallback.

One env object may carry `n_envs` independent copies (a batch on one GPU).
With n_envs == 1 (default) `reset/step` return numpy arrays / Python scalars
exactly like the reference's gym env; with n_envs > 1 they return torch CUDA
tensors [E, obs_dim], [E], [E] that alias the engine's buffers (DLPack).
"""
import math
import time
from copy import deepcopy

import numpy as np

from . import engine as EN

VEHICLE_LENGTH = 5.0          # flow/core/kernel/network/base.py:10
RING_JUNCTION_LENGTH = 0.1    # flow/networks/ring.py:197
SPEED_MODES = {"aggressive": 0, "obey_safe_speed": 1, "no_collide": 7,
               "right_of_way": 25, "all_checks": 31}
LC_MODES = {"no_lc_safe": 512, "no_lc_aggressive": 0, "sumo_default": 1621,
            "no_lat_collide": 512, "strategic": 1621, "aggressive": 0}


class FatalFlowError(Exception):
    """flow/utils/exceptions.py: raised where the reference aborts a run."""

    def __init__(self, msg='Fatal error in Flow'):
        super().__init__(msg)


# =============================================================================
# controllers — constructor signatures of flow/controllers/*.py
# =============================================================================
class _Controller:
    """Parameter carrier. `ENGINE` = FS_CTRL_* id, `ORDER` = positional order
    of the engine's p[] table for this law."""
    ENGINE = None
    ORDER = ()
    DEFAULTS = {}

    def __init__(self, veh_id=None, car_following_params=None, time_delay=0.0,
                 noise=0, fail_safe=None, **kw):
        unknown = set(kw) - set(self.DEFAULTS) - {'delay'}
        if unknown:
            raise TypeError('%s() got unexpected arguments %s'
                            % (type(self).__name__, sorted(unknown)))
        self.veh_id = veh_id
        self.car_following_params = car_following_params
        self.delay = kw.pop('delay', time_delay)
        self.noise = noise
        self.fail_safe = fail_safe
        for k, v in self.DEFAULTS.items():
            setattr(self, k, kw.get(k, v))

    def engine_params(self):
        return [getattr(self, k) for k in self.ORDER]


class SimCarFollowingController(_Controller):
    """flow/controllers/car_following_models.py:585-590 (simulator drives)."""
    ENGINE = EN.CTRL_SIM


class RLController(_Controller):
    """flow/controllers/rlcontroller.py: accel comes from the action vector."""
    ENGINE = EN.CTRL_RL


class IDMController(_Controller):
    """car_following_models.py:437-482"""
    ENGINE = EN.CTRL_IDM
    ORDER = ('v0', 'T', 'a', 'b', 'delta', 's0')
    DEFAULTS = dict(v0=30, T=1, a=1, b=1.5, delta=4, s0=2)


class OVMController(_Controller):
    """car_following_models.py:282-328"""
    ENGINE = EN.CTRL_OVM
    ORDER = ('alpha', 'beta', 'h_st', 'h_go', 'v_max')
    DEFAULTS = dict(alpha=1, beta=1, h_st=2, h_go=15, v_max=30)


class LinearOVM(_Controller):
    """car_following_models.py:359-397"""
    ENGINE = EN.CTRL_LINEAR_OVM
    ORDER = ('v_max', 'adaptation', 'h_st')
    DEFAULTS = dict(v_max=30, adaptation=0.65, h_st=5)


class BCMController(_Controller):
    """car_following_models.py:125-176"""
    ENGINE = EN.CTRL_BCM
    ORDER = ('k_d', 'k_v', 'k_c', 'd_des', 'v_des')
    DEFAULTS = dict(k_d=1, k_v=1, k_c=1, d_des=1, v_des=8)


class CFMController(BCMController):
    """car_following_models.py:49-88"""
    ENGINE = EN.CTRL_CFM


class LACController(_Controller):
    """car_following_models.py:205-245"""
    ENGINE = EN.CTRL_LAC
    ORDER = ('k_1', 'k_2', 'h', 'tau', 'a')
    DEFAULTS = dict(k_1=0.3, k_2=0.4, h=1, tau=0.1, a=0)


class GippsController(_Controller):
    """car_following_models.py:538-582"""
    ENGINE = EN.CTRL_GIPPS
    ORDER = ('v0', 'acc', 'b', 'b_l', 's0', 'tau')
    DEFAULTS = dict(v0=30, acc=1.5, b=-1, b_l=-1, s0=2, tau=1)


class FollowerStopper(_Controller):
    """velocity_controllers.py:24-116: delay 1.0 and the safe_velocity
    fail-safe are forced by the constructor (:31-33)."""
    ENGINE = EN.CTRL_FOLLOWER_STOPPER
    ORDER = ('v_des',)
    DEFAULTS = dict(v_des=15, danger_edges=None)

    def __init__(self, veh_id=None, car_following_params=None, **kw):
        super().__init__(veh_id, car_following_params, **kw)
        self.delay, self.fail_safe = 1.0, 'safe_velocity'
        if self.danger_edges:
            raise NotImplementedError('danger_edges: ring/highway have no such edges')


class NonLocalFollowerStopper(FollowerStopper):
    """velocity_controllers.py:119-164 (v_des = mean speed of the env)."""
    ENGINE = EN.CTRL_NONLOCAL_FOLLOWER_STOPPER
    ORDER = ()


class PISaturation(_Controller):
    """velocity_controllers.py:167-240 (delay 1.0, no fail-safe)."""
    ENGINE = EN.CTRL_PI_SATURATION

    def __init__(self, veh_id=None, car_following_params=None):
        super().__init__(veh_id, car_following_params)
        self.delay = 1.0


class SimLaneChangeController:
    """lane_change_controllers.py: SUMO decides (mode 512: never)."""
    ENGINE = EN.LC_SIM

    def __init__(self, veh_id=None, lane_change_params=None):
        self.veh_id = veh_id


class StaticLaneChanger(SimLaneChangeController):
    """lane_change_controllers.py: always direction 0."""
    ENGINE = EN.LC_STATIC


class ContinuousRouter:
    """routing_controllers.py: re-routes at the end of the loop; a closed
    ring needs no routing in the engine (one continuous coordinate)."""

    def __init__(self, veh_id=None, router_params=None):
        self.veh_id = veh_id


_ACCEL_BY_NAME = {c.__name__: c for c in (
    SimCarFollowingController, RLController, IDMController, OVMController, LinearOVM,
    BCMController, CFMController, LACController, GippsController, FollowerStopper,
    NonLocalFollowerStopper, PISaturation)}
_LC_BY_NAME = {c.__name__: c for c in (SimLaneChangeController, StaticLaneChanger)}


def _resolve(cls, table, what):
    """Map a controller class (ours or the reference's, by name) to ours."""
    name = cls if isinstance(cls, str) else getattr(cls, '__name__', None)
    if name not in table:
        raise NotImplementedError(
            '%s %r cannot run on the B200 engine: only the built-in laws are compiled '
            '(%s)' % (what, name, ', '.join(sorted(table))))
    return table[name]


# =============================================================================
# params — flow/core/params.py
# =============================================================================
class SumoCarFollowingParams:
    """flow/core/params.py:785-902"""

    def __init__(self, speed_mode='right_of_way', accel=2.6, decel=4.5, sigma=0.5, tau=1.0,
                 min_gap=2.5, max_speed=30, speed_factor=1.0, speed_dev=0.1, impatience=0.5,
                 car_follow_model='IDM', **kwargs):
        old = dict(minGap='min_gap', maxSpeed='max_speed', speedFactor='speed_factor',
                   speedDev='speed_dev', carFollowModel='car_follow_model')
        loc = dict(min_gap=min_gap, max_speed=max_speed, speed_factor=speed_factor,
                   speed_dev=speed_dev, car_follow_model=car_follow_model)
        for k, v in kwargs.items():
            if k not in old:
                raise TypeError('unexpected argument %r' % k)
            loc[old[k]] = v   # deprecated camelCase spellings
        self.controller_params = {
            'accel': accel, 'decel': decel, 'sigma': sigma, 'tau': tau,
            'minGap': loc['min_gap'], 'maxSpeed': loc['max_speed'],
            'speedFactor': loc['speed_factor'], 'speedDev': loc['speed_dev'],
            'impatience': impatience, 'carFollowModel': loc['car_follow_model']}
        if isinstance(speed_mode, str) and speed_mode in SPEED_MODES:
            speed_mode = SPEED_MODES[speed_mode]
        elif not isinstance(speed_mode, (int, float)):
            speed_mode = SPEED_MODES['obey_safe_speed']
        self.speed_mode = speed_mode


class SumoLaneChangeParams:
    """flow/core/params.py:904-1067. The engine only reads `lane_change_mode`; the SUMO
    model parameters are carried so that flow_params.json stays interchangeable."""
    _BASE = (('lcStrategic', 'lc_strategic', 1.0), ('lcCooperative', 'lc_cooperative', 1.0),
             ('lcSpeedGain', 'lc_speed_gain', 1.0), ('lcKeepRight', 'lc_keep_right', 1.0))
    _SUBLANE = (('lcLookaheadLeft', 'lc_look_ahead_left', 2.0),
                ('lcSpeedGainRight', 'lc_speed_gain_right', 1.0), ('lcSublane', 'lc_sublane', 1.0),
                ('lcPushy', 'lc_pushy', 0), ('lcPushyGap', 'lc_pushy_gap', 0.6),
                ('lcAssertive', 'lc_assertive', 1), ('lcAccelLat', 'lc_accel_lat', 1.0))

    def __init__(self, lane_change_mode='no_lat_collide', model='LC2013', **kwargs):
        if model not in ('LC2013', 'SL2015'):
            model = 'LC2013'
        rows = self._BASE + (self._SUBLANE if model == 'SL2015' else ())
        known = {k for r in self._BASE + self._SUBLANE for k in r[:2]}
        unknown = set(kwargs) - known
        if unknown:
            raise TypeError('unexpected arguments %s' % sorted(unknown))
        self.controller_params = {'laneChangeModel': model}
        for sumo_key, key, default in rows:   # camelCase spellings are the deprecated aliases
            self.controller_params[sumo_key] = str(kwargs.get(sumo_key, kwargs.get(key, default)))
        if isinstance(lane_change_mode, str) and lane_change_mode in LC_MODES:
            lane_change_mode = LC_MODES[lane_change_mode]
        elif not isinstance(lane_change_mode, (int, float)):
            lane_change_mode = LC_MODES['no_lat_collide']
        self.lane_change_mode = lane_change_mode


class VehicleParams:
    """flow/core/params.py:199-362"""

    def __init__(self):
        self.ids = []
        self.initial = []
        self.types = []
        self.type_parameters = {}
        self.minGap = {}
        self.num_vehicles = 0
        self.num_rl_vehicles = 0
        self.num_types = 0
        self._type_of = {}

    def add(self, veh_id, length=None,
            acceleration_controller=(SimCarFollowingController, {}),
            lane_change_controller=(SimLaneChangeController, {}),
            routing_controller=None, initial_speed=0, num_vehicles=0,
            car_following_params=None, lane_change_params=None, color=None):
        cf = car_following_params or SumoCarFollowingParams()
        lc = lane_change_params or SumoLaneChangeParams()
        type_params = dict(cf.controller_params)
        type_params.update(lc.controller_params)
        row = dict(acceleration_controller=acceleration_controller,
                   lane_change_controller=lane_change_controller,
                   routing_controller=routing_controller, initial_speed=initial_speed,
                   car_following_params=cf, lane_change_params=lc)
        self.type_parameters[veh_id] = dict(row)
        if length:
            type_params['length'] = length
            self.type_parameters[veh_id]['length'] = length
        if color:
            type_params['color'] = color
            self.type_parameters[veh_id]['color'] = color
        self.initial.append(dict(row, veh_id=veh_id, num_vehicles=num_vehicles))
        self.minGap[veh_id] = type_params['minGap']
        is_rl = getattr(acceleration_controller[0], '__name__', '') == 'RLController'
        for i in range(num_vehicles):
            vid = '%s_%d' % (veh_id, i)
            self.ids.append(vid)
            self._type_of[vid] = veh_id
            self.num_vehicles += 1
            self.num_rl_vehicles += int(is_rl)
        self.num_types += 1
        self.types.append({'veh_id': veh_id, 'type_params': type_params})

    def get_type(self, veh_id):
        return self._type_of[veh_id]


class SimParams:
    """flow/core/params.py:364-422"""

    def __init__(self, sim_step=0.1, render=False, restart_instance=False, emission_path=None,
                 save_render=False, sight_radius=25, show_radius=False, pxpm=2,
                 force_color_update=False):
        self.sim_step, self.render = sim_step, render
        self.restart_instance, self.emission_path = restart_instance, emission_path
        self.save_render, self.sight_radius = save_render, sight_radius
        self.pxpm, self.show_radius = pxpm, show_radius
        self.force_color_update = force_color_update


class SumoParams(SimParams):
    """flow/core/params.py:510-618"""

    def __init__(self, port=None, sim_step=0.1, emission_path=None, lateral_resolution=None,
                 no_step_log=True, render=False, save_render=False, sight_radius=25,
                 show_radius=False, pxpm=2, force_color_update=False, overtake_right=False,
                 seed=None, restart_instance=False, print_warnings=True, start_at_load=True,
                 teleport_time=-1, num_clients=1, color_by_speed=False, use_ballistic=False):
        super().__init__(sim_step, render, restart_instance, emission_path, save_render,
                         sight_radius, show_radius, pxpm, force_color_update)
        self.port, self.lateral_resolution = port, lateral_resolution
        self.no_step_log, self.seed = no_step_log, seed
        self.overtake_right, self.print_warnings = overtake_right, print_warnings
        self.start_at_load, self.teleport_time = start_at_load, teleport_time
        self.num_clients, self.color_by_speed = num_clients, color_by_speed
        self.use_ballistic = use_ballistic


class EnvParams:
    """flow/core/params.py:620-671"""

    def __init__(self, additional_params=None, horizon=float('inf'), warmup_steps=0,
                 sims_per_step=1, evaluate=False, clip_actions=True):
        self.additional_params = additional_params if additional_params is not None else {}
        self.horizon, self.warmup_steps = horizon, warmup_steps
        self.sims_per_step, self.evaluate = sims_per_step, evaluate
        self.clip_actions = clip_actions

    def get_additional_param(self, key):
        return self.additional_params[key]


class InFlows:
    """flow/core/params.py:1070-1221 (the list lives in the same name-mangled
    attribute as the reference's, so a flow_params.json round-trips both ways)"""

    def __init__(self):
        self.__flows = []

    def add(self, edge, veh_type, vehs_per_hour=None, probability=None, period=None,
            depart_lane='first', depart_speed=0, name='flow', begin=1, end=86400,
            number=None, **kwargs):
        given = [x is not None for x in (vehs_per_hour, probability, period)]
        if sum(given) > 1:
            raise ValueError('define exactly one of vehs_per_hour, probability, period')
        if probability is not None and not 0 <= probability <= 1:
            raise ValueError('Inflow.add called with parameter \'probability\' set to {}, but '
                             'probability should be between 0 and 1'.format(probability))
        new = dict(name='%s_%d' % (name, len(self.__flows)), vtype=veh_type, edge=edge,
                   departLane=depart_lane, departSpeed=depart_speed, begin=begin, end=end)
        new.update(kwargs)
        if vehs_per_hour is not None:
            new['vehsPerHour'] = vehs_per_hour
        if period is not None:
            new['period'] = period
        if probability is not None:
            new['probability'] = probability
        if number is not None:
            new['number'] = number
        self.__flows.append(new)

    def sort(self, key):
        self.__flows.sort(key=key)

    def get(self):
        return self.__flows


class NetParams:
    """flow/core/params.py:674-711"""

    def __init__(self, inflows=None, osm_path=None, template=None, additional_params=None):
        self.inflows = inflows or InFlows()
        self.osm_path, self.template = osm_path, template
        self.additional_params = additional_params or {}


class InitialConfig:
    """flow/core/params.py:714-782"""

    def __init__(self, shuffle=False, spacing='uniform', min_gap=0, perturbation=0.0, x0=0,
                 bunching=0, lanes_distribution=float('inf'), edges_distribution='all',
                 additional_params=None):
        self.shuffle, self.spacing, self.min_gap = shuffle, spacing, min_gap
        self.perturbation, self.x0, self.bunching = perturbation, x0, bunching
        self.lanes_distribution = lanes_distribution
        self.edges_distribution = edges_distribution
        self.additional_params = additional_params or dict()


class TrafficLightParams:
    """flow/core/params.py:29-196 — out of scope (rings/highways have none)."""

    def __init__(self, baseline=False):
        self.baseline = baseline

    def get_properties(self):
        return {}


# =============================================================================
# networks — flow/networks/ring.py, highway.py (closed form, no netconvert)
# =============================================================================
class Network:
    ADDITIONAL = {}
    TOPOLOGY = EN.TOPO_RING

    def __init__(self, name, vehicles, net_params, initial_config=None, traffic_lights=None,
                 detector_params=None):
        for p in self.ADDITIONAL:
            if p not in net_params.additional_params:
                raise KeyError('Network parameter "{}" not supplied'.format(p))
        self.orig_name = name
        self.name = name
        self.vehicles = vehicles
        self.net_params = net_params
        self.initial_config = initial_config or InitialConfig()
        self.traffic_lights = traffic_lights or TrafficLightParams()
        self.detector_params = detector_params


RING_NET_PARAMS = {'length': 230, 'lanes': 1, 'speed_limit': 30, 'resolution': 40}
HIGHWAY_NET_PARAMS = {'length': 1000, 'lanes': 4, 'speed_limit': 30, 'num_edges': 1}


class RingNetwork(Network):
    """flow/networks/ring.py:20-216: four equal edges bottom/right/top/left
    joined by 0.1 m junctions."""
    ADDITIONAL = RING_NET_PARAMS
    TOPOLOGY = EN.TOPO_RING

    def edge_starts(self, length=None):
        L = float(self.net_params.additional_params['length'] if length is None else length)
        j = RING_JUNCTION_LENGTH
        return [('bottom', 0.0), ('right', 0.25 * L + j), ('top', 0.5 * L + 2 * j),
                ('left', 0.75 * L + 3 * j)]


class HighwayNetwork(Network):
    """flow/networks/highway.py:20-170"""
    ADDITIONAL = HIGHWAY_NET_PARAMS
    TOPOLOGY = EN.TOPO_HIGHWAY


MERGE_NET_PARAMS = {'merge_length': 100, 'pre_merge_length': 200, 'post_merge_length': 100,
                    'merge_lanes': 1, 'highway_lanes': 1, 'speed_limit': 30}


class MergeNetwork(Network):
    """flow/networks/merge.py:27-221: a highway (inflow_highway -> left -> center) joined by an
    on-ramp (inflow_merge -> bottom -> center), single lane each."""
    ADDITIONAL = MERGE_NET_PARAMS
    TOPOLOGY = EN.TOPO_MERGE


FIGURE_EIGHT_NET_PARAMS = {'radius_ring': 30, 'lanes': 1, 'speed_limit': 30, 'resolution': 40}


class FigureEightNetwork(Network):
    """flow/networks/figure_eight.py:22-263: two rings joined by an intersection - one closed loop
    that passes the centre junction twice. One lane."""
    ADDITIONAL = FIGURE_EIGHT_NET_PARAMS
    TOPOLOGY = EN.TOPO_FIGURE8


_NET_BY_NAME = {'RingNetwork': RingNetwork, 'HighwayNetwork': HighwayNetwork,
                'MergeNetwork': MergeNetwork, 'FigureEightNetwork': FigureEightNetwork,
                'LoopScenario': RingNetwork, 'RingScenario': RingNetwork,
                'HighwayScenario': HighwayNetwork, 'MergeScenario': MergeNetwork,
                'Figure8Scenario': FigureEightNetwork, 'FigureEightScenario': FigureEightNetwork}


# =============================================================================
# gym.spaces.Box stand-in (gym is not a dependency of the engine)
# =============================================================================
class Box:
    def __init__(self, low, high, shape=None, dtype=np.float32):
        if shape is None:
            low, high = np.asarray(low, dtype=dtype), np.asarray(high, dtype=dtype)
            shape = low.shape
        else:
            low = np.full(shape, low, dtype=dtype)
            high = np.full(shape, high, dtype=dtype)
        self.low, self.high, self.shape, self.dtype = low, high, tuple(shape), np.dtype(dtype)
        self._rng = np.random.default_rng()

    def seed(self, s=None):
        self._rng = np.random.default_rng(s)

    def sample(self):
        lo = np.where(np.isfinite(self.low), self.low, -1.0)
        hi = np.where(np.isfinite(self.high), self.high, 1.0)
        return self._rng.uniform(lo, hi).astype(self.dtype)

    def contains(self, x):
        x = np.asarray(x)
        return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

    def __repr__(self):
        return 'Box%s' % (self.shape,)


def _box(*a, **k):
    try:  # hand RLlib a real gym space when gym is installed
        from gym.spaces.box import Box as GymBox
        return GymBox(*a, **k)
    except Exception:
        return Box(*a, **k)


# =============================================================================
# flow_params -> engine configuration
# =============================================================================
class EngineSpec:
    """Everything fs_create / fs_set_types need, derived on the host."""

    def __init__(self, types, type_of, kw, veh_ids, type_names, capacity=None):
        self.types, self.type_of, self.kw = types, type_of, kw
        self.veh_ids, self.type_names, self.capacity = veh_ids, type_names, capacity


def _vtype_row(ini, lc_env):
    cls, kwargs = ini['acceleration_controller']
    ctrl = _resolve(cls, _ACCEL_BY_NAME, 'acceleration controller')(
        veh_id=ini['veh_id'], car_following_params=ini['car_following_params'],
        **dict(kwargs or {}))
    lc_cls = _resolve(ini['lane_change_controller'][0], _LC_BY_NAME, 'lane-change controller')
    cf = ini['car_following_params'].controller_params
    lc = lc_cls.ENGINE
    if lc_env and ctrl.ENGINE == EN.CTRL_RL:
        lc = EN.LC_RL   # LaneChangeAccelEnv._apply_rl_actions, lane_change_accel.py:132-154
    if ctrl.fail_safe not in EN.FAILSAFE:
        raise ValueError('unknown fail_safe %r' % (ctrl.fail_safe,))
    length = ini.get('length') or VEHICLE_LENGTH
    return EN.make_vtype(ctrl.ENGINE, ctrl.engine_params(), fail_safe=ctrl.fail_safe,
                         noise=ctrl.noise, delay=ctrl.delay, max_accel=cf['accel'],
                         max_decel=cf['decel'], length=length, sim_tau=cf['tau'],
                         sim_min_gap=cf['minGap'], sim_max_speed=cf['maxSpeed'],
                         lc_controller=lc,
                         # setSpeedMode at departure (vehicle/traci.py:343-347): the restated
                         # clamp on the speeds Flow commands (include/flowstep.h fs_vtype)
                         speed_mode=int(ini['car_following_params'].speed_mode))


def _inflow_rows(network, names, types, rl_ok, arms=None):
    """InFlows.get() -> [(type, vehs_per_hour, depart_speed, begin, end, probability, arm)]
    (flow/core/params.py:1135-1236 holds SUMO's <flow> attributes)."""
    flows = network.net_params.inflows.get()
    if len(flows) > 4:
        raise NotImplementedError('more than four inflows')
    rows = []
    for f in flows:
        rate, prob = f.get('vehsPerHour'), float(f.get('probability') or 0.0)
        begin, end = float(f['begin']), float(f.get('end') or 0.0)
        if rate is None and f.get('period') is not None:
            rate = 3600.0 / float(f['period'])
        if f.get('number') is not None and end > begin:
            # SUMO: `number` vehicles, evenly spaced over [begin, end) when no rate is given,
            # else the flow ends after `number` vehicles
            if rate is None and not prob:
                rate = 3600.0 * float(f['number']) / (end - begin)
            elif rate is not None:
                end = min(end, begin + float(f['number']) * 3600.0 / float(rate))
        if rate is None and not prob:
            raise ValueError('inflow %r: one of vehs_per_hour, probability, period, number is '
                             'needed' % (f.get('name'),))
        if isinstance(f['departSpeed'], str):
            raise NotImplementedError('depart_speed=%r: SUMO picks the speed; only numeric '
                                      'departure speeds are compiled' % (f['departSpeed'],))
        if f['vtype'] not in names:
            raise KeyError('inflow of unknown vehicle type %r' % (f['vtype'],))
        t = names.index(f['vtype'])
        if types[t].controller == EN.CTRL_RL and not rl_ok:
            raise NotImplementedError('an RL inflow needs MultiAgentHighwayPOEnv')
        arm = 0
        if arms is not None:
            if f['edge'] not in arms:
                raise NotImplementedError('inflow on edge %r: vehicles enter at %s'
                                          % (f['edge'], ' / '.join(arms)))
            arm = arms[f['edge']]
        rows.append((t, float(rate or 0.0), float(f['departSpeed']), begin, end, prob, arm))
    return rows


def compile_spec(env_cls, env_params, sim_params, network, seed=None):
    """Build the engine configuration an env of class `env_cls` needs.
    Raises the reference's exceptions for the reference's error cases."""
    ap = env_params.additional_params
    for p in env_cls.ADDITIONAL:   # accel.py:70-73 etc.
        if p not in ap:
            raise KeyError('Environment parameter "{}" not supplied'.format(p))
    veh, net, ini = network.vehicles, network.net_params.additional_params, network.initial_config
    if ini.spacing not in ('uniform', 'random'):
        raise NotImplementedError('InitialConfig(spacing=%r): only "uniform" and "random" placements '
                                  'exist here ("custom" belongs to networks that are not built)'
                                  % (ini.spacing,))
    if ini.shuffle and len([r for r in veh.initial if r['num_vehicles'] > 0]) > 1:
        # random.shuffle(initial_ids) moves every vehicle TYPE to another start position in every
        # env; the engine's type table is per vehicle index, shared by the batch. With one type
        # (what the reference's own shuffle test and non_rl/highway.py use) only the names move.
        raise NotImplementedError('InitialConfig(shuffle=True) with several vehicle types')
    if ini.edges_distribution != 'all':
        raise NotImplementedError('edges_distribution other than "all"')
    sort_on = bool(ap.get('sort_vehicles')) and env_cls.KIND in (
        EN.ENV_ACCEL, EN.ENV_LANE_CHANGE_ACCEL, EN.ENV_LANE_CHANGE_ACCEL_PO)
    if sort_on and network.TOPOLOGY != EN.TOPO_RING:
        raise NotImplementedError('sort_vehicles=True on the open highway (slots are re-packed)')
    lc_env = env_cls.KIND in (EN.ENV_LANE_CHANGE_ACCEL, EN.ENV_LANE_CHANGE_ACCEL_PO)
    merge = network.TOPOLOGY == EN.TOPO_MERGE
    if merge != (env_cls.KIND == EN.ENV_MERGE_PO):
        raise NotImplementedError('MergePOEnv runs on MergeNetwork, and only it does')
    n_lanes = 1 if merge else int(net['lanes'])
    if merge and (int(net['merge_lanes']) != 1 or int(net['highway_lanes']) != 1):
        raise NotImplementedError('merge: one lane on the highway and on the ramp only')
    if merge and (ini.spacing != 'uniform' or ini.perturbation > 0 or ini.shuffle):
        raise NotImplementedError('merge: the even initial placement only')
    types, names, type_of, ids = [], [], [], []
    for t, row in enumerate(veh.initial):
        lcp = row.get('lane_change_params')
        mode = getattr(lcp, 'lane_change_mode', 512)
        if n_lanes > 1 and isinstance(mode, (int, float)) and int(mode) & 0xFF:
            # bits 0-7 = SUMO's autonomous strategic / cooperative / speed-gain / keep-right
            # changes (LC2013 / SL2015): not restated, and never silently dropped
            raise NotImplementedError(
                'lane_change_mode %r of "%s" lets SUMO change lanes on its own; only modes without '
                'autonomous changes (512 "no_lat_collide", 0 "aggressive") are compiled'
                % (mode, row['veh_id']))
        types.append(_vtype_row(row, lc_env))
        names.append(row['veh_id'])
        type_of += [t] * row['num_vehicles']
        ids += ['%s_%d' % (row['veh_id'], i) for i in range(row['num_vehicles'])]
    if len(types) > 16:
        raise NotImplementedError('more than 16 vehicle types')
    horizon = env_params.horizon
    kw = dict(
        topology=network.TOPOLOGY, env_kind=env_cls.KIND, n_lanes=n_lanes,
        sim_step=float(sim_params.sim_step), speed_limit=float(net['speed_limit']),
        x0=float(ini.x0), bunching=float(ini.bunching), init_min_gap=float(ini.min_gap),
        lanes_distribution=int(min(ini.lanes_distribution, 1 << 30)),
        warmup_steps=int(env_params.warmup_steps),
        horizon=int(horizon) if math.isfinite(horizon) else (1 << 30),
        sims_per_step=int(env_params.sims_per_step),
        clip_actions=int(bool(env_params.clip_actions)),
        target_velocity=float(ap.get('target_velocity', 10.0)),
        act_lo=-abs(float(ap.get('max_decel', 4.5))), act_hi=float(ap.get('max_accel', 2.6)),
        lane_change_duration=int(ap.get('lane_change_duration', 5)),
        # AccelEnv.compute_reward is the only reader of env_params.evaluate on this path
        # (accel.py:111-112); sorted_ids: accel.py:150-169
        evaluate=int(bool(env_params.evaluate) and
                     env_cls.KIND in (EN.ENV_ACCEL, EN.ENV_MA_HIGHWAY_PO)),
        sort_vehicles=int(sort_on),
        use_ballistic=int(bool(getattr(sim_params, 'use_ballistic', False))),
        seed=int(seed if seed is not None else (getattr(sim_params, 'seed', None) or 0)))
    capacity = None
    if network.TOPOLOGY == EN.TOPO_RING:
        ring = ap.get('ring_length')
        L = float(net['length'])
        lo, hi = (float(ring[0]), float(ring[1])) if ring else (L, L)
        kw.update(length_lo=lo, length_hi=hi, junction_len=RING_JUNCTION_LENGTH)
        if env_cls.KIND in (EN.ENV_WAVE_ATTENUATION_PO, EN.ENV_MA_WAVE_ATTENUATION_PO) and ring:
            kw['obs_max_length'] = hi        # wave_attenuation.py:253-256
        # network/base.py:599-605: no room -> the reference aborts the run
        lanes_d = int(min(max(ini.lanes_distribution, 1), n_lanes))
        room = lo * lanes_d - lanes_d * max(ini.bunching, 0) - \
            len(ids) * (max(ini.min_gap, 0) + VEHICLE_LENGTH)
        if room < 0:
            raise FatalFlowError('There is not enough space to place all vehicles in the '
                                 'network.')
    elif network.TOPOLOGY == EN.TOPO_FIGURE8:
        if env_cls.KIND != EN.ENV_ACCEL or n_lanes != 1 or sort_on:
            raise NotImplementedError('figure eight: AccelEnv on one lane, sort_vehicles False')
        if ini.spacing != 'uniform' or ini.perturbation > 0:
            raise NotImplementedError('figure eight: the even initial placement only')
        from . import placement as PL
        try:      # network/base.py:599-605: no room -> the reference aborts the run
            PL.even_positions(*PL.figure8_edges(float(net['radius_ring'])), ini.x0, ini.min_gap,
                              ini.bunching, len(ids))
        except ValueError as exc:
            raise FatalFlowError(str(exc))
        kw.update(fig8_radius=float(net['radius_ring']))
    elif merge:
        pre, ramp, post = (float(net[k]) for k in ('pre_merge_length', 'merge_length',
                                                   'post_merge_length'))
        rows = _inflow_rows(network, names, types, rl_ok=True,
                            arms={'inflow_highway': 0, 'inflow_merge': 1})
        if any(r[4] < 86400 or r[5] for r in rows):
            raise NotImplementedError('merge: periodic inflows (vehs_per_hour / period) without an '
                                      'end only')
        kw.update(merge_pre=pre, merge_ramp=ramp, merge_post=post,
                  max_agents=int(ap['num_rl']), inflows=[r[:4] + (0.0, 0.0, r[6]) for r in rows])
        from . import placement as PL
        real, internal = PL.merge_edges(pre, ramp, post)
        try:      # network/base.py:599-605: no room -> the reference aborts the run
            PL.even_positions(real, internal, ini.x0, ini.min_gap, ini.bunching, len(ids))
        except ValueError as exc:
            raise FatalFlowError(str(exc))
        # slots: both arms and the merged road bumper to bumper is the physical maximum
        road = 2 * PL.MERGE_INFLOW_EDGE_LEN + pre + ramp + post + 45.2
        capacity = min(256, max(64, int(math.ceil(road / (VEHICLE_LENGTH + 2.5) / 32) * 32)))
        if len(ids) > capacity:
            raise NotImplementedError('merge: more than %d vehicles' % capacity)
    else:
        if int(net['lanes']) != 1 or int(net.get('num_edges', 1)) != 1:
            raise NotImplementedError('highway: single lane, single edge only')
        ma = env_cls.KIND == EN.ENV_MA_HIGHWAY_PO
        kw.update(length_lo=float(net['length']), length_hi=float(net['length']),
                  junction_len=0.0, env_kind=EN.ENV_MA_HIGHWAY_PO if ma else EN.ENV_HIGHWAY)
        rows = [r[:6] for r in _inflow_rows(network, names, types, rl_ok=ma)]
        if len(rows) == 1 and not rows[0][5] and rows[0][4] >= 86400:   # the single <flow> of fs_config
            kw.update(inflow_type=rows[0][0], inflow_vehs_per_hour=rows[0][1],
                      inflow_depart_speed=rows[0][2], inflow_begin=rows[0][3])
        elif rows:             # fs_set_inflows
            kw['inflows'] = rows
        # slots: one vehicle per (length + minGap) of road is the physical maximum
        capacity = max(32, int(math.ceil(float(net['length']) / (VEHICLE_LENGTH + 1.0) / 32) * 32))
        capacity = min(capacity, 256)
        if ma:
            kw['max_agents'] = int(min(capacity, ap.get('max_agents', env_cls.MAX_AGENTS)))
    return EngineSpec(types, type_of, kw, ids, names, capacity)


# =============================================================================
# kernel facade: env.k.vehicle / env.k.network / env.k.simulation
# =============================================================================
class _VehicleKernel:
    """The KernelVehicle getters the reference's envs, rewards and Experiment
    call (flow/core/kernel/vehicle/traci.py:427-689), answered from the
    engine's device buffers for ONE env of the batch (index `env.view_env`).
    A host copy is fetched lazily once per step."""

    def __init__(self, env):
        self._env = env
        self._cache = None

    def _snap(self):
        if self._cache is None:
            e, i = self._env.engine, self._env.view_env
            g = lambda t: None if t is None else t[:, i].cpu().numpy()  # noqa: E731
            self._cache = c = dict(pos=g(e.pos), vel=g(e.vel), lane=g(e.lane), head=g(e.headway),
                                   lead=g(e.leader), foll=g(e.follower), act=g(e.active),
                                   C=float(e.length[i]), names=None)
            if e.veh_id is not None and hasattr(self._env, 'vehicle_name'):
                # vehicles come and go: every occupied slot answers to the name SUMO would use
                vid, c['vtype'] = e.veh_id[i].cpu().numpy(), e.veh_type[i].cpu().numpy()
                c['slot_names'] = [self._env.vehicle_name(v) if c['act'][k] and v >= 0 else None
                                   for k, v in enumerate(vid)]
                c['names'] = {n: k for k, n in enumerate(c['slot_names']) if n is not None}
        return self._cache

    def _invalidate(self):
        self._cache = None

    def _idx(self, veh_id):
        names = self._snap()['names']
        return (self._env._index if names is None else names).get(veh_id, -1)

    def _vtype(self, i):
        s = self._snap()
        t = s['vtype'][i] if s['names'] is not None else self._env.spec.type_of[i]
        return self._env.spec.types[int(t)], self._env.spec.type_names[int(t)]

    def _get(self, key, veh_id, error):
        if isinstance(veh_id, (list, tuple, np.ndarray)):
            return [self._get(key, v, error) for v in veh_id]
        i = self._idx(veh_id)
        s = self._snap()
        if i < 0 or (s['act'] is not None and not s['act'][i]):
            return error
        return s[key][i].item()

    # ids ------------------------------------------------------------------
    def get_ids(self):
        s, ids = self._snap(), self._env.spec.veh_ids
        if s['names'] is not None:
            return list(s['names'])
        if self._env.spec.capacity is None:
            return list(ids)
        return [v for i, v in enumerate(ids) if s['act'][i]]

    def get_rl_ids(self):
        s = self._snap()
        if s['names'] is not None:     # kept sorted by the reference (vehicle/traci.py:259,366)
            return sorted(n for n, k in s['names'].items()
                          if self._vtype(k)[0].controller == EN.CTRL_RL)
        return [v for v in self._env.spec.veh_ids if self._env._is_rl[self._env._index[v]]]

    def get_human_ids(self):
        s = self._snap()
        if s['names'] is not None:
            return [n for n, k in s['names'].items() if self._vtype(k)[0].controller != EN.CTRL_RL]
        return [v for v in self._env.spec.veh_ids if not self._env._is_rl[self._env._index[v]]]

    def get_controlled_ids(self):
        sim = (EN.CTRL_SIM, EN.CTRL_RL)
        return [v for v in self.get_ids() if self._vtype(self._idx(v))[0].controller not in sim]

    @property
    def num_vehicles(self):
        return len(self.get_ids())

    @property
    def num_rl_vehicles(self):
        return len(self.get_rl_ids())

    def get_type(self, veh_id):
        return self._vtype(self._idx(veh_id))[1]

    # state ----------------------------------------------------------------
    def get_speed(self, veh_id, error=-1001):
        return self._get('vel', veh_id, error)

    def _merge_lengths(self):
        net = self._env.net_params.additional_params
        return (float(net['pre_merge_length']), float(net['merge_length']),
                float(net['post_merge_length']))

    def _arm_pos(self, veh_id):
        """merge: (arm, metres along it) of a vehicle, None when it is not in the network"""
        p = self._get('pos', veh_id, None)
        return None if p is None else (int(self._get('lane', veh_id, 0)), p)

    def _fig8_radius(self):
        return float(self._env.net_params.additional_params['radius_ring'])

    def get_x_by_id(self, veh_id):
        if self._env.network.TOPOLOGY == EN.TOPO_FIGURE8:
            # node slots report their start, everything else start + position (network/traci.py:273-289)
            if isinstance(veh_id, (list, tuple, np.ndarray)):
                return [self.get_x_by_id(v) for v in veh_id]
            from . import placement as PL
            s = self._get('pos', veh_id, None)
            if s is None:
                return -1001
            name, p = PL.figure8_edge_of(s, self._fig8_radius())
            return s - p if name.endswith('_0') and not name.startswith(':center') else (s - p) + p
        if self._env.network.TOPOLOGY == EN.TOPO_MERGE:
            if isinstance(veh_id, (list, tuple, np.ndarray)):
                return [self.get_x_by_id(v) for v in veh_id]
            from . import placement as PL
            ap = self._arm_pos(veh_id)
            return -1001 if ap is None else PL.merge_x(ap[0], ap[1], *self._merge_lengths())
        return self._get('pos', veh_id, -1001)

    def get_lane(self, veh_id, error=-1001):
        if self._env.network.TOPOLOGY == EN.TOPO_MERGE:    # the lane buffer holds the arm
            if isinstance(veh_id, (list, tuple, np.ndarray)):
                return [self.get_lane(v, error) for v in veh_id]
            return error if self._get('pos', veh_id, None) is None else 0
        return self._get('lane', veh_id, error)

    def get_length(self, veh_id, error=-1001):
        if isinstance(veh_id, (list, tuple, np.ndarray)):
            return [self.get_length(v, error) for v in veh_id]
        i = self._idx(veh_id)
        return error if i < 0 else float(self._vtype(i)[0].length)

    def get_edge(self, veh_id, error=''):
        """Edge name from the loop coordinate (ring.py:194-216 edge starts;
        a junction slot reports the internal edge ':name_0')."""
        if isinstance(veh_id, (list, tuple, np.ndarray)):
            return [self.get_edge(v, error) for v in veh_id]
        x = self._get('pos', veh_id, None)
        if x is None:
            return error
        if self._env.network.TOPOLOGY == EN.TOPO_FIGURE8:
            from . import placement as PL
            return PL.figure8_edge_of(x, self._fig8_radius())[0]
        if self._env.network.TOPOLOGY == EN.TOPO_MERGE:
            from . import placement as PL
            return PL.merge_edge_of(self._arm_pos(veh_id)[0], x, *self._merge_lengths())[0]
        return self._env._edge_of(x)[0]

    def get_position(self, veh_id, error=-1001):
        """Position relative to the start of the current edge."""
        if isinstance(veh_id, (list, tuple, np.ndarray)):
            return [self.get_position(v, error) for v in veh_id]
        x = self._get('pos', veh_id, None)
        if x is None:
            return error
        if self._env.network.TOPOLOGY == EN.TOPO_FIGURE8:
            from . import placement as PL
            return PL.figure8_edge_of(x, self._fig8_radius())[1]
        if self._env.network.TOPOLOGY == EN.TOPO_MERGE:
            from . import placement as PL
            return PL.merge_edge_of(self._arm_pos(veh_id)[0], x, *self._merge_lengths())[1]
        return x - self._env._edge_of(x)[1]

    def get_headway(self, veh_id, error=-1001):
        self._env._need_views()
        return self._get('head', veh_id, error)

    def _neighbour(self, key, veh_id, error):
        self._env._need_views()
        if isinstance(veh_id, (list, tuple, np.ndarray)):
            return [self._neighbour(key, v, error) for v in veh_id]
        j = self._get(key, veh_id, None)
        if j is None:
            return error          # unknown vehicle
        if j < 0:
            return None                                       # nobody ahead / behind
        s = self._snap()
        return self._env.spec.veh_ids[j] if s['names'] is None else s['slot_names'][j]

    def get_leader(self, veh_id, error=''):
        return self._neighbour('lead', veh_id, error)

    def get_follower(self, veh_id, error=''):
        return self._neighbour('foll', veh_id, error)

    # rates (vehicle/traci.py:493-505) -------------------------------------------
    def _rate(self, row, time_span):
        e = self._env
        if e.engine.counters is None or e.spec.capacity is None:
            return 0
        hist = e._count_hist          # cumulative (departed, arrived), entry 0 = at reset
        span = min(len(hist) - 1, int(time_span / e.sim_step))
        if span <= 0:
            return 0
        return 3600.0 * (hist[-1][row] - hist[-1 - span][row]) / (span * e.sim_step)

    def get_inflow_rate(self, time_span):
        return self._rate(0, time_span)

    def get_outflow_rate(self, time_span):
        return self._rate(1, time_span)


class _NetworkKernel:
    """flow/core/kernel/network/traci.py:273-325 in closed form."""

    def __init__(self, env):
        self._env = env
        self.network = env.network

    def length(self):
        return float(self._env.engine.length[self._env.view_env])

    def non_internal_length(self):
        if self._env.network.TOPOLOGY == EN.TOPO_RING:
            return self.length() - 4 * RING_JUNCTION_LENGTH
        if self._env.network.TOPOLOGY == EN.TOPO_MERGE:
            return self.length() - (2 * 0.1 + 22.5)
        if self._env.network.TOPOLOGY == EN.TOPO_FIGURE8:
            return self.length() - (2 * 6.2 + 4 * 0.28)
        return self.length()

    def max_speed(self):
        return float(self._env.network.net_params.additional_params['speed_limit'])

    def speed_limit(self, edge=None):
        return self.max_speed()

    def num_lanes(self, edge=None):
        return int(self._env.network.net_params.additional_params.get('lanes', 1))

    def get_edge_list(self):
        if self._env.network.TOPOLOGY == EN.TOPO_RING:
            return ['bottom', 'right', 'top', 'left']
        if self._env.network.TOPOLOGY == EN.TOPO_MERGE:
            return ['inflow_highway', 'left', 'inflow_merge', 'bottom', 'center']
        if self._env.network.TOPOLOGY == EN.TOPO_FIGURE8:
            return ['bottom', 'top', 'right', 'left', 'upper_ring', 'lower_ring']
        return ['highway_0']

    def get_junction_list(self):
        if self._env.network.TOPOLOGY == EN.TOPO_RING:
            return [':right_0', ':top_0', ':left_0', ':bottom_0']
        if self._env.network.TOPOLOGY == EN.TOPO_MERGE:
            return [':left', ':bottom', ':center']
        if self._env.network.TOPOLOGY == EN.TOPO_FIGURE8:
            return [':bottom_0', ':center_1', ':top_0', ':right_0', ':center_0', ':left_0']
        return []

    def edge_length(self, edge):
        if self._env.network.TOPOLOGY == EN.TOPO_MERGE:
            net = self._env.network.net_params.additional_params
            return {'inflow_highway': 100.0, 'inflow_merge': 100.0, ':left': 0.1, ':bottom': 0.1,
                    ':center': 22.5, 'left': float(net['pre_merge_length']),
                    'bottom': float(net['merge_length']),
                    'center': float(net['post_merge_length'])}[edge]
        if self._env.network.TOPOLOGY == EN.TOPO_FIGURE8:
            from . import placement as PL
            r = float(self._env.network.net_params.additional_params['radius_ring'])
            if edge.startswith(':'):
                return 6.2 if edge.startswith(':center') else 0.28
            return {n: ln for n, _, ln in PL.figure8_edges(r)[0]}[edge]
        if edge.startswith(':'):
            return RING_JUNCTION_LENGTH
        n = len(self.get_edge_list())
        return self.non_internal_length() / n


class _SimulationKernel:
    def __init__(self, env):
        self._env = env

    def check_collision(self):
        return bool(self._env.engine.crash[self._env.view_env].item())

    def close(self):
        pass


class _Kernel:
    def __init__(self, env):
        self.vehicle = _VehicleKernel(env)
        self.network = _NetworkKernel(env)
        self.simulation = _SimulationKernel(env)
        self.traffic_light = None
        self.detector = None

    def close(self):
        pass


# =============================================================================
# environments — flow/envs/base.py + flow/envs/ring/*.py
# =============================================================================
class Env:
    """Base env: the reference's `Env` (flow/envs/base.py) with the per-vehicle
    loop replaced by fused CUDA launches."""
    KIND = EN.ENV_ACCEL
    ADDITIONAL = {}

    def __init__(self, env_params, sim_params, network, simulator='b200', n_envs=1,
                 dtype=None, device=0, env_id_base=0, seed=None, record_views=None,
                 lc_corrected=False):
        if simulator not in ('b200', 'traci'):
            # 'traci' is accepted so unmodified flow_params run; 'aimsun' is out of scope
            raise FatalFlowError('simulator %r is not supported by the B200 engine' % simulator)
        if getattr(sim_params, 'render', False) not in (False, None, 0):
            raise FatalFlowError('Mode %r is not supported (headless engine: no renderer)'
                                 % (sim_params.render,))
        self.env_params, self.sim_params = env_params, sim_params
        self.network, self.simulator = network, simulator
        self.initial_vehicles = network.vehicles
        self.net_params, self.initial_config = network.net_params, network.initial_config
        self.sim_step = sim_params.sim_step
        self.n_envs = int(n_envs)
        self.view_env = 0
        self.spec = compile_spec(type(self), env_params, sim_params, network, seed=seed)
        self.dtype = dtype or ('f64' if self.n_envs == 1 else 'f32')
        if record_views is None:
            record_views = self.n_envs == 1
        self.engine = EN.Engine(self.spec.types, self.spec.type_of, self.n_envs,
                                dtype=self.dtype, device=device, capacity=self.spec.capacity,
                                env_id_base=env_id_base, record_views=int(bool(record_views)),
                                lc_corrected=int(bool(lc_corrected)), **self.spec.kw)
        if self.spec.capacity is not None:   # highway: vehicles come and go -> slot ids
            self._initial_ids = list(self.spec.veh_ids)
            self.spec.veh_ids = ['slot_%d' % i for i in range(self.engine.n_veh)]
            self.spec.type_of = list(self.spec.type_of) + [self.spec.kw.get('inflow_type', 0)] * \
                (self.engine.n_veh - len(self.spec.type_of))
        self._index = {v: i for i, v in enumerate(self.spec.veh_ids)}
        self._is_rl = [self.spec.types[t].controller == EN.CTRL_RL for t in self.spec.type_of]
        self._count_hist = []
        self.k = _Kernel(self)
        self.time_counter = 0
        self.step_counter = 0
        self.state = None
        self.obs_var_labels = []
        # placements the reference draws from host random streams (flow_b200/placement.py)
        ic = self.initial_config
        self._host_placement = ic.spacing == 'random' or ic.perturbation > 0 or bool(ic.shuffle) \
            or network.TOPOLOGY == EN.TOPO_MERGE     # five unequal edges: placed on the host
        self._placed = None
        if self._host_placement and self.spec.kw.get('length_lo') != self.spec.kw.get('length_hi'):
            # WaveAttenuationEnv.reset builds a NEW InitialConfig(bunching=50, min_gap=0) for every
            # ring length it draws (wave_attenuation.py:171-174): the options never take effect
            import warnings
            warnings.warn('InitialConfig(perturbation / spacing="random" / shuffle) has no effect on '
                          'an env that redraws its ring length at reset, as in the reference')
            self._host_placement = False
        if self._host_placement:
            if self.spec.capacity is not None and (ic.spacing == 'random' or ic.perturbation > 0) \
                    and len(set(self.spec.type_of[:len(self._initial_ids)])) > 1:
                raise NotImplementedError('perturbed / random start positions on the highway with '
                                          'several vehicle types (slots are kept head first)')
            # Env.__init__ -> setup_initial_state (base.py:183-199): one shuffle + one draw here,
            # so that the host streams stand where the reference's stand after construction
            self._draw_host_placement()

    # spaces -----------------------------------------------------------------
    @property
    def action_space(self):
        ap, n = self.env_params.additional_params, self.initial_vehicles.num_rl_vehicles
        return _box(low=-abs(ap['max_decel']), high=ap['max_accel'], shape=(n,),
                    dtype=np.float32)

    @property
    def observation_space(self):
        return _box(low=0, high=1, shape=(2 * self.initial_vehicles.num_vehicles,),
                    dtype=np.float32)

    # helpers ----------------------------------------------------------------
    def _need_views(self):
        """Leader / headway / follower views cost bandwidth, so a batched env only
        keeps them on request (record_views=True); a single env always has them."""
        if self.engine.headway is None:
            raise RuntimeError('get_headway / get_leader / get_follower need '
                               'record_views=True on a batched env')

    def _draw_host_placement(self):
        """setup_initial_state (flow/envs/base.py:268-292): shuffle the ids, then generate the
        start positions - with the reference's own host streams (placement.py). Leaves the
        engine placed (stage FS_RESET_PLACE done) and the drawn state in self._placed."""
        from . import placement as PL
        ic, eng, ring = self.initial_config, self.engine, self.network.TOPOLOGY == EN.TOPO_RING
        if ic.shuffle:
            if self.spec.capacity is not None:
                self._initial_ids = PL.shuffled(self._initial_ids)
            else:
                self.spec.veh_ids = PL.shuffled(self.spec.veh_ids)
                self._index = {v: i for i, v in enumerate(self.spec.veh_ids)}
                self.rl_ids = [v for i, v in enumerate(self.spec.veh_ids) if self._is_rl[i]]
        eng.reset_stage(eng.RESET_PLACE)
        self._placed = None
        if ic.spacing != 'random' and not ic.perturbation > 0:
            return                                    # shuffle only: the even placement stands
        pos = eng.pos.double().cpu().numpy().copy()   # [V or slots, E]
        lane = eng.lane.cpu().numpy().copy()
        circ = eng.length.double().cpu().numpy()
        lanes = int(self.net_params.additional_params['lanes'])
        n = len(self._initial_ids) if self.spec.capacity is not None else eng.n_veh
        for e in range(self.n_envs):
            edges = PL.ring_edges(circ[e] - 4 * RING_JUNCTION_LENGTH) if ring else \
                PL.highway_edges(circ[e])
            if ic.spacing == 'random':
                try:
                    xs, ls = PL.random_positions(edges, lanes, ic.lanes_distribution, ic.min_gap,
                                                 ic.bunching, n)
                except ValueError as exc:
                    raise FatalFlowError(str(exc))
            else:
                order = list(range(n)) if ring else sorted(range(n), key=lambda i: pos[i, e])
                xs = PL.perturb([pos[i, e] for i in order], edges, ic.perturbation)
                ls = [int(lane[i, e]) for i in order]
            if not ring:      # highway slots are kept head first (largest x in slot 0)
                by_x = sorted(range(n), key=lambda i: -xs[i])
                xs, ls = [xs[i] for i in by_x], [ls[i] for i in by_x]
            pos[:n, e], lane[:n, e] = xs, ls
        self._placed = (pos, lane)

    def _reset_host_placed(self):
        """Env.reset with host-drawn placements: with shuffle the reference redraws everything at
        every reset (base.py:466-467), otherwise it re-adds the vehicles at the positions drawn
        once at construction (base.py:497-530)."""
        eng = self.engine
        if self.initial_config.shuffle:
            self._draw_host_placement()
        else:
            eng.reset_stage(eng.RESET_PLACE)
        if self._placed is not None:
            eng.set_state(pos=self._placed[0], lane=self._placed[1])
        return eng.reset_stage(eng.RESET_WARMUP)

    def _edge_of(self, x):
        """(edge name, edge start) of loop coordinate x."""
        if self.network.TOPOLOGY != EN.TOPO_RING:
            return 'highway_0', 0.0
        L = self.k.network.non_internal_length()
        starts = self.network.edge_starts(L)
        j = RING_JUNCTION_LENGTH
        internal = [(':right_0', 0.25 * L), (':top_0', 0.5 * L + j),
                    (':left_0', 0.75 * L + 2 * j), (':bottom_0', L + 3 * j)]
        best = max((s for s in starts + internal if s[1] <= x + 1e-12), key=lambda s: s[1])
        return best

    def _out(self, obs, rew, done):
        if self.n_envs == 1:
            return (obs[0].double().cpu().numpy(), float(rew[0]), bool(done[0]), {})
        return obs, rew, done, {}

    def _to_engine_actions(self, rl_actions):
        """gym layout ([act_dim] or [E, act_dim]) -> engine layout [act_dim, E]."""
        import torch
        if rl_actions is None:
            return None
        if isinstance(rl_actions, dict):
            raise NotImplementedError('multi-agent action dicts')
        td = torch.float64 if self.dtype == 'f64' else torch.float32
        if isinstance(rl_actions, torch.Tensor):
            a = rl_actions.to(device='cuda:%d' % self.engine.device, dtype=td)
        else:
            a_np = np.asarray(rl_actions, dtype=np.float64)
            self._check_host_actions(a_np)
            a = torch.as_tensor(a_np, dtype=td).to('cuda:%d' % self.engine.device)
        a = a.reshape(self.n_envs, self.engine.act_dim) if a.numel() == \
            self.n_envs * self.engine.act_dim else None
        if a is None:
            raise ValueError('rl_actions must have n_envs*act_dim = %d elements'
                             % (self.n_envs * self.engine.act_dim))
        return a.t().contiguous()

    def _check_host_actions(self, a):
        pass

    # gym API ----------------------------------------------------------------
    def reset(self):
        """flow/envs/base.py:414-560: new episode (ring length draw, even
        placement, warm-up steps with rl_actions=None) for every env."""
        self.time_counter = 0
        obs = self._reset_host_placed() if self._host_placement else self.engine.reset()
        self._count_hist = []
        self._log_counts()
        self.k.vehicle._invalidate()
        self.state = obs
        return obs[0].double().cpu().numpy() if self.n_envs == 1 else obs

    def step(self, rl_actions):
        """flow/envs/base.py:294-412 for every env of the batch."""
        a = self._to_engine_actions(rl_actions)
        obs, rew, done = self.engine.step(a, k=1)
        self.time_counter += self.env_params.sims_per_step
        self.step_counter += self.env_params.sims_per_step
        self.k.vehicle._invalidate()
        self._log_counts()
        self.state = obs
        return self._out(obs, rew, done)

    def _log_counts(self):
        if self.engine.counters is not None and self.spec.capacity is not None:
            c = self.engine.counters[:, self.view_env].cpu().numpy()
            self._count_hist.append((int(c[1]), int(c[2])))

    def rollout(self, actions, obs_seq=None, reward_seq=None, done_seq=None):
        """K fused steps in ONE launch (no reference counterpart: the reason the
        engine exists). actions: device tensor [K, act_dim, E] or None + K."""
        k = int(actions.shape[0])
        self.engine.step(actions, k=k, obs_seq=obs_seq, reward_seq=reward_seq,
                         done_seq=done_seq)
        self.time_counter += k * self.env_params.sims_per_step
        self.k.vehicle._invalidate()
        return obs_seq, reward_seq, done_seq

    def clip_actions(self, rl_actions=None):
        """flow/envs/base.py:566-597 (host-side helper; the engine clips too)."""
        if rl_actions is None:
            return None
        sp = self.action_space
        return np.clip(rl_actions, a_min=sp.low, a_max=sp.high)

    def terminate(self):
        self.engine.close()

    close = terminate

    def render(self, *a, **k):
        return None


class AccelEnv(Env):
    """flow/envs/ring/accel.py:25-165"""
    KIND = EN.ENV_ACCEL
    ADDITIONAL = {'max_accel': 3, 'max_decel': 3, 'target_velocity': 10, 'sort_vehicles': False}

    @property
    def sorted_ids(self):
        return self.k.vehicle.get_ids()


class WaveAttenuationEnv(Env):
    """flow/envs/ring/wave_attenuation.py:33-210"""
    KIND = EN.ENV_WAVE_ATTENUATION
    ADDITIONAL = {'max_accel': 1, 'max_decel': 1, 'ring_length': [220, 270]}


class WaveAttenuationPOEnv(WaveAttenuationEnv):
    """flow/envs/ring/wave_attenuation.py:213-281"""
    KIND = EN.ENV_WAVE_ATTENUATION_PO

    @property
    def observation_space(self):
        return _box(low=-float('inf'), high=float('inf'), shape=(3,), dtype=np.float32)


class LaneChangeAccelEnv(AccelEnv):
    """flow/envs/ring/lane_change_accel.py:29-165"""
    KIND = EN.ENV_LANE_CHANGE_ACCEL
    ADDITIONAL = {'max_accel': 3, 'max_decel': 3, 'lane_change_duration': 5,
                  'target_velocity': 10, 'sort_vehicles': False}

    @property
    def action_space(self):
        ap, n = self.env_params.additional_params, self.initial_vehicles.num_rl_vehicles
        return _box(np.array([-abs(ap['max_decel']), -1] * n),
                    np.array([ap['max_accel'], 1] * n), dtype=np.float32)

    @property
    def observation_space(self):
        return _box(low=0, high=1, shape=(3 * self.initial_vehicles.num_vehicles,),
                    dtype=np.float32)

    def _check_host_actions(self, a):
        # TraCIVehicle.apply_lane_change, vehicle/traci.py:973-975
        d = a.reshape(-1, 2 * sum(t == EN.CTRL_RL for t in (
            self.spec.types[i].controller for i in self.spec.type_of)))[:, 1::2]
        if not np.isin(d, (-1, 0, 1)).all():
            raise ValueError('Direction values for lane changes may only be: {-1, 0, 1}.')


class LaneChangeAccelPOEnv(LaneChangeAccelEnv):
    """flow/envs/ring/lane_change_accel.py:164-269. The observation is what the
    reference's get_state RETURNS (it returns inside its loop over RL vehicles): the
    per-lane [headways | tailways | leader speeds | follower speeds] block of the first
    RL vehicle, zeros for the others and one ego speed - 4*n_rl*lanes + 1 numbers, while
    `observation_space` declares 4*n_rl*lanes + n_rl exactly like the reference. Pass
    `lc_corrected=True` to the env for the full, normalised observation."""
    KIND = EN.ENV_LANE_CHANGE_ACCEL_PO

    def __init__(self, env_params, sim_params, network, simulator='b200', **kw):
        super().__init__(env_params, sim_params, network, simulator, **kw)
        self.num_lanes = int(network.net_params.additional_params['lanes'])
        self.visible = []

    @property
    def observation_space(self):
        n = self.initial_vehicles.num_rl_vehicles
        return _box(low=0, high=1, shape=(4 * n * self.num_lanes + n,), dtype=np.float32)


class MultiAgentWaveAttenuationPOEnv(Env):
    """flow/envs/multiagent/ring/wave_attenuation.py:143-312 +
    flow/envs/multiagent/base.py:31-225 (MultiEnv): every RL vehicle is an agent.
    `step` takes {rl_id: action} (or None) and returns dicts keyed by rl id:
    obs (3,), the SHARED reward, done (+ '__all__'), info. Batched envs get
    tensors [E, 3] / [E] per key.

    Deviation, documented: the reference's MultiEnv discards the collision flag
    (`crash = 0`, base.py:174) and lets SUMO's teleporting continue the run; the
    engine keeps its collision rule (env frozen, reward 0, done['__all__'])."""
    KIND = EN.ENV_MA_WAVE_ATTENUATION_PO
    ADDITIONAL = {'max_accel': 1, 'max_decel': 1, 'ring_length': [220, 270]}

    def __init__(self, env_params, sim_params, network, simulator='b200', **kw):
        super().__init__(env_params, sim_params, network, simulator, **kw)
        # action slot r belongs to the r-th RL vehicle in vehicle-index order
        self.rl_ids = [v for i, v in enumerate(self.spec.veh_ids) if self._is_rl[i]]

    @property
    def action_space(self):
        ap = self.env_params.additional_params
        return _box(low=-abs(ap['max_decel']), high=ap['max_accel'], shape=(1,), dtype=np.float32)

    @property
    def observation_space(self):
        return _box(low=-5, high=5, shape=(3,), dtype=np.float32)

    OBS_PER_AGENT = 3
    REWARD_WITHOUT_ACTIONS = 0     # compute_reward returns 0 (not a dict) when rl_actions is None

    def _split(self, obs):
        w = self.OBS_PER_AGENT
        if self.n_envs == 1:
            o = obs[0].double().cpu().numpy().reshape(len(self.rl_ids), w)
            return {v: o[r] for r, v in enumerate(self.rl_ids)}
        o = obs.view(self.n_envs, len(self.rl_ids), w)
        return {v: o[:, r] for r, v in enumerate(self.rl_ids)}

    def reset(self):
        obs = super().reset()
        return self._split(self.state)

    def step(self, rl_actions):
        a = None
        if rl_actions:   # MultiEnv._apply_rl_actions: `if rl_actions:` (empty dict = no command)
            missing = [v for v in self.rl_ids if v not in rl_actions]
            if missing:
                raise KeyError('no action for agent(s) %s' % missing)
            if self.n_envs == 1:
                a = np.array([float(np.asarray(rl_actions[v]).reshape(-1)[0]) for v in self.rl_ids])
            else:
                import torch
                a = torch.stack([torch.as_tensor(rl_actions[v]).reshape(self.n_envs)
                                 for v in self.rl_ids], dim=1)
        obs, rew, done, _ = super().step(a)
        o = self._split(self.state)
        if self.n_envs == 1:
            # wave attenuation: compute_reward returns 0 (not a dict) when rl_actions is None
            # (:223-224); accel: always the dict
            r = {v: rew for v in self.rl_ids} \
                if (rl_actions or self.REWARD_WITHOUT_ACTIONS is None) else 0
        else:
            r = {v: rew for v in self.rl_ids}
        d = {v: False if self.n_envs == 1 else done.new_zeros(done.shape) for v in self.rl_ids}
        d['__all__'] = done
        return o, r, d, {v: {} for v in self.rl_ids}


class MultiAgentAccelPOEnv(MultiAgentWaveAttenuationPOEnv):
    """flow/envs/multiagent/ring/accel.py:87-226 on a single-lane ring: every RL vehicle
    observes (x/L, v/vmax, (v_lead - v)/vmax, (x_lead - x - own length)/L, (v - v_follow)/vmax,
    headway(follower)/L) - the leader distance is NOT wrapped, as in the reference - and all
    agents share rewards.desired_velocity. Same collision deviation as the class above."""
    KIND = EN.ENV_MA_ACCEL_PO
    ADDITIONAL = {'max_accel': 1, 'max_decel': 1, 'target_velocity': 20}
    OBS_PER_AGENT = 6
    REWARD_WITHOUT_ACTIONS = None

    def __init__(self, env_params, sim_params, network, simulator='b200', **kw):
        super().__init__(env_params, sim_params, network, simulator, **kw)
        self.leader, self.follower = [], []

    @property
    def observation_space(self):
        return _box(low=-5, high=5, shape=(6,), dtype=np.float32)


class MultiAgentHighwayPOEnv(Env):
    """flow/envs/multiagent/highway.py:14-196 + MultiEnv (flow/envs/multiagent/base.py:116-225) on
    the single-lane, single-edge HighwayNetwork: every RL vehicle currently IN the network is an
    agent; agents arrive through RL inflows and leave at the end of the edge.

    One env (n_envs == 1, the gym seam): `step({veh_id: [accel]})` -> dicts keyed by vehicle id
    ("rl_0" for initial vehicles, "<inflow name>.<k>" for inflow vehicles, as SUMO names them):
    obs (5,), per-agent reward, done (+ '__all__'), info; an RL vehicle that left the network
    in this step reports done = True, reward 0 and observation None (base.py:214-217). Actions are
    not clipped (MultiEnv.clip_actions returns its input); an agent without an entry keeps the
    simulator's model. Batched envs: `step(tensor [E, max_agents])` (NaN = no command) ->
    (obs [E, A, 5], reward [E, A], done [E], {'agent_id': [E, A]}), row a = a-th RL vehicle
    counted from the downstream end.

    Deviations, documented: the collision rule stays on (see MultiAgentWaveAttenuationPOEnv);
    the reference's config (multiagent_highway.py) runs on HighwayRampsNetwork with three lanes
    and SUMO lane changing, which is not restated - this is the single-lane straight edge."""
    KIND = EN.ENV_MA_HIGHWAY_PO
    ADDITIONAL = {'max_accel': 1, 'max_decel': 1, 'target_velocity': 25}
    MAX_AGENTS = 16     # observation / action rows per env (additional_params['max_agents'])

    def __init__(self, env_params, sim_params, network, simulator='b200', **kw):
        super().__init__(env_params, sim_params, network, simulator, **kw)
        self._flow_names = [f['name'] for f in network.net_params.inflows.get()]
        self._present = []

    @property
    def action_space(self):
        ap = self.env_params.additional_params
        return _box(low=-abs(ap['max_decel']), high=ap['max_accel'], shape=(1,), dtype=np.float32)

    @property
    def observation_space(self):
        return _box(low=-float('inf'), high=float('inf'), shape=(5,), dtype=np.float32)

    def vehicle_name(self, vid):
        """engine vehicle id -> the id SUMO would give the vehicle"""
        vid = int(vid)
        if vid < (1 << 24):
            return self._initial_ids[vid]
        return '%s.%d' % (self._flow_names[(vid >> 24) - 1], vid & 0xFFFFFF)

    def _agent_rows(self):
        ids = self.engine.agent_id[self.view_env].cpu().numpy()
        return [(a, self.vehicle_name(v)) for a, v in enumerate(ids) if v >= 0]

    def get_rl_ids(self):
        return [name for _, name in self._agent_rows()]

    def _obs_dict(self, rows):
        o = self.engine.agent_obs[self.view_env].double().cpu().numpy()
        return {name: o[a].copy() for a, name in rows}

    def reset(self):
        super().reset()
        if self.n_envs > 1:
            return self.engine.agent_obs
        self._present = self._agent_rows()
        return self._obs_dict(self._present)

    def step(self, rl_actions):
        if self.n_envs > 1:
            _, _, done, _ = super().step(rl_actions)
            return self.engine.agent_obs, self.engine.agent_reward, done, \
                {'agent_id': self.engine.agent_id}
        a = None
        if rl_actions:   # `if rl_actions:` highway.py:88 - None and {} both mean "no command"
            a = np.full(self.engine.max_agents, np.nan)
            for row, name in self._present:
                if name in rl_actions:
                    a[row] = float(np.asarray(rl_actions[name], dtype=np.float64).reshape(-1)[0])
        before = [name for _, name in self._present]
        _, _, done, _ = super().step(a)
        self._present = rows = self._agent_rows()
        obs = self._obs_dict(rows)
        d = {name: False for _, name in rows}
        d['__all__'] = done
        rew = {}
        if rl_actions is not None:   # compute_reward: {} during warm-up (highway.py:141-142)
            r = self.engine.agent_reward[self.view_env].double().cpu().numpy()
            rew = {name: float(r[a]) for a, name in rows}
        now = set(obs)
        for name in before:          # get_arrived_rl_ids, base.py:214-217
            if name not in now:
                d[name], rew[name], obs[name] = True, 0, None
        return obs, rew, d, {name: {} for _, name in rows}


class MergePOEnv(Env):
    """flow/envs/merge.py:24-220 on MergeNetwork (one lane each way). At most `num_rl` RL vehicles
    are controlled at a time: action k drives the k-th vehicle of `rl_veh`; vehicles that leave
    free their row, RL vehicles that entered later wait for one in order of arrival. Per row the
    observation holds own speed, the speed difference and bumper-to-bumper distance to the leader
    and to the follower, scaled by the speed limit and the network length; the reward is
    rewards.desired_velocity minus a penalty for time headways under one second.

    Vehicle names are the ones SUMO gives: "<type>_<i>" for the initial vehicles,
    "<inflow name>.<k>" for inflow vehicles.

    Restated simulator behaviour, documented in DESIGN.md: both arms share one axis measured from
    the merge point; a ramp vehicle and a highway vehicle see each other from `merge_zone` before
    it on (default: the ramp edge + the 22.5 m junction); a scheduled vehicle enters when the gap
    to the last vehicle of its arm is the model's desired gap at the departure speed, flows are
    served earliest due first. The reference removes from `rl_veh` while iterating over it, which
    skips one element for a step when a controlled vehicle leaves; not reproduced."""
    KIND = EN.ENV_MERGE_PO
    ADDITIONAL = {'max_accel': 3, 'max_decel': 3, 'target_velocity': 25, 'num_rl': 5}

    def __init__(self, env_params, sim_params, network, simulator='b200', **kw):
        super().__init__(env_params, sim_params, network, simulator, **kw)
        self.num_rl = int(env_params.additional_params['num_rl'])
        self._flow_names = [f['name'] for f in network.net_params.inflows.get()]
        self.leader, self.follower = [], []

    @property
    def action_space(self):
        ap = self.env_params.additional_params
        return _box(low=-abs(ap['max_decel']), high=ap['max_accel'], shape=(self.num_rl,),
                    dtype=np.float32)

    @property
    def observation_space(self):
        return _box(low=0, high=1, shape=(5 * self.num_rl,), dtype=np.float32)

    def vehicle_name(self, vid):
        """engine vehicle id -> the id SUMO would give the vehicle"""
        vid = int(vid)
        if vid < (1 << 24):
            return self._initial_ids[vid]
        return '%s.%d' % (self._flow_names[(vid >> 24) - 1], vid & 0xFFFFFF)

    @property
    def rl_veh(self):
        """names of the controlled vehicles, row order (merge.py:60-63)"""
        ids = self.engine.agent_id[self.view_env].cpu().numpy()
        return [self.vehicle_name(v) for v in ids if v >= 0]

    def _draw_host_placement(self):
        """gen_even_start_pos over the five edges (placement.even_positions), identical in every
        env of the batch; initial vehicles take the positions in the order they were added."""
        from . import placement as PL
        eng, ic, net = self.engine, self.initial_config, self.net_params.additional_params
        pre, ramp, post = (float(net[k]) for k in ('pre_merge_length', 'merge_length',
                                                   'post_merge_length'))
        eng.reset_stage(eng.RESET_PLACE)
        real, internal = PL.merge_edges(pre, ramp, post)
        where, _ = PL.even_positions(real, internal, ic.x0, ic.min_gap, ic.bunching,
                                     len(self._initial_ids))
        pos = np.zeros((eng.n_veh, self.n_envs), eng.np_dtype)
        arm = np.zeros((eng.n_veh, self.n_envs), np.uint8)
        for i, (edge, p) in enumerate(where):
            arm[i], pos[i] = PL.merge_arm_position(edge, p, pre)
        self._placed = (pos, arm)


_ENV_BY_NAME = {c.__name__: c for c in (AccelEnv, WaveAttenuationEnv, WaveAttenuationPOEnv,
                                        LaneChangeAccelEnv, LaneChangeAccelPOEnv,
                                        MultiAgentWaveAttenuationPOEnv, MultiAgentAccelPOEnv,
                                        MultiAgentHighwayPOEnv, MergePOEnv)}


# =============================================================================
# registry + experiment
# =============================================================================
_registered = set()


def make_create_env(params, version=0, render=None, **engine_kw):
    """flow/utils/registry.py:13-134. `params` is the reference's flow_params
    dict (its own param / controller / env / network classes are recognised by
    name). Extra keywords (n_envs, dtype, device, env_id_base, seed) size the
    batch. Returns (create_env, env_name)."""
    env_cls, net_cls = params['env_name'], params['network']
    base = env_cls if isinstance(env_cls, str) else env_cls.__name__
    while '%s-v%d' % (base, version) in _registered:
        version += 1
    env_name = '%s-v%d' % (base, version)
    env_ours = _resolve(env_cls, _ENV_BY_NAME, 'environment')
    net_ours = _resolve(net_cls, _NET_BY_NAME, 'network')

    def create_env(*_):
        sim_params = deepcopy(params['sim'])
        vehicles = deepcopy(params['veh'])
        network = net_ours(name=params['exp_tag'], vehicles=vehicles, net_params=params['net'],
                           initial_config=params.get('initial') or InitialConfig(),
                           traffic_lights=params.get('tls'),
                           detector_params=params.get('detector_params'))
        sim_params.render = render or sim_params.render
        _registered.add(env_name)
        return env_ours(env_params=params['env'], sim_params=sim_params, network=network,
                        simulator='b200', **engine_kw)

    return create_env, env_name


class Experiment:
    """flow/core/experiment.py:17-200: run `num_runs` rollouts of one env and
    collect returns / mean speeds / outflows."""

    def __init__(self, flow_params, custom_callables=None, **engine_kw):
        self.custom_callables = custom_callables or {}
        create_env, _ = make_create_env(flow_params, **engine_kw)
        self.env = create_env()

    def run(self, num_runs, rl_actions=None, convert_to_csv=False):
        env = self.env
        num_steps = env.env_params.horizon
        if convert_to_csv and env.sim_params.emission_path is None:
            raise ValueError('The experiment was run with convert_to_csv set to True, but no '
                             'emission file will be generated.')
        rec = None
        if convert_to_csv:   # the reference converts SUMO's emission xml; here rows come from the engine
            from .emission import EmissionRecorder
            rec = EmissionRecorder(env)
        info = {'returns': [], 'velocities': [], 'outflows': []}
        info.update({k: [] for k in self.custom_callables})
        if rl_actions is None:
            def rl_actions(*_):
                return None
        t, steps = time.time(), 0
        for i in range(num_runs):
            ret, vel = 0, []
            custom = {k: [] for k in self.custom_callables}
            state = env.reset()
            if rec is not None and i == num_runs - 1:
                rec.rows, rec._waiting = [], {}
                rec.record()
            for _ in range(int(num_steps)):
                state, reward, done, _i = env.step(rl_actions(state))
                steps += 1
                if rec is not None and i == num_runs - 1:   # like SUMO's file: the last run
                    rec.record()
                vel.append(np.mean(env.k.vehicle.get_speed(env.k.vehicle.get_ids())))
                ret += reward
                for k, f in self.custom_callables.items():
                    custom[k].append(f(env))
                if np.all(np.asarray(done.cpu() if hasattr(done, 'cpu') else done)):
                    break
            info['returns'].append(ret)
            info['velocities'].append(np.mean(vel))
            info['outflows'].append(env.k.vehicle.get_outflow_rate(int(500)))
            for k in custom:
                info[k].append(np.mean(custom[k]))
            print('Round {0}, return: {1}'.format(i, ret))
        for k in info:
            print('Average, std {}: {}, {}'.format(k, np.mean(info[k]), np.std(info[k])))
        dt = time.time() - t
        print('Total time:', dt)
        print('steps/second:', steps / dt if dt > 0 else float('inf'))
        if rec is not None:
            import os
            rec.write_csv(os.path.join(env.sim_params.emission_path,
                                       '{0}-emission.csv'.format(env.network.name)))
        env.terminate()
        return info
