"""Drive the REAL reference controllers / fail-safes / rewards / observation
functions (imported from /root/reference via ref_shim) on a ring whose
simulator part (Euler move + leader search, i.e. what SUMO does) is the
oracle's restatement. Used by make_golden.py to produce tests/golden/*.npz and
by tests/test_oracle_vs_reference.py (skipped where the reference is absent).
"""
import numpy as np

from . import ref_shim

ref_shim.install()

from flow.core.params import SumoCarFollowingParams, EnvParams  # noqa: E402
from flow.core.kernel.vehicle.traci import TraCIVehicle  # noqa: E402
from flow.core.kernel.network.base import BaseKernelNetwork  # noqa: E402
from flow.core.params import InitialConfig  # noqa: E402
from flow.core import rewards as ref_rewards  # noqa: E402
from flow.envs.ring.wave_attenuation import WaveAttenuationEnv, \
    WaveAttenuationPOEnv  # noqa: E402
from flow.envs.ring.accel import AccelEnv  # noqa: E402
from flow.envs.multiagent.ring.accel import MultiAgentAccelPOEnv  # noqa: E402
from flow.envs.multiagent.ring.wave_attenuation import \
    MultiAgentWaveAttenuationPOEnv  # noqa: E402
import flow.controllers as ref_ctrl  # noqa: E402


class _Net:
    """Closed-form stand-in for TraCIKernelNetwork on a ring: supplies the
    tables the reference's own gen_even_start_pos reads."""

    def __init__(self, length, lanes, speed_limit, junction=0.1):
        self._length = float(length)
        self.lanes = lanes
        self.speed = speed_limit
        j = junction
        L = self._length
        self.edgestarts = [("bottom", 0), ("right", 0.25 * L + j),
                           ("top", 0.5 * L + 2 * j), ("left", 0.75 * L + 3 * j)]
        self.internal_edgestarts = [(":right_0", 0.25 * L), (":top_0", 0.5 * L + j),
                                    (":left_0", 0.75 * L + 2 * j),
                                    (":bottom_0", L + 3 * j)]
        self.total_edgestarts = sorted(self.edgestarts + self.internal_edgestarts,
                                       key=lambda t: t[1])
        self.total_edgestarts_dict = dict(self.total_edgestarts)
        self.circ = L + 4 * j

    def get_edge_list(self):
        return ["bottom", "right", "top", "left"]

    def num_lanes(self, edge):
        return self.lanes

    def edge_length(self, edge):
        return self._length / 4 if edge[0] != ':' else 0.1

    def non_internal_length(self):
        return self._length

    def length(self):
        return self.circ

    def max_speed(self):
        return self.speed

    def get_edge(self, x):  # flow/core/kernel/network/traci.py:267-271
        for (edge, start_pos) in reversed(self.total_edgestarts):
            if x >= start_pos:
                return edge, x - start_pos

    def get_x(self, edge, position):
        return self.total_edgestarts_dict[edge] + position

    # connection tables of the generated ring net (edge -> junction -> edge, lane kept),
    # what TraCIKernelNetwork.next_edge / prev_edge return (network/traci.py:347-359)
    _ORDER = ["bottom", ":right_0", "right", ":top_0", "top", ":left_0", "left", ":bottom_0"]

    def get_junction_list(self):
        return [":right_0", ":top_0", ":left_0", ":bottom_0"]

    def get_lane_list(self):
        return ['%s_%d' % (e, l) for e in self._ORDER for l in range(self.lanes)]

    def next_edge(self, edge, lane):
        return [(self._ORDER[(self._ORDER.index(edge) + 1) % 8], lane)]

    def prev_edge(self, edge, lane):
        return [(self._ORDER[(self._ORDER.index(edge) - 1) % 8], lane)]

    gen_even_start_pos = BaseKernelNetwork.gen_even_start_pos
    gen_random_start_pos = BaseKernelNetwork.gen_random_start_pos
    _get_start_pos_util = BaseKernelNetwork._get_start_pos_util


def ref_start_positions(length, num_vehicles, lanes=1, **init_kw):
    """Loop coordinates + lanes from the reference's own gen_even_start_pos (or
    gen_random_start_pos for spacing='random'); these read the host RNG streams
    (`random`, `np.random`) the caller has seeded."""
    net = _Net(length, lanes, 30)
    ic = InitialConfig(**init_kw)
    if ic.spacing == 'random':
        pos, lane = net.gen_random_start_pos(ic, num_vehicles)
    else:
        pos, lane = net.gen_even_start_pos(ic, num_vehicles)
    return [net.get_x(e, p) for e, p in pos], list(lane)


class _EdgeNet:
    """Stand-in for TraCIKernelNetwork on a network of the REFERENCE (its own specify_edges /
    specify_edge_starts / specify_internal_edge_starts tables, nominal edge lengths): what the
    reference's gen_even_start_pos reads."""

    def __init__(self, network):
        self.edgestarts = list(network.edge_starts)
        self.internal_edgestarts = list(network.internal_edge_starts)
        self.total_edgestarts = sorted(self.edgestarts + self.internal_edgestarts,
                                       key=lambda t: t[1])
        self.total_edgestarts_dict = dict(self.total_edgestarts)
        self._len = {e['id']: float(e['length']) for e in network.edges}
        lanes = {t['id']: int(t['numLanes']) for t in network.types}
        self._lanes = {e['id']: lanes[e['type']] for e in network.edges}

    def get_edge_list(self):
        return list(self._len)

    def num_lanes(self, edge):
        return self._lanes[edge]

    def edge_length(self, edge):
        return self._len[edge]

    def non_internal_length(self):
        return sum(self._len.values())

    get_edge = _Net.get_edge
    get_x = _Net.get_x
    gen_even_start_pos = BaseKernelNetwork.gen_even_start_pos
    _get_start_pos_util = BaseKernelNetwork._get_start_pos_util


def ref_merge_start_positions(pre, merge, post, num_vehicles, **init_kw):
    """[(edge, position)] from the reference's gen_even_start_pos on its own MergeNetwork."""
    from flow.networks.merge import MergeNetwork
    from flow.core.params import NetParams, VehicleParams
    net = MergeNetwork('merge', VehicleParams(), NetParams(additional_params={
        'merge_length': merge, 'pre_merge_length': pre, 'post_merge_length': post,
        'merge_lanes': 1, 'highway_lanes': 1, 'speed_limit': 30}))
    pos, lane = _EdgeNet(net).gen_even_start_pos(InitialConfig(**init_kw), num_vehicles)
    return [(e, float(p)) for e, p in pos], list(lane)


def ref_figure8_start_positions(radius, num_vehicles, **init_kw):
    """[(edge, position)] from the reference's gen_even_start_pos on its own FigureEightNetwork."""
    from flow.networks.figure_eight import FigureEightNetwork
    from flow.core.params import NetParams, VehicleParams
    net = FigureEightNetwork('figure8', VehicleParams(), NetParams(additional_params={
        'radius_ring': radius, 'lanes': 1, 'speed_limit': 30, 'resolution': 40}))
    en = _EdgeNet(net)
    pos, lane = en.gen_even_start_pos(InitialConfig(**init_kw), num_vehicles)
    return [(e, float(p)) for e, p in pos], [float(en.get_x(e, p)) for e, p in pos]


class _Veh:
    """The getter surface of TraCIVehicle the controllers/rewards call,
    string ids in, backed by arrays."""

    def __init__(self, h):
        self.h = h

    def _i(self, vid):
        return self.h.index[vid]

    def get_ids(self):
        return list(self.h.ids)

    def get_rl_ids(self):
        return [self.h.ids[i] for i in self.h.rl_idx]

    def get_human_ids(self):
        return [v for i, v in enumerate(self.h.ids) if i not in self.h.rl_idx]

    @property
    def num_vehicles(self):
        return len(self.h.ids)

    @property
    def num_rl_vehicles(self):
        return len(self.h.rl_idx)

    def get_speed(self, vid, error=-1001):
        if isinstance(vid, (list, np.ndarray)):
            return [self.get_speed(v, error) for v in vid]
        if vid is None or vid not in self.h.index:
            return error
        return self.h.vel[self._i(vid)]

    def get_headway(self, vid, error=-1001):
        if vid is None or vid not in self.h.index:
            return error
        return self.h.headway[self._i(vid)]

    def get_leader(self, vid, error=""):
        j = self.h.leader[self._i(vid)]
        return None if j < 0 else self.h.ids[j]

    def get_follower(self, vid, error=""):
        j = self.h.follower[self._i(vid)]
        return None if j < 0 else self.h.ids[j]

    def get_edge(self, vid, error=""):
        # restated continuous loop: one logical edge, never '' or ':...'
        return "ring"

    def get_position(self, vid, error=-1001):
        return self.h.pos[self._i(vid)]

    def get_x_by_id(self, vid):
        return self.h.pos[self._i(vid)]

    def get_length(self, vid, error=-1001):
        return 5.0

    def get_lane(self, vid, error=-1001):
        return self.h.lane[self._i(vid)]

    def set_observed(self, vid):
        pass


class _SlowDownSink:
    """kernel_api.vehicle stand-in: records what apply_acceleration commands."""

    def __init__(self):
        self.cmd = {}

    def slowDown(self, vid, speed, duration):
        self.cmd[vid] = speed


class RefRing:
    """Ring env whose controllers, fail-safes, apply_acceleration, rewards and
    observations are the reference's code; only the simulator move and the
    leader search (SUMO's job) are the oracle restatement (oracle/sim.py)."""

    def __init__(self, groups, length, env_kind='accel', dt=0.1, lanes=1,
                 speed_limit=30, init_kw=None, env_additional=None,
                 warmup_steps=0, horizon=3000, clip_actions=False, index_shortcut=False,
                 use_speed_mode=False):
        """groups: list of (name, controller_class_or_'rl', ctrl_kwargs,
        SumoCarFollowingParams kwargs, count)."""
        self.dt = self.sim_step = dt
        # bench.py's CPU arm: let the restated SUMO part use the single-lane leader = i+1
        # shortcut (same answer, oracle/sim.py RingSim.update) so that the time measured is the
        # reference's own Python, not the O(V^2) restatement of SUMO's C++ leader search
        self.index_shortcut = index_shortcut
        self.ids, self.ctrl, self.rl_idx = [], [], []
        self.sim_ctrl = {}
        self.speed_mode, self.cf_limits = [], []
        for name, cls, kw, cf_kw, n in groups:
            for q in range(n):
                vid = '%s_%d' % (name, q)
                cf = SumoCarFollowingParams(**cf_kw)
                # restated simulator default model: the reference's IDM
                # law fed with the type's SUMO params. It drives un-commanded RL vehicles and
                # is the "safe speed" of the restated speed mode (bit 0) of every vehicle.
                cp = cf.controller_params
                self.sim_ctrl[len(self.ids)] = ref_ctrl.IDMController(
                    vid, v0=min(float(cp['maxSpeed']), speed_limit),
                    T=float(cp['tau']), a=float(cp['accel']),
                    b=float(cp['decel']), delta=4, s0=float(cp['minGap']),
                    car_following_params=cf)
                # the reference's params object carries the mode (params.py:841, default 25);
                # the harness applies it only on request (goldens made before the restated
                # clamp existed stay what they were)
                self.speed_mode.append(int(cf.speed_mode) if use_speed_mode else 0)
                self.cf_limits.append((float(cp['accel']), float(cp['decel'])))
                if cls == 'rl':
                    self.rl_idx.append(len(self.ids))
                    self.ctrl.append(None)
                else:
                    self.ctrl.append(cls(vid, car_following_params=cf, **kw))
                self.ids.append(vid)
        self.index = {v: i for i, v in enumerate(self.ids)}
        self.lanes = lanes
        self.speed_limit = speed_limit
        self.init_kw = init_kw or {}
        self.env_kind = env_kind
        self.env_params = EnvParams(
            horizon=horizon, warmup_steps=warmup_steps,
            clip_actions=clip_actions, additional_params=env_additional or {})
        self.k = type('K', (), {})()
        self.k.vehicle = _Veh(self)
        self.sink = _SlowDownSink()
        self.reset(length)

    # what SUMO would do -------------------------------------------------
    def _update(self):
        from oracle.sim import RingSim
        # borrow the restated leader search
        shim = type('S', (), {})()
        shim.num_vehicles = len(self.ids)
        # golden generation: always the general per-lane search, never the shortcut
        shim.n_lanes = self.lanes if self.index_shortcut else 0
        shim.lane, shim.pos, shim.circ = self.lane, self.pos, self.net.circ
        from oracle.controllers import VType
        shim.types, shim.type_of = [VType()], [0] * len(self.ids)
        RingSim.update(shim)
        self.leader, self.follower, self.headway = \
            shim.leader, shim.follower, shim.headway

    def reset(self, length):
        self.net = _Net(length, self.lanes, self.speed_limit)
        self.k.network = self.net
        ic = InitialConfig(**self.init_kw)
        pos, lane = self.net.gen_even_start_pos(ic, len(self.ids))
        self.pos = [float(self.net.get_x(e, p)) for e, p in pos]
        self.lane = list(lane)
        self.vel = [0.0] * len(self.ids)
        self.accel = [0.0] * len(self.ids)
        self.time_counter = 0
        self._update()

    def step(self, rl_actions):
        n = len(self.ids)
        self.time_counter += 1
        acc = [None] * n
        for i in range(n):
            if self.ctrl[i] is not None:
                acc[i] = self.ctrl[i].get_action(self)
        if rl_actions is not None:
            space_lo = -abs(self.env_params.additional_params.get('max_decel', 1e9))
            space_hi = self.env_params.additional_params.get('max_accel', 1e9)
            clipped = np.clip(rl_actions, a_min=space_lo, a_max=space_hi)
            self._assign_rl(acc, clipped)
        else:
            for i in self.rl_idx:
                acc[i] = self.sim_ctrl[i].get_accel(self)
        self._before_move()
        self.accel_cmd = list(acc)
        acc = self._speed_mode(acc, rl_actions is not None)
        # reference apply_acceleration (vehicle/traci.py:952-963), unbound
        stub = type('V', (), {})()
        stub.sim_step = self.dt
        stub.get_ids = lambda: self.ids
        stub.get_speed = lambda vid: self.vel[self.index[vid]]
        stub.kernel_api = type('A', (), {})()
        stub.kernel_api.vehicle = self.sink
        self.sink.cmd = {}
        TraCIVehicle.apply_acceleration(stub, self.ids, acc)
        for i, vid in enumerate(self.ids):
            self.accel[i] = self.accel_cmd[i]      # what the controller commanded
            self.vel[i] = self.sink.cmd[vid]
            self.pos[i] = (self.pos[i] + self.vel[i] * self.dt) % self.net.circ
        self._update()
        crash = any(h < 0 for h in self.headway)
        obs = self.get_state()
        acts = rl_actions
        reward = self.compute_reward(acts, crash)
        return obs, reward, crash

    def _assign_rl(self, acc, clipped):
        for q, i in enumerate(self.rl_idx):
            acc[i] = float(clipped[q])

    def _speed_mode(self, acc, rl_commanded):
        """RESTATED SUMO (what the simulator makes of a commanded speed, see
        oracle/controllers.speed_mode_accel): bit 0 caps it by the simulator model's own
        acceleration - evaluated with the reference's IDMController on the SUMO params -
        bit 1 / bit 2 bound it by the type's accel / decel. Vehicles the simulator drives
        itself (un-commanded RL) are left alone."""
        out = list(acc)
        for i, m in enumerate(self.speed_mode):
            if not m or (i in self.rl_idx and not rl_commanded):
                continue
            a = out[i]
            if m & 1:
                a = min(a, self.sim_ctrl[i].get_accel(self))
            if m & 2:
                a = min(a, self.cf_limits[i][0])
            if m & 4:
                a = max(a, -self.cf_limits[i][1])
            out[i] = a
        return out

    def _before_move(self):      # Env.additional_command, base.py:365
        pass

    def get_state(self):
        if self.env_kind == 'ma_wave_po':
            # dict keyed by rl id -> rows in get_rl_ids() order (= action-slot order)
            d = MultiAgentWaveAttenuationPOEnv.get_state(self)
            return np.concatenate([d[v] for v in self.k.vehicle.get_rl_ids()])
        if self.env_kind == 'ma_accel_po':
            # the reference's get_state keeps its own `leader` / `follower` lists on the env:
            # hand it a view that shares the kernel but not this harness's arrays
            view = type('EnvView', (), {})()
            view.k, view.env_params = self.k, self.env_params
            d = MultiAgentAccelPOEnv.get_state(view)
            return np.concatenate([d[v] for v in self.k.vehicle.get_rl_ids()])
        if self.env_kind == 'wave_po':
            return WaveAttenuationPOEnv.get_state(self)
        if self.env_kind == 'wave':
            return WaveAttenuationEnv.get_state(self)
        self.sorted_ids = self.ids
        return AccelEnv.get_state(self)

    def compute_reward(self, rl_actions, fail):
        if self.env_kind == 'ma_wave_po':
            acts = None if rl_actions is None else \
                dict(zip(self.k.vehicle.get_rl_ids(), [float(a) for a in rl_actions]))
            r = MultiAgentWaveAttenuationPOEnv.compute_reward(self, acts, fail=fail)
            if isinstance(r, dict):   # shared reward: every agent gets the same number
                vals = list(r.values())
                assert all(x == vals[0] for x in vals)
                return vals[0]
            return r
        if self.env_kind == 'ma_accel_po':
            r = MultiAgentAccelPOEnv.compute_reward(self, rl_actions, fail=fail)
            vals = list(r.values())
            assert all(x == vals[0] for x in vals)
            return vals[0]
        if self.env_kind in ('wave', 'wave_po'):
            return WaveAttenuationEnv.compute_reward(self, rl_actions, fail=fail)
        return AccelEnv.compute_reward(self, rl_actions, fail=fail)


class _SortedMixin:
    """additional_params['sort_vehicles'] = True with the reference's own
    AccelEnv.sorted_ids / _get_abs_position / additional_command / _apply_rl_actions;
    `finish_reset` is AccelEnv.reset's tail (accel.py:178-181), to be called where the
    reference's reset would return (after the warm-up steps)."""
    sorted_ids = AccelEnv.sorted_ids
    _get_abs_position = AccelEnv._get_abs_position

    def _init_keys(self):
        self.absolute_position = dict()      # accel.py:77-78: made once, survive resets
        self.prev_pos = dict()

    def finish_reset(self):
        for vid in self.k.vehicle.get_ids():
            self.absolute_position[vid] = self.k.vehicle.get_x_by_id(vid)
            self.prev_pos[vid] = self.k.vehicle.get_x_by_id(vid)


class RefRingSorted(_SortedMixin, RefRing):
    def __init__(self, *a, **kw):
        self._init_keys()
        add = dict(kw.pop('env_additional', None) or {})
        add['sort_vehicles'] = True
        super().__init__(*a, env_additional=add, **kw)

    def _assign_rl(self, acc, clipped):
        got = {}
        self.k.vehicle.apply_acceleration = lambda ids, a: got.update(zip(ids, a))
        AccelEnv._apply_rl_actions(self, clipped)
        for vid, a in got.items():
            acc[self.index[vid]] = float(a)

    def _before_move(self):
        AccelEnv.additional_command(self)

    def get_state(self):
        return AccelEnv.get_state(self)


class RefRingLC(RefRing):
    """Multi-lane ring driven by the reference's LaneChangeAccelEnv code:
    get_state, compute_reward, _apply_rl_actions, TraCIVehicle.apply_lane_change
    and TraCIVehicle.get_last_lc (with its headway bug) run unmodified; only the
    execution of a requested change and the move are the oracle restatement."""

    def __init__(self, *a, lane_change_duration=5, **kw):
        from flow.envs.ring.lane_change_accel import LaneChangeAccelEnv
        from flow.envs.ring.lane_change_accel import LaneChangeAccelPOEnv
        self.LC = LaneChangeAccelPOEnv if kw.get('env_kind') == 'lc_po' else LaneChangeAccelEnv
        kw.setdefault('env_kind', 'lc')
        self.requests = {}
        super().__init__(*a, **kw)
        self.env_params.additional_params.setdefault('lane_change_duration',
                                                     lane_change_duration)
        self.env_params.additional_params.setdefault('sort_vehicles', False)
        veh = self.k.vehicle
        h = self

        class _TV:  # what TraCIVehicle methods touch
            sim_step = self.dt
            kernel_api = type('A', (), {})()
            master_kernel = type('M', (), {})()
            prev_last_lc = {}

            def get_ids(s):
                return h.ids

            def get_rl_ids(s):
                return veh.get_rl_ids()

            def get_speed(s, vid):
                return veh.get_speed(vid)

            def get_lane(s, vid):
                return veh.get_lane(vid)

            def get_edge(s, vid):
                return 'ring'

            def get_headway(s, vid, error=-1001):
                return veh.get_headway(vid, error)

            def get_length(s, vid, error=-1001):
                return 5.0

        class _TVPO(_TV):
            def get_edge(s, vid):
                return h.net.get_edge(h.pos[h.index[vid]])[0]

            def get_position(s, vid, error=-1001):
                return h.net.get_edge(h.pos[h.index[vid]])[1]

        for name in ('_multi_lane_headways', '_multi_lane_headways_util', '_next_edge_leaders',
                     '_prev_edge_followers', 'set_lane_headways', 'set_lane_tailways',
                     'set_lane_leaders', 'set_lane_followers', 'get_lane_headways',
                     'get_lane_tailways', 'get_lane_leaders', 'get_lane_followers'):
            setattr(_TVPO, name, getattr(TraCIVehicle, name))

        tv = _TVPO() if self.env_kind == 'lc_po' else _TV()
        tv.kernel_api.vehicle = type('KV', (), {})()
        tv.kernel_api.vehicle.slowDown = self.sink.slowDown
        tv.kernel_api.vehicle.changeLane = \
            lambda vid, lane, dur: self.requests.__setitem__(vid, lane)
        tv.master_kernel.network = self.net
        self.tv = tv
        veh.apply_acceleration = lambda ids, acc: TraCIVehicle.apply_acceleration(tv, ids, acc)
        veh.apply_lane_change = lambda ids, direction: TraCIVehicle.apply_lane_change(
            tv, ids, direction)
        veh.get_last_lc = lambda vid, error=-1001: TraCIVehicle.get_last_lc(tv, vid, error)
        if not isinstance(getattr(type(self), 'sorted_ids', None), property):
            self.sorted_ids = self.ids
        if self.env_kind == 'lc_po':
            for name in ('get_lane_headways', 'get_lane_tailways', 'get_lane_leaders',
                         'get_lane_followers'):
                setattr(veh, name, getattr(tv, name))
            self.num_lanes = self.lanes        # LaneChangeAccelPOEnv.__init__
            self.visible = []

    def _sync_tv(self):
        self.tv._TraCIVehicle__rl_ids = self.k.vehicle.get_rl_ids()
        self.tv._TraCIVehicle__vehicles = {
            vid: {'headway': self.headway[i], 'last_lc': -float('inf')}
            for i, vid in enumerate(self.ids)}

    def step(self, rl_actions):
        n = len(self.ids)
        self.time_counter += 1
        self._sync_tv()
        acc = [None] * n
        for i in range(n):
            if self.ctrl[i] is not None:
                acc[i] = self.ctrl[i].get_action(self)
        self.sink.cmd = {}
        self.requests = {}
        TraCIVehicle.apply_acceleration(self.tv, self.ids, acc)   # humans (base.py:335)
        if rl_actions is not None:
            lo = [-abs(self.env_params.additional_params['max_decel']), -1] * len(self.rl_idx)
            hi = [self.env_params.additional_params['max_accel'], 1] * len(self.rl_idx)
            clipped = np.clip(np.array(rl_actions, dtype=float), lo, hi)
            self.LC._apply_rl_actions(self, clipped)
        else:
            for i in self.rl_idx:
                a = self.sim_ctrl[i].get_accel(self)
                self.sink.cmd[self.ids[i]] = max(self.vel[i] + a * self.dt, 0)
                acc[i] = a
        # restated SUMO: execute requested changes (no overlap rule), then move
        new_lane = list(self.lane)
        for vid, target in self.requests.items():
            i = self.index[vid]
            ok = True
            for j in range(n):
                if j == i or self.lane[j] != target:
                    continue
                fwd = (self.pos[j] - self.pos[i]) % self.net.circ
                back = (self.pos[i] - self.pos[j]) % self.net.circ
                if fwd - 5.0 < 0 or back - 5.0 < 0:
                    ok = False
            if ok:
                new_lane[i] = target
        self.lane = new_lane
        for i, vid in enumerate(self.ids):
            self.vel[i] = self.sink.cmd[vid]
            self.pos[i] = (self.pos[i] + self.vel[i] * self.dt) % self.net.circ
        self._update()
        self._sync_tv()
        if self.env_kind == 'lc_po':
            self.tv._multi_lane_headways()        # k.update, vehicle/traci.py:255-256
        crash = any(h < 0 for h in self.headway)
        obs = self.LC.get_state(self)
        reward = self.LC.compute_reward(self, rl_actions, fail=crash)
        return obs, reward, crash


class RefRingLCSorted(_SortedMixin, RefRingLC):
    """LaneChangeAccelEnv with sort_vehicles: its get_state / _apply_rl_actions read
    sorted_ids, its additional_command (lane_change_accel.py:156-161) keeps no
    bookkeeping, so the keys are the positions AccelEnv.reset stored."""

    def __init__(self, *a, **kw):
        self._init_keys()
        add = dict(kw.pop('env_additional', None) or {})
        add['sort_vehicles'] = True
        super().__init__(*a, env_additional=add, **kw)


class RefHighwayMA:
    """MultiAgentHighwayPOEnv.get_state / compute_reward of the REFERENCE
    (flow/envs/multiagent/highway.py:97-172, with rewards.desired_velocity) evaluated on
    the state of the restated open highway (oracle/highway.py: SUMO's move, insertion and
    arrival are the restatement). Vehicle ids are str(engine id)."""

    def __init__(self, sim, target_velocity, evaluate=False):
        from flow.envs.multiagent.highway import MultiAgentHighwayPOEnv
        self._cls = MultiAgentHighwayPOEnv
        self.sim = sim
        self.env_params = EnvParams(evaluate=evaluate, additional_params={
            'max_accel': 1, 'max_decel': 1, 'target_velocity': target_velocity})
        h = self

        class V:
            def _s(self, vid):
                return h.sim.ids.index(int(vid))

            def get_ids(self):
                return [str(i) for i in h.sim.ids]

            def get_rl_ids(self):
                return [str(h.sim.ids[s]) for s in h.sim.rl_slots()]

            @property
            def num_vehicles(self):
                return len(h.sim.ids)

            def get_speed(self, vid, error=-1001):
                if isinstance(vid, (list, np.ndarray)):
                    return [self.get_speed(v, error) for v in vid]
                return h.sim.vel[self._s(vid)]

            def get_headway(self, vid, error=-1001):
                return h.sim.headway[self._s(vid)]

            def get_leader(self, vid, error=''):
                s = self._s(vid)
                return None if s == 0 else str(h.sim.ids[s - 1])   # vehicle/traci.py:234-238

            def get_follower(self, vid, error=''):
                s = self._s(vid)
                return None if s + 1 >= len(h.sim.ids) else str(h.sim.ids[s + 1])

        class Nw:
            def max_speed(self):
                return h.sim.speed_limit

            def length(self):
                return h.sim.length

        self.k = type('K', (), {})()
        self.k.vehicle, self.k.network = V(), Nw()

    def get_state(self):
        return {int(k): np.asarray(v, float) for k, v in self._cls.get_state(self).items()}

    def compute_reward(self, rl_actions, fail=False):
        r = self._cls.compute_reward(self, rl_actions, fail=fail)
        return {int(k): float(v) for k, v in r.items()}


class RefMerge:
    """MergePOEnv of the REFERENCE (flow/envs/merge.py:106-220: _apply_rl_actions,
    additional_command, get_state, compute_reward with rewards.desired_velocity) run on the state
    of the restated merge network (oracle/merge.py: SUMO's move, junction, insertion and arrival
    are the restatement). Vehicle ids are zero-padded decimal strings of the engine ids."""

    def __init__(self, sim, max_accel=3, max_decel=3, evaluate=False):
        import collections
        from flow.envs.merge import MergePOEnv
        self._cls = MergePOEnv
        self.sim = sim
        self.num_rl = sim.num_rl
        self.rl_queue, self.rl_veh = collections.deque(), []
        self.leader, self.follower = [], []
        self.env_params = EnvParams(evaluate=evaluate, additional_params={
            'max_accel': max_accel, 'max_decel': max_decel, 'target_velocity': sim.target_velocity,
            'num_rl': sim.num_rl})
        h = self
        name = lambda i: '%012d' % i          # noqa: E731

        class V:
            def _s(self, vid):
                try:
                    return h.sim.ids.index(int(vid))
                except ValueError:
                    return -1

            def get_ids(self):
                return [name(i) for i in h.sim.ids]

            def get_rl_ids(self):     # kept sorted by the reference (vehicle/traci.py:259,366)
                return sorted(name(h.sim.ids[s]) for s in range(len(h.sim.ids)) if h.sim.is_rl(s))

            @property
            def num_vehicles(self):
                return len(h.sim.ids)

            def get_speed(self, vid, error=-1001):
                if isinstance(vid, (list, np.ndarray)):
                    return [self.get_speed(v, error) for v in vid]
                s = self._s(vid)
                return error if s < 0 else h.sim.vel[s]

            def get_headway(self, vid, error=-1001):
                s = self._s(vid)
                return error if s < 0 else h.sim.headway[s]

            def get_leader(self, vid, error=''):
                s = self._s(vid)
                if s < 0:
                    return error
                return None if h.sim.leader[s] < 0 else name(h.sim.ids[h.sim.leader[s]])

            def get_follower(self, vid, error=''):
                s = self._s(vid)
                if s < 0:
                    return error
                return None if h.sim.follower[s] < 0 else name(h.sim.ids[h.sim.follower[s]])

            def get_x_by_id(self, vid):
                s = self._s(vid)
                return h.sim.g.x_abs(h.sim.arm[s], h.sim.pos[s])

            def get_length(self, vid, error=-1001):
                return h.sim.types[h.sim.type_of[self._s(vid)]].length

            def set_observed(self, vid):
                pass

            def apply_acceleration(self, vid, acc):
                h.commands[int(vid)] = float(acc)

        class Nw:
            def max_speed(self):
                return h.sim.speed_limit

            def length(self):
                return h.sim.g.net_length

        self.k = type('K', (), {})()
        self.k.vehicle, self.k.network = V(), Nw()
        self.commands = {}

    def step(self, rl_actions):
        """Env.step (flow/envs/base.py:294-412) around the restated simulator."""
        sim = self.sim
        for _ in range(sim.sims_per_step):
            self.commands = {}
            if rl_actions is not None:
                self._cls._apply_rl_actions(self, rl_actions)
            sim._sim_step(None, commanded_ids=self.commands,
                          additional_command=lambda: self._cls.additional_command(self))
            if sim.crash:
                break
        sim.done = sim.crash or sim.time_counter >= sim.sims_per_step * (sim.warmup_steps + sim.horizon)
        obs = np.asarray(self._cls.get_state(self), dtype=float)
        rew = float(self._cls.compute_reward(self, rl_actions, fail=sim.crash))
        return obs, rew, sim.done



class RefFigure8:
    """AccelEnv on FigureEightNetwork with the REFERENCE's code wherever the reference has code:
    the controllers' get_action (incl. the internal-edge guard, base_controller.py:96-99, fed with
    SUMO-style edge ids), AccelEnv._apply_rl_actions / get_state / compute_reward, get_x through
    TraCIKernelNetwork.get_x on the network's own edge-start tables. SUMO's part (move, loop and
    crossing leaders, collision, the model that drives un-commanded vehicles) is the restatement
    oracle/figure8.py, which this harness steps through its `commanded` hook."""

    def __init__(self, sim, groups, max_accel=3, max_decel=3, clip_actions=True, evaluate=False):
        """groups: [(name, controller class or 'rl', ctrl kwargs, SumoCarFollowingParams kwargs,
        count)] in the order of sim.type_of."""
        from flow.networks.figure_eight import FigureEightNetwork
        from flow.core.params import NetParams, VehicleParams
        from flow.core.kernel.network.traci import TraCIKernelNetwork
        self.sim = sim
        net = FigureEightNetwork('figure8', VehicleParams(), NetParams(additional_params={
            'radius_ring': sim.g.r, 'lanes': 1, 'speed_limit': sim.speed_limit, 'resolution': 40}))
        self.ids, self.ctrl, self.rl_idx = [], [], []
        for name, cls, kw, cf_kw, n in groups:
            for q in range(n):
                vid = '%s_%d' % (name, q)
                if cls == 'rl':
                    self.rl_idx.append(len(self.ids))
                    self.ctrl.append(None)
                else:
                    self.ctrl.append(cls(vid, car_following_params=SumoCarFollowingParams(**cf_kw), **kw))
                self.ids.append(vid)
        self.index = {v: i for i, v in enumerate(self.ids)}
        self.sim_step = sim.dt
        self.env_params = EnvParams(
            horizon=sim.horizon, clip_actions=clip_actions, evaluate=evaluate, additional_params={
                'max_accel': max_accel, 'max_decel': max_decel, 'target_velocity': sim.target_velocity,
                'sort_vehicles': False})
        h = self

        class Nw:
            edgestarts = list(net.edge_starts)
            internal_edgestarts = list(net.internal_edge_starts)
            total_edgestarts = sorted(edgestarts + internal_edgestarts, key=lambda t: t[1])
            total_edgestarts_dict = dict(total_edgestarts)
            internal_edgestarts_dict = dict(internal_edgestarts)
            get_x = TraCIKernelNetwork.get_x

            def max_speed(self):
                return h.sim.speed_limit

            def length(self):
                return h.sim.g.L

        class V(_Veh):
            def get_edge(self, vid, error=""):
                # SUMO's ids: a junction's internal edges are ':<node>_<k>'. The centre junction's
                # two passes are ':center_1' / ':center_0' (the keys the network lists); the four
                # plain nodes carry one internal edge each, ':<node>_0'
                name = h.sim.g.segment(h.sim.pos[self._i(vid)])[0]
                return name + '_0' if name in h.sim.g.start_only else name

            def get_position(self, vid, error=-1001):
                s = h.sim.pos[self._i(vid)]
                return s - h.sim.g.segment(s)[1]

            def get_x_by_id(self, vid):
                return h.k.network.get_x(self.get_edge(vid), self.get_position(vid))

            def get_length(self, vid, error=-1001):
                return h.sim.length_of(self._i(vid))

            def apply_acceleration(self, veh_ids, acc):
                for vid, a in zip(veh_ids, acc):
                    if a is not None:
                        h.commanded[self._i(vid)] = float(a)

        self.k = type('K', (), {})()
        self.k.vehicle, self.k.network = V(self), Nw()
        self.commanded = {}
        self._view()

    # arrays the _Veh getters read: what the controllers are meant to see
    def _view(self):
        from oracle.figure8 import _ControlView
        sim = self.sim
        self.vel, self.pos, self.lane = sim.vel, sim.pos, [0] * len(self.ids)
        self.leader, self.follower = sim.leader, sim.follower
        self.headway = _ControlView(sim).headway

    sorted_ids = property(lambda self: list(self.ids))

    def step(self, rl_actions):
        """Env.step (flow/envs/base.py:294-412) around the restated simulator."""
        sim = self.sim
        if rl_actions is not None and self.env_params.clip_actions:
            rl_actions = np.clip(rl_actions, -abs(self.env_params.additional_params['max_decel']),
                                 self.env_params.additional_params['max_accel'])
        for _ in range(sim.sims_per_step):
            self.commanded = {}
            self._view()
            for i, c in enumerate(self.ctrl):        # base.py:329-339
                if c is not None:
                    a = c.get_action(self)
                    if a is not None:
                        self.commanded[i] = float(a)
            if rl_actions is not None:
                AccelEnv._apply_rl_actions(self, rl_actions)
            sim._sim_step(None, commanded=self.commanded)
            if sim.crash:
                break
        sim.done = sim.crash or sim.time_counter >= sim.sims_per_step * (sim.warmup_steps + sim.horizon)
        self._view()
        obs = np.asarray(AccelEnv.get_state(self), dtype=float)
        rew = float(AccelEnv.compute_reward(self, rl_actions, fail=sim.crash))
        return obs, rew, sim.done
