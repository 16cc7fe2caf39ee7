"""Python handle on one libflowstep engine (one per GPU).

Device buffers stay owned by the C library; they are exposed to Python as
torch tensors through DLPack capsules built here from the raw pointers
fs_buffers() returns (zero copy, never freed by torch).
"""
import ctypes as C

import numpy as np

from . import _lib

# ---- constants mirrored from include/flowstep.h ----------------------------
F32, F64 = 0, 1
TOPO_RING, TOPO_HIGHWAY, TOPO_MERGE, TOPO_FIGURE8 = 0, 1, 2, 3
ENV_ACCEL, ENV_WAVE_ATTENUATION, ENV_WAVE_ATTENUATION_PO, \
    ENV_LANE_CHANGE_ACCEL, ENV_LANE_CHANGE_ACCEL_PO, ENV_HIGHWAY, \
    ENV_MA_WAVE_ATTENUATION_PO, ENV_MA_ACCEL_PO, ENV_MA_HIGHWAY_PO, ENV_MERGE_PO = range(10)
CTRL_SIM, CTRL_IDM, CTRL_OVM, CTRL_LINEAR_OVM, CTRL_BCM, CTRL_CFM, CTRL_LAC, \
    CTRL_GIPPS, CTRL_FOLLOWER_STOPPER, CTRL_NONLOCAL_FOLLOWER_STOPPER, \
    CTRL_PI_SATURATION, CTRL_RL = range(12)
FAILSAFE = {None: 0, 'instantaneous': 1, 'safe_velocity': 2}
LC_SIM, LC_STATIC, LC_RL = 0, 1, 2
ENVF_EVALUATE, ENVF_SORT_VEHICLES, ENVF_BALLISTIC = 1, 2, 4


class FlowStepError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__('libflowstep error %d: %s' % (code, msg))
        self.code = code


# ---- DLPack capsule from a raw device pointer ---------------------------------
# The DLManagedTensor is built and freed inside libflowstep (fs_dlpack_wrap): no
# Python callback sits in a tensor's destructor, so tensors may outlive the
# interpreter's teardown order safely.
_DL_CODES = {'f32': (2, 32), 'f64': (2, 64), 'u8': (1, 8), 'i32': (0, 32)}
_PyCapsule_New = C.pythonapi.PyCapsule_New
_PyCapsule_New.restype = C.py_object
_PyCapsule_New.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p]


def dlpack_capsule(ptr, shape, dtype, device_id):
    """DLPack ('dltensor') capsule for device memory the engine owns."""
    code, bits = _DL_CODES[dtype]
    shp = (C.c_int64 * len(shape))(*shape)
    m = _lib.load().fs_dlpack_wrap(C.c_void_p(ptr), device_id, code, bits, len(shape), shp, None)
    if not m:
        raise ValueError('fs_dlpack_wrap rejected ptr=%r shape=%r' % (ptr, shape))
    return _PyCapsule_New(m, b'dltensor', None)


def _as_tensor(ptr, shape, dtype, device_id):
    import torch
    if not ptr:
        return None
    return torch.from_dlpack(dlpack_capsule(ptr, shape, dtype, device_id))


class VType(dict):
    """Keyword bag for one fs_vtype row (see include/flowstep.h)."""


def make_vtype(controller, p=(), fail_safe=None, noise=0.0, delay=0.0,
               max_accel=2.6, max_decel=4.5, length=5.0, sim_tau=1.0,
               sim_min_gap=2.5, sim_max_speed=30.0, lc_controller=LC_SIM, speed_mode=0):
    """speed_mode: SumoCarFollowingParams.speed_mode as an int (flow/core/params.py:12-18);
    0 = 'aggressive' (Flow's commanded speed is taken as is), the reference's default is 25."""
    t = _lib.FsVType()
    t.speed_mode = int(speed_mode)
    t.controller = controller
    t.fail_safe = FAILSAFE[fail_safe] if not isinstance(fail_safe, int) else fail_safe
    t.lc_controller = lc_controller
    for i, v in enumerate(p):
        t.p[i] = float(v)
    t.noise, t.delay = float(noise), float(delay)
    t.max_accel, t.max_decel, t.length = float(max_accel), abs(float(max_decel)), float(length)
    t.sim_tau, t.sim_min_gap, t.sim_max_speed = float(sim_tau), float(sim_min_gap), float(sim_max_speed)
    return t


class Engine:
    """One batched step engine on one GPU. Not thread-safe."""

    def __init__(self, types, type_of_vehicle, n_envs, dtype='f32', device=0,
                 use_torch_stream=True, capacity=None, **cfg):
        self.lib = _lib.load()
        c = _lib.FsConfig()
        c.abi_version = _lib.FS_ABI_VERSION
        c.device = device
        c.dtype = F64 if dtype in ('f64', F64, 'float64') else F32
        c.n_envs = n_envs
        if capacity is not None:   # highway: slots; the table lists the initial vehicles
            cfg.setdefault('n_initial', len(type_of_vehicle))
            type_of_vehicle = list(type_of_vehicle) + [0] * (capacity - len(type_of_vehicle))
        c.n_veh = len(type_of_vehicle)
        defaults = dict(topology=TOPO_RING, env_kind=ENV_ACCEL, n_lanes=1, env_id_base=0,
                        seed=0, sim_step=0.1, length_lo=230.0, length_hi=230.0,
                        junction_len=0.1, speed_limit=30.0, x0=0.0, bunching=0.0,
                        init_min_gap=0.0, lanes_distribution=1 << 30, warmup_steps=0,
                        horizon=3000, sims_per_step=1, clip_actions=0,
                        lane_change_duration=5, target_velocity=10.0, act_lo=-4.5,
                        act_hi=2.6, obs_max_length=0.0, inflow_vehs_per_hour=0.0,
                        inflow_depart_speed=0.0, inflow_type=0, collision_min_gap=0,
                        record_views=0, lc_corrected=0, n_initial=0, inflow_begin=1.0,
                        evaluate=0, sort_vehicles=0, use_ballistic=0, max_agents=0,
                        merge_pre=0.0, merge_ramp=0.0, merge_post=0.0, merge_zone=0.0,
                        fig8_radius=0.0, fig8_zone=0.0)
        # [(vtype, vehs_per_hour, depart_speed, begin[, end[, probability[, arm]]]), ...]
        inflows = cfg.pop('inflows', None)
        unknown = set(cfg) - set(defaults)
        if unknown:
            raise TypeError('unknown engine config keys: %s' % sorted(unknown))
        defaults.update(cfg)
        # EnvParams.evaluate / additional_params['sort_vehicles'] -> fs_config.env_flags
        defaults['env_flags'] = (ENVF_EVALUATE if defaults.pop('evaluate') else 0) | \
            (ENVF_SORT_VEHICLES if defaults.pop('sort_vehicles') else 0) | \
            (ENVF_BALLISTIC if defaults.pop('use_ballistic') else 0)
        for k, v in defaults.items():
            setattr(c, k, v)
        self.config = c
        self.dtype = 'f64' if c.dtype == F64 else 'f32'
        self.np_dtype = np.float64 if c.dtype == F64 else np.float32
        self.device = device
        self.n_envs, self.n_veh = n_envs, c.n_veh
        h = C.c_void_p()
        rc = self.lib.fs_create(C.byref(c), C.byref(h))
        if rc != 0:
            raise FlowStepError(rc, self.lib.fs_last_error(None).decode())
        self.h = h
        self._use_torch_stream = use_torch_stream
        self._stream = None
        self._bind_stream()
        arr = (_lib.FsVType * len(types))(*types)
        tov = np.ascontiguousarray(type_of_vehicle, dtype=np.uint8)
        self._check(self.lib.fs_set_types(self.h, arr, len(types), tov.ctypes.data))
        if inflows is not None:
            rows = (_lib.FsInflow * max(1, len(inflows)))()
            for r, row in zip(rows, inflows):
                vt, rate, speed, begin = row[:4]
                r.vtype, r.vehs_per_hour, r.depart_speed, r.begin = \
                    int(vt), float(rate), float(speed), float(begin)
                r.end = float(row[4]) if len(row) > 4 else 0.0
                r.probability = float(row[5]) if len(row) > 5 else 0.0
                r.arm = int(row[6]) if len(row) > 6 else 0
            self._check(self.lib.fs_set_inflows(self.h, rows, len(inflows)))
        v = _lib.FsViews()
        self._check(self.lib.fs_buffers(self.h, C.byref(v)))
        self.views = v
        self.obs_dim, self.act_dim = v.obs_dim, v.act_dim
        E, V, dt = n_envs, c.n_veh, self.dtype
        T = lambda p, shape, d=dt: _as_tensor(p, shape, d, device)  # noqa: E731

        def VE(p, d=dt):
            """logical [V, E] view whatever the engine's memory order is"""
            if not p:
                return None
            if v.env_stride == 1:
                return T(p, (V, E), d)
            return T(p, (E, V), d).t()
        self.env_major = v.env_stride != 1
        self.pos, self.vel = VE(v.pos), VE(v.vel)
        self.lane, self.active = VE(v.lane, 'u8'), VE(v.active, 'u8')
        self.headway, self.accel = VE(v.headway), VE(v.accel)
        self.leader, self.follower = VE(v.leader, 'i32'), VE(v.follower, 'i32')
        self.length = T(v.length, (E,))
        self.obs, self.reward = T(v.obs, (E, self.obs_dim)), T(v.reward, (E,))
        self.done, self.crash = T(v.done, (E,), 'u8'), T(v.crash, (E,), 'u8')
        self.time_counter = T(v.time_counter, (E,), 'i32')
        self.counters = T(v.counters, (3, E), 'i32')   # highway: in network, departed, arrived
        self.agent_obs = self.agent_reward = self.agent_id = self.veh_id = self.veh_type = None
        if c.env_kind in (ENV_MA_HIGHWAY_PO, ENV_MERGE_PO):
            av = _lib.FsAgentViews()
            self._check(self.lib.fs_agents(self.h, C.byref(av)))
            A = av.max_agents
            self.max_agents = A
            self.agent_obs = T(av.agent_obs, (E, A, av.obs_per_agent))
            self.agent_reward = T(av.agent_reward, (E, A)) if av.agent_reward else None
            self.agent_id = T(av.agent_id, (E, A), 'i32')
            self.veh_id = T(av.veh_id, (E, V), 'i32')       # per highway slot, head first
            self.veh_type = T(av.veh_type, (E, V), 'u8')
        self._fs_step = self.lib.fs_step

    def _bind_stream(self):
        """Run on torch's CURRENT stream of the engine's device: re-read on every call, the
        engine is re-bound (fs_set_stream waits for the old stream) when the caller moved to
        another one, e.g. inside `with torch.cuda.stream(s)`."""
        if not self._use_torch_stream:
            return
        import torch
        s = torch.cuda.current_stream(self.device).cuda_stream
        if s != self._stream:
            self._check(self.lib.fs_set_stream(self.h, C.c_void_p(s)))
            self._stream = s

    def _dev_ptr(self, t, name, numel=None, dtype=None):
        """data_ptr of a tensor that must live on this engine's GPU (None passes through)."""
        if t is None:
            return None
        if not t.is_cuda or t.device.index != self.device:
            raise ValueError('%s must be a tensor on cuda:%d, got %s' % (name, self.device, t.device))
        if not t.is_contiguous():
            raise ValueError('%s must be contiguous' % name)
        if dtype is not None and t.dtype != dtype:
            raise ValueError('%s must have dtype %s, got %s' % (name, dtype, t.dtype))
        if numel is not None and t.numel() != numel:
            raise ValueError('%s must have %d elements, got %d' % (name, numel, t.numel()))
        return t.data_ptr()

    def _check(self, rc):
        if rc != 0:
            raise FlowStepError(rc, self.lib.fs_last_error(self.h).decode())

    _VIEWS = ('pos', 'vel', 'lane', 'active', 'headway', 'accel', 'leader', 'follower', 'length',
              'obs', 'reward', 'done', 'crash', 'time_counter', 'counters', 'agent_obs',
              'agent_reward', 'agent_id', 'veh_id', 'veh_type')

    def close(self):
        """Free the engine. The tensors it handed out alias memory the library owns: the
        attributes are dropped here so that nothing of this object can reach freed memory;
        tensors a caller kept elsewhere (e.g. the return value of reset()) dangle after close."""
        if getattr(self, 'h', None):
            for name in self._VIEWS:
                if hasattr(self, name):
                    setattr(self, name, None)
            self._staging = None
            self.lib.fs_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- reset / step -------------------------------------------------------
    def reset(self, mask=None, lengths=None):
        self._bind_stream()
        m = None if mask is None else np.ascontiguousarray(mask, dtype=np.uint8)
        ln = None if lengths is None else np.ascontiguousarray(lengths, dtype=np.float64)
        self._check(self.lib.fs_reset(self.h, None if m is None else m.ctypes.data,
                                      None if ln is None else ln.ctypes.data))
        return self.obs

    RESET_PLACE, RESET_WARMUP = 1, 2

    def reset_stage(self, stages, mask=None, lengths=None):
        """fs_reset_stage: RESET_PLACE (ring length, even placement, counters) and / or
        RESET_WARMUP (the warm-up steps on the current state); between the two the caller may
        replace positions / lanes with set_state (host-drawn placements, flow_api)."""
        self._bind_stream()
        m = None if mask is None else np.ascontiguousarray(mask, dtype=np.uint8)
        ln = None if lengths is None else np.ascontiguousarray(lengths, dtype=np.float64)
        self._check(self.lib.fs_reset_stage(self.h, None if m is None else m.ctypes.data,
                                            None if ln is None else ln.ctypes.data, int(stages)))
        return self.obs

    def step(self, actions=None, k=1, obs_seq=None, reward_seq=None, done_seq=None):
        """k fused steps with DEVICE tensors (any DLPack producer).
        actions: [k, act_dim, E] (or [act_dim, E] / [E] when k == 1)."""
        import torch
        self._bind_stream()
        want = torch.float64 if self.dtype == 'f64' else torch.float32
        a_ptr = None
        if actions is not None:
            if not isinstance(actions, torch.Tensor):
                actions = torch.from_dlpack(actions)
            if not actions.is_cuda or actions.device.index != self.device:
                raise ValueError('actions must be on cuda:%d (host arrays: step_host), got %s'
                                 % (self.device, actions.device))
            if actions.dtype != want or not actions.is_contiguous():
                actions = actions.to(want).contiguous()     # on the current stream, like the launch
            if actions.numel() != k * self.act_dim * self.n_envs:
                raise ValueError('actions must have k*act_dim*n_envs = %d elements, got %d'
                                 % (k * self.act_dim * self.n_envs, actions.numel()))
            a_ptr = actions.data_ptr()
        E = self.n_envs
        po = self._dev_ptr(obs_seq, 'obs_seq', k * E * self.obs_dim, want)
        pr = self._dev_ptr(reward_seq, 'reward_seq', k * E, want)
        pd = self._dev_ptr(done_seq, 'done_seq', k * E, torch.uint8)
        self._check(self.lib.fs_step(self.h, a_ptr, k, po, pr, pd))
        return self.obs, self.reward, self.done

    def step_ptr(self, actions_ptr, k, obs_ptr=0, reward_ptr=0, done_ptr=0):
        """Same as step() with raw DEVICE addresses (ints; 0 = NULL): the
        low-overhead path for callers that slice big pre-allocated buffers."""
        rc = self._fs_step(self.h, actions_ptr or None, k, obs_ptr or None,
                           reward_ptr or None, done_ptr or None)
        if rc != 0:
            self._check(rc)

    def step_unfused_ptr(self, actions_ptr, k, obs_ptr=0, reward_ptr=0, done_ptr=0):
        """k steps as k single-step launches issued from C / replayed from a CUDA graph
        (fs_step_unfused): the variant whose state round-trips through HBM every step."""
        rc = self.lib.fs_step_unfused(self.h, actions_ptr or None, k, obs_ptr or None,
                                      reward_ptr or None, done_ptr or None)
        if rc != 0:
            self._check(rc)

    def step_host(self, actions, k=1, want_obs=True):
        """k fused steps with HOST numpy buffers; returns (obs, reward, done) arrays
        of every step ([k,E,obs_dim], [k,E], [k,E])."""
        E = self.n_envs
        a = None
        if actions is not None:
            a = np.ascontiguousarray(actions, dtype=self.np_dtype)
            if a.size != k * self.act_dim * E:
                raise ValueError('actions must have k*act_dim*n_envs elements')
        obs = np.empty((k, E, self.obs_dim), self.np_dtype) if want_obs else None
        rew = np.empty((k, E), self.np_dtype)
        done = np.empty((k, E), np.uint8)
        self._check(self.lib.fs_step_host(
            self.h, None if a is None else a.ctypes.data, k,
            None if obs is None else obs.ctypes.data, rew.ctypes.data, done.ctypes.data))
        return obs, rew, done

    def step_host_noalloc(self, actions, k=1, want_obs=True):
        """step_host() through fs_step_host_noalloc: same results, device staging in the
        library's static scratch array (no cudaMalloc in the call; see include/flowstep.h for
        the size limit and the sharing rule)."""
        E = self.n_envs
        a = None
        if actions is not None:
            a = np.ascontiguousarray(actions, dtype=self.np_dtype)
            if a.size != k * self.act_dim * E:
                raise ValueError('actions must have k*act_dim*n_envs elements')
        obs = np.empty((k, E, self.obs_dim), self.np_dtype) if want_obs else None
        rew = np.empty((k, E), self.np_dtype)
        done = np.empty((k, E), np.uint8)
        self._check(self.lib.fs_step_host_noalloc(
            self.h, None if a is None else a.ctypes.data, k,
            None if obs is None else obs.ctypes.data, rew.ctypes.data, done.ctypes.data))
        return obs, rew, done

    def host_staging(self, k=1):
        """The engine's pinned host block for closed-loop calls of k <= 8 steps
        (fs_host_staging): numpy views (actions [k, act_dim, E], obs [k, E, obs_dim],
        reward [k, E], done [k, E]) plus their raw addresses. Writing actions into the view
        and calling step_staged() costs no host-side copy; the outputs are overwritten by the
        next call."""
        hv = _lib.FsHostViews()
        self._check(self.lib.fs_host_staging(self.h, int(k), C.byref(hv)))
        E, ct = self.n_envs, (C.c_double if self.dtype == 'f64' else C.c_float)

        def arr(ptr, shape, ctype):
            n = int(np.prod(shape))
            if n == 0:
                return np.empty(shape, self.np_dtype)
            return np.ctypeslib.as_array((ctype * n).from_address(ptr)).reshape(shape)
        self._staging = dict(
            k=hv.k, ptr=(hv.actions, hv.obs, hv.reward, hv.done),
            actions=arr(hv.actions, (hv.k, self.act_dim, E), ct),
            obs=arr(hv.obs, (hv.k, E, self.obs_dim), ct),
            reward=arr(hv.reward, (hv.k, E), ct),
            done=arr(hv.done, (hv.k, E), C.c_uint8))
        return self._staging

    def step_staged(self, k=1, with_actions=True):
        """k <= 8 steps on the actions already written into host_staging()['actions'];
        blocks until obs / reward / done of those steps are in the staging views."""
        st = getattr(self, '_staging', None)
        if st is None or st['k'] < k:
            st = self.host_staging(max(k, 1))
        pa, po, pr, pd = st['ptr']
        rc = self.lib.fs_step_host(self.h, pa if (with_actions and self.act_dim) else None, k,
                                   po, pr, pd)
        if rc != 0:
            self._check(rc)
        return st

    def set_state(self, pos=None, vel=None, lane=None, lengths=None):
        f = lambda x, d: None if x is None else np.ascontiguousarray(x, dtype=d)  # noqa: E731
        pos, vel = f(pos, self.np_dtype), f(vel, self.np_dtype)
        lane, lengths = f(lane, np.uint8), f(lengths, np.float64)
        q = lambda x: None if x is None else x.ctypes.data  # noqa: E731
        self._check(self.lib.fs_set_state(self.h, q(pos), q(vel), q(lane), q(lengths)))

    def stats(self, reset=False):
        out = (C.c_double * 8)()
        self._check(self.lib.fs_stats(self.h, out, 8, int(reset)))
        keys = ['sum_reward', 'sum_speed', 'sum_speed_sq', 'vehicle_steps', 'env_steps', 'crashes']
        return dict(zip(keys, list(out)[:6]))

    def sync(self):
        self._check(self.lib.fs_sync(self.h))

    @property
    def launch_count(self):
        return int(self.lib.fs_launch_count(self.h))
