"""Build and load libflowstep.so (the C-ABI in include/flowstep.h) with ctypes.

The library is built IN-TREE (flow_b200/libflowstep.so) with nvcc for sm_100a
only. There is no fallback: if the library is missing or does not load, every
product entry point raises.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)
SO_PATH = os.path.join(_HERE, 'libflowstep.so')
CSRC = os.path.join(_HERE, 'csrc')
SOURCES = ['flowstep.cu', 'fs_fast_f32.cu', 'fs_fast_f64.cu']
HEADERS = ['fs_common.cuh', 'fs_ring.cuh', 'fs_ring_fast.cuh', 'fs_ring_hot.cuh', 'fs_reset.cuh', 'fs_launch.h',
           'fs_multilane.cuh', 'fs_highway.cuh', 'fs_sort.cuh']

FS_NPARAM = 8
FS_MAX_TYPES = 16


class FsVType(C.Structure):
    _fields_ = [('controller', C.c_int32), ('fail_safe', C.c_int32),
                ('lc_controller', C.c_int32), ('reserved', C.c_int32),
                ('p', C.c_double * FS_NPARAM),
                ('noise', C.c_double), ('delay', C.c_double),
                ('max_accel', C.c_double), ('max_decel', C.c_double),
                ('length', C.c_double),
                ('sim_tau', C.c_double), ('sim_min_gap', C.c_double),
                ('sim_max_speed', C.c_double)]


class FsConfig(C.Structure):
    _fields_ = [('abi_version', C.c_int32), ('device', C.c_int32),
                ('dtype', C.c_int32), ('topology', C.c_int32),
                ('env_kind', C.c_int32), ('n_envs', C.c_int32),
                ('n_veh', C.c_int32), ('n_lanes', C.c_int32),
                ('env_id_base', C.c_int64), ('seed', C.c_uint64),
                ('sim_step', C.c_double),
                ('length_lo', C.c_double), ('length_hi', C.c_double),
                ('junction_len', C.c_double), ('speed_limit', C.c_double),
                ('x0', C.c_double), ('bunching', C.c_double),
                ('init_min_gap', C.c_double),
                ('lanes_distribution', C.c_int32),
                ('warmup_steps', C.c_int32), ('horizon', C.c_int32),
                ('sims_per_step', C.c_int32), ('clip_actions', C.c_int32),
                ('lane_change_duration', C.c_int32),
                ('target_velocity', C.c_double),
                ('act_lo', C.c_double), ('act_hi', C.c_double),
                ('obs_max_length', C.c_double),
                ('inflow_vehs_per_hour', C.c_double),
                ('inflow_depart_speed', C.c_double),
                ('inflow_type', C.c_int32), ('collision_min_gap', C.c_int32),
                ('record_views', C.c_int32), ('lc_corrected', C.c_int32),
                ('n_initial', C.c_int32), ('env_flags', C.c_int32),
                ('inflow_begin', C.c_double),
                ('max_agents', C.c_int32), ('reserved0', C.c_int32)]


FS_ABI_VERSION = 2
FS_MAX_INFLOWS = 4


class FsInflow(C.Structure):
    _fields_ = [('vtype', C.c_int32), ('reserved', C.c_int32),
                ('vehs_per_hour', C.c_double), ('depart_speed', C.c_double),
                ('begin', C.c_double), ('end', C.c_double), ('probability', C.c_double)]


class FsAgentViews(C.Structure):
    _fields_ = [('agent_obs', C.c_void_p), ('agent_reward', C.c_void_p),
                ('agent_id', C.c_void_p), ('veh_id', C.c_void_p),
                ('veh_type', C.c_void_p),
                ('max_agents', C.c_int32), ('obs_per_agent', C.c_int32)]


class FsViews(C.Structure):
    _fields_ = [('pos', C.c_void_p), ('vel', C.c_void_p), ('lane', C.c_void_p),
                ('active', C.c_void_p), ('headway', C.c_void_p),
                ('leader', C.c_void_p), ('follower', C.c_void_p),
                ('accel', C.c_void_p), ('length', C.c_void_p),
                ('obs', C.c_void_p), ('reward', C.c_void_p),
                ('done', C.c_void_p), ('crash', C.c_void_p),
                ('time_counter', C.c_void_p), ('counters', C.c_void_p),
                ('n_envs', C.c_int32), ('n_veh', C.c_int32),
                ('n_lanes', C.c_int32), ('obs_dim', C.c_int32),
                ('act_dim', C.c_int32), ('dtype', C.c_int32),
                ('veh_stride', C.c_int32), ('env_stride', C.c_int32)]


# every symbol include/flowstep.h declares: (name, restype, argtypes)
SYMBOLS = [
    ('fs_abi_version', C.c_int, []),
    ('fs_create', C.c_int, [C.POINTER(FsConfig), C.POINTER(C.c_void_p)]),
    ('fs_destroy', None, [C.c_void_p]),
    ('fs_last_error', C.c_char_p, [C.c_void_p]),
    ('fs_set_types', C.c_int, [C.c_void_p, C.POINTER(FsVType), C.c_int, C.c_void_p]),
    ('fs_reset', C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    ('fs_step', C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    ('fs_step_unfused', C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    ('fs_step_host', C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    ('fs_step_host_async', C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    ('fs_buffers', C.c_int, [C.c_void_p, C.POINTER(FsViews)]),
    ('fs_dlpack_wrap', C.c_void_p, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    ('fs_set_state', C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    ('fs_stats', C.c_int, [C.c_void_p, C.POINTER(C.c_double), C.c_int, C.c_int]),
    ('fs_sync', C.c_int, [C.c_void_p]),
    ('fs_stream', C.c_void_p, [C.c_void_p]),
    ('fs_launch_count', C.c_int64, [C.c_void_p]),
    ('fs_set_stream', C.c_int, [C.c_void_p, C.c_void_p]),
    ('fs_set_inflows', C.c_int, [C.c_void_p, C.POINTER(FsInflow), C.c_int]),
    ('fs_agents', C.c_int, [C.c_void_p, C.POINTER(FsAgentViews)]),
]

NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3',
              '-ftz=true', '-prec-sqrt=false',  # fp32 mode only (tolerance 1e-4); fp64 is untouched
              '-std=c++17', '-Xcompiler', '-fPIC']


def _stale():
    if not os.path.exists(SO_PATH):
        return True
    t = os.path.getmtime(SO_PATH)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS]
    deps.append(os.path.join(ROOT, 'include', 'flowstep.h'))
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """Compile libflowstep.so for sm_100a with nvcc (cross-compiles without a GPU)."""
    if not force and not _stale():
        return SO_PATH
    nvcc = os.environ.get('NVCC', 'nvcc')
    bdir = os.path.join(_HERE, 'build')
    os.makedirs(bdir, exist_ok=True)
    extra = ['-Xptxas', '-v'] if verbose else []
    if os.environ.get('FS_HOT_EXPERIMENTS'):   # timing variants of the hot kernel (FS_HOT_VAR)
        extra.append('-DFS_HOT_EXPERIMENTS')
    procs = []
    for src in SOURCES:   # one nvcc per translation unit, in parallel
        obj = os.path.join(bdir, src.replace('.cu', '.o'))
        cmd = [nvcc] + NVCC_FLAGS + extra + ['-c', '-o', obj, os.path.join(CSRC, src)]
        procs.append((cmd, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE,
                                                  stderr=subprocess.PIPE, text=True)))
    objs = []
    for cmd, obj, p in procs:
        out, err = p.communicate()
        if p.returncode != 0:
            raise RuntimeError('nvcc failed:\n%s\n%s' % (' '.join(cmd), err[-4000:]))
        if verbose:
            print(err)
        objs.append(obj)
    cmd = [nvcc, '-shared', '-gencode', 'arch=compute_100a,code=sm_100a', '-o', SO_PATH] + objs
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError('link failed:\n%s\n%s' % (' '.join(cmd), res.stderr[-4000:]))
    return SO_PATH


_lib = None


def load():
    """Load the library; raises (never falls back) when it is unavailable."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        try:
            build()
        except Exception as exc:  # no nvcc on this machine
            raise RuntimeError(
                'libflowstep.so is not built (run python -c "import '
                '__graft_entry__ as g; g.build()"): %s' % exc)
    lib = C.CDLL(SO_PATH)
    for name, restype, argtypes in SYMBOLS:
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype = restype
        fn.argtypes = argtypes
    if lib.fs_abi_version() != FS_ABI_VERSION:
        raise RuntimeError('libflowstep ABI version mismatch')
    _lib = lib
    return lib
