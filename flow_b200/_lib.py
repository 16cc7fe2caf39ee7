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
SOURCES = ['flowstep.cu', 'fs_fast_f32.cu', 'fs_fast_f64.cu', 'fs_ring_f32.cu', 'fs_ring_f64.cu',
           'fs_ml_f32.cu', 'fs_ml_f64.cu', 'fs_hw_f32.cu', 'fs_hw_f64.cu', 'fs_mg_f32.cu', 'fs_mg_f64.cu',
           'fs_f8_f32.cu', 'fs_f8_f64.cu']

FS_NPARAM = 8
FS_MAX_TYPES = 16


class FsVType(C.Structure):
    _fields_ = [('controller', C.c_int32), ('fail_safe', C.c_int32),
                ('lc_controller', C.c_int32), ('speed_mode', C.c_int32),
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
                ('max_agents', C.c_int32), ('reserved0', C.c_int32),
                ('merge_pre', C.c_double), ('merge_ramp', C.c_double),
                ('merge_post', C.c_double), ('merge_zone', C.c_double),
                ('fig8_radius', C.c_double), ('fig8_zone', C.c_double)]


FS_ABI_VERSION = 5
FS_MAX_INFLOWS = 4
FS_NOALLOC_SCRATCH_BYTES = 4 * 1024 * 1024


class FsInflow(C.Structure):
    _fields_ = [('vtype', C.c_int32), ('arm', C.c_int32),
                ('vehs_per_hour', C.c_double), ('depart_speed', C.c_double),
                ('begin', C.c_double), ('end', C.c_double), ('probability', C.c_double)]


class FsAgentViews(C.Structure):
    _fields_ = [('agent_obs', C.c_void_p), ('agent_reward', C.c_void_p),
                ('agent_id', C.c_void_p), ('veh_id', C.c_void_p),
                ('veh_type', C.c_void_p),
                ('max_agents', C.c_int32), ('obs_per_agent', C.c_int32)]


class FsHostViews(C.Structure):
    _fields_ = [('actions', C.c_void_p), ('obs', C.c_void_p), ('reward', C.c_void_p),
                ('done', C.c_void_p),
                ('k', C.c_int32), ('n_envs', C.c_int32), ('obs_dim', C.c_int32),
                ('act_dim', C.c_int32), ('dtype', C.c_int32), ('reserved', C.c_int32)]


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
    ('fs_reset_stage', C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
    ('fs_step', C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    ('fs_step_unfused', C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    ('fs_step_host', C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    ('fs_host_staging', C.c_int, [C.c_void_p, C.c_int, C.POINTER(FsHostViews)]),
    ('fs_step_host_noalloc', C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
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


def _deps(path, seen=None):
    """`path` plus every header it includes (transitively, quoted includes only)."""
    import re
    seen = set() if seen is None else seen
    if path in seen or not os.path.exists(path):
        return seen
    seen.add(path)
    for inc in re.findall(r'^\s*#include\s+"([^"]+)"', open(path).read(), re.M):
        _deps(os.path.normpath(os.path.join(os.path.dirname(path), inc)), seen)
    return seen


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


HASH_PATH = SO_PATH + '.srchash'


def _src_hash():
    """Content hash of every source / header / compiler flag the library is built from
    (mtimes do not survive the snapshot copy to the GPU box; contents do)."""
    import hashlib
    deps = set()
    for src in SOURCES:
        _deps(os.path.join(CSRC, src), deps)
    h = hashlib.sha256(' '.join(NVCC_FLAGS).encode())
    for d in sorted(deps):
        h.update(os.path.relpath(d, ROOT).encode())
        h.update(open(d, 'rb').read())
    return h.hexdigest()


def _stale():
    """True when libflowstep.so was not built from the sources that are in the tree now."""
    if not os.path.exists(SO_PATH) or not os.path.exists(HASH_PATH):
        return True
    return open(HASH_PATH).read().strip() != _src_hash()


def build(force=False, verbose=False):
    """Compile libflowstep.so for sm_100a with nvcc (cross-compiles without a GPU): one
    translation unit per kernel family and dtype, compiled in parallel; only units whose
    sources or headers changed are recompiled unless `force`. The library is written to a
    temporary name and renamed, and the whole build holds a file lock, so concurrent ranks
    never load a half-written file."""
    if not force and not _stale():
        return SO_PATH
    import fcntl
    nvcc = os.environ.get('NVCC', 'nvcc')
    bdir = os.path.join(_HERE, 'build')
    os.makedirs(bdir, exist_ok=True)
    lock = open(os.path.join(bdir, '.lock'), 'w')
    fcntl.flock(lock, fcntl.LOCK_EX)
    try:
        if not force and not _stale():      # another process built it while we waited
            return SO_PATH
        extra = ['-Xptxas', '-v'] if verbose else []
        procs, objs = [], []
        for src in SOURCES:
            path = os.path.join(CSRC, src)
            obj = os.path.join(bdir, src.replace('.cu', '.o'))
            objs.append(obj)
            if not force and not _newer(obj, _deps(path)):
                continue
            cmd = [nvcc] + NVCC_FLAGS + extra + ['-c', '-o', obj, path]
            procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE,
                                                stderr=subprocess.PIPE, text=True)))
        failed = None
        for cmd, p in procs:
            out, err = p.communicate()
            if p.returncode != 0 and failed is None:
                failed = 'nvcc failed:\n%s\n%s' % (' '.join(cmd), err[-4000:])
            if verbose:
                print(err)
        if failed:
            raise RuntimeError(failed)
        tmp = SO_PATH + '.tmp.%d' % os.getpid()
        cmd = [nvcc, '-shared', '-gencode', 'arch=compute_100a,code=sm_100a', '-o', tmp] + objs
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError('link failed:\n%s\n%s' % (' '.join(cmd), res.stderr[-4000:]))
        os.replace(tmp, SO_PATH)
        with open(HASH_PATH, 'w') as f:
            f.write(_src_hash())
        return SO_PATH
    finally:
        fcntl.flock(lock, fcntl.LOCK_UN)
        lock.close()


_lib = None


def load():
    """Load the library; raises (never falls back) when it is unavailable."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH) or _stale():
        # edited sources must never run against an old binary; without nvcc (a box that only
        # received the built library) a missing library is fatal, a stale one is reported
        try:
            build()
        except Exception as exc:
            if not os.path.exists(SO_PATH):
                raise RuntimeError(
                    'libflowstep.so is not built (run python -c "import '
                    '__graft_entry__ as g; g.build()"): %s' % exc)
            import shutil
            if shutil.which(os.environ.get('NVCC', 'nvcc')):
                raise
            import warnings
            warnings.warn('libflowstep.so is older than its sources and nvcc is absent: %s' % exc)
    lib = C.CDLL(SO_PATH)
    for name, restype, argtypes in SYMBOLS:
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype = restype
        fn.argtypes = argtypes
    if lib.fs_abi_version() != FS_ABI_VERSION:
        raise RuntimeError('libflowstep ABI version mismatch')
    _lib = lib
    return lib
