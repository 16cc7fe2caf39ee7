"""GPU: fs_step_host_noalloc (include/flowstep.h; no counterpart in the reference)
against fs_step_host through the C-ABI: same bytes on two calls
in a row, oversize calls refused with a clear message, nothing of an earlier call leaks.
Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _pair(dtype, E, seed):
    from flow_b200 import workloads as W
    from flow_b200.vec_env import B200VecEnv
    types, tov, kw = W.ring_c2(noise=0.2)
    kw['warmup_steps'] = 25
    a = B200VecEnv(types, tov, E, dtype=dtype, seed=seed, **kw)
    b = B200VecEnv(types, tov, E, dtype=dtype, seed=seed, **kw)
    a.reset(); b.reset()
    return a, b


def _call(fn, eng, act, k):
    """one call of an fs_step_host-shaped entry point with caller-owned numpy buffers"""
    E, nd = eng.n_envs, eng.np_dtype
    a = np.ascontiguousarray(act, dtype=nd)
    obs = np.full((k, E, eng.obs_dim), 7.0, nd)
    rew = np.full((k, E), 7.0, nd)
    done = np.full((k, E), 7, np.uint8)
    rc = fn(eng.h, a.ctypes.data, k, obs.ctypes.data, rew.ctypes.data, done.ctypes.data)
    assert rc == 0, eng.lib.fs_last_error(eng.h)
    return obs, rew, done


@pytest.mark.parametrize('dtype', ['f32', 'f64'])
def test_noalloc_host_step_equals_fs_step_host_on_consecutive_calls(dtype):
    E = 300
    a, b = _pair(dtype, E, seed=11)
    lib = a.engine.lib
    assert lib.fs_step_host_noalloc(b.engine.h, None, 0, None, None, None) == 0   # set-up only
    rng = np.random.default_rng(4)
    # two calls in a row (the second smaller than the first: its results sit where the first
    # call's were), then a third that is larger again
    for k in (12, 5, 20):
        act = rng.uniform(-1, 1, size=(k, 1, E))
        ref = _call(lib.fs_step_host, a.engine, act, k)
        got = _call(lib.fs_step_host_noalloc, b.engine, act, k)
        for x, y in zip(ref, got):
            assert np.array_equal(x, y), k          # same kernels, same inputs: bit for bit
        assert not (got[2] == 7).any()
    import torch
    assert torch.equal(a.engine.pos, b.engine.pos) and torch.equal(a.engine.vel, b.engine.vel)
    assert torch.equal(a.engine.time_counter, b.engine.time_counter)


def test_noalloc_rejects_calls_beyond_the_static_scratch():
    from flow_b200 import _lib
    E = 4096
    a, b = _pair('f32', E, seed=3)
    lib, eng = b.engine.lib, b.engine
    per_step = E * (4 + 12 + 4 + 1)
    k_big = _lib.FS_NOALLOC_SCRATCH_BYTES // per_step + 2
    act = np.zeros((k_big, 1, E), np.float32)
    obs = np.empty((k_big, E, 3), np.float32); rew = np.empty((k_big, E), np.float32)
    done = np.empty((k_big, E), np.uint8)
    tc0 = eng.time_counter.clone()
    rc = lib.fs_step_host_noalloc(eng.h, act.ctypes.data, k_big, obs.ctypes.data, rew.ctypes.data,
                                  done.ctypes.data)
    assert rc == -1                                                    # FS_EINVAL
    msg = lib.fs_last_error(eng.h).decode()
    assert 'FS_NOALLOC_SCRATCH_BYTES' in msg and str(_lib.FS_NOALLOC_SCRATCH_BYTES) in msg
    import torch
    assert torch.equal(eng.time_counter, tc0)                          # nothing ran
    # the largest call that fits still works and equals fs_step_host
    k_ok = 40
    actk = np.random.default_rng(0).uniform(-1, 1, size=(k_ok, 1, E)).astype(np.float32)
    ref = _call(lib.fs_step_host, a.engine, actk, k_ok)
    got = _call(lib.fs_step_host_noalloc, eng, actk, k_ok)
    for x, y in zip(ref, got):
        assert np.array_equal(x, y)
