"""CPU: libflowstep.so builds/loads and exports every symbol include/flowstep.h
declares; argument validation that needs no GPU."""
import ctypes as C
import os
import re

import pytest

from flow_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, 'include', 'flowstep.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return set(re.findall(r'\b(fs_[a-z_]+)\s*\(', src))


def test_library_builds_and_exports_header_symbols():
    _lib.build()
    lib = _lib.load()
    declared = header_symbols()
    bound = {name for name, _, _ in _lib.SYMBOLS}
    assert declared == bound, declared ^ bound
    for name in declared:
        assert hasattr(lib, name)
    assert lib.fs_abi_version() == _lib.FS_ABI_VERSION == 5


def test_struct_layouts_match_header():
    # sizes computed from the C declarations (no padding surprises)
    assert C.sizeof(_lib.FsVType) == 4 * 4 + 8 * 8 + 8 * 8
    assert C.sizeof(_lib.FsConfig) % 8 == 0
    assert _lib.FsConfig.env_id_base.offset == 32
    assert _lib.FsConfig.sim_step.offset == 48


def test_create_rejects_bad_arguments_without_gpu():
    lib = _lib.load()
    h = C.c_void_p()
    c = _lib.FsConfig()
    assert lib.fs_create(C.byref(c), C.byref(h)) == -1       # abi_version 0
    assert b'abi_version' in lib.fs_last_error(None)
    c.abi_version = _lib.FS_ABI_VERSION
    c.n_envs, c.n_veh, c.n_lanes, c.sim_step = 4, 1000, 1, 0.1
    assert lib.fs_create(C.byref(c), C.byref(h)) == -1       # n_veh > 256
    c.n_veh = 22
    c.dtype = 7
    assert lib.fs_create(C.byref(c), C.byref(h)) == -1


def test_no_cpu_fallback():
    """Without a CUDA device the product path must fail loudly."""
    import torch
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    lib = _lib.load()
    h = C.c_void_p()
    c = _lib.FsConfig()
    c.abi_version, c.n_envs, c.n_veh, c.n_lanes, c.sim_step = _lib.FS_ABI_VERSION, 4, 22, 1, 0.1
    assert lib.fs_create(C.byref(c), C.byref(h)) == -2       # FS_ECUDA
    assert b'no CPU path' in lib.fs_last_error(None)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, 'flow_b200')
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                txt = open(os.path.join(d, f)).read()
                assert 'import oracle' not in txt and 'from oracle' not in txt, f
