"""flow_b200 — B200-native batched step engine behind Flow's Env.reset/step API.

The compute path is libflowstep.so (hand-written CUDA for sm_100a, C-ABI in
include/flowstep.h); this package is the Python host side: ctypes binding
(`_lib`), engine handle with DLPack views (`engine`), the batched env
(`vec_env`) and the mirror of Flow's params/controllers/envs/registry API.
"""
__version__ = '0.1.0'
