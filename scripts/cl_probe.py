"""Where the closed-loop host step spends its time (c2, 4096 envs, fp32)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from flow_b200 import workloads as W
from flow_b200.vec_env import B200VecEnv

def mk():
    types, tov, kw = W.ring_c2(noise=0.2)
    env = B200VecEnv(types, tov, 4096, dtype='f32', seed=1, **kw)
    env.reset()
    return env

def t(fn, n=2000):
    for _ in range(200): fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e6

for zc in ('0', '1', '2', '3'):
    os.environ['FS_CL_MODE'] = zc
    env = mk(); e = env.engine
    st = e.host_staging(1)
    pa, po, pr, pd = st['ptr']
    a = np.random.uniform(-1, 1, 4096).astype(np.float32)
    lib, h = e.lib, e.h
    print('FS_CL_MODE=%s' % zc)
    print('  C fs_step_host(k=1), staging pointers : %.2f us' % t(lambda: lib.fs_step_host(h, pa, 1, po, pr, pd)))
    print('  python B200VecEnv.step_host(a)        : %.2f us' % t(lambda: env.step_host(a)))
    env.close()
env = mk(); e = env.engine
act = torch.rand(1, 4096, device='cuda') * 2 - 1
ap = act.data_ptr()
print('fs_step(k=1) device, no sync            : %.2f us' % t(lambda: e.step_ptr(ap, 1)))
print('fs_step(k=1) device + fs_sync           : %.2f us' % t(lambda: (e.step_ptr(ap, 1), e.sync())))
ev = torch.cuda.Event()
print('torch empty kernel + sync               : %.2f us' % t(lambda: (act.add_(0.0), torch.cuda.synchronize())))
