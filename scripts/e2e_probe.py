"""Where does the host-buffer path spend its time? (development probe)"""
import time
import torch
from flow_b200.vec_env import B200VecEnv
from flow_b200 import workloads as W

E, F, N = 4096, 100, 20
types, tov, kw = W.ring_c2(noise=0.2)
kw['horizon'] = 100000


def run(tag, stream=None, per_call_wait=False, out_small=False):
    ctx = torch.cuda.stream(stream) if stream is not None else torch.cuda.stream(torch.cuda.current_stream())
    with ctx:
        env = B200VecEnv(types, tov, E, dtype='f32', seed=1, **kw)
        env.reset()
        act = (torch.rand(N * F, 1, E) * 2 - 1).pin_memory()
        n_out = F if out_small else N * F
        obs = torch.empty(n_out, E, 3).pin_memory()
        rew = torch.empty(n_out, E).pin_memory()
        done = torch.empty(n_out, E, dtype=torch.uint8).pin_memory()
        for rep in range(3):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for c in range(N):
                s = c * F
                o = 0 if out_small else s
                env.rollout_host_async(act[s:s + F], obs[o:o + F], rew[o:o + F], done[o:o + F])
                if per_call_wait:
                    env.wait()
            env.wait()
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        print('%-40s %.3f ms/call  %.2f GB/s' % (tag, 1e3 * dt / N, N * F * E * 21 / dt / 1e9))
        env.close()


def blocking():
    env = B200VecEnv(types, tov, E, dtype='f32', seed=1, **kw)
    env.reset()
    act = (torch.rand(N * F, 1, E) * 2 - 1).pin_memory()
    for rep in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for c in range(N):
            env.rollout_host(act[c * F:(c + 1) * F])
        dt = time.perf_counter() - t0
    print('%-40s %.3f ms/call  %.2f GB/s' % ('blocking fs_step_host', 1e3 * dt / N, N * F * E * 21 / dt / 1e9))


def copies():
    d = torch.empty(64 << 20, dtype=torch.uint8, device='cuda')
    h = torch.empty(64 << 20, dtype=torch.uint8).pin_memory()
    for name, fn in (('D2H', lambda: h.copy_(d, non_blocking=True)), ('H2D', lambda: d.copy_(h, non_blocking=True))):
        fn(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(10):
            fn()
        torch.cuda.synchronize()
        print('%s 64 MiB: %.1f GB/s' % (name, 10 * 64 * 1.048576e-3 / (time.perf_counter() - t0)))


copies()
blocking()
run('async, wait after every call', per_call_wait=True)
run('async, 20 calls in flight')
run('async, 20 in flight, small out buffer', out_small=True)
run('async, in flight, own torch stream', stream=torch.cuda.Stream())
