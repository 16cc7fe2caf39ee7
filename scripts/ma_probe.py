import numpy as np, torch
from flow_b200.vec_env import B200VecEnv
from flow_b200 import workloads as W
for dtype in ('f32', 'f64'):
    for noise in (0.0, 0.2):
        for warm in (0, 750):
            types, tov, kw = W.ring_ma(noise=noise)
            kw['warmup_steps'] = warm
            env = B200VecEnv(types, tov, 256, dtype=dtype, seed=1, **kw)
            env.reset()
            e = env.engine
            c0 = float(e.crash.float().mean())
            td = torch.float32 if dtype == 'f32' else torch.float64
            act = torch.zeros(100, 2, 256, device='cuda', dtype=td)
            env.rollout(act)
            hmin = None
            print(dtype, 'noise', noise, 'warm', warm, 'crash after reset %.3f after 100 steps %.3f' % (c0, float(e.crash.float().mean())),
                  'tc', int(e.time_counter[0]), 'v mean %.3f' % float(e.vel.mean()), 'vmax %.2f' % float(e.vel.max()))
            env.close()
