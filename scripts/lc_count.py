"""How often lanes change in the c3 workload (bench actions): lane-change events per env-step."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from flow_b200 import workloads as W
from flow_b200.vec_env import B200VecEnv
types, tov, kw = W.ring_c3(noise=0.2)
E, K = 2048, 3000
env = B200VecEnv(types, tov, E, dtype='f32', seed=1234, **kw)
env.reset()
gen = torch.Generator(device='cuda'); gen.manual_seed(4321)
ev = 0
for s in range(0, K, 10):
    act = torch.rand(10, 6, E, device='cuda', generator=gen) * 2 - 1
    act[:, 0::2] = act[:, 0::2] * 0.6 - 0.4
    act[:, 1::2] = torch.round(act[:, 1::2] * 1.49)
    l0 = env.engine.lane.clone()
    env.rollout(act)
    ev += int((env.engine.lane != l0).sum())
print('c3: >= %d lane changes in %d env-steps (%.4f per env-step), crashed %d envs'
      % (ev, E * K, ev / (E * K), int(env.engine.crash.sum())))
