"""N > 1 host logic on CPU: world_size-2 gloo processes shard the env range and
reduce their rollout statistics exactly as bench.py does over NCCL."""
import os
import socket
import subprocess
import sys

from flow_b200 import sharding as SH

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import json, os, sys
sys.path.insert(0, sys.argv[1])
import torch.distributed as dist
from flow_b200 import sharding as SH
rank, world, _ = SH.env_rank_world()
dist.init_process_group('gloo', init_method='tcp://127.0.0.1:%s' % os.environ['MASTER_PORT'],
                        rank=rank, world_size=world)
base, n = SH.shard_envs(1001, rank, world)
# per-rank "rollout statistics": every env contributes its global index so the
# reduced sums are checkable in closed form
ids = range(base, base + n)
st = dict(sum_reward=float(sum(ids)), sum_speed=2.0 * n * 22, sum_speed_sq=4.0 * n * 22,
          vehicle_steps=22.0 * n, env_steps=float(n), crashes=float(rank))
red, t = SH.reduce_stats(st, time_ms=10.0 + rank)
if rank == 0:
    print(json.dumps(dict(red=red, t=t, base=base, n=n)))
dist.destroy_process_group()
'''


def test_shard_envs_partition():
    for total, world in ((65536, 8), (1001, 2), (7, 4), (3, 8)):
        blocks = [SH.shard_envs(total, r, world) for r in range(world)]
        assert blocks[0][0] == 0 and sum(n for _, n in blocks) == total
        for (b0, n0), (b1, _) in zip(blocks, blocks[1:]):
            assert b1 == b0 + n0
        assert max(n for _, n in blocks) - min(n for _, n in blocks) <= 1
    assert SH.weak_shard(4096, 3) == (3 * 4096, 4096)


def test_reduce_is_identity_without_a_group():
    st = dict(zip(SH.STAT_KEYS, [1.0, 2.0, 3.0, 4.0, 5.0, 6.0]))
    red, t = SH.reduce_stats(st, time_ms=1.5)
    assert red == st and t == 1.5
    s = SH.summarize(dict(sum_reward=10.0, sum_speed=44.0, sum_speed_sq=88.0, vehicle_steps=22.0,
                          env_steps=1.0, crashes=0.0))
    assert s['mean_speed'] == 2.0 and s['speed_std'] == 0.0 and s['mean_reward'] == 10.0


def test_world_size_2_gloo_reduction():
    import json
    s = socket.socket(); s.bind(('127.0.0.1', 0)); port = s.getsockname()[1]; s.close()
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE='2', LOCAL_RANK=str(r),
                   MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, '-c', WORKER, ROOT], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=300) for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    got = json.loads(outs[0][0].strip().splitlines()[-1])
    assert (got['base'], got['n']) == (0, 501)
    red = got['red']
    assert red['sum_reward'] == sum(range(1001))          # every global env exactly once
    assert red['env_steps'] == 1001 and red['vehicle_steps'] == 22 * 1001
    assert red['crashes'] == 1.0 and got['t'] == 11.0     # SUM of stats, MAX of time
