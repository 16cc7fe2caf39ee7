"""Multi-GPU plumbing: environments are independent, so the path shards with
NO data-path collective (in the reference every env is its own
SUMO process). One process per GPU owns a contiguous block of the global env
index range; `env_id_base` keys the engine's Philox streams by GLOBAL env id, so
ring lengths, noise and therefore results do not depend on the GPU count. The
only exchange is a reduction of the small rollout-statistics vector
(`Experiment.run`'s returns / velocities bookkeeping,
flow/core/experiment.py:123-176) over torch.distributed — NCCL on GPUs, gloo in
the CPU tests.
"""
import os

STAT_KEYS = ('sum_reward', 'sum_speed', 'sum_speed_sq', 'vehicle_steps', 'env_steps', 'crashes')


def env_rank_world():
    """(rank, world, local_rank) from the torchrun environment (1-process default)."""
    return (int(os.environ.get('RANK', '0')), int(os.environ.get('WORLD_SIZE', '1')),
            int(os.environ.get('LOCAL_RANK', '0')))


def bind_to_gpu_cpus(local_rank):
    """Pin this process to the host cores NVML reports as closest to its GPU, BEFORE any
    pinned staging buffer is allocated: first-touch then places the buffers on the GPU's own
    NUMA node and eight ranks do not push their 45 GB/s of results across the socket link.
    Returns the number of cores kept (0: NVML unavailable or nothing changed)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(int(local_rank))
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return 0


def shard_envs(total_envs, rank, world):
    """Contiguous block of the global env range owned by `rank`:
    returns (env_id_base, n_local). Blocks differ by at most one env."""
    if not 0 <= rank < world:
        raise ValueError('rank %d outside world of %d' % (rank, world))
    q, r = divmod(int(total_envs), world)
    n = q + (1 if rank < r else 0)
    base = rank * q + min(rank, r)
    return base, n


def weak_shard(envs_per_gpu, rank):
    """Weak scaling: every rank carries `envs_per_gpu` envs."""
    return rank * int(envs_per_gpu), int(envs_per_gpu)


def reduce_stats(stats, device='cpu', time_ms=None):
    """Sum the engine's rollout statistics over all ranks; `time_ms` (if given)
    is reduced with MAX — a multi-GPU step is as slow as its slowest rank.
    Returns (stats summed, time_ms max). Without an initialised process group
    this is the identity."""
    import torch
    import torch.distributed as dist
    v = torch.tensor([float(stats[k]) for k in STAT_KEYS], dtype=torch.float64, device=device)
    t = torch.tensor([0.0 if time_ms is None else float(time_ms)], dtype=torch.float64,
                     device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(v, op=dist.ReduceOp.SUM)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    out = dict(zip(STAT_KEYS, [float(x) for x in v.tolist()]))
    return out, (None if time_ms is None else float(t[0]))


def summarize(stats):
    """Aggregates the reference's Experiment prints (mean return per env-step,
    mean speed, speed std) from a reduced statistics dict."""
    vs, es = max(stats['vehicle_steps'], 1.0), max(stats['env_steps'], 1.0)
    mean_v = stats['sum_speed'] / vs
    var = max(stats['sum_speed_sq'] / vs - mean_v * mean_v, 0.0)
    return dict(mean_reward=stats['sum_reward'] / es, mean_speed=mean_v, speed_std=var ** 0.5,
                crash_rate=stats['crashes'] / es)
