#!/usr/bin/env python
"""bench.py — env-steps/s and vehicle-steps/s of the batched ring step path.

  python bench.py --gpus N --steps K --warmup W [--workload c2|c5] [--fuse F]
  python bench.py --impl reference ...      # CPU arm: the oracle port on host cores

A "step" is one Env.step of EVERY environment of the batch (E per GPU).
Workload c2 (default, BASELINE.json configs[1]): WaveAttenuationPOEnv ring,
21 IDM + 1 RL vehicle, ring length ~ U{220..270}, 750 warm-up steps at reset,
random RL actions, E = 4096 envs per GPU. c5: same with E = 65536 per GPU.

Timed numbers
  value    vehicle-steps/s, whole job, inputs (actions) resident in HBM, the
           K steps issued as fused launches of F steps (per-step obs/reward/
           done of every env are written for every step);
  e2e      same metric through B200VecEnv.rollout_host: HOST pinned action
           buffers in, obs/reward/done of every step back on the host, the
           copies inside the timed region;
  single_step  K launches of one step each, issued by fs_step_unfused and replayed from a
           CUDA graph (the HBM-bound variant); single_step_per_call = the same from a Python
           loop of fs_step(k=1) calls (the closed-loop gym cadence).
All device timings use CUDA events on the stream the kernels run on; N > 1 is
one process per GPU with a barrier and the max over ranks.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

_emit = print
ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# algorithmic bytes: pos+vel read+write 16 B per
# vehicle-step, plus per env-step: ring length 4 + action 4 read, obs 12 +
# reward 4 + done 1 written = 25 B  ->  22*16 + 25 = 377 B per env-step.
BYTES_PER_ENV_STEP = 22 * 16 + 25
N_VEH = 22


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=3000)
    ap.add_argument('--warmup', type=int, default=300)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--workload', default='c2', choices=['c2', 'c5', 'ma', 'c3', 'c4', 'mah'])
    ap.add_argument('--fuse', type=int, default=0,
                    help='steps per fused launch; 0 = per workload: 100 (what BASELINE.json states for '
                         'c5), 500 for the 4096-env rings c2 / ma, whose 100-step launch is short '
                         'enough (53 us) for its ramp and drain to cost 10 %% (measured: 1.51e11 at '
                         '100, 1.61e11 at 250, 1.66e11 at 500, 1.69e11 at 1000 vehicle-steps/s)')
    ap.add_argument('--dtype', default='f32', choices=['f32', 'f64'])
    ap.add_argument('--noise', type=float, default=0.2,
                    help='IDM accel noise std (0.2 in singleagent_ring.py; 0 = parity setting)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--envs', type=int, default=0, help='development: override envs per GPU')
    ap.add_argument('--only-fused', action='store_true',
                    help='development: time only the fused device-resident pass')
    ap.add_argument('--cpu-envs-per-core', type=int, default=2)
    ap.add_argument('--scaling', default='weak', choices=['weak', 'strong'],
                    help="weak (default, the driver's contract): the workload's envs on EVERY GPU; "
                         'strong: the same total split over the GPUs (SURVEY 8d c5)')
    return ap.parse_args()


# ------------------------------------------------------------- CPU arm (oracle)
def _cpu_worker(args):
    """One env, `steps` Env.steps of workload c2 on the oracle port (scalar
    Python, like the reference's per-vehicle loop). Returns seconds."""
    seed, steps = args
    import numpy as np
    from oracle import sim as OS
    from oracle import controllers as OC
    rng = np.random.default_rng(seed)
    types = [OC.VType(OC.IDM, OC.idm_params(), sim_min_gap=0.0), OC.VType(OC.RL)]
    length = int(rng.integers(220, 271))
    env = OS.RingSim(types, [0] * 21 + [1], length, bunching=50.0,
                     env_kind=OS.ENV_WAVE_ATTENUATION_PO, warmup_steps=0, horizon=10 ** 9,
                     act_lo=-1, act_hi=1, obs_max_length=270.0)
    acts = rng.uniform(-1, 1, size=steps)
    t0 = time.perf_counter()
    for k in range(steps):
        _, _, done = env.step([acts[k]])
        if done:
            env.reset(length, run_warmup=False)
    return time.perf_counter() - t0


def cpu_baseline(steps_per_env, envs_per_core, cores=None):
    """Oracle port on all host cores, one env per worker process at a time
    (how RLlib runs the reference). Returns dict for the JSON line."""
    import multiprocessing as mp
    cores = cores or os.cpu_count() or 1
    jobs = [(1000 + i, steps_per_env) for i in range(cores * envs_per_core)]
    ctx = mp.get_context('spawn')
    with ctx.Pool(cores) as pool:
        secs = pool.map(_cpu_worker, jobs, chunksize=1)
    # process start-up / imports are excluded: every worker times only its
    # stepping loop; `cores` workers run concurrently, so the aggregate rate
    # is cores x the mean per-worker rate.
    rate = cores * statistics.mean(steps_per_env / t for t in secs)
    return dict(value=rate * N_VEH, unit='vehicle-steps/s', cores=cores, kind='port',
                env_steps_per_s=rate,
                sample='%d envs x %d steps of workload c2 on oracle/sim.py (scalar Python port '
                       'of the reference loop), %d worker processes' % (len(jobs), steps_per_env, cores))


_REF_ENV = {}


def _ref_init(seed_base):
    """Pool initializer of the reference arm: one persistent env per worker."""
    import numpy as np
    from oracle import sim as OS
    from oracle import controllers as OC
    rng = np.random.default_rng(seed_base + os.getpid())
    types = [OC.VType(OC.IDM, OC.idm_params(), sim_min_gap=0.0), OC.VType(OC.RL)]
    length = int(rng.integers(220, 271))
    _REF_ENV['env'] = OS.RingSim(types, [0] * 21 + [1], length, bunching=50.0,
                                 env_kind=OS.ENV_WAVE_ATTENUATION_PO, warmup_steps=0,
                                 horizon=10 ** 9, act_lo=-1, act_hi=1, obs_max_length=270.0)
    _REF_ENV['rng'], _REF_ENV['length'] = rng, length


def _ref_advance(n):
    """Advance this worker's env by n Env.steps; returns seconds spent stepping."""
    env, rng, length = _REF_ENV['env'], _REF_ENV['rng'], _REF_ENV['length']
    acts = rng.uniform(-1, 1, size=n)
    t0 = time.perf_counter()
    for k in range(n):
        _, _, done = env.step([acts[k]])
        if done:
            env.reset(length, run_warmup=False)
    return time.perf_counter() - t0


def run_reference(a):
    """CPU arm: the oracle port of the reference loop on ALL host cores, one env
    per worker process (how RLlib runs the reference). One "step" of this arm is
    a bounded sample: every core advances its env by `chunk` Env.steps, with
    `chunk` calibrated so that the whole --steps/--warmup run takes about a
    minute of wall clock whatever K and W are."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    import multiprocessing as mp
    cores = os.cpu_count() or 1
    ctx = mp.get_context('spawn')
    with ctx.Pool(cores, initializer=_ref_init, initargs=(977,)) as pool:
        cal = pool.map(_ref_advance, [300] * cores, chunksize=1)       # calibration, untimed
        per_core = 300.0 / statistics.mean(cal)
        chunk = int(max(1, min(3000, per_core * 60.0 / max(a.warmup + a.steps, 1))))
        for _ in range(a.warmup):
            pool.map(_ref_advance, [chunk] * cores, chunksize=1)
        t0 = time.perf_counter()
        for _ in range(a.steps):
            pool.map(_ref_advance, [chunk] * cores, chunksize=1)
        wall = time.perf_counter() - t0
    env_steps = float(a.steps) * chunk * cores
    val = env_steps * N_VEH / wall
    sample = ('%d worker processes (all host cores) x 1 env each, %d Env.steps per worker per '
              'bench step, %d bench steps, workload c2 on oracle/sim.py (scalar Python port of the '
              'reference per-vehicle loop; SUMO/TraCI unavailable on this box)'
              % (cores, chunk, a.steps))
    _emit(json.dumps({
        'impl': 'reference', 'metric': 'vehicle-steps/s', 'value': val, 'unit': 'vehicle-steps/s',
        'n_gpus': a.gpus, 'steps': a.steps, 'warmup': a.warmup, 'ms_per_step': 1e3 * wall / a.steps,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64',
        'data': 'synthetic', 'env_steps_per_s': env_steps / wall,
        'config': {'workload': 'c2: WaveAttenuationPOEnv ring 21 IDM + 1 RL, random RL actions; '
                               'CPU arm: one env per host core', 'env_steps_per_bench_step': chunk * cores},
        'cpu_baseline': dict(value=val, unit='vehicle-steps/s', cores=cores, kind='port',
                             sample=sample),
        'e2e': {'value': val, 'unit': 'vehicle-steps/s', 'h2d_bytes_per_step': 0,
                'd2h_bytes_per_step': 0},
        'gpu_launches': 0}))


# ------------------------------------------------------------------ B200 arm
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed regions."""
    Q = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,'
         'clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile('w+', suffix='.csv', delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(
                ['nvidia-smi', '-i', str(gpu_index), '--query-gpu=' + self.Q,
                 '--format=csv,noheader,nounits', '-lms', '100'],
                stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self):
        out = {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': [], 'samples': 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for line in self.f:
            c = [x.strip() for x in line.split(',')]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for n, v in zip(names, c[5:9]):
                if v.lower().startswith('active'):
                    reasons.add(n)
        self.f.close()
        os.unlink(self.f.name)
        if sm:
            # "under load" = upper half of the samples (idle gaps between phases excluded)
            hi = sorted(sm)[len(sm) // 2:]
            out.update(sm_mhz=statistics.median(hi), sm_max_mhz=max(mx), samples=len(sm))
        out['reasons'] = sorted(reasons)
        return out


def measured_peak():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        try:
            return float(json.load(open(p))['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
        except Exception:
            pass
    return 6650.0, 'fallback (B200_PROFILING.md)'


def run_b200(a):
    import torch
    import torch.distributed as dist
    from flow_b200.vec_env import B200VecEnv
    from flow_b200.workloads import WORKLOADS

    from flow_b200 import sharding as SH
    rank, world, local = SH.env_rank_world()
    if world != a.gpus and world > 1:
        raise SystemExit('--gpus %d but WORLD_SIZE=%d' % (a.gpus, world))
    torch.cuda.set_device(local)
    numa_cores = SH.bind_to_gpu_cpus(local) if world > 1 else 0
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))

    wl = WORKLOADS[a.workload]
    fuse = a.fuse or (500 if a.workload in ('c2', 'ma') else 100)
    E, K, W, F = (a.envs or wl['n_envs']), a.steps, a.warmup, max(1, min(fuse, a.steps))
    types, tov, kw = wl['builder'](noise=a.noise)
    kw['horizon'] = max(kw['horizon'], K)
    if a.scaling == 'strong':
        env_id_base, E = SH.shard_envs(E, rank, world)   # the workload's envs split over the GPUs
    else:
        env_id_base, E = SH.weak_shard(E, rank)          # weak scaling: E envs on every GPU
    env = B200VecEnv(types, tov, E, dtype=a.dtype, device=local, env_id_base=env_id_base,
                     seed=1234, **kw)
    eng = env.engine
    td = torch.float32 if a.dtype == 'f32' else torch.float64
    ts = 4 if a.dtype == 'f32' else 8
    gen = torch.Generator(device='cuda'); gen.manual_seed(4321 + rank)
    act = torch.rand(K, max(env.act_dim, 1), E, device='cuda', dtype=td, generator=gen) * 2 - 1
    if a.workload == 'c3':
        # rows [accel, direction] per RL vehicle: gentle accelerations, directions in {-1, 0, 1}
        # braking-biased like 'ma': an RL car with zero-mean accelerations rams its leader
        act[:, 0::2] = act[:, 0::2] * 0.6 - 0.4
        act[:, 1::2] = torch.round(act[:, 1::2] * 1.49)
    if a.workload == 'mah':
        # random accelerations make every agent ram its leader within a few hundred steps (the
        # observation does not even hold the agent's own gap): the agents get "no command" rows
        # (NaN), i.e. the simulator model, which still runs the ranking / observation / reward path
        act.fill_(float('nan'))
    if a.workload == 'ma':
        # agents that ignore their leaders in stop-and-go traffic ram them within ~100 steps
        # (zero or zero-mean accelerations: > 90 % of the rings); braking-biased random
        # actions U(-1, 0.2) keep the rings alive for the timed pass
        act = act * 0.6 - 0.4
    obs_seq = torch.empty(K, E, env.obs_dim, device='cuda', dtype=td)
    rew_seq = torch.empty(K, E, device='cuda', dtype=td)
    done_seq = torch.empty(K, E, device='cuda', dtype=torch.uint8)
    act_host = act.cpu().pin_memory()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # raw device addresses of step s inside the big buffers (no per-launch torch slicing)
    pa, po_, pr, pd = act.data_ptr(), obs_seq.data_ptr(), rew_seq.data_ptr(), done_seq.data_ptr()
    sa, so, sr, sd = E * env.act_dim * ts, E * env.obs_dim * ts, E * ts, E

    def fused_pass(n_steps):
        for s in range(0, n_steps, F):
            eng.step_ptr(pa + s * sa, min(F, n_steps - s), po_ + s * so, pr + s * sr, pd + s * sd)

    def single_pass(n_steps):
        # K single-step launches issued by libflowstep itself; the second call with the same
        # buffers (the timed one: warm-up covers the same K steps) replays them from a CUDA graph
        eng.step_unfused_ptr(pa, n_steps, po_, pr, pd)

    def single_call_pass(n_steps):
        for s in range(n_steps):
            eng.step_ptr(pa + s * sa, 1, po_ + s * so, pr + s * sr, pd + s * sd)

    host_out = {}

    def host_bufs(n_steps):
        """pinned result buffers, allocated once (outside every timed region)"""
        if not host_out:
            n = max(K, W)
            host_out.update(obs=torch.empty(n, E, env.obs_dim, dtype=td).pin_memory(),
                            rew=torch.empty(n, E, dtype=td).pin_memory(),
                            done=torch.empty(n, E, dtype=torch.uint8).pin_memory())
        return host_out['obs'], host_out['rew'], host_out['done']

    if not a.only_fused:
        host_bufs(K)

    def host_pass(n_steps):
        """e2e, open loop: pinned host actions in, every step's obs/reward/done back in
        pinned host memory; calls are pipelined (fs_step_host_async), one wait at the end."""
        o, r, d = host_bufs(n_steps)
        for s in range(0, n_steps, F):
            k = min(F, n_steps - s)
            env.rollout_host_async(act_host[s:s + k], o[s:s + k], r[s:s + k], d[s:s + k])
        env.wait()

    def host_sync_pass(n_steps):
        """e2e, closed-loop cadence: every call blocks until its results are on the host."""
        for s in range(0, n_steps, F):
            k = min(F, n_steps - s)
            env.rollout_host(act_host[s:s + k])

    def host_single_pass(n_steps):
        for s in range(n_steps):
            env.rollout_host(act_host[s:s + 1])

    def timed(fn, n_steps, warm=None):
        """reset -> W warm-up steps -> reset -> barrier -> time n_steps -> barrier.
        Returns (ms max over ranks, vehicle-steps summed over ranks, launches on this rank)."""
        env.reset()
        fn(warm or min(W, n_steps))
        env.reset()
        eng.stats(reset=True)
        barrier()
        l0 = eng.launch_count
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn(n_steps)
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        st = eng.stats()
        # rollout statistics summed, time max'ed over ranks (NCCL all-reduce of 7 doubles)
        st, ms = SH.reduce_stats(st, device='cuda', time_ms=ms)
        agg = SH.summarize(st)
        return dict(ms=ms, vehicle_steps=st['vehicle_steps'], env_steps=st['env_steps'],
                    crashes=st['crashes'], mean_reward=agg['mean_reward'],
                    mean_speed=agg['mean_speed'], launches=eng.launch_count - l0)

    sampler = ClockSampler(local) if rank == 0 else None
    fused = timed(fused_pass, K)
    if a.only_fused:
        single = single_call = host = host_sync = host1 = fused
    else:
        single = timed(single_pass, K, warm=K)
        single_call = timed(single_call_pass, K)
        # host passes warm up over the whole pinned buffers (first DMA into a pinned page is slow)
        host = timed(host_pass, K, warm=K)
        host_sync = timed(host_sync_pass, K, warm=K)
        host1 = timed(host_single_pass, min(K, 1000))
    clocks = sampler.stop() if sampler else None
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    def vps(r):
        return r['vehicle_steps'] / (r['ms'] * 1e-3)

    peak, peak_src = measured_peak()
    n_launch = fused['launches']
    launch_ms = fused['ms'] / max(n_launch, 1)
    # per-GPU algorithmic bytes of one fused launch / its duration
    # SURVEY 8d per env-step: state round trip 16 B x V + ring length 4 + actions + obs + reward 4 + done 1 (f32)
    nv = eng.n_veh            # vehicles (ring) or slots (highway) whose state makes the round trip
    # + lane / slot-type byte read and written (+ the 4-byte vehicle id on the multi-agent highway)
    extra = {'c3': 2, 'c4': 2, 'mah': 10}.get(a.workload, 0)
    bpe = (nv * 16 + 4 + 4 * env.act_dim + 4 * env.obs_dim + 4) * (ts / 4) + 1 + nv * extra
    alg_bytes = bpe * (fused['env_steps'] / world) / max(n_launch, 1)
    achieved = alg_bytes / (launch_ms * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, 'profiles', 'ncu_traffic.json')
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get('%s_%s_fuse%d' % (a.workload, a.dtype, F))
        except Exception:
            traffic = None
    h2d = F * E * env.act_dim * ts
    d2h = F * E * (env.obs_dim * ts + ts + 1)
    line = {
        'metric': 'vehicle-steps/s', 'value': vps(fused), 'unit': 'vehicle-steps/s',
        'n_gpus': world, 'steps': K, 'warmup': W, 'ms_per_step': fused['ms'] / K,
        'higher_is_better': True, 'scaling': a.scaling, 'vs_baseline': None, 'dtype': a.dtype,
        'data': 'synthetic',
        'env_steps_per_s': fused['env_steps'] / (fused['ms'] * 1e-3),
        'config': {'workload': '%s: %s' % (a.workload, wl['name']), 'envs_per_gpu': E,
                   'vehicles_per_env': nv, 'fuse': F, 'idm_noise': a.noise, 'parallelism': 'envs sharded, %d GPU(s), '
                   'no data-path collective' % world,
                   'l2': 'streamed actions/obs/reward/done %.0f MB per pass > 126 MB L2 is %s; '
                         'vehicle state is register-resident inside a fused launch'
                         % ((act.numel() + obs_seq.numel() + rew_seq.numel()) * ts / 1e6 + done_seq.numel() / 1e6,
                            str((act.numel() + obs_seq.numel() + rew_seq.numel()) * ts + done_seq.numel() > 126e6)),
                   'host_cores_bound_to_gpu_numa_node': numa_cores,
                   'counted': 'vehicle-steps actually advanced (engine counter); frozen (crashed) '
                              'envs are not counted'},
        'crashed_env_frac': fused['crashes'] / (E * world),
        'mean_reward': fused['mean_reward'], 'mean_speed': fused['mean_speed'],
        'gpu_launches': n_launch,
        'single_step': {'value': vps(single), 'unit': 'vehicle-steps/s',
                        'ms_per_step': single['ms'] / K, 'launches': single['launches'],
                        'api': 'fs_step_unfused: K one-step launches (state through HBM every '
                               'step) replayed from one CUDA graph'},
        'single_step_per_call': {'value': vps(single_call), 'unit': 'vehicle-steps/s',
                                 'ms_per_step': single_call['ms'] / K,
                                 'api': 'fs_step(k=1) called K times from Python (ctypes)'},
        'e2e': {'value': vps(host), 'unit': 'vehicle-steps/s', 'h2d_bytes_per_step': h2d / F,
                'd2h_bytes_per_step': d2h / F, 'api': 'B200VecEnv.rollout_host_async -> '
                'fs_step_host_async, %d steps per call, pinned host buffers, upload / kernel / '
                'download of consecutive calls overlapped, one wait at the end' % F,
                'ms_per_step': host['ms'] / K},
        'e2e_blocking': {'value': vps(host_sync), 'unit': 'vehicle-steps/s',
                         'ms_per_step': host_sync['ms'] / K,
                         'api': 'B200VecEnv.rollout_host -> fs_step_host, %d steps per call, every '
                                'call waits for its results (closed-loop cadence)' % F},
        'e2e_single_step': {'value': vps(host1), 'unit': 'vehicle-steps/s',
                            'ms_per_step': host1['ms'] / min(K, 1000),
                            'api': 'B200VecEnv.rollout_host, 1 step per call'},
        'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s',
                     'frac': achieved / peak, 'traffic': traffic, 'peak_source': peak_src,
                     'kernel': '%s<%s> (flow_b200/csrc/), %d fused steps per launch, threads/env = %s'
                               % ({'ma': 'ring1_fast_kernel', 'c3': 'ring_ml_kernel', 'c4': 'highway_kernel',
                                   'mah': 'highway_kernel'}.get(
                                   a.workload, 'ring1_hot_kernel'), a.dtype, F,
                                  os.environ.get('FS_RING_SHAPE', 'auto')),
                     'launch_ms': launch_ms,
                     'note': 'achieved = ALGORITHMIC bytes (SURVEY 8d: %d B per env-step = state round '
                             'trip + action/obs/reward/done, the single-step figure) x env-steps per '
                             'launch / launch duration. The fused launch keeps the state in registers, '
                             'so its real DRAM traffic (`traffic`, ncu) is far below that and frac may '
                             'exceed 1: it says how the fused path compares with the best a '
                             'one-step-per-launch HBM-bound kernel could do. The fused kernel itself is '
                             'bound by instruction issue (profiles/: issue-active 51 %% on c2, 74 %% on '
                             'c5; `issue` below).' % int(bpe)},
        # what the fused kernel must move at least: action in, obs/reward/done out per env-step,
        # state once per launch
        'roofline_fused_stream': (lambda b: {
            'bound': 'hbm', 'unit': 'GB/s', 'peak': peak, 'bytes_per_launch': b,
            'achieved': b / (launch_ms * 1e-3) / 1e9, 'frac': b / (launch_ms * 1e-3) / 1e9 / peak})(
                (fused['env_steps'] / world / max(n_launch, 1)) * (env.act_dim * ts + env.obs_dim * ts + ts + 1)
                + E * (2 * nv * 2 * ts + ts)),
        'issue': {'unit': 'warp-instructions per vehicle-step (static SASS count of the step loop, '
                          'scripts/sass_loops.py)', 'c5_shape1': 46.9, 'c2_shape8': 89.0,
                  'note': 'B200: 148 SMs x 4 schedulers x 1 warp-instruction/cycle x 1.965 GHz'},
        'roofline_single_step': {
            'bound': 'hbm', 'unit': 'GB/s', 'peak': peak,
            'achieved': bpe * E / (single['ms'] / K * 1e-3) / 1e9,
            'frac': bpe * E / (single['ms'] / K * 1e-3) / 1e9 / peak},
        'clocks': clocks,
    }
    if world == 1 and not a.no_cpu_baseline:
        line['cpu_baseline'] = cpu_baseline(10000, a.cpu_envs_per_core)
    _emit(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    a = parse()
    # stdout carries exactly ONE JSON line: library chatter on fd 1 (e.g. NCCL's version
    # banner) is sent to stderr, the line is written to the saved descriptor at the end
    sys.stdout.flush()
    real_out = os.dup(1)
    os.dup2(2, 1)
    global _emit

    def _emit(line):
        os.write(real_out, (line + '\n').encode())
    if a.impl == 'reference':
        run_reference(a)
    else:
        run_b200(a)


if __name__ == '__main__':
    main()
