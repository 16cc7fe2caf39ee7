#!/usr/bin/env python
"""bench.py — env-steps/s and vehicle-steps/s of the batched step path (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W [--workload c2|c5|c3|c4|ma|mah|mg|f8] [--dtype f32|f64]
  python bench.py --impl reference ...     # CPU arm: the reference's own code on the host cores

A "step" is one Env.step of EVERY environment of the batch (E per GPU). Workload c2 (default,
BASELINE.json configs[1]): WaveAttenuationPOEnv ring, 21 IDM + 1 RL vehicle, ring length
~ U{220..270}, 750 warm-up steps at reset, random RL actions, E = 4096 envs per GPU.

How the K steps are timed. One K-step pass of c2 lasts K x 0.4 us — far
below what a CUDA-event bracket resolves (the driver runs K = 20). The pass is therefore
repeated R times back to back, R taken from an untimed calibration so that the bracket holds
>= --target-ms (100 ms) of GPU work, the same R on every rank (all-reduce MAX). The R x K
Env.steps are issued exactly like a long rollout: fused launches of `fuse` steps (a constant of
the workload, NOT derived from K), the envs reset (ring length redrawn, 750 warm-up steps) every
`horizon` = 3000 steps as Experiment.run does between rollouts. The resets are enqueued on
the same stream BETWEEN the event pairs, so the GPU never idles and the timed region is the sum
of the per-episode brackets. ms_per_step = that sum / (R x K). Actions, observations, rewards and
done flags stream through device ring buffers larger than L2, every step's outputs are written.

Timed numbers
  value        vehicle-steps/s, whole job (all ranks), actions resident in HBM, fused launches;
  single_step  one launch per Env.step (state through memory every step), issued by
               fs_step_unfused and replayed from a CUDA graph; single_step_per_call = the same
               from a Python loop of fs_step(k=1);
  e2e          same metric through B200VecEnv with HOST pinned buffers, copies inside the timed
               region: `value` = open-loop pipelined rollout (fs_step_host_async), `blocking` =
               every `fuse`-step call waits for its results, `closed_loop` = ONE Env.step per
               call, results on the host before the next call (what Env.step(policy(obs)) sees);
  legs         (default run only) c5 = 65 536 envs per GPU fused + single-step, f64 = c2 in the
               reference's own arithmetic, strong = c5's 65 536 envs split over the GPUs (N > 1).
All device timings are CUDA events on the stream the kernels run on; N > 1 is one process per
GPU with a barrier on both sides and the max over ranks.
"""
import argparse
import json
import math
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

N_VEH = 22
L2_BYTES = 126e6
# FP32 issue roofline of the fused variant: 148 SMs x 128 FP32 lanes x 2 (FMA) x
# 1.965 GHz; FP64: 64 lanes per SM
SM_COUNT, SM_GHZ = 148, 1.965
PEAK_TFLOPS = {'f32': SM_COUNT * 128 * 2 * SM_GHZ / 1e3, 'f64': SM_COUNT * 64 * 2 * SM_GHZ / 1e3}
# counted flop per vehicle-step (IDM + move + headway + reward = 30; Philox +
# Box-Muller amortised over 4 normals = 25 more when the humans carry noise)
FLOP_IDM, FLOP_NOISE = 30, 25
# speed-mode bit 0 (the reference's default mode 25 has it): a second IDM evaluation (the restated
# simulator model) + min: |h| clamp 2 + dv 1 + s* 6 + (v/v0)^4 3 + (s*/h)^2 2 + a(1-.-.) 3 + min 1
FLOP_SAFE = 18
DEFAULT_FUSE = {'c2': 500, 'ma': 500}          # else 100 (what BASELINE.json states for c5)


def parse(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=3000)
    ap.add_argument('--warmup', type=int, default=300)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--workload', default='c2', choices=['c2', 'c5', 'ma', 'c3', 'c4', 'mah', 'mg', 'f8'])
    ap.add_argument('--fuse', type=int, default=0,
                    help='steps per fused launch; 0 = per workload: 100 (what BASELINE.json states '
                         'for c5), 500 for the 4096-env rings c2 / ma (their 100-step launch lasts '
                         '~50 us, 10 %% of it ramp and drain)')
    ap.add_argument('--dtype', default='f32', choices=['f32', 'f64'])
    ap.add_argument('--noise', type=float, default=0.2,
                    help='IDM accel noise std (0.2 in singleagent_ring.py; 0 = parity setting)')
    ap.add_argument('--speed-mode', type=int, default=25,
                    help="SumoCarFollowingParams.speed_mode of every type (25 = 'right_of_way', the "
                         "reference's default, which every shipped config leaves in place; 0 = "
                         "'aggressive': no restated SUMO clamp on commanded speeds)")
    ap.add_argument('--target-ms', type=float, default=100.0,
                    help='GPU work inside every timed bracket (the K-step pass is repeated)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-extra-legs', action='store_true',
                    help='skip the c5 / f64 / strong sub-legs of the default run')
    ap.add_argument('--envs', type=int, default=0, help='development: override envs per GPU')
    ap.add_argument('--only-fused', action='store_true',
                    help='development / profiling: time only the fused device-resident pass')
    ap.add_argument('--reps', type=int, default=0, help='development: fix R instead of calibrating')
    ap.add_argument('--cpu-seconds', type=float, default=12.0,
                    help='wall-clock budget of the cpu_baseline sample')
    ap.add_argument('--scaling', default='weak', choices=['weak', 'strong'],
                    help="weak (default, the driver's contract): the workload's envs on EVERY GPU; "
                         'strong: the same total split over the GPUs (SURVEY 8d c5)')
    return ap.parse_args(argv)


# ------------------------------------------------------------------ planning helpers (CPU-testable)
def plan_reps(k_steps, est_ms_per_step, target_ms, max_total=4_000_000):
    """R such that R x K steps hold >= target_ms of work at the calibrated per-step time."""
    if est_ms_per_step <= 0:
        return 1
    r = int(math.ceil(target_ms / (k_steps * est_ms_per_step)))
    return max(1, min(r, max(1, max_total // max(k_steps, 1))))


def round_reps(reps, k_steps, chunk):
    """Smallest R' >= reps with R' x K a multiple of `chunk` (no partial launch in the pass)."""
    m = chunk // math.gcd(chunk, max(k_steps, 1))
    return int(math.ceil(reps / m)) * m


def plan_ring(fuse, k_steps, act_bytes_per_step, out_bytes_per_step, cap_bytes=12e9):
    """Steps held by the device ring buffers: a multiple of `fuse`, at least one K-step pass,
    large enough that the streamed inputs alone exceed L2 (or, with no action input, the
    outputs do), capped by `cap_bytes` in total."""
    need = max(fuse, k_steps)
    if act_bytes_per_step > 0:
        need = max(need, int(math.ceil(1.07 * L2_BYTES / act_bytes_per_step)))
    need = max(need, int(math.ceil(2.0 * L2_BYTES / max(act_bytes_per_step + out_bytes_per_step, 1))))
    cap = int(cap_bytes // max(act_bytes_per_step + out_bytes_per_step, 1))
    need = min(need, max(cap, fuse))
    return int(math.ceil(need / fuse)) * fuse


def plan_episodes(total_steps, horizon, chunk):
    """[(episode, [(cursor-relative step, k), ...])]: `total_steps` cut into episodes of
    `horizon` steps (a reset precedes each), every episode into calls of <= chunk steps."""
    eps, done = [], 0
    while done < total_steps:
        n = min(horizon, total_steps - done)
        calls, s = [], 0
        while s < n:
            k = min(chunk, n - s)
            calls.append((s, k))
            s += k
        eps.append((n, calls))
        done += n
    return eps


# ------------------------------------------------------------- CPU arm
def _ref_root():
    """Where the reference's package can be imported from on this machine: the read-only
    checkout (build container) or the offline install baseline/_ref (travels to the GPU box)."""
    for p in (os.environ.get('FLOW_REFERENCE'), '/root/reference', os.path.join(ROOT, 'baseline', '_ref')):
        if p and os.path.isdir(os.path.join(p, 'flow', 'controllers')):
            return p
    return None


_CPU = {}


def _cpu_init(kind, seed_base, noise, speed_mode):
    """Pool initializer: one persistent c2 env per worker process.
    kind 'reference': the reference's OWN IDMController.get_action (noise + fail-safe),
    TraCIVehicle.apply_acceleration, WaveAttenuationPOEnv.get_state / compute_reward
    (flow/envs/base.py:294-412's per-vehicle loop) driven by tests/golden/ref_harness.RefRing;
    only SUMO's part (Euler move + leader search) is the oracle restatement.
    kind 'port': oracle/sim.py, the scalar Python restatement of the same loop."""
    import numpy as np
    rng = np.random.default_rng(seed_base + os.getpid())
    length = int(rng.integers(220, 271))
    if kind == 'reference':
        os.environ['FLOW_REFERENCE'] = _ref_root()
        from tests.golden.ref_harness import RefRing, ref_ctrl
        env = RefRing([('human', ref_ctrl.IDMController, {'noise': noise},
                        {'min_gap': 0, 'speed_mode': speed_mode}, 21),
                       ('rl', 'rl', {}, {'speed_mode': speed_mode}, 1)], length, 'wave_po',
                      init_kw=dict(bunching=50, min_gap=0),
                      env_additional={'max_accel': 1, 'max_decel': 1, 'ring_length': [220, 270]},
                      index_shortcut=True, use_speed_mode=True)

        def advance(acts):
            for a in acts:
                _, _, crash = env.step([a])
                if crash:
                    env.reset(length)
    else:
        from oracle import sim as OS
        from oracle import controllers as OC
        types = [OC.VType(OC.IDM, OC.idm_params(), noise=noise, sim_min_gap=0.0, speed_mode=speed_mode),
                 OC.VType(OC.RL, speed_mode=speed_mode)]
        env = OS.RingSim(types, [0] * 21 + [1], length, bunching=50.0,
                         env_kind=OS.ENV_WAVE_ATTENUATION_PO, warmup_steps=0, horizon=10 ** 9,
                         act_lo=-1, act_hi=1, obs_max_length=270.0,
                         noise_fn=(lambda step, veh: rng.normal()) if noise > 0 else None)

        def advance(acts):
            for a in acts:
                _, _, done = env.step([a])
                if done:
                    env.reset(length, run_warmup=False)
    _CPU['advance'], _CPU['rng'] = advance, rng


def _cpu_advance(n):
    """Advance this worker's env by n Env.steps; returns seconds spent stepping."""
    acts = _CPU['rng'].uniform(-1, 1, size=n)
    t0 = time.perf_counter()
    _CPU['advance'](acts)
    return time.perf_counter() - t0


def _cpu_kind():
    return 'reference' if _ref_root() else 'port'


class CpuArm:
    """All host cores, one env per worker process (how RLlib runs the reference)."""

    def __init__(self, kind, noise, speed_mode, cores=None):
        import multiprocessing as mp
        self.kind, self.cores = kind, cores or os.cpu_count() or 1
        self.pool = mp.get_context('spawn').Pool(self.cores, initializer=_cpu_init,
                                                 initargs=(kind, 977, noise, speed_mode))
        cal = self.pool.map(_cpu_advance, [60] * self.cores, chunksize=1)   # untimed: imports, first steps
        self.per_core = 60.0 / statistics.mean(cal)                          # Env.steps/s per worker

    def run(self, chunk, n_passes):
        """n_passes x (every worker advances chunk Env.steps); returns (env-steps, seconds)."""
        t0 = time.perf_counter()
        for _ in range(n_passes):
            self.pool.map(_cpu_advance, [chunk] * self.cores, chunksize=1)
        return float(n_passes) * chunk * self.cores, time.perf_counter() - t0

    def close(self):
        self.pool.close()
        self.pool.join()

    def describe(self, chunk, n_passes):
        what = {'reference': "the reference's own IDMController.get_action / apply_acceleration / "
                             'WaveAttenuationPOEnv.get_state / compute_reward (imported from %s) on the '
                             'numpy ring harness tests/golden/ref_harness.RefRing; SUMO part (move + '
                             'leader search) = oracle restatement; SUMO/TraCI unavailable on this box'
                             % os.path.relpath(_ref_root() or '?', ROOT),
                'port': 'oracle/sim.py (scalar Python restatement of the reference per-vehicle loop)'}[self.kind]
        return ('%d worker processes (all host cores) x 1 env each, %d Env.steps per worker per pass, '
                '%d passes, workload c2 on %s' % (self.cores, chunk, n_passes, what))


def cpu_baseline(seconds, noise, speed_mode):
    """Bounded sample of workload c2 on the host cores: the reference-code harness
    (kind 'reference') when the reference package is importable, and the oracle port beside it."""
    out = None
    for kind in ([_cpu_kind()] if _cpu_kind() == 'port' else ['reference', 'port']):
        arm = CpuArm(kind, noise, speed_mode)
        budget = seconds * (0.7 if kind == 'reference' else 0.3) if _cpu_kind() == 'reference' else seconds
        chunk = int(max(20, min(3000, arm.per_core * budget)))
        steps, wall = arm.run(chunk, 1)
        arm.close()
        d = dict(value=steps * N_VEH / wall, unit='vehicle-steps/s', cores=arm.cores, kind=kind,
                 env_steps_per_s=steps / wall, sample=arm.describe(chunk, 1))
        if out is None:
            out = d
        else:
            out['port'] = d
    return out


def workload_name(wl):
    from flow_b200.workloads import WORKLOADS
    return '%s: %s' % (wl, WORKLOADS[wl]['name'])


def run_reference(a):
    """CPU arm. One "step" of this arm is a bounded sample of the same workload: every host core
    advances its env by `chunk` Env.steps, `chunk` calibrated so that the whole --steps/--warmup
    run takes about a minute of wall clock whatever K and W are."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    kind = _cpu_kind()
    arm = CpuArm(kind, a.noise, a.speed_mode)
    chunk = int(max(1, min(3000, arm.per_core * 60.0 / max(a.warmup + a.steps, 1))))
    arm.run(chunk, a.warmup)
    env_steps, wall = arm.run(chunk, a.steps)
    arm.close()
    val = env_steps * N_VEH / wall
    sample = arm.describe(chunk, a.steps)
    _emit(json.dumps({
        'impl': 'reference', 'metric': 'vehicle-steps/s', 'value': val, 'unit': 'vehicle-steps/s',
        'n_gpus': a.gpus, 'steps': a.steps, 'warmup': a.warmup, 'ms_per_step': 1e3 * wall / a.steps,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64',
        'data': 'synthetic', 'env_steps_per_s': env_steps / wall,
        'config': {'workload': workload_name('c2'), 'vehicles_per_env': N_VEH, 'idm_noise': a.noise,
                   'speed_mode': a.speed_mode,
                   'envs_per_gpu': arm.cores, 'env_steps_per_bench_step': chunk * arm.cores,
                   'note': 'CPU arm: one env per host core; a bench step = every core advances its env '
                           'by %d Env.steps (bounded sample of the same workload)' % chunk},
        'cpu_baseline': dict(value=val, unit='vehicle-steps/s', cores=arm.cores, kind=kind,
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


def ncu_profile(workload, dtype):
    """profiles/ncu_traffic.json: per workload_dtype the DRAM bytes per env-step and the issue
    counters of the last ncu --set full capture of the fused kernel (scripts/make_profiles.py)."""
    tp = os.path.join(ROOT, 'profiles', 'ncu_traffic.json')
    try:
        return json.load(open(tp)).get('%s_%s' % (workload, dtype))
    except Exception:
        return None


KERNEL_OF = {'ma': 'ring1_fast_kernel', 'c3': 'ring_ml_kernel', 'c4': 'highway_kernel',
             'mah': 'highway_kernel', 'mg': 'merge_kernel', 'f8': 'fig8_kernel'}


class Leg:
    """One workload x dtype x sharding on this rank: engine, device ring buffers, timed passes."""

    def __init__(self, a, ctx, workload, dtype, scaling, k_steps):
        import torch
        from flow_b200.vec_env import B200VecEnv
        from flow_b200.workloads import WORKLOADS
        from flow_b200 import sharding as SH
        self.a, self.ctx, self.workload, self.dtype, self.scaling = a, ctx, workload, dtype, scaling
        self.torch, self.SH = torch, SH
        rank, world, local = ctx['rank'], ctx['world'], ctx['local']
        wl = WORKLOADS[workload]
        self.wl = wl
        self.K = k_steps
        self.F = a.fuse or DEFAULT_FUSE.get(workload, 100)
        E = (a.envs if (a.envs and workload == a.workload) else 0) or wl['n_envs']
        types, tov, kw = wl['builder'](noise=a.noise, speed_mode=a.speed_mode)
        self.H = int(kw['horizon'])
        if scaling == 'strong':
            env_id_base, E = SH.shard_envs(E, rank, world)   # the workload's envs split over the GPUs
        else:
            env_id_base, E = SH.weak_shard(E, rank)          # weak scaling: E envs on every GPU
        self.E = E
        self.env = B200VecEnv(types, tov, E, dtype=dtype, device=local, env_id_base=env_id_base,
                              seed=1234, **kw)
        env = self.env
        self.eng = env.engine
        self.td = torch.float32 if dtype == 'f32' else torch.float64
        self.ts = ts = 4 if dtype == 'f32' else 8
        self.A = env.act_dim
        self.act_b = E * self.A * ts                          # bytes per Env.step of the batch
        self.out_b = E * (env.obs_dim * ts + ts + 1)
        self.NB = NB = plan_ring(self.F, min(k_steps, self.H), self.act_b, self.out_b)
        gen = torch.Generator(device='cuda'); gen.manual_seed(4321 + rank)
        act = torch.rand(NB, max(self.A, 1), E, device='cuda', dtype=self.td, generator=gen) * 2 - 1
        if workload == 'c3':
            # rows [accel, direction] per RL vehicle: gentle accelerations, directions in {-1, 0, 1}
            # braking-biased like 'ma': an RL car with zero-mean accelerations rams its leader
            act[:, 0::2] = act[:, 0::2] * 0.6 - 0.4
            act[:, 1::2] = torch.round(act[:, 1::2] * 1.49)
        if workload == 'mah':
            # random accelerations make every agent ram its leader within a few hundred steps (the
            # observation does not even hold the agent's own gap): the agents get "no command" rows
            # (NaN), i.e. the simulator model, which still runs the ranking / observation / reward path
            act.fill_(float('nan'))
        if workload == 'ma':
            # agents that ignore their leaders in stop-and-go traffic ram them within ~100 steps;
            # braking-biased random actions U(-1, 0.2) keep the rings alive for the timed pass
            act = act * 0.6 - 0.4
        self.act = act
        self.obs_seq = torch.empty(NB, E, env.obs_dim, device='cuda', dtype=self.td)
        self.rew_seq = torch.empty(NB, E, device='cuda', dtype=self.td)
        self.done_seq = torch.empty(NB, E, device='cuda', dtype=torch.uint8)
        # raw device addresses of ring slot s (no per-launch torch slicing)
        self.p = (act.data_ptr(), self.obs_seq.data_ptr(), self.rew_seq.data_ptr(), self.done_seq.data_ptr())
        self.st = (self.act_b, E * env.obs_dim * ts, E * ts, E)
        self.cursor = 0
        self.host = None

    # ---- ring-buffer cursor: a call of k <= F steps starting at a multiple of F never wraps
    def _slot(self, k, align):
        c = self.cursor
        if c % align:
            c += align - c % align
        if c + k > self.NB:
            c = 0
        self.cursor = c + k
        return c

    def _ptrs(self, s):
        return tuple(p + s * st for p, st in zip(self.p, self.st))

    # ---- the ways of issuing k Env.steps
    def call_fused(self, k):
        pa, po, pr, pd = self._ptrs(self._slot(k, self.F))
        self.eng.step_ptr(pa if self.A else 0, k, po, pr, pd)

    def call_single_graph(self, k):
        # k single-step launches issued by libflowstep; a repeated call with the same buffers is
        # replayed from ONE CUDA graph, so this leg re-uses slot 0 (k * out_b bytes; the vehicle
        # state these launches stream is L2-resident at these batch sizes anyway: 0.7 MB c2, 11.5 MB c5)
        pa, po, pr, pd = self._ptrs(0)
        self.eng.step_unfused_ptr(pa if self.A else 0, k, po, pr, pd)

    def call_single_python(self, k):
        for _ in range(k):
            pa, po, pr, pd = self._ptrs(self._slot(1, 1))
            self.eng.step_ptr(pa if self.A else 0, 1, po, pr, pd)

    def _host_bufs(self):
        """pinned host ring of 4 x F steps, allocated once (outside every timed region)"""
        if self.host is None:
            t, n = self.torch, 4 * self.F
            self.host = dict(
                n=n, cur=0,
                act=self.act[:n].cpu().pin_memory(),
                obs=t.empty(n, self.E, self.env.obs_dim, dtype=self.td).pin_memory(),
                rew=t.empty(n, self.E, dtype=self.td).pin_memory(),
                done=t.empty(n, self.E, dtype=t.uint8).pin_memory())
        return self.host

    def _host_slot(self, k):
        h = self._host_bufs()
        c = h['cur']
        if c + k > h['n']:
            c = 0
        h['cur'] = c + k
        return h, c

    def call_host_async(self, k):
        h, c = self._host_slot(k)
        self.env.rollout_host_async(h['act'][c:c + k], h['obs'][c:c + k], h['rew'][c:c + k],
                                    h['done'][c:c + k])

    def call_host_blocking(self, k):
        h, c = self._host_slot(k)
        self.env.rollout_host(h['act'][c:c + k])

    def call_host_closed_loop(self, k):
        h, c = self._host_slot(k)
        if 'act_np' not in h:
            h['act_np'] = h['act'].numpy()
        a, step = h['act_np'], self.env.step_host
        for s in range(c, c + k):
            step(a[s] if self.A else None)

    # ---- timing
    def barrier(self):
        import torch.distributed as dist
        self.torch.cuda.synchronize()
        if self.ctx['world'] > 1:
            dist.barrier()
            self.torch.cuda.synchronize()

    def _run(self, call, total, chunk, drain=None, timed=True):
        """`total` Env.steps as episodes of H steps (reset before each, outside the brackets),
        each episode in calls of <= chunk steps. Returns (sum of bracket ms, launches inside)."""
        torch, eng = self.torch, self.eng
        pairs, launches = [], 0
        for n, calls in plan_episodes(total, self.H, chunk):
            self.env.reset()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            l0 = eng.launch_count
            e0.record()
            for _, k in calls:
                call(k)
            if drain:
                drain()
            e1.record()
            launches += eng.launch_count - l0
            pairs.append((e0, e1))
        torch.cuda.synchronize()
        return sum(e0.elapsed_time(e1) for e0, e1 in pairs), launches

    def timed(self, call, chunk, k_steps=None, drain=None, target_ms=None, max_total=4_000_000):
        """calibrate -> R -> W warm-up steps -> barrier -> R x K steps -> barrier; time max'ed and
        rollout statistics summed over ranks. Returns the result dict (same on every rank)."""
        import torch.distributed as dist
        a, torch, eng = self.a, self.torch, self.eng
        K = k_steps or self.K
        target = a.target_ms if target_ms is None else target_ms
        cal_steps = max(chunk, min(self.H, 4 * chunk)) if chunk > 1 else 64
        self._run(call, cal_steps, chunk, drain)                       # first touch, graph capture
        ms_cal, _ = self._run(call, cal_steps, chunk, drain)
        R = a.reps or plan_reps(K, ms_cal / cal_steps, target, max_total)
        if self.ctx['world'] > 1:
            r = torch.tensor([R], device='cuda', dtype=torch.int64)
            dist.all_reduce(r, op=dist.ReduceOp.MAX)
            R = int(r.item())
        R = round_reps(R, K, chunk)          # R x K = a whole number of `chunk`-step calls
        # the W untimed warm-up steps (whole calls: a partial one would re-capture the CUDA graph
        # of the single-step leg inside the timed region)
        self._run(call, min(self.H, int(math.ceil(max(a.warmup, 1) / chunk)) * chunk), chunk, drain)
        eng.stats(reset=True)
        self.barrier()
        ms, launches = self._run(call, R * K, chunk, drain)
        self.barrier()
        st = eng.stats()
        # rollout statistics summed, time max'ed over ranks (NCCL all-reduce of 7 doubles)
        st, ms = self.SH.reduce_stats(st, device='cuda', time_ms=ms)
        agg = self.SH.summarize(st)
        return dict(ms=ms, reps=R, steps=R * K, vehicle_steps=st['vehicle_steps'],
                    env_steps=st['env_steps'], crashes=st['crashes'],
                    mean_reward=agg['mean_reward'], mean_speed=agg['mean_speed'],
                    launches=launches, vps=st['vehicle_steps'] / (ms * 1e-3),
                    ms_per_step=ms / (R * K))

    def close(self):
        self.env.close()
        self.act = self.obs_seq = self.rew_seq = self.done_seq = self.host = None
        self.torch.cuda.empty_cache()

    # ---- rooflines
    def flop_per_vehicle_step(self):
        return FLOP_IDM + (FLOP_NOISE if self.a.noise > 0 else 0) + \
            (FLOP_SAFE if self.a.speed_mode & 1 else 0)

    def bytes_per_env_step(self, active_per_env):
        """SURVEY 8d single-step algorithmic bytes: state round trip 4 x T per (active) vehicle
        (+ lane / slot-type byte r/w, + the 4-byte vehicle id on the multi-agent highway), ring
        length + actions in, obs + reward + done out."""
        ts, env = self.ts, self.env
        extra = {'c3': 2, 'c4': 2, 'mah': 10, 'mg': 20}.get(self.workload, 0)   # mg: arm, type, id, seq r/w
        return active_per_env * (4 * ts + extra) + ts + ts * self.A + ts * env.obs_dim + ts + 1

    def roofline_fused(self, r):
        world = self.ctx['world']
        prof = ncu_profile(self.workload, self.dtype) or {}
        peak = PEAK_TFLOPS[self.dtype]
        per_gpu_vps = r['vps'] / world
        ach = per_gpu_vps * self.flop_per_vehicle_step() / 1e12
        env_steps_per_launch = r['env_steps'] / world / max(r['launches'], 1)
        tr = prof.get('dram_bytes_per_env_step')
        return {
            'bound': 'fp32-issue' if self.dtype == 'f32' else 'fp64-issue',
            'achieved': ach, 'peak': peak, 'unit': 'TFLOP/s', 'frac': ach / peak,
            'traffic': None if tr is None else tr * env_steps_per_launch,
            'traffic_note': 'ncu dram__bytes_read.sum + dram__bytes_write.sum of one fused launch '
                            '(profiles/ncu_traffic.json: bytes per env-step) x env-steps per launch here',
            'peak_source': '148 SMs x %d lanes x 2 (FMA) x 1.965 GHz (SURVEY 8d; B200_PROFILING.md has no '
                           'measured non-tensor figure)' % (128 if self.dtype == 'f32' else 64),
            'flop_per_vehicle_step': self.flop_per_vehicle_step(),
            'kernel': '%s<%s> (flow_b200/csrc/), %d fused steps per launch'
                      % (KERNEL_OF.get(self.workload, 'ring1_hot_kernel'), self.dtype, self.F),
            'launch_ms': r['ms'] / max(r['launches'], 1),
            'ncu': {k: prof.get(k) for k in ('issue_active_pct', 'warps_active_pct',
                                             'warp_inst_per_env_step', 'thread_inst_per_vehicle_step',
                                             'fma_pipe_pct', 'alu_pipe_pct', 'xu_pipe_pct',
                                             'capture') if k in prof} or None,
            'note': 'fused K-step launch: vehicle state lives in registers / shared memory for the whole '
                    'launch, DRAM carries only actions in and obs/reward/done out (`traffic`), so the '
                    'bound is instruction issue (SURVEY 8d second bullet): achieved = vehicle-steps/s '
                    'per GPU x counted flop per vehicle-step (IDM + move + headway + reward 30, '
                    'Philox + Box-Muller noise 25, speed-mode safe speed = second IDM 18). The HBM reading belongs to the one-step-per-launch '
                    'variant: `roofline_single_step`.'}

    def roofline_single(self, r, active_per_env):
        peak, src = measured_peak()
        per_gpu_env_steps_per_s = r['env_steps'] / self.ctx['world'] / (r['ms'] * 1e-3)
        bpe = self.bytes_per_env_step(active_per_env)
        ach = bpe * per_gpu_env_steps_per_s / 1e9
        return {'bound': 'hbm', 'unit': 'GB/s', 'peak': peak, 'peak_source': src, 'achieved': ach,
                'frac': ach / peak, 'bytes_per_env_step': bpe,
                'note': 'one launch per Env.step: ALGORITHMIC bytes (SURVEY 8d: state read + written '
                        'every step + ring length / action in, obs / reward / done out) per env-step x '
                        'env-steps/s. At %d envs the state (%.1f MB) is L2-resident between launches, '
                        'so real DRAM traffic is lower; the step time is launch ramp + one dependent '
                        'pass over the state.' % (self.E, self.E * active_per_env * 2 * self.ts / 1e6)}


def run_leg(a, ctx, workload, dtype, scaling, k_steps, modes):
    """Build the leg, run the requested timed passes, return plain dicts (same on all ranks)."""
    leg = Leg(a, ctx, workload, dtype, scaling, k_steps)
    F, out = leg.F, {}
    out['fused'] = leg.timed(leg.call_fused, F)
    if 'single' in modes:
        # graph of F single-step launches replayed; per-step cost ~6-8 us
        out['single'] = leg.timed(leg.call_single_graph, F, target_ms=a.target_ms / 2)
    if 'single_call' in modes:
        out['single_call'] = leg.timed(leg.call_single_python, F, target_ms=a.target_ms / 2)
    if 'host' in modes:
        out['host'] = leg.timed(leg.call_host_async, F, drain=leg.env.wait)
        out['host_blocking'] = leg.timed(leg.call_host_blocking, F, target_ms=a.target_ms / 2)
        out['host_closed'] = leg.timed(leg.call_host_closed_loop, F, target_ms=a.target_ms / 2)
    fused = out['fused']
    sps = int(getattr(leg.eng.config, 'sims_per_step', 1) or 1)
    # vehicles in an env: vehicle-steps per Env.step / simulation steps per Env.step
    active = fused['vehicle_steps'] / max(fused['env_steps'], 1.0) / sps
    info = dict(workload=workload_name(workload), envs_per_gpu=leg.E, dtype=dtype, fuse=F,
                scaling=scaling, obs_dim=leg.env.obs_dim, act_dim=leg.A, ring_steps=leg.NB,
                active_vehicles_per_env=active, n_slots=leg.eng.n_veh, horizon=leg.H,
                sims_per_step=sps,
                h2d_bytes_per_step=leg.act_b, d2h_bytes_per_step=leg.out_b,
                streamed_mb_per_ring_cycle=leg.NB * (leg.act_b + leg.out_b) / 1e6)
    roof = dict(fused=leg.roofline_fused(fused))
    if 'single' in out:
        roof['single'] = leg.roofline_single(out['single'], active)
    leg.close()
    return out, info, roof


def brief(r):
    return {'value': r['vps'], 'unit': 'vehicle-steps/s', 'ms_per_step': r['ms_per_step'],
            'reps': r['reps'], 'timed_region_ms': r['ms'], 'env_steps_per_s': r['env_steps'] / (r['ms'] * 1e-3),
            'launches': r['launches']}


def run_b200(a):
    import torch
    import torch.distributed as dist
    from flow_b200 import sharding as SH

    rank, world, local = SH.env_rank_world()
    if world != a.gpus and world > 1:
        raise SystemExit('--gpus %d but WORLD_SIZE=%d' % (a.gpus, world))
    torch.cuda.set_device(local)
    numa_cores = SH.bind_to_gpu_cpus(local) if world > 1 else 0
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    ctx = dict(rank=rank, world=world, local=local)
    K, W = a.steps, a.warmup

    sampler = ClockSampler(local) if rank == 0 else None
    modes = () if a.only_fused else ('single', 'single_call', 'host')
    res, info, roof = run_leg(a, ctx, a.workload, a.dtype, a.scaling, K, modes)
    legs = {}
    if not (a.only_fused or a.no_extra_legs) and a.workload == 'c2' and a.dtype == 'f32' \
            and a.scaling == 'weak' and not a.envs:
        def sub(name, wl, dt, sc, md):
            r, i, rf = run_leg(a, ctx, wl, dt, sc, K, md)
            d = brief(r['fused'])
            d.update(config=i, roofline=rf['fused'],
                     crashed_env_frac=r['fused']['crashes'] / max(
                         i['envs_per_gpu'] * world * math.ceil(r['fused']['steps'] / float(i['horizon'])), 1))
            if 'single' in r:
                d['single_step'] = brief(r['single'])
                d['roofline_single_step'] = rf['single']
            legs[name] = d
        sub('c5', 'c5', 'f32', 'weak', ('single',))
        sub('f64', 'c2', 'f64', 'weak', ())
        if world > 1:
            sub('strong', 'c5', 'f32', 'strong', ())
    clocks = sampler.stop() if sampler else None
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    fused = res['fused']
    E = info['envs_per_gpu']
    line = {
        'metric': 'vehicle-steps/s', 'value': fused['vps'], 'unit': 'vehicle-steps/s',
        'n_gpus': world, 'steps': K, 'warmup': W, 'ms_per_step': fused['ms_per_step'],
        'higher_is_better': True, 'scaling': a.scaling, 'vs_baseline': None, 'dtype': a.dtype,
        'data': 'synthetic',
        'env_steps_per_s': fused['env_steps'] / (fused['ms'] * 1e-3),
        'config': {'workload': info['workload'], 'envs_per_gpu': E,
                   'vehicles_per_env': info['n_slots'], 'idm_noise': a.noise,
                   'speed_mode': a.speed_mode, 'fuse': info['fuse'], 'horizon': info['horizon'],
                   'sims_per_step': info['sims_per_step'],
                   'reps': fused['reps'], 'timed_region_ms': fused['ms'],
                   'env_steps_timed_per_env': fused['steps'],
                   'timing': 'the K-step pass repeated `reps` times back to back (calibrated so the '
                             'bracket holds >= %.0f ms of GPU work, same R on every rank); issued as a '
                             'long rollout: fused launches of `fuse` steps, envs reset (+ warm-up) every '
                             '%d steps (the horizon) BETWEEN the CUDA-event pairs on the same stream; '
                             'ms_per_step = sum of brackets / (reps x steps)' % (a.target_ms, info['horizon']),
                   'parallelism': 'envs sharded, %d GPU(s), no data-path collective' % world,
                   'l2': 'actions / obs / reward / done stream through device ring buffers of %d steps = '
                         '%.0f MB per cycle (> 126 MB L2; actions alone %.0f MB); vehicle state is '
                         'register-resident inside a fused launch'
                         % (info['ring_steps'], info['streamed_mb_per_ring_cycle'],
                            info['ring_steps'] * info['h2d_bytes_per_step'] / 1e6),
                   'host_cores_bound_to_gpu_numa_node': numa_cores,
                   'counted': 'vehicle-steps actually advanced (engine counter); frozen (crashed) '
                              'envs are not counted'},
        'crashed_env_frac': fused['crashes'] / max(E * world * math.ceil(fused['steps'] / float(info['horizon'])), 1),   # per episode
        'mean_reward': fused['mean_reward'], 'mean_speed': fused['mean_speed'],
        'active_vehicles_per_env': info['active_vehicles_per_env'],
        'gpu_launches': fused['launches'],
        'roofline': roof['fused'],
        'clocks': clocks,
    }
    if not a.only_fused:
        single, single_call = res['single'], res['single_call']
        host, hb, hc = res['host'], res['host_blocking'], res['host_closed']
        line['single_step'] = dict(brief(single), api='fs_step_unfused: one launch per Env.step (state '
                                   'through memory every step), %d launches replayed from one CUDA graph'
                                   % info['fuse'])
        line['single_step_per_call'] = dict(brief(single_call),
                                            api='fs_step(k=1) called from Python (ctypes), device buffers')
        line['roofline_single_step'] = roof['single']
        line['e2e'] = {
            'value': host['vps'], 'unit': 'vehicle-steps/s',
            'h2d_bytes_per_step': info['h2d_bytes_per_step'], 'd2h_bytes_per_step': info['d2h_bytes_per_step'],
            'ms_per_step': host['ms_per_step'], 'reps': host['reps'], 'timed_region_ms': host['ms'],
            'api': 'open loop: B200VecEnv.rollout_host_async -> fs_step_host_async, %d steps per call, '
                   'pinned host buffers, upload / kernel / download of consecutive calls overlapped, '
                   'one wait per episode' % info['fuse'],
            'blocking': dict(brief(hb), api='B200VecEnv.rollout_host -> fs_step_host, %d steps per call, '
                             'every call waits for its results' % info['fuse']),
            'closed_loop': dict(brief(hc), api='B200VecEnv.step_host -> fs_step_host(k=1): ONE Env.step per '
                                'call, host action copied into the engine\'s pinned staging block, upload -> '
                                'kernel -> one download replayed as one CUDA graph, obs / reward / done on '
                                'the host before the next call (the gym cadence of Env.step(policy(obs)))')}
    if legs:
        line['legs'] = legs
    if world == 1 and not a.no_cpu_baseline and not a.only_fused:
        line['cpu_baseline'] = cpu_baseline(a.cpu_seconds, a.noise, a.speed_mode)
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
