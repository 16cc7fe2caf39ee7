// fs_ring_hot.cuh — the c2 / c5 kernel: single-lane ring of V = 22 vehicles,
// 21 IDM humans of one type followed by ONE RL vehicle at index V-1, observed
// by WaveAttenuationPOEnv, no fail-safe, no views. This is
// exactly what examples/exp_configs/rl/singleagent/singleagent_ring.py builds;
// anything else runs on ring1_fast_kernel (fs_ring_fast.cuh) or ring1_kernel.
//
// Same thread mapping and step semantics as ring1_fast_kernel (G threads per
// env, VPT vehicles per thread in registers for the K fused steps), with
// everything that kernel decides at run time per vehicle or per step fixed at
// compile time:
//   * V, the RL vehicle's thread (RL_G) and slot (RL_SLOT) are constants, so
//     the per-vehicle "is this one RL?" tests, mask shifts and the alternate
//     action loads disappear; the RL thread is also the writer of obs/reward/
//     done, and the leader of vehicle V-1 is vehicle 0 = the exchanged value;
//   * kAct (RL actions present) is a template flag: warm-up launches carry no
//     action loads, reward arithmetic or clipping;
//   * one-vehicle-per-thread shapes (VPT <= 4) run the step loop in blocks of 4
//     steps aligned to the env's time counter, unrolled: the four normals of a
//     Philox call and the four prefetched actions of a block are addressed with
//     compile-time indices (no register rotation, no per-step phase tests);
//   * Philox4x32-10 runs with a compile-time key schedule (seed and global env
//     id travel in the counter words), so the 20 key additions per call fold
//     into immediates;
//   * the steps-advanced statistic is the time-counter difference.
// Reference lines restated: see fs_ring.cuh header (same step semantics).
#pragma once
#include "fs_ring_fast.cuh"

namespace fs {

struct PhiloxCK {
  // counter = (c0,c1,c2,c3) -> 4 x u32, fixed key (0xA4093822, 0x299F31D0)
  static __device__ __forceinline__ void gen(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                             uint32_t (&out)[4]) {
#pragma unroll
    for (int i = 0; i < 10; ++i) {
      const uint32_t k0 = 0xA4093822u + (uint32_t)i * 0x9E3779B9u;
      const uint32_t k1 = 0x299F31D0u + (uint32_t)i * 0xBB67AE85u;
      unsigned long long p0, p1;   // one IMAD.WIDE.U32 each (a C++ 64-bit product adds a dead carry)
      asm("mul.wide.u32 %0, %1, %2;" : "=l"(p0) : "r"(c0), "r"(0xD2511F53u));
      asm("mul.wide.u32 %0, %1, %2;" : "=l"(p1) : "r"(c2), "r"(0xCD9E8D57u));
      const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
      const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
      c1 = (uint32_t)p1; c3 = (uint32_t)p0; c0 = n0; c2 = n2;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
  }
};

// four standard normals (Box-Muller on both pairs of one Philox call)
template <typename T>
__device__ __forceinline__ void normal4_ck(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                           T (&z)[4]) {
  uint32_t r[4];
  PhiloxCK::gen(c0, c1, c2, c3, r);
  if (sizeof(T) == 4) {
    const float k = 2.3283064365386963e-10f;  // 2^-32
    const float u0 = fminf(fmaf((float)r[0], k, k), 1.0f), u2 = fminf(fmaf((float)r[2], k, k), 1.0f);
    const float ra = sqrtf(-1.3862943611198906f * __log2f(u0));   // sqrt(-2 ln u)
    const float rb = sqrtf(-1.3862943611198906f * __log2f(u2));
    float s0, c0f, s1, c1f;
    const float k2pi = 1.4629180792671596e-09f;   // 2 pi 2^-32
    __sincosf((float)r[1] * k2pi, &s0, &c0f);
    __sincosf((float)r[3] * k2pi, &s1, &c1f);
    z[0] = (T)(ra * c0f); z[1] = (T)(ra * s0); z[2] = (T)(rb * c1f); z[3] = (T)(rb * s1);
  } else {
    const double k = 1.0 / 4294967296.0;
    const double ra = ::sqrt(-2.0 * ::log(((double)r[0] + 1.0) * k));
    const double rb = ::sqrt(-2.0 * ::log(((double)r[2] + 1.0) * k));
    double s0, c0d, s1, c1d;
    sincos(6.283185307179586 * ((double)r[1] * k), &s0, &c0d);
    sincos(6.283185307179586 * ((double)r[3] * k), &s1, &c1d);
    z[0] = (T)(ra * c0d); z[1] = (T)(ra * s0); z[2] = (T)(rb * c1d); z[3] = (T)(rb * s1);
  }
}

constexpr int kHotV = 22;

// sum of `part` over the G threads of each env of the warp; every lane of the
// warp must call it (full-mask collectives). One env per warp in fp32: a single
// integer REDUX in 2^-20 m/s fixed point (|sum| < 2000 m/s); otherwise an
// xor-shuffle tree over the group's lanes (log2 G steps).
template <typename T, int G>
__device__ __forceinline__ T env_sum_full(T part, int es) {
  constexpr int EPW = 32 / G;
  if (G == 1) return part;
  if (sizeof(T) == 4 && EPW == 1) {
    const int q = __float2int_rn((float)part * 1048576.0f);
    return (T)((float)__reduce_add_sync(0xffffffffu, q) * 9.5367431640625e-07f);
  }
#pragma unroll
  for (int o = EPW; o < 32; o <<= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
  return part;
}

// kSafe: some type carries a SUMO speed mode (the reference's default, 'right_of_way' = 25, has
// bit 0 set): the speed Flow commands is capped by the restated simulator model's own next
// speed (speed_mode_fast, fs_ring_fast.cuh) - a second IDM evaluation per vehicle and step
template <typename T, int VPT, int G, bool kNoise, bool kAct, bool kSafe>
__global__ void __launch_bounds__(128)
ring1_hot_kernel(const __grid_constant__ FastP<T> P, const RingBufs<T> B, const StepArgs<T> A) {
  typedef Num<T> N;
  constexpr int V = kHotV;
  constexpr int EPW = 32 / G;
  constexpr unsigned FULL = 0xffffffffu;
  constexpr bool kF32 = sizeof(T) == 4;
  constexpr bool kExact = VPT * G == V;
  constexpr int RL_G = (V - 1) / VPT, RL_SLOT = (V - 1) % VPT;  // thread / slot of vehicle V-1
  constexpr bool kTimeNoise = VPT <= 4;   // 4 normals of a call spread over 4 consecutive steps
  constexpr int NQ = kTimeNoise ? VPT : (VPT + 3) / 4;
  constexpr int UNR = kTimeNoise ? 4 : 1;  // steps per unrolled block
  // Several envs share a warp (1 < G < 32): the step body runs for every lane
  // and frozen / not-selected envs are handled by predication, so that every
  // collective is a full-warp one (a partial-mask REDUX / VOTE compiles to a
  // loop over mask groups). One env per warp or per thread: plain branches.
  constexpr bool kPred = G > 1 && G < 32;
  static_assert(VPT * G >= V, "shape does not hold 22 vehicles");

  // Programmatic dependent launch (fs_step_unfused issues its one-step launches with
  // programmatic stream serialization): let the next launch of the stream start being
  // scheduled right away, and do not touch memory before the previous one has completed and
  // flushed. Both are no-ops for an ordinary launch.
  asm volatile("griddepcontrol.launch_dependents;");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const int lane = threadIdx.x & 31;
  const int g = (G == 1) ? 0 : lane / EPW;
  const int es = (G == 1) ? lane : lane % EPW;
  const int E = P.n_envs;
  const int e_raw = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * EPW + es;
  const bool valid = e_raw < E && (A.mask == nullptr || A.mask[e_raw < E ? e_raw : 0] != 0);
  if (!__any_sync(FULL, valid)) return;   // warp-uniform: all 32 lanes stay together below
  const int e = e_raw < E ? e_raw : E - 1;
  unsigned gmask = 0;
#pragma unroll
  for (int q = 0; q < G; ++q) gmask |= 1u << (es + q * EPW);

  const int base = g * VPT;
  const int n_mine = kExact ? VPT : max(0, min(VPT, V - base));
  const bool is_rl = g == RL_G;   // owns vehicle V-1: RL vehicle, writer of obs / reward / done
  const int lane_next = ((g >= RL_G) ? 0 : g + 1) * EPW + es;
  const int K = A.k_steps;

  IdmRegs<T> pm = P.pa;
  T nz = P.noise;
  if (VPT == 1 && is_rl) { pm = P.pb; nz = T(0); }
  // simulator model behind the speed mode for slot RL_SLOT: the RL type's on the RL thread
  IdmRegs<T> psim_rl = P.ps;
  if (kSafe && is_rl) psim_rl = P.pb;
  const T len = P.len, dt = P.dt;
  const T C = B.circ[e];
  const T crash_thr = P.collision_min_gap ? P.min_gap : T(0);
  const T inv_obs_len = P.inv_obs_len > T(0) ? P.inv_obs_len : T(1) / C;
  const T obs_len = P.obs_max_length > T(0) ? P.obs_max_length : C;
  const T inv_C = T(1) / C;

  T x[VPT], v[VPT], h[VPT];
  FS_SLOTS(i) {
    const bool ok = kExact || i < n_mine;
    x[i] = ok ? B.pos[(size_t)(base + i) * E + e] : T(0);
    v[i] = ok ? B.vel[(size_t)(base + i) * E + e] : T(0);
  }
  // first vehicle of the next thread = leader of my last vehicle (vehicle 0 for the RL thread)
  T x_nf = (G == 1) ? x[0] : __shfl_sync(FULL, x[0], lane_next);
  T v_nf = (G == 1) ? v[0] : __shfl_sync(FULL, v[0], lane_next);
  auto update_headways = [&]() {  // k.update: vehicle/traci.py:218-250
    FS_SLOTS(i) {
      T xl;
      if (i + 1 < VPT) xl = (kExact || i + 1 < n_mine) ? x[i + 1 < VPT ? i + 1 : 0] : x_nf;
      else xl = x_nf;
      h[i] = fwd_dist(xl, x[i], C) - len;
    }
  };
  update_headways();

  int tc = B.time_counter[e];
  const int tc0 = tc;
  const int episode = B.episode[e];
  bool done = A.ignore_done ? false : (B.done[e] != 0);
  bool crash = B.crash[e] != 0;
  const long long gid = P.env_id_base + e;
  T reward = B.reward[e];
  // Philox counter words 2, 3: global env id and seed; word 1: episode / vehicle block
  const uint32_t pc2 = (uint32_t)gid;
  const uint32_t pc3 = (uint32_t)(gid >> 32) ^ (uint32_t)P.seed ^ ((uint32_t)(P.seed >> 32) * 0x9E3779B9u);
  const uint32_t pc1 = (uint32_t)(episode * 256 + base) | (kTimeNoise ? 0u : 0x80000000u);

  const T* act_p = A.actions + e;  // action of step s: act_p[s * E] (act_dim == 1)
  auto ld_act = [&](int s) -> T {  // clamped: slots outside [0, K) are fetched but never used
    s = s < 0 ? 0 : (s >= K ? K - 1 : s);
    return __ldg(act_p + (size_t)s * E);
  };

  float st_rew = 0.f, st_v = 0.f, st_v2 = 0.f;
  int st_cr = 0;
  // per-step outputs: a missing sequence is redirected to the env's last-step
  // slot with stride 0 (it ends up holding the last step's value, which is what
  // the epilogue stores there anyway), so one predicate guards all stores
  const bool seq_on = A.obs_seq != nullptr || A.reward_seq != nullptr || A.done_seq != nullptr;
  const bool wr = valid && is_rl;
  T* const obs_b = A.obs_seq ? A.obs_seq + (size_t)e * 3 : B.obs + (size_t)e * 3;
  T* const rew_b = A.reward_seq ? A.reward_seq + e : B.reward + e;
  uint8_t* const done_b = A.done_seq ? A.done_seq + e : B.done + e;
  const unsigned obs_st = A.obs_seq ? (unsigned)E * 3u : 0u;   // elements per step
  const unsigned rew_st = A.reward_seq ? (unsigned)E : 0u, done_st = A.done_seq ? (unsigned)E : 0u;

  auto po_obs = [&](T& o0, T& o1, T& o2) {  // WaveAttenuationPOEnv.get_state, wave_attenuation.py:248-269
    const T xr = x[RL_SLOT], vr = v[RL_SLOT];
    if (kF32) { o0 = vr * T(1.0 / 15.0); o1 = (v_nf - vr) * T(1.0 / 15.0); }
    else { o0 = N::div(vr, T(15)); o1 = N::div(v_nf - vr, T(15)); }
    o2 = kF32 ? fwd_dist(x_nf, xr, C) * inv_obs_len : N::div(fwd_dist(x_nf, xr, C), obs_len);
  };
  // One pass per phase of the 4-step noise block (tc & 3) that occurs in this
  // warp: a pass advances the envs of that phase through all K steps with the
  // block loop aligned to their time counters; the other envs are predicated
  // off. All envs of a warp share the phase unless they were reset at different
  // times, so normally exactly one pass runs.
  int my_phase = kTimeNoise ? (tc & 3) : 0;
  if (kTimeNoise && G < 32) {
    // frozen envs (done: their time counter stopped at some other phase) ride
    // along with the first live phase of the warp instead of costing a pass
    unsigned live_ph = 0;
#pragma unroll
    for (int ph = 0; ph < 4; ++ph)
      if (__any_sync(FULL, !done && my_phase == ph)) live_ph |= 1u << ph;
    if (done) my_phase = live_ph ? __ffs(live_ph) - 1 : 0;
  }
#pragma unroll 1
  for (int ph = 0; ph < (kTimeNoise ? 4 : 1); ++ph) {
    const bool sel = my_phase == ph;
    if (kTimeNoise && !__any_sync(FULL, sel)) continue;
    T ac[UNR], an[UNR];
    int j0 = ph;        // first (possibly partial) block starts at slot j0
    int kb = -j0;       // launch-relative step of the block's slot 0
    // running output pointers of the block's slot 0 (a 64-bit add per block instead of a
    // 64-bit multiply per step and stream; slots before step 0 are never dereferenced)
    T* obs_p = obs_b + (ptrdiff_t)kb * (ptrdiff_t)obs_st;
    T* rew_p = rew_b + (ptrdiff_t)kb * (ptrdiff_t)rew_st;
    uint8_t* done_p = done_b + (ptrdiff_t)kb * (ptrdiff_t)done_st;
#pragma unroll
    for (int j = 0; j < UNR; ++j) { ac[j] = T(0); an[j] = (kAct && is_rl) ? ld_act(kb + j) : T(0); }

#pragma unroll 1
    while (kb < K) {
      if (kAct) {
#pragma unroll
        for (int j = 0; j < UNR; ++j) ac[j] = an[j];
        if (is_rl) {   // actions of the next block, requested one block ahead
          if (kb >= 0 && kb + 2 * UNR <= K) {   // whole next block inside the launch: no clamping
            const T* q = act_p + (size_t)(unsigned)(kb + UNR) * (unsigned)E;
#pragma unroll
            for (int j = 0; j < UNR; ++j) an[j] = __ldg(q + (size_t)((unsigned)j * (unsigned)E));
          } else {
#pragma unroll
            for (int j = 0; j < UNR; ++j) an[j] = ld_act(kb + UNR + j);
          }
        }
      }
      T zn[kNoise ? NQ : 1][4];
      if (kNoise && kTimeNoise && (kPred || !done)) {
#pragma unroll
        for (int q = 0; q < NQ; ++q) normal4_ck<T>((uint32_t)(tc >> 2), pc1 + q, pc2, pc3, zn[q]);
      }
#pragma unroll
      for (int j = 0; j < UNR; ++j) {
        if (j >= j0 && kb + j < K) {    // warp-uniform
          const bool live = sel && !done;
          if (kPred || live) {
            if (kNoise && !kTimeNoise) {  // base_controller.py:109-110
#pragma unroll
              for (int q = 0; q < NQ; ++q) normal4_ck<T>((uint32_t)tc, pc1 + 4 * q, pc2, pc3, zn[q]);
            }
            // ---- controllers -> apply_acceleration -> Euler move (all read the old state)
            const T v_lead_last = v_nf;
            T act_abs = T(0);
            FS_SLOTS(i) {
              const bool ok = (kExact || i < n_mine) && (!kPred || live);
              const T vi = v[i], hi = h[i];
              T vl;
              if (i + 1 < VPT) vl = (kExact || i + 1 < n_mine) ? v[i + 1 < VPT ? i + 1 : 0] : v_lead_last;
              else vl = v_lead_last;
              // controller law and, with a speed mode (kSafe), the simulator model whose next
              // speed caps what Flow commands (restated SUMO, speed_mode_fast): for the RL
              // vehicle's slot the model of the RL type on the RL thread (psim_rl)
              T a, a_sim = T(0);
              if (kSafe && kF32) {
                float af, asf;
                idm_pair_f32((float)vi, (float)vl, (float)hi, reinterpret_cast<const IdmRegs<float>&>(pm),
                             reinterpret_cast<const IdmRegs<float>&>(i == RL_SLOT ? psim_rl : P.ps), af, asf);
                a = (T)af; a_sim = (T)asf;
              } else {
                a = idm_fast<T, true>(vi, vl, hi, pm);
                if (kSafe) a_sim = idm_fast<T, true>(vi, vl, hi, i == RL_SLOT ? psim_rl : P.ps);
              }
              if (kNoise) {
                const T z = kTimeNoise ? zn[kNoise ? i : 0][j] : zn[kNoise ? (i >> 2) : 0][i & 3];
                a = add(a, mul(VPT == 1 ? nz : P.noise, z));
              }
              if (kSafe) {   // SUMO's speed mode on the human's commanded speed
                if (P.safe_a) a = N::min(a, a_sim);
                a = N::max(N::min(a, P.cap_a), P.floor_a);
              }
              if (i == RL_SLOT) {
                if (kAct) {  // Env.apply_rl_actions: Box clip, then apply_acceleration (base.py:599-615)
                  const T raw = ac[j];
                  T a_c = N::min(N::max(raw, P.act_lo), P.act_hi);
                  act_abs = N::abs(P.clip_actions ? a_c : raw);
                  if (kSafe) {
                    if (P.safe_b) a_c = N::min(a_c, a_sim);
                    a_c = N::max(N::min(a_c, P.cap_b), P.floor_b);
                  }
                  a = is_rl ? a_c : a;
                } else {  // un-commanded RL vehicle: restated simulator default model, no speed mode
                  if (is_rl) a = idm_fast<T, true>(vi, vl, hi, P.pb);
                }
              }
              const T vn = N::max(add(vi, mul(a, dt)), T(0));   // vehicle/traci.py:961-962
              const T xn = wrap_fwd(add(x[i], mul(vn, dt)), C);  // simulator Euler move
              if (ok) { v[i] = vn; x[i] = xn; }
            }
            if (G > 1) { x_nf = __shfl_sync(FULL, x[0], lane_next); v_nf = __shfl_sync(FULL, v[0], lane_next); }
            else { x_nf = x[0]; v_nf = v[0]; }
            update_headways();
            bool cr = false;
            T sv = T(0);
            float s2 = 0.f;
            FS_SLOTS(i) if (kExact || i < n_mine) {
              cr |= h[i] < crash_thr;
              sv = add(sv, v[i]);
              s2 = fmaf((float)v[i], (float)v[i], s2);
            }
            if (!kPred || live) { st_v += (float)sv; st_v2 += s2; }
            if (G > 1) cr = (__ballot_sync(FULL, cr) & gmask) != 0u;
            T r = T(0);
            if (kAct) {  // WaveAttenuationEnv.compute_reward, wave_attenuation.py:113-139
              sv = env_sum_full<T, G>(sv, es);
              r = kF32 ? sv * P.rew_scale : N::div(mul(T(4), N::div(sv, P.n_veh_t)), T(20));
              if (act_abs > T(0)) r = add(r, mul(T(4), T(0) - act_abs));
              r = cr ? T(0) : r;
            }
            if (!kPred || live) {
              crash = cr;
              tc += 1;
              done = crash || tc >= P.t_end;
              reward = r;
              st_rew += (float)r; st_cr += cr ? 1 : 0;
            } else if (sel) {
              reward = T(0);
            }
          } else if (sel) {
            reward = T(0);
          }
          // ---- per-step outputs
          if (seq_on) {
            if (wr && sel) {
              T o0, o1, o2; po_obs(o0, o1, o2);
              T* o = obs_p + (unsigned)j * obs_st;
              o[0] = o0; o[1] = o1; o[2] = o2;
              rew_p[(unsigned)j * rew_st] = reward;
              done_p[(unsigned)j * done_st] = done ? 1 : 0;
            }
          }
        }
      }
      kb += UNR;
      j0 = 0;
      obs_p += (unsigned)UNR * obs_st; rew_p += (unsigned)UNR * rew_st; done_p += (unsigned)UNR * done_st;
    }
  }

  if (valid) {
    T* ol = B.obs + (size_t)e * 3;
    if (is_rl) { T o0, o1, o2; po_obs(o0, o1, o2); ol[0] = o0; ol[1] = o1; ol[2] = o2; }
    FS_SLOTS(i) if (kExact || i < n_mine) {
      B.pos[(size_t)(base + i) * E + e] = x[i];
      B.vel[(size_t)(base + i) * E + e] = v[i];
    }
    if (is_rl) {
      B.time_counter[e] = tc;
      B.done[e] = done ? 1 : 0;
      B.crash[e] = crash ? 1 : 0;
      B.reward[e] = reward;
    }
  }
  if (A.accumulate_stats) {
    // stats[]: sum reward, sum v, sum v^2, vehicle-steps, env-steps, crashes
    const bool st_on = valid;
    const int adv = st_on ? tc - tc0 : 0;
    double vals[6] = {(st_on && is_rl) ? (double)st_rew : 0.0, st_on ? (double)st_v : 0.0,
                      st_on ? (double)st_v2 : 0.0, (double)(adv * n_mine),
                      is_rl ? (double)adv : 0.0, (st_on && is_rl) ? (double)st_cr : 0.0};
    substep_stats(A.accumulate_stats, vals);
    double* sp = stat_stripe(B.stats);
#pragma unroll
    for (int q = 0; q < 6; ++q) {
      double s = vals[q];
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(FULL, s, o);
      if (lane == 0 && s != 0.0) atomicAdd(&sp[q], s);
    }
  }
}

// ---- host-side dispatch (one translation unit per dtype instantiates these)
template <typename T, int VPT, int G>
static cudaError_t launch_hot_shape(bool noise, bool act, int grid, int block, cudaStream_t st,
                                    const FastP<T>& c, const RingBufs<T>& b, const StepArgs<T>& a) {
  auto go = [&](void (*kern)(const FastP<T>, const RingBufs<T>, const StepArgs<T>)) {
    if (a.pdl) {   // programmatic dependent launch: overlaps this launch's ramp with the previous tail
      cudaLaunchConfig_t lc = {};
      lc.gridDim = dim3(grid); lc.blockDim = dim3(block); lc.dynamicSmemBytes = 0; lc.stream = st;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      at[0].val.programmaticStreamSerializationAllowed = 1;
      lc.attrs = at; lc.numAttrs = 1;
      cudaLaunchKernelEx(&lc, kern, c, b, a);
    } else {
      kern<<<grid, block, 0, st>>>(c, b, a);
    }
  };
  const bool safe = c.any_speed_mode != 0;
#define FS_HOT(NZ, AC) \
  do { if (safe) go(ring1_hot_kernel<T, VPT, G, NZ, AC, true>); \
       else go(ring1_hot_kernel<T, VPT, G, NZ, AC, false>); } while (0)
  if (noise) { if (act) FS_HOT(true, true); else FS_HOT(true, false); }
  else { if (act) FS_HOT(false, true); else FS_HOT(false, false); }
#undef FS_HOT
  return cudaGetLastError();
}

// shape = threads per env (1, 2, 4, 8, 16, 32)
template <typename T>
static cudaError_t launch_hot22(int shape, bool noise, int n_envs, cudaStream_t st,
                                const FastP<T>& c, const RingBufs<T>& b, const StepArgs<T>& a) {
  const int epw = 32 / shape;
  const int warps = (n_envs + epw - 1) / epw;
  const int wpb = warps <= 148 * 8 ? 1 : (warps <= 148 * 32 ? 2 : 4);
  const int grid = (warps + wpb - 1) / wpb, block = wpb * 32;
  const bool act = a.actions != nullptr;
  switch (shape) {
    case 1: return launch_hot_shape<T, 22, 1>(noise, act, grid, block, st, c, b, a);
    case 2: return launch_hot_shape<T, 11, 2>(noise, act, grid, block, st, c, b, a);
    case 4: return launch_hot_shape<T, 6, 4>(noise, act, grid, block, st, c, b, a);
    case 8: return launch_hot_shape<T, 3, 8>(noise, act, grid, block, st, c, b, a);
    case 16: return launch_hot_shape<T, 2, 16>(noise, act, grid, block, st, c, b, a);
    default: return launch_hot_shape<T, 1, 32>(noise, act, grid, block, st, c, b, a);
  }
}

}  // namespace fs
