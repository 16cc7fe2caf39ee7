// fs_ring_fast.cuh — the hot kernel: single-lane ring where every vehicle
// follows the IDM law (IDMController humans of ONE type + RL vehicles of one
// type, which is what examples/exp_configs/non_rl/ring.py and
// rl/singleagent/singleagent_ring.py build), V <= 32.
//
// Thread mapping as in fs_ring.cuh (G threads own one env, VPT vehicles per
// thread, all state in registers for the K fused steps of a launch), tuned
// for issue rate:
//   * everything uniform is precomputed on the host into FastP (a ~200 B
//     kernel argument): IDM block of the human / RL type, RL bit mask, obs and
//     reward constants. No per-vehicle table gathers, no prologue loops;
//   * ONE exchange per step: after the move the group's first (x, v) is
//     shuffled to the previous thread; it feeds the headway of this step, the
//     partial observation of this step and the leader speed of the next step;
//   * the RL vehicle's thread writes obs / reward / done itself (it already
//     holds its leader's state), so the PO observation costs no shuffles;
//   * reward sums: fp32 uses ONE integer warp reduction (REDUX) of the speeds
//     in 2^-20 m/s fixed point (abs error < 3e-6 on the sum) instead of a
//     shuffle tree; the collision flag is one warp vote; fp64 keeps the tree;
//   * noise: Philox4x32-10 + Box-Muller gives 4 normals per call; a vehicle
//     consumes them over 4 CONSECUTIVE steps (counter = step/4, component =
//     step%4), so the cost is 1/4 call per vehicle-step for every shape;
//   * RL actions are prefetched in blocks of 4 steps, one block ahead (the only
//     global load in the loop would otherwise sit on the step's critical path);
//   * delta == 4, noise, fail-safes, views and kExact (V == VPT*G) are
//     template flags.
// Reference lines restated: see fs_ring.cuh header (same step semantics).
#pragma once
#include "fs_common.cuh"
#include "fs_ring.cuh"

namespace fs {

template <typename T>
struct IdmRegs {  // one type's effective IDM block, register resident
  T v0, inv_v0, Tt, a, tsab, inv_tsab, delta, s0;
  __host__ __device__ void load(const TypeP<T>& t) {
    v0 = t.e_v0; inv_v0 = t.e_inv_v0; Tt = t.e_T; a = t.e_a; tsab = t.e_tsab;
    inv_tsab = t.e_inv_tsab; delta = t.e_delta; s0 = t.e_s0;
  }
};

// the simulator-model block of a type whatever its controller (for the speed mode's safe speed)
template <typename T>
static inline void load_sim_block(IdmRegs<T>& r, const TypeP<T>& t) {
  r.v0 = t.sim_v0; r.inv_v0 = t.sim_inv_v0; r.Tt = t.sim_tau; r.a = t.max_accel;
  r.tsab = t.sim_two_sqrt_ab; r.inv_tsab = t.sim_inv_two_sqrt_ab; r.delta = T(4); r.s0 = t.sim_min_gap;
}
// speed-mode bits of a type as kernel constants: bit 0 -> safe, bits 1 / 2 -> cap / floor
// (+-inf when the bit is off, so they are applied unconditionally)
template <typename T>
static inline void speed_mode_consts(const TypeP<T>& t, int& safe, T& cap, T& floor) {
  safe = t.speed_mode & 1;
  cap = (t.speed_mode & 2) ? t.max_accel : (T)INFINITY;
  floor = (t.speed_mode & 4) ? T(0) - t.max_decel : T(0) - (T)INFINITY;
}

// Host-precomputed constants of the hot kernel (built by make_fastp()).
template <typename T>
struct FastP {
  int n_envs, n_veh, env_kind, obs_dim, act_dim, n_rl;
  int t_end, clip_actions, collision_min_gap, fail_safe;
  unsigned rl_mask;   // bit i: vehicle i takes its acceleration from the action vector
  int rl_veh;         // vehicle of action slot 0 (WaveAttenuationPO observes it)
  long long env_id_base;
  unsigned long long seed;
  T dt, len, noise, delay, min_gap, target, act_lo, act_hi;
  T obs_max_length, speed_limit, max_cost, inv_nrl, n_veh_t;
  T rew_scale;        // 4 / (20 V): WaveAttenuation reward per unit of summed speed (fp32 path)
  T inv_obs_len;      // 1 / obs_max_length, 0 when the ring's own length normalises
  T inv_speed_limit;
  IdmRegs<T> pa, pb;  // human type / un-commanded RL type (simulator default model)
  IdmRegs<T> ps;      // simulator model of the HUMAN type (safe speed of its speed mode)
  T cap_a, floor_a, cap_b, floor_b;   // speed-mode bits 1 / 2 of the two types (+-inf when off)
  int safe_a, safe_b;                 // speed-mode bit 0
  int any_speed_mode;
};

template <typename T>
static inline FastP<T> make_fastp(const Cfg<T>& c) {
  FastP<T> p;
  memset(&p, 0, sizeof(p));
  p.n_envs = c.n_envs; p.n_veh = c.n_veh; p.env_kind = c.env_kind; p.obs_dim = c.obs_dim;
  p.act_dim = c.act_dim; p.n_rl = c.n_rl;
  p.t_end = c.sims_per_step * (c.warmup_steps + c.horizon);
  p.clip_actions = c.clip_actions; p.collision_min_gap = c.collision_min_gap;
  int ta = 0, tb = 0;
  for (int i = 0; i < c.n_veh; ++i) {
    if (c.rl_index[i] >= 0) { p.rl_mask |= 1u << i; tb = c.type_of[i]; if (c.rl_index[i] == 0) p.rl_veh = i; }
    else ta = c.type_of[i];
  }
  p.env_id_base = c.env_id_base; p.seed = c.seed;
  p.dt = c.dt; p.len = c.types[ta].length; p.noise = c.types[ta].noise;
  p.delay = c.types[ta].delay; p.min_gap = c.types[ta].sim_min_gap;
  p.fail_safe = c.types[ta].fail_safe;
  p.target = c.target_velocity; p.act_lo = c.act_lo; p.act_hi = c.act_hi;
  p.obs_max_length = c.obs_max_length; p.speed_limit = c.speed_limit;
  {  // ||[target]*V||, rewards.py:52-53 (same summation order as numpy's for V < 8 pairwise block)
    T mc2 = T(0);
    const T t2 = c.target_velocity * c.target_velocity;
    for (int i = 0; i < c.n_veh; ++i) mc2 = mc2 + t2;
    p.max_cost = (T)sqrt((double)mc2);
  }
  p.inv_nrl = T(1) / T(c.n_rl > 0 ? c.n_rl : 1);
  p.n_veh_t = T(c.n_veh);
  p.rew_scale = T(4) / (T(20) * T(c.n_veh));
  p.inv_obs_len = c.obs_max_length > T(0) ? T(1) / c.obs_max_length : T(0);
  p.inv_speed_limit = T(1) / c.speed_limit;
  p.pa.load(c.types[ta]); p.pb.load(c.types[tb]);
  load_sim_block(p.ps, c.types[ta]);
  speed_mode_consts(c.types[ta], p.safe_a, p.cap_a, p.floor_a);
  speed_mode_consts(c.types[tb], p.safe_b, p.cap_b, p.floor_b);
  if (c.types[ta].ctrl != FS_CTRL_IDM) {   // SimCarFollowing humans: the simulator drives, no mode
    p.safe_a = 0; p.cap_a = (T)INFINITY; p.floor_a = T(0) - (T)INFINITY;
  }
  p.any_speed_mode = (c.types[ta].ctrl == FS_CTRL_IDM && c.types[ta].speed_mode != 0) ||
                     (c.n_rl > 0 && c.types[tb].speed_mode != 0);
  return p;
}

// four standard normals from one Philox4x32-10 call (Box-Muller on both pairs)
template <typename T>
__device__ __forceinline__ void normal4(unsigned long long seed, long long env_gid, uint32_t c0,
                                        uint32_t c1, T (&z)[4]) {
  uint32_t r[4];
  Philox::gen(c0, c1, (uint32_t)env_gid, (uint32_t)(env_gid >> 32), (uint32_t)seed,
              (uint32_t)(seed >> 32), r);
  if (sizeof(T) == 4) {
    const float k = 2.3283064365386963e-10f;  // 2^-32
    // u in (0,1]: (r + 1) * 2^-32 evaluated as fma so r = 2^32-1 maps to 1
    const float u0 = fminf(fmaf((float)r[0], k, k), 1.0f), u2 = fminf(fmaf((float)r[2], k, k), 1.0f);
    // sqrt(-2 ln u) = sqrt(-2 ln2 * log2 u)
    const float ra = sqrtf(-1.3862943611198906f * __log2f(u0));
    const float rb = sqrtf(-1.3862943611198906f * __log2f(u2));
    float s0, c0f, s1, c1f;
    __sincosf(6.2831853071795865f * ((float)r[1] * k), &s0, &c0f);
    __sincosf(6.2831853071795865f * ((float)r[3] * k), &s1, &c1f);
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

// IDM with the exponent handled at compile time (car_following_models.py:464-482)
template <typename T, bool kDelta4>
__device__ __forceinline__ T idm_fast(T v, T v_lead, T h, const IdmRegs<T>& p) {
  typedef Num<T> N;
  if (N::abs(h) < T(1e-3)) h = T(1e-3);
  T q;
  if (sizeof(T) == 8) q = N::div(mul(v, v - v_lead), p.tsab);
  else q = v * (v - v_lead) * p.inv_tsab;
  const T s_star = add(p.s0, N::max(T(0), add(mul(v, p.Tt), q)));
  const T r = (sizeof(T) == 8) ? N::div(v, p.v0) : v * p.inv_v0;
  T pw;
  if (kDelta4) { const T r2 = mul(r, r); pw = mul(r2, r2); }
  else pw = N::pow(r, p.delta);
  const T sh = N::div(s_star, h);
  return mul(p.a, (T(1) - pw) - mul(sh, sh));
}

// speed_mode_accel (fs_common.cuh) with the type's constants in registers / kernel arguments
template <typename T>
__device__ __forceinline__ T speed_mode_fast(T a, T v, T v_l, T h, bool has_lead,
                                             const IdmRegs<T>& sim, int safe, T cap, T floor) {
  typedef Num<T> N;
  if (safe)
    a = N::min(a, idm_accel<T>(v, v_l, h, has_lead, sim.v0, sim.inv_v0, sim.Tt, sim.a, sim.tsab,
                               sim.inv_tsab, sim.delta, sim.s0));
  return N::max(N::min(a, cap), floor);
}

// fp32 only: the controller's IDM and the simulator model's IDM (the safe speed of the speed
// mode) on the same (v, v_lead, h), sharing the |h| clamp, ONE reciprocal of h and v (v - v_lead);
// both with delta = 4. The values of two idm_fast<float, true> calls up to fp32 rounding (x / h there
// is x * rcp(h) too; (v / v0)^4 is v^4 * (1 / v0^4) here), about 9 instructions instead of 16 for
// the second law and no second MUFU.
__device__ __forceinline__ void idm_pair_f32(float v, float v_lead, float h, const IdmRegs<float>& p,
                                             const IdmRegs<float>& q, float& a_p, float& a_q,
                                             bool has_lead = true) {
  if (fabsf(h) < 1e-3f) h = 1e-3f;
  const float ih = __fdividef(1.0f, h);
  const float vv = v * (v - v_lead);
  // no leader: s* = 0 (car_following_models.py:474-480)
  const float sp = has_lead ? p.s0 + fmaxf(0.0f, v * p.Tt + vv * p.inv_tsab) : 0.0f;
  const float sq = has_lead ? q.s0 + fmaxf(0.0f, v * q.Tt + vv * q.inv_tsab) : 0.0f;
  // (v / v0)^4 of both laws from ONE v^4: the 1 / v0^4 are loop-invariant constants
  const float v2 = v * v, v4 = v2 * v2;
  const float cp = (p.inv_v0 * p.inv_v0) * (p.inv_v0 * p.inv_v0);
  const float cq = (q.inv_v0 * q.inv_v0) * (q.inv_v0 * q.inv_v0);
  const float hp = sp * ih, hq = sq * ih;
  a_p = p.a * (fmaf(-v4, cp, 1.0f) - hp * hp);
  a_q = q.a * (fmaf(-v4, cq, 1.0f) - hq * hq);
}

// sum over the threads of one env group. fp32: integer REDUX in 2^-20 fixed
// point (|v| < 90 m/s per vehicle, V <= 22); fp64: xor-shuffle tree.
template <typename T, int G>
__device__ __forceinline__ T env_sum(T part, unsigned gmask) {
  if (G == 1) return part;
  if (sizeof(T) == 4) {
    const int q = __float2int_rn((float)part * 1048576.0f);
    return (T)((float)__reduce_add_sync(gmask, q) * 9.5367431640625e-07f);
  }
  return group_sum(part, gmask, 32 / G);
}


// kFS: fail-safe and / or speed mode compiled in; kFail: the fail-safe itself (false: speed mode only)
template <typename T, int VPT, int G, bool kExact, bool kNoise, bool kFS, bool kViews, bool kFail = true>
__global__ void __launch_bounds__(128)
ring1_fast_kernel(const __grid_constant__ FastP<T> P, const RingBufs<T> B, const StepArgs<T> A) {
  typedef Num<T> N;
  constexpr int EPW = 32 / G;
  constexpr bool kF32 = sizeof(T) == 4;
  constexpr bool kTimeNoise = VPT <= 4;     // 4 normals of a call spread over 4 steps
  constexpr int NQ = kTimeNoise ? VPT : (VPT + 3) / 4;
  const int lane = threadIdx.x & 31;
  const int g = (G == 1) ? 0 : lane / EPW;
  const int es = (G == 1) ? lane : lane % EPW;
  const int E = P.n_envs, V = P.n_veh;
  const int e_raw = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * EPW + es;
  const bool valid = e_raw < E && (A.mask == nullptr || A.mask[e_raw < E ? e_raw : 0] != 0);
  if (!__any_sync(0xffffffffu, valid)) return;
  const int e = e_raw < E ? e_raw : E - 1;
  unsigned gmask = 0;
#pragma unroll
  for (int q = 0; q < G; ++q) gmask |= 1u << (es + q * EPW);

  const int base = g * VPT;
  const int n_mine = kExact ? VPT : max(0, min(VPT, V - base));
  const int g_next = (base + n_mine >= V) ? 0 : g + 1;
  const int lane_next = g_next * EPW + es;
  const unsigned my_rl = (P.rl_mask >> base) & ((VPT >= 32) ? 0xffffffffu : ((1u << VPT) - 1u));

  // the thread that owns the observed RL vehicle writes obs / reward / done
  const int wbase = P.n_rl > 0 ? (P.rl_veh / VPT) * VPT : 0;
  const int rs = P.n_rl > 0 ? P.rl_veh - wbase : 0;
  const bool is_writer = base == wbase;
  const bool rl_tail = rs == (kExact ? VPT : max(0, min(VPT, V - wbase))) - 1;

  const bool have_act = A.actions != nullptr;
  const bool po = P.env_kind == FS_ENV_WAVE_ATTENUATION_PO;
  const bool accel_env = P.env_kind == FS_ENV_ACCEL;
  const bool ma = P.env_kind == FS_ENV_MA_WAVE_ATTENUATION_PO;
  const int K = A.k_steps;

  // IDM block of my vehicles. One vehicle per thread: picked once (an
  // un-commanded RL vehicle follows the restated simulator default model).
  IdmRegs<T> pm = P.pa;
  T noise_m = P.noise;
  if (VPT == 1 && (my_rl & 1u)) { pm = P.pb; noise_m = T(0); }
  const T len = P.len, dt = P.dt;
  const T C = B.circ[e];
  const T crash_thr = P.collision_min_gap ? P.min_gap : T(0);
  const T inv_obs_len = P.inv_obs_len > T(0) ? P.inv_obs_len : T(1) / C;
  const T obs_len = P.obs_max_length > T(0) ? P.obs_max_length : C;
  const T inv_C = T(1) / C;

  T x[VPT], v[VPT], h[VPT], acc[kViews ? VPT : 1];
  FS_SLOTS(j) {
    const bool ok = kExact || j < n_mine;
    x[j] = ok ? B.pos[(size_t)(base + j) * E + e] : T(0);
    v[j] = ok ? B.vel[(size_t)(base + j) * E + e] : T(0);
    if (kViews) acc[kViews ? j : 0] = T(0);
  }
  // exchange: first vehicle of the next thread = leader of my last vehicle
  T x_nf = (G == 1) ? x[0] : shfl(x[0], lane_next, gmask);
  T v_nf = (G == 1) ? v[0] : shfl(v[0], lane_next, gmask);
  auto update_headways = [&]() {  // k.update: vehicle/traci.py:218-250
    FS_SLOTS(j) {
      T xl;
      if (j + 1 < VPT) xl = (kExact || j + 1 < n_mine) ? x[j + 1 < VPT ? j + 1 : 0] : x_nf;
      else xl = x_nf;
      h[j] = fwd_dist(xl, x[j], C) - len;
    }
  };
  update_headways();

  int tc = B.time_counter[e];
  const int episode = B.episode[e];
  bool done = A.ignore_done ? false : (B.done[e] != 0);
  bool crash = B.crash[e] != 0;
  const long long gid = P.env_id_base + e;
  T reward = B.reward[e];

  // RL actions of my first RL vehicle: blocks of 4 steps, fetched one block ahead
  const int rl_first = my_rl ? __ffs(my_rl) - 1 : -1;
  const bool pf = have_act && rl_first >= 0;
  const T* act_p = A.actions;
  size_t act_stride = 0;
  if (pf) {
    act_p = A.actions + (size_t)__popc(P.rl_mask & ((1u << (base + rl_first)) - 1u)) * E + e;
    act_stride = (size_t)P.act_dim * E;
  }
  T ac0 = T(0), ac1 = T(0), ac2 = T(0), ac3 = T(0);
  T an0 = T(0), an1 = T(0), an2 = T(0), an3 = T(0);
  if (pf) {
    if (0 < K) an0 = __ldg(act_p);
    if (1 < K) an1 = __ldg(act_p + act_stride);
    if (2 < K) an2 = __ldg(act_p + 2 * act_stride);
    if (3 < K) an3 = __ldg(act_p + 3 * act_stride);
  }

  float st_rew = 0.f, st_v = 0.f, st_v2 = 0.f;
  int st_adv = 0, st_cr = 0;
  T zn[kNoise ? NQ : 1][4];
  int zleft = 0;

  auto po_obs = [&](T& o0, T& o1, T& o2) {  // WaveAttenuationPOEnv.get_state, wave_attenuation.py:248-269
    T xr, vr, xl, vl;
    if (rl_tail) {
      xr = kExact ? x[VPT - 1] : pick(x, rs); vr = kExact ? v[VPT - 1] : pick(v, rs);
      xl = x_nf; vl = v_nf;
    } else {
      xr = pick(x, rs); vr = pick(v, rs);
      xl = pick(x, rs + 1 < VPT ? rs + 1 : 0); vl = pick(v, rs + 1 < VPT ? rs + 1 : 0);
    }
    o0 = N::div(vr, T(15)); o1 = N::div(vl - vr, T(15));
    o2 = kF32 ? fwd_dist(xl, xr, C) * inv_obs_len : N::div(fwd_dist(xl, xr, C), obs_len);
  };
  // MultiAgentWaveAttenuationPOEnv.get_state (multiagent/ring/wave_attenuation.py:195-213):
  // every RL vehicle: own speed, speed difference to its leader, bumper-to-bumper headway
  auto ma_obs = [&](T* o) {
    FS_SLOTS(j) if ((kExact || j < n_mine) && ((my_rl >> j) & 1u)) {
      const int r = __popc(P.rl_mask & ((1u << (base + j)) - 1u));
      T vl;
      if (j + 1 < VPT) vl = (kExact || j + 1 < n_mine) ? v[j + 1 < VPT ? j + 1 : 0] : v_nf;
      else vl = v_nf;
      o[3 * r] = N::div(v[j], T(15));
      o[3 * r + 1] = N::div(vl - v[j], T(15));
      o[3 * r + 2] = kF32 ? h[j] * inv_obs_len : N::div(h[j], obs_len);
    }
  };
  auto full_obs = [&](T* o) {  // AccelEnv / WaveAttenuationEnv.get_state, accel.py:116-123
    FS_SLOTS(j) if (kExact || j < n_mine) {
      o[base + j] = kF32 ? v[j] * P.inv_speed_limit : N::div(v[j], P.speed_limit);
      o[V + base + j] = kF32 ? x[j] * inv_C : N::div(x[j], C);
    }
  };

  // per-step output cursors (null = not requested / not mine)
  const bool w_obs = valid && A.obs_seq != nullptr && (!po || is_writer);
  const bool w_rew = valid && is_writer && A.reward_seq != nullptr;
  const bool w_done = valid && is_writer && A.done_seq != nullptr;
  T* obs_k = A.obs_seq + (w_obs ? (size_t)e * P.obs_dim : 0);
  T* rew_k = A.reward_seq + (w_rew ? e : 0);
  uint8_t* done_k = A.done_seq + (w_done ? e : 0);
  const size_t obs_stride = (size_t)E * P.obs_dim;
  const bool st_on = A.accumulate_stats && valid;
  // a second law is evaluated next to the controller's (see the step body)
  const bool two_laws = (kFS && P.any_speed_mode) || P.n_rl > 0;

  for (int k = 0; k < K; ++k) {
    // ---- RL action of this step (block of 4 promoted every 4th step; next block requested)
    T a_cur = T(0);
    if (have_act) {
      if ((k & 3) == 0) {
        ac0 = an0; ac1 = an1; ac2 = an2; ac3 = an3;
        if (pf) {
          const T* q = act_p + (size_t)(k + 4) * act_stride;
          if (k + 4 < K) an0 = __ldg(q);
          if (k + 5 < K) an1 = __ldg(q + act_stride);
          if (k + 6 < K) an2 = __ldg(q + 2 * act_stride);
          if (k + 7 < K) an3 = __ldg(q + 3 * act_stride);
        }
      }
      a_cur = ac0; ac0 = ac1; ac1 = ac2; ac2 = ac3;
    }
    if (!done) {
      // ---- noise of this step (base_controller.py:109-110)
      T zs[kNoise ? VPT : 1];
      if (kNoise) {
        if (kTimeNoise) {
          if (zleft == 0) {
#pragma unroll
            for (int q = 0; q < NQ; ++q)
              normal4<T>(P.seed, gid, (uint32_t)(tc >> 2), (uint32_t)(episode * 256 + base + q), zn[q]);
            int sk = tc & 3;   // launch started inside a 4-step block: drop the used components
            zleft = 4 - sk;
            while (sk-- > 0) {
#pragma unroll
              for (int q = 0; q < NQ; ++q) { zn[q][0] = zn[q][1]; zn[q][1] = zn[q][2]; zn[q][2] = zn[q][3]; }
            }
          }
          --zleft;
#pragma unroll
          for (int q = 0; q < NQ; ++q) {
            zs[kNoise ? q : 0] = zn[q][0];
            zn[q][0] = zn[q][1]; zn[q][1] = zn[q][2]; zn[q][2] = zn[q][3];
          }
        } else {
#pragma unroll
          for (int q = 0; q < NQ; ++q)
            normal4<T>(P.seed, gid, (uint32_t)tc,
                       0x80000000u | (uint32_t)(episode * 256 + base + 4 * q), zn[q]);
          FS_SLOTS(j) zs[kNoise ? j : 0] = zn[kNoise ? (j >> 2) : 0][j & 3];
        }
      }
      // ---- controllers -> apply_acceleration -> Euler move (all read the old state)
      T act_abs_sum = T(0);
      const T v_lead_last = v_nf;  // leader speed (old state) of my last vehicle
      FS_SLOTS(j) {
        const bool ok = kExact || j < n_mine;
        const bool rlj = (my_rl >> j) & 1u;
        const T vi = v[j], hi = h[j];
        T vl;
        if (j + 1 < VPT) vl = (kExact || j + 1 < n_mine) ? v[j + 1 < VPT ? j + 1 : 0] : v_lead_last;
        else vl = v_lead_last;
        // fp32: ONE evaluation of two laws per vehicle - the controller's and law q, which is
        // the simulator model behind a human's speed mode (P.ps) or, in an RL vehicle's slot,
        // the RL type's own model (P.pb: it drives the vehicle while it has no command and is
        // the safe speed of a commanded one). fp64 keeps the separate evaluations, whose
        // operation order is the reference's.
        T a, a_q = T(0), a_rl;
        if (kF32 && two_laws) {
          IdmRegs<float> q;
          const IdmRegs<float>& qa = reinterpret_cast<const IdmRegs<float>&>(P.ps);
          const IdmRegs<float>& qb = reinterpret_cast<const IdmRegs<float>&>(P.pb);
          q.s0 = rlj ? qb.s0 : qa.s0; q.Tt = rlj ? qb.Tt : qa.Tt; q.inv_tsab = rlj ? qb.inv_tsab : qa.inv_tsab;
          q.inv_v0 = rlj ? qb.inv_v0 : qa.inv_v0; q.a = rlj ? qb.a : qa.a;
          float af, aqf;
          idm_pair_f32((float)vi, (float)vl, (float)hi, reinterpret_cast<const IdmRegs<float>&>(pm), q, af, aqf);
          a = (T)af; a_q = (T)aqf;
          a_rl = a_q;
        } else {
          a = idm_fast<T, true>(vi, vl, hi, pm);
          a_rl = a;
          if (VPT > 1 && rlj && !have_act) a_rl = idm_fast<T, true>(vi, vl, hi, P.pb);
        }
        if (kNoise) a = add(a, mul(VPT == 1 ? noise_m : P.noise, zs[kNoise ? j : 0]));
        if (kFS && kFail) {  // base_controller.py:113-116
          if (P.fail_safe == FS_FAILSAFE_INSTANTANEOUS)
            a = failsafe_instantaneous<T>(a, vi, hi, true, V, dt);
          else if (P.fail_safe == FS_FAILSAFE_SAFE_VELOCITY)
            a = failsafe_safe_velocity<T>(a, vi, vl, hi, V, dt, P.delay);
        }
        T a_h = a;          // what moves a human (after SUMO's speed mode, restated)
        if (kFS && P.any_speed_mode) {
          if (kF32) {
            if (P.safe_a) a_h = N::min(a_h, a_q);
            a_h = N::max(N::min(a_h, P.cap_a), P.floor_a);
          } else {
            a_h = speed_mode_fast<T>(a, vi, vl, hi, true, P.ps, P.safe_a, P.cap_a, P.floor_a);
          }
        }
        T a_rl_eff = a_rl;
        if (have_act) {
          // Env.apply_rl_actions: Box clip, then apply_acceleration (base.py:599-615)
          T raw = a_cur;
          if (VPT > 1 && rlj && j != rl_first)
            raw = A.actions[((size_t)k * P.act_dim +
                             __popc(P.rl_mask & ((1u << (base + j)) - 1u))) * E + e];
          a_rl = N::min(N::max(raw, P.act_lo), P.act_hi);
          const T aa = N::abs(P.clip_actions ? a_rl : raw);
          act_abs_sum = rlj ? add(act_abs_sum, aa) : act_abs_sum;
          a_rl_eff = a_rl;
          if (kFS && P.any_speed_mode) {
            if (kF32) {
              if (P.safe_b) a_rl_eff = N::min(a_rl_eff, a_q);
              a_rl_eff = N::max(N::min(a_rl_eff, P.cap_b), P.floor_b);
            } else {
              a_rl_eff = speed_mode_fast<T>(a_rl, vi, vl, hi, true, P.pb, P.safe_b, P.cap_b, P.floor_b);
            }
          }
        }
        a = rlj ? a_rl : a;                       // commanded (the `accel` view)
        const T a_mv = rlj ? a_rl_eff : a_h;
        const T vn = N::max(add(vi, mul(a_mv, dt)), T(0));   // vehicle/traci.py:961-962
        const T xn = wrap_fwd(add(x[j], mul(vn, dt)), C);  // simulator Euler move
        if (ok) { v[j] = vn; x[j] = xn; if (kViews) acc[kViews ? j : 0] = a; }
      }
      if (G > 1) { x_nf = shfl(x[0], lane_next, gmask); v_nf = shfl(v[0], lane_next, gmask); }
      else { x_nf = x[0]; v_nf = v[0]; }
      T h_old[VPT];
      FS_SLOTS(j) h_old[j] = h[j];
      update_headways();
      bool cr = false;
      T sv = T(0), sq = T(0);
      float s2 = 0.f;
      const T half_lap = T(0.5) * C;
      FS_SLOTS(j) if (kExact || j < n_mine) {
        cr |= h[j] < crash_thr;
        cr |= (h[j] - h_old[j]) > half_lap;   // went through its leader inside the step
        sv = add(sv, v[j]);
        if (accel_env) { const T d = v[j] - P.target; sq = add(sq, mul(d, d)); }
        if (st_on) s2 = fmaf((float)v[j], (float)v[j], s2);
      }
      if (st_on) { st_v += (float)sv; st_v2 += s2; st_adv += 1; }
      if (G > 1) cr = __any_sync(gmask, cr);
      crash = cr;
      tc += 1;
      done = crash || tc >= P.t_end;
      if (accel_env) {  // rewards.desired_velocity, flow/core/rewards.py:6-59
        if (G > 1) sq = group_sum(sq, gmask, EPW);
        const T cost = N::sqrt(sq);
        const T eps = T(1.1920928955078125e-07);
        reward = crash ? T(0) : N::div(N::max(P.max_cost - cost, T(0)), add(P.max_cost, eps));
      } else if (have_act) {  // WaveAttenuationEnv.compute_reward, wave_attenuation.py:113-139
        sv = env_sum<T, G>(sv, gmask);
        if (G > 1 && P.n_rl > 1) act_abs_sum = group_sum(act_abs_sum, gmask, EPW);
        T r = kF32 ? sv * P.rew_scale : N::div(mul(T(4), N::div(sv, P.n_veh_t)), T(20));
        const T mean_act = mul(act_abs_sum, P.inv_nrl);
        if (mean_act > T(0)) r = add(r, mul(T(4), T(0) - mean_act));
        reward = crash ? T(0) : r;
      } else {
        reward = T(0);
      }
      if (st_on && is_writer) { st_rew += (float)reward; st_cr += crash ? 1 : 0; }
    } else {
      reward = T(0);
    }
    // ---- per-step outputs
    if (w_obs) {
      if (po) { T o0, o1, o2; po_obs(o0, o1, o2); obs_k[0] = o0; obs_k[1] = o1; obs_k[2] = o2; }
      else if (ma) ma_obs(obs_k);
      else full_obs(obs_k);
      obs_k += obs_stride;
    }
    if (w_rew) { *rew_k = reward; rew_k += E; }
    if (w_done) { *done_k = done ? 1 : 0; done_k += E; }
  }

  if (valid) {
    T* ol = B.obs + (size_t)e * P.obs_dim;
    if (po) {
      if (is_writer) { T o0, o1, o2; po_obs(o0, o1, o2); ol[0] = o0; ol[1] = o1; ol[2] = o2; }
    } else if (ma) ma_obs(ol);
    else full_obs(ol);
    FS_SLOTS(j) if (kExact || j < n_mine) {
      B.pos[(size_t)(base + j) * E + e] = x[j];
      B.vel[(size_t)(base + j) * E + e] = v[j];
    }
    if (is_writer) {
      B.time_counter[e] = tc;
      B.done[e] = done ? 1 : 0;
      B.crash[e] = crash ? 1 : 0;
      B.reward[e] = reward;
    }
    if (kViews && A.write_views) {
      FS_SLOTS(j) if (kExact || j < n_mine) {
        const int i = base + j;
        const size_t o = (size_t)i * E + e;
        B.v_headway[o] = h[j];
        B.v_leader[o] = (i + 1 < V) ? i + 1 : 0;
        B.v_follower[o] = (i > 0) ? i - 1 : V - 1;
        if (K > 0) B.v_accel[o] = acc[kViews ? j : 0];
      }
    }
  }
  if (A.accumulate_stats) {
    // stats[]: sum reward, sum v, sum v^2, vehicle-steps, env-steps, crashes
    double vals[6] = {(double)st_rew, (double)st_v, (double)st_v2, (double)(st_adv * n_mine),
                      is_writer ? (double)st_adv : 0.0, (double)st_cr};
    substep_stats(A.accumulate_stats, vals);
#pragma unroll
    for (int q = 0; q < 6; ++q) {
      double s = vals[q];
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (lane == 0 && s != 0.0) atomicAdd(&stat_stripe(B.stats)[q], s);
    }
  }
}

// ---- host-side dispatch (one translation unit per dtype instantiates these)
template <typename T, int VPT, int G, bool kExact>
static cudaError_t launch_fast_shape(int variant, int grid, int block, cudaStream_t st,
                                     const FastP<T>& c, const RingBufs<T>& b, const StepArgs<T>& a) {
  switch (variant) {
    case 0: ring1_fast_kernel<T, VPT, G, kExact, false, false, false><<<grid, block, 0, st>>>(c, b, a); break;
    case 1: ring1_fast_kernel<T, VPT, G, kExact, true, false, false><<<grid, block, 0, st>>>(c, b, a); break;
    case 2: ring1_fast_kernel<T, VPT, G, kExact, true, true, false><<<grid, block, 0, st>>>(c, b, a); break;
    case 4: ring1_fast_kernel<T, VPT, G, kExact, true, true, false, false><<<grid, block, 0, st>>>(c, b, a); break;
    default: ring1_fast_kernel<T, VPT, G, kExact, true, true, true><<<grid, block, 0, st>>>(c, b, a); break;
  }
  return cudaGetLastError();
}

// shape = threads per env (1, 2, 4, 8, 16, 32) for V = 22; variant: 0 plain,
// 1 noise, 2 noise + fail-safe + speed mode, 3 = 2 + views (accel / headway / leader / follower),
// 4 = noise + speed mode without the fail-safe code (what the reference's default mode 25 needs)
template <typename T>
static cudaError_t launch_fast22(int shape, int variant, int n_envs, cudaStream_t st,
                                 const FastP<T>& c, const RingBufs<T>& b, const StepArgs<T>& a) {
  const int epw = 32 / shape;
  const int warps = (n_envs + epw - 1) / epw;
  // spread small batches over all 148 SMs with one-warp blocks; fatter blocks
  // once every SM has several warps anyway
  const int wpb = warps <= 148 * 8 ? 1 : (warps <= 148 * 32 ? 2 : 4);
  const int grid = (warps + wpb - 1) / wpb, block = wpb * 32;
  switch (shape) {
    case 1: return launch_fast_shape<T, 22, 1, true>(variant, grid, block, st, c, b, a);
    case 2: return launch_fast_shape<T, 11, 2, true>(variant, grid, block, st, c, b, a);
    case 4: return launch_fast_shape<T, 6, 4, false>(variant, grid, block, st, c, b, a);
    case 8: return launch_fast_shape<T, 3, 8, false>(variant, grid, block, st, c, b, a);
    case 16: return launch_fast_shape<T, 2, 16, false>(variant, grid, block, st, c, b, a);
    default: return launch_fast_shape<T, 1, 32, false>(variant, grid, block, st, c, b, a);
  }
}

}  // namespace fs
