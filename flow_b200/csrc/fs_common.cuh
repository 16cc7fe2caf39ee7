// fs_common.cuh — device-side parameter tables, math helpers and the
// car-following / fail-safe laws shared by every step kernel.
//
// Reference (parthjaggi/flow) lines each law restates are cited at the
// function. T is float (FS_F32) or double (FS_F64). In double every formula
// keeps the reference's operation order so results agree to ~1 ulp; in float
// divisions become reciprocal multiplies / fast intrinsics (tolerance 1e-4).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/flowstep.h"

#define FS_MAX_VEH 256  // vehicles (ring) or slots (highway) per env
// Rollout statistics are accumulated with fire-and-forget double atomics. All
// warps of a launch hitting the SAME 6 addresses serialise in one L2 slice
// (~25 us for 4096 warps), so every warp adds into one of FS_STAT_STRIPES
// copies of the 8-double vector; fs_stats() sums the copies on the host.
#define FS_STAT_STRIPES 64

namespace fs {

template <typename T>
struct TypeP {
  int ctrl, fail_safe, lc, speed_mode;   // speed_mode: SUMO bits 0-2, see speed_mode_accel
  T p[FS_NPARAM];
  T noise, delay, max_accel, max_decel, length;
  T sim_tau, sim_min_gap, sim_v0;  // sim_v0 = min(maxSpeed, speed_limit)
  T two_sqrt_ab;      // IDM: 2*sqrt(a*b)
  T sim_two_sqrt_ab;  // simulator-default IDM: 2*sqrt(accel*decel)
  T inv_v0, sim_inv_v0, inv_two_sqrt_ab, sim_inv_two_sqrt_ab;  // f32 fast path
  // effective IDM block: the law's own parameters for FS_CTRL_IDM, the
  // simulator-default ones for FS_CTRL_SIM / FS_CTRL_RL (used by the fast kernel)
  T e_v0, e_inv_v0, e_T, e_a, e_tsab, e_inv_tsab, e_delta, e_s0;
};

template <typename T>
struct Cfg {
  int n_envs, n_veh, n_lanes, env_kind;
  int n_types, warmup_steps, horizon, sims_per_step;
  int clip_actions, lane_change_duration, act_dim, obs_dim;
  int n_rl, collision_min_gap, any_noise, any_follower;
  long long env_id_base;
  unsigned long long seed;
  T dt, inv_dt, junction_len, speed_limit, inv_speed_limit;
  T x0, bunching, init_min_gap;
  T target_velocity, act_lo, act_hi, obs_max_length;
  T length_lo, length_hi;
  int lanes_distribution;
  // FS_TOPO_MERGE: arm lengths up to the merge point (highway / ramp), the common edge, the
  // cross-arm look-ahead zone, the edge lengths they were built from (flowstep.cu fill_cfg)
  T mg_LH, mg_LM, mg_post, mg_zone, mg_pre, mg_ramp;
  // FS_TOPO_FIGURE8: loop length, the two junction boxes, junction length, look-ahead zone and
  // the starts of the 12 segments in driving order (even index = internal), fs::f8_geometry
  T f8_L, f8_cA, f8_cB, f8_J, f8_zone;
  T f8_start[12];
  int16_t evaluate;   // EnvParams.evaluate (AccelEnv reward = mean speed)
  int16_t ballistic;  // SumoParams.use_ballistic: x += (v_old + v_new)/2 dt instead of v_new dt
  TypeP<T> types[FS_MAX_TYPES];
  uint8_t type_of[FS_MAX_VEH];
  int16_t rl_index[FS_MAX_VEH];  // vehicle -> action slot, -1 if not RL
  int16_t pi_index[FS_MAX_VEH];  // vehicle -> PISaturation history slot, -1 if none
};

// FigureEightNetwork laid out as ONE loop (flow/networks/figure_eight.py:66-73, 217-263), in
// double and in the operation order of oracle/figure8.py Figure8Geometry. Segments in driving
// order: ':bottom' bottom ':center_1' top ':top' upper_ring ':right' right ':center_0' left
// ':left' lower_ring (even index = internal edge).
// Python's built-in sum() over floats (CPython >= 3.12: Neumaier-compensated; the reference's
// _get_start_pos_util and non_internal_length call it): running sum t, compensation c, t + c.
struct PySum {
  double t = 0.0, c = 0.0;
  __host__ __device__ void push(double x) {
    const double n = t + x;
    c += fabs(t) >= fabs(x) ? (t - n) + x : (x - n) + t;
    t = n;
  }
  __host__ __device__ double value() const { return c != 0.0 ? t + c : t; }
};

struct F8Geo {
  double start[12], len[12];
  double L, cA, cB, J, non_internal;
  double aimsun;   // start of the network's 'right_to_left' entry (see fig8_reset_kernel)
};
inline F8Geo f8_geometry(double r) {
  F8Geo g;
  const double rel = r * 3.141592653589793 / 2.0, il = 2 * r, jl = 2.9 + 3.3 * 1, sp = 0.28,
               re3 = 3 * rel;
  const double st[12] = {0.0, sp, il / 2 + sp, il / 2 + jl + sp, il + jl + sp, il + jl + 2 * sp,
                         il + re3 + jl + 2 * sp, il + re3 + jl + 3 * sp,
                         3.0 / 2 * il + re3 + jl + 3 * sp, 3.0 / 2 * il + re3 + 2 * jl + 3 * sp,
                         2 * il + re3 + 2 * jl + 3 * sp, 2 * il + re3 + 2 * jl + 4 * sp};
  // edge LENGTHS of the two ring edges are 3 r pi / 2 (specify_edges), their starts use
  // 3 * (r pi / 2) (specify_edge_starts): an ulp apart, both kept as the reference has them
  const double rl = 3 * r * 3.141592653589793 / 2.0;
  const double ln[12] = {sp, r, jl, r, sp, rl, sp, r, jl, r, sp, rl};
  for (int k = 0; k < 12; ++k) { g.start[k] = st[k]; g.len[k] = ln[k]; }
  g.L = 2 * il + 2 * re3 + 2 * jl + 4 * sp;
  g.cA = il / 2 + sp;
  g.cB = 3.0 / 2 * il + re3 + jl + 3 * sp;
  g.J = jl;
  g.aimsun = +jl + 3 * sp;
  // get_edge_list order of the harness = specify_edges order: bottom, top, right, left,
  // upper_ring, lower_ring, summed like Python's sum()
  PySum ni;
  const int order[6] = {1, 3, 7, 9, 5, 11};
  for (int q = 0; q < 6; ++q) ni.push(ln[order[q]]);
  g.non_internal = ni.value();
  return g;
}

// ---------------------------------------------------------------- helpers
__device__ __forceinline__ double* stat_stripe(double* stats) {
  const unsigned w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  return stats + (w & (FS_STAT_STRIPES - 1)) * 8;
}

// StepArgs.accumulate_stats == 2: an INTERMEDIATE simulation sub-step of an Env.step
// (EnvParams.sims_per_step > 1). The reference computes the reward and counts the Env.step once,
// after the last sub-step or after the one that collided (flow/envs/base.py:325-410): speeds,
// vehicle-steps and crashes are kept, the reward is not summed and the env-step is counted only
// where the env collided in this sub-step (vals: reward, v, v^2, vehicle-steps, env-steps, crashes).
__device__ __forceinline__ void substep_stats(int mode, double* vals) {
  if (mode == 2) { vals[0] = 0.0; vals[4] = vals[5]; }
}

template <typename T> struct Num;
template <> struct Num<float> {
  static __device__ __forceinline__ float div(float a, float b) { return __fdividef(a, b); }
  static __device__ __forceinline__ float sqrt(float a) { return sqrtf(a); }
  static __device__ __forceinline__ float pow(float a, float b) { return __powf(a, b); }
  static __device__ __forceinline__ float cospi(float a) { return ::cospif(a); }
  static __device__ __forceinline__ float abs(float a) { return fabsf(a); }
  static __device__ __forceinline__ float max(float a, float b) { return fmaxf(a, b); }
  static __device__ __forceinline__ float min(float a, float b) { return fminf(a, b); }
  static __device__ __forceinline__ float nan() { return __int_as_float(0x7fc00000); }
};
template <> struct Num<double> {
  static __device__ __forceinline__ double div(double a, double b) { return a / b; }
  static __device__ __forceinline__ double sqrt(double a) { return ::sqrt(a); }
  static __device__ __forceinline__ double pow(double a, double b) { return ::pow(a, b); }
  static __device__ __forceinline__ double cospi(double a) { return ::cos(3.141592653589793 * a); }
  static __device__ __forceinline__ double abs(double a) { return fabs(a); }
  static __device__ __forceinline__ double max(double a, double b) { return fmax(a, b); }
  static __device__ __forceinline__ double min(double a, double b) { return fmin(a, b); }
  static __device__ __forceinline__ double nan() { return __longlong_as_double(0x7ff8000000000000LL); }
};

// No fused multiply-add in double: Python evaluates a*b+c with two roundings,
// and the 1e-9 parity bar is easier to hold when we do the same.
__device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float mul(float a, float b) { return a * b; }
__device__ __forceinline__ float add(float a, float b) { return a + b; }

// wrap x into [0, C) given x in [0, 2C): equals Python's x % C bit for bit
template <typename T>
__device__ __forceinline__ T wrap_fwd(T x, T C) { return x >= C ? x - C : x; }
// (a - b) mod C for a, b in [0, C): equals Python's (a - b) % C bit for bit
template <typename T>
__device__ __forceinline__ T fwd_dist(T a, T b, T C) {
  T d = a - b;
  return d < T(0) ? d + C : d;
}

// The simulator move (restated SUMO): Euler x += v_new dt by default,
// the ballistic update x += (v_old + v_new)/2 dt with SumoParams(use_ballistic=True)
// (flow/core/kernel/simulation/traci.py:95-96 only hands SUMO the switch).
template <typename T>
__device__ __forceinline__ T move_dist(T v_old, T v_new, T dt, int ballistic) {
  return ballistic ? mul(mul(T(0.5), add(v_old, v_new)), dt) : mul(v_new, dt);
}

// ------------------------------------------------------------ Philox4x32-10
struct Philox {
  static __device__ __forceinline__ void round(uint32_t (&c)[4], uint32_t (&k)[2]) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    uint32_t hi0 = __umulhi(M0, c[0]), lo0 = M0 * c[0];
    uint32_t hi1 = __umulhi(M1, c[2]), lo1 = M1 * c[2];
    uint32_t n0 = hi1 ^ c[1] ^ k[0], n2 = hi0 ^ c[3] ^ k[1];
    c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
    k[0] += 0x9E3779B9u; k[1] += 0xBB67AE85u;
  }
  // counter = (c0,c1,c2,c3), key = (k0,k1) -> 4 x u32
  static __device__ __forceinline__ void gen(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                             uint32_t k0, uint32_t k1, uint32_t (&out)[4]) {
    uint32_t c[4] = {c0, c1, c2, c3}, k[2] = {k0, k1};
#pragma unroll
    for (int i = 0; i < 10; ++i) round(c, k);
    out[0] = c[0]; out[1] = c[1]; out[2] = c[2]; out[3] = c[3];
  }
};

// Standard normal for (env, vehicle, step): Box-Muller on two Philox words.
// Keyed by the GLOBAL env id so results do not depend on the GPU count.
template <typename T>
__device__ __forceinline__ T normal_sample(unsigned long long seed, long long env_gid,
                                           int veh, int step) {
  uint32_t r[4];
  Philox::gen((uint32_t)step, (uint32_t)veh, (uint32_t)env_gid, (uint32_t)(env_gid >> 32),
              (uint32_t)seed, (uint32_t)(seed >> 32), r);
  // u1 in (0,1], u2 in [0,1)
  double u1 = ((double)r[0] + 1.0) * (1.0 / 4294967296.0);
  double u2 = (double)r[1] * (1.0 / 4294967296.0);
  if (sizeof(T) == 4) {
    float rad = sqrtf(-2.0f * __logf((float)u1));
    return (T)(rad * cospif(2.0f * (float)u2));
  }
  return (T)(::sqrt(-2.0 * ::log(u1)) * ::cos(6.283185307179586 * u2));
}

// Four standard normals from ONE Philox call (Box-Muller on both word pairs), all in T's
// own precision: what the warp-per-env kernels use, one call per thread and step.
template <typename T>
__device__ __forceinline__ void normal4(unsigned long long seed, long long env_gid, int idx,
                                        int step, T* z) {
  uint32_t r[4];
  Philox::gen((uint32_t)step, (uint32_t)idx, (uint32_t)env_gid, (uint32_t)(env_gid >> 32),
              (uint32_t)seed, (uint32_t)(seed >> 32), r);
  if (sizeof(T) == 4) {
    const float k = 2.3283064365386963e-10f;  // 2^-32; u in (0, 1]
    const float u0 = fminf(fmaf((float)r[0], k, k), 1.0f), u2 = fminf(fmaf((float)r[2], k, k), 1.0f);
    const float ra = sqrtf(-1.3862943611198906f * __log2f(u0));   // sqrt(-2 ln u)
    const float rb = sqrtf(-1.3862943611198906f * __log2f(u2));
    float s0, c0, s1, c1;
    const float k2pi = 1.4629180792671596e-09f;   // 2 pi 2^-32
    __sincosf((float)r[1] * k2pi, &s0, &c0);
    __sincosf((float)r[3] * k2pi, &s1, &c1);
    z[0] = (T)(ra * c0); z[1] = (T)(ra * s0); z[2] = (T)(rb * c1); z[3] = (T)(rb * s1);
  } else {
    const double k = 1.0 / 4294967296.0;
    const double ra = ::sqrt(-2.0 * ::log(((double)r[0] + 1.0) * k));
    const double rb = ::sqrt(-2.0 * ::log(((double)r[2] + 1.0) * k));
    double s0, c0, s1, c1;
    sincos(6.283185307179586 * ((double)r[1] * k), &s0, &c0);
    sincos(6.283185307179586 * ((double)r[3] * k), &s1, &c1);
    z[0] = (T)(ra * c0); z[1] = (T)(ra * s0); z[2] = (T)(rb * c1); z[3] = (T)(rb * s1);
  }
}

// ------------------------------------------------------------ controller laws
// IDM, flow/controllers/car_following_models.py:464-482
template <typename T>
__device__ __forceinline__ T idm_accel(T v, T v_lead, T h, bool has_leader, T v0, T inv_v0,
                                       T Tt, T a, T two_sqrt_ab, T inv_two_sqrt_ab, T delta,
                                       T s0) {
  typedef Num<T> N;
  if (N::abs(h) < T(1e-3)) h = T(1e-3);
  T s_star = T(0);
  if (has_leader) {
    T q;
    if (sizeof(T) == 8) q = N::div(mul(v, v - v_lead), two_sqrt_ab);
    else q = v * (v - v_lead) * inv_two_sqrt_ab;
    s_star = add(s0, N::max(T(0), add(mul(v, Tt), q)));
  }
  T r = (sizeof(T) == 8) ? N::div(v, v0) : v * inv_v0;
  T pw;
  if (delta == T(4)) { T r2 = mul(r, r); pw = mul(r2, r2); }
  else pw = N::pow(r, delta);
  T sh = N::div(s_star, h);
  return mul(a, (T(1) - pw) - mul(sh, sh));
}

// get_accel for vehicle with type tp. v: own speed, lead/foll: indices validity,
// v_l: leader speed (-1001 if none), h: headway, v_f / h_f: follower speed and
// follower's headway (-1001 if none), mean_v: mean speed over the env.
// st0/st1: controller state words (LAC: a ; PISaturation: v_cmd, running sum).
template <typename T>
__device__ __forceinline__ T ctrl_accel(const TypeP<T>& tp, int ctrl, T v, bool has_lead, T v_l,
                                        T h, bool has_foll, T v_f, T h_f, T mean_v, T dt,
                                        T inv_dt, T& st0) {
  typedef Num<T> N;
  const T* p = tp.p;
  switch (ctrl) {
    case FS_CTRL_IDM:
      return idm_accel<T>(v, v_l, h, has_lead, p[0], tp.inv_v0, p[1], p[2], tp.two_sqrt_ab,
                          tp.inv_two_sqrt_ab, p[4], p[5]);
    case FS_CTRL_SIM:
    case FS_CTRL_RL:
      // restated simulator default model: IDM fed with the
      // type's SumoCarFollowingParams, flow/core/params.py:842-846
      return idm_accel<T>(v, v_l, h, has_lead, tp.sim_v0, tp.sim_inv_v0, tp.sim_tau,
                          tp.max_accel, tp.sim_two_sqrt_ab, tp.sim_inv_two_sqrt_ab, T(4),
                          tp.sim_min_gap);
    case FS_CTRL_OVM: {  // car_following_models.py:308-328
      if (!has_lead) return tp.max_accel;
      T h_dot = v_l - v, v_h;
      if (h <= p[2]) v_h = T(0);
      else if (h < p[3]) {
        T c;
        if (sizeof(T) == 8)  // reference order: cos(pi * (h - h_st) / (h_go - h_st))
          c = (T)::cos((double)N::div(mul(T(3.141592653589793), h - p[2]), p[3] - p[2]));
        else
          c = N::cospi(N::div(h - p[2], p[3] - p[2]));
        v_h = mul(p[4] / T(2), T(1) - c);
      } else v_h = p[4];
      return add(mul(p[0], v_h - v), mul(p[1], h_dot));
    }
    case FS_CTRL_LINEAR_OVM: {  // car_following_models.py:383-397
      const T alpha = T(1.689);
      T v_h;
      if (h < p[2]) v_h = T(0);
      else if (h <= add(p[2], N::div(p[0], alpha))) v_h = mul(alpha, h - p[2]);
      else v_h = p[0];
      return N::div(v_h - v, p[1]);
    }
    case FS_CTRL_CFM:  // car_following_models.py:76-88
      if (!has_lead) return tp.max_accel;
      return add(add(mul(p[0], h - p[3]), mul(p[1], v_l - v)), mul(p[2], p[4] - v));
    case FS_CTRL_BCM: {  // car_following_models.py:152-176
      if (!has_lead) return tp.max_accel;
      return add(add(mul(p[0], h - h_f), mul(p[1], (v_l - v) - (v - v_f))),
                 mul(p[2], p[4] - v));
    }
    case FS_CTRL_LAC: {  // car_following_models.py:232-245 ; st0 = a
      T ex = (h - tp.length) - mul(p[2], v);
      T ev = v_l - v;
      T u = add(mul(p[0], ex), mul(p[1], ev));
      T a_dot = add(-N::div(st0, p[3]), N::div(u, p[3]));
      st0 = add(mul(a_dot, dt), st0);
      return st0;
    }
    case FS_CTRL_GIPPS: {  // car_following_models.py:567-582
      T v0 = p[0], acc = p[1], b = p[2], b_l = p[3], s0 = p[4], tau = p[5];
      T r = N::div(v, v0);
      T v_acc = add(v, mul(mul(mul(mul(T(2.5), acc), tau), T(1) - r),
                           N::sqrt(add(T(0.025), r))));
      T inner = (mul(T(2), h - s0) - mul(tau, v)) - N::div(mul(v_l, v_l), b_l);
      T v_safe = add(mul(tau, b),
                     N::sqrt(mul(mul(tau, tau), mul(b, b)) - mul(b, inner)));
      T v_next = N::min(N::min(v_acc, v_safe), v0);
      return N::div(v_next - v, dt);
    }
    case FS_CTRL_FOLLOWER_STOPPER:
    case FS_CTRL_NONLOCAL_FOLLOWER_STOPPER: {  // velocity_controllers.py:75-116, :122-164
      T v_des = (ctrl == FS_CTRL_NONLOCAL_FOLLOWER_STOPPER) ? mean_v : p[0];
      T v_cmd;
      if (!has_lead) v_cmd = v_des;
      else {
        T dvm = N::min(v_l - v, T(0));
        T dv2 = mul(dvm, dvm);
        T dx1 = add(T(4.5), mul(T(1) / (T(2) * T(1.5)), dv2));
        T dx2 = add(T(5.25), mul(T(1) / (T(2) * T(1.0)), dv2));
        T dx3 = add(T(6.0), mul(T(1) / (T(2) * T(0.5)), dv2));
        T vv = N::min(N::max(v_l, T(0)), v_des);
        if (h <= dx1) v_cmd = T(0);
        else if (h <= dx2) v_cmd = N::div(mul(vv, h - dx1), dx2 - dx1);
        else if (h <= dx3) v_cmd = add(vv, N::div(mul(v_des - v, h - dx2), dx3 - dx2));
        else v_cmd = v_des;
      }
      return N::div(v_cmd - v, dt);
    }
    default:
      return T(0);
  }
}

// What SUMO makes of a speed commanded through slowDown (apply_acceleration,
// flow/core/kernel/vehicle/traci.py:952-963) under the vehicle's speed mode (setSpeedMode at
// departure, vehicle/traci.py:343-347; SPEED_MODES flow/core/params.py:12-18: aggressive 0,
// obey_safe_speed 1, no_collide 7, right_of_way 25 = the default, all_checks 31).
// RESTATED SUMO, "parity unpinned": bit 0 (regard safe speed) caps the commanded speed by the
// car-following model's own next speed, i.e. the restated simulator model = IDM fed with the
// type's SumoCarFollowingParams (carFollowModel "IDM" is the default, params.py:833); bit 1 caps
// the acceleration at the type's `accel`, bit 2 floors it at -`decel`; bits 3-4 (right of way,
// red lights) have nothing to act on on a ring / straight highway. Written on accelerations:
// max(v + min(a, a_sim) dt, 0) = min(max(v + a dt, 0), max(v + a_sim dt, 0)).
template <typename T>
__device__ __forceinline__ T speed_mode_accel(const TypeP<T>& tp, T a, T v, T v_l, T h,
                                              bool has_lead) {
  typedef Num<T> N;
  const int m = tp.speed_mode;
  if (m & 1)
    a = N::min(a, idm_accel<T>(v, v_l, h, has_lead, tp.sim_v0, tp.sim_inv_v0, tp.sim_tau,
                               tp.max_accel, tp.sim_two_sqrt_ab, tp.sim_inv_two_sqrt_ab, T(4),
                               tp.sim_min_gap));
  if (m & 2) a = N::min(a, tp.max_accel);
  if (m & 4) a = N::max(a, T(0) - tp.max_decel);
  return a;
}

// base_controller.py:120-169
template <typename T>
__device__ __forceinline__ T failsafe_instantaneous(T a, T v, T h, bool has_lead, int n_veh,
                                                    T dt) {
  if (n_veh == 1 || !has_lead) return a;
  T next_vel = add(v, mul(a, dt));
  if (next_vel > T(0)) {
    T lim = add(add(mul(dt, next_vel), mul(v, T(1e-3))), mul(mul(T(0.5), v), dt));
    if (h < lim) return Num<T>::div(-v, dt);
  }
  return a;
}

// base_controller.py:171-236
template <typename T>
__device__ __forceinline__ T failsafe_safe_velocity(T a, T v, T v_l, T h, int n_veh, T dt,
                                                    T delay) {
  if (n_veh == 1) return a;
  T dv = v_l - v;
  T v_safe = add(Num<T>::div(mul(T(2), h), dt), dv) - mul(v, mul(T(2), delay));
  if (add(v, mul(a, dt)) > v_safe) {
    if (v_safe > T(0)) return Num<T>::div(v_safe - v, dt);
    return Num<T>::div(-v, dt);
  }
  return a;
}

}  // namespace fs
