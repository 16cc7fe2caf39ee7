// fs_ring_fast.cuh — the hot kernel: single-lane ring where every vehicle
// follows the IDM law (IDMController humans of ONE type + RL vehicles of one
// type, which is what examples/exp_configs/non_rl/ring.py and
// rl/singleagent/singleagent_ring.py build). Same thread mapping as
// fs_ring.cuh (G threads per env, VPT register-resident vehicles per thread),
// but with everything that is uniform hoisted out of the K-step loop:
//   * the human type's IDM parameters live in registers (no per-vehicle
//     constant-bank gathers), RL vehicles are a bit mask;
//   * delta == 4 is a template flag (two multiplies instead of pow);
//   * noise (Philox4x32-10, one call per 4 vehicles, Box-Muller on both
//     outputs) and the fail-safes are template flags;
//   * kExact (V == VPT*G) removes all padding predicates.
// Reference lines restated: see fs_ring.cuh header (same step semantics).
#pragma once
#include "fs_common.cuh"
#include "fs_ring.cuh"

namespace fs {

// four standard normals for (env, vehicle group q = vehicle/4, step)
template <typename T>
__device__ __forceinline__ void normal4(unsigned long long seed, long long env_gid, int q,
                                        int step, T (&z)[4]) {
  uint32_t r[4];
  Philox::gen((uint32_t)step, (uint32_t)q, (uint32_t)env_gid, (uint32_t)(env_gid >> 32),
              (uint32_t)seed, (uint32_t)(seed >> 32), r);
  if (sizeof(T) == 4) {
    const float k = 2.3283064365386963e-10f;  // 2^-32
    const float u0 = ((float)r[0] + 1.0f) * k, u2 = ((float)r[2] + 1.0f) * k;
    const float ra = sqrtf(-2.0f * __logf(fminf(u0, 1.0f)));
    const float rb = sqrtf(-2.0f * __logf(fminf(u2, 1.0f)));
    float s0, c0, s1, c1;
    __sincosf(6.2831853071795865f * ((float)r[1] * k), &s0, &c0);
    __sincosf(6.2831853071795865f * ((float)r[3] * k), &s1, &c1);
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

template <typename T>
struct IdmRegs {  // one type's effective IDM block, register resident
  T v0, inv_v0, Tt, a, tsab, inv_tsab, delta, s0;
  __device__ __forceinline__ void load(const TypeP<T>& t) {
    v0 = t.e_v0; inv_v0 = t.e_inv_v0; Tt = t.e_T; a = t.e_a; tsab = t.e_tsab;
    inv_tsab = t.e_inv_tsab; delta = t.e_delta; s0 = t.e_s0;
  }
};

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

template <typename T, int VPT, int G, bool kExact, bool kNoise, bool kFS, bool kViews>
__global__ void __launch_bounds__(128)
ring1_fast_kernel(const __grid_constant__ Cfg<T> cfg, const RingBufs<T> B, const StepArgs<T> A) {
  typedef Num<T> N;
  constexpr int EPW = 32 / G;
  constexpr int NQ = (VPT + 3) / 4 + (G > 1 ? 1 : 0);  // Philox calls covering my vehicles
  const int lane = threadIdx.x & 31;
  const int g = (G == 1) ? 0 : lane / EPW;
  const int es = (G == 1) ? lane : lane % EPW;
  const int E = cfg.n_envs, V = cfg.n_veh;
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

  // uniform tables -> registers
  unsigned rl_mask = 0;
  int type_a = 0, type_b = 0;
  for (int i = 0; i < V; ++i) {
    if (cfg.rl_index[i] >= 0) { rl_mask |= 1u << i; type_b = cfg.type_of[i]; }
    else type_a = cfg.type_of[i];
  }
  const unsigned my_rl = rl_mask >> base;
  IdmRegs<T> pa;
  pa.load(cfg.types[type_a]);
  const T len = cfg.types[type_a].length;
  const T noise_a = cfg.types[type_a].noise, delay_a = cfg.types[type_a].delay;
  const int fs_a = cfg.types[type_a].fail_safe;

  T x[VPT], v[VPT], h[VPT], acc[kViews ? VPT : 1];
  const T C = B.circ[e];
  const T dt = cfg.dt;
  FS_SLOTS(j) {
    const bool ok = kExact || j < n_mine;
    x[j] = ok ? B.pos[(size_t)(base + j) * E + e] : T(0);
    v[j] = ok ? B.vel[(size_t)(base + j) * E + e] : T(0);
    if (kViews) acc[kViews ? j : 0] = T(0);
  }
  auto update_headways = [&]() {  // k.update: vehicle/traci.py:218-250
    const T x_nf = (G == 1) ? x[0] : shfl(x[0], lane_next, gmask);
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
  const int t_end = cfg.sims_per_step * (cfg.warmup_steps + cfg.horizon);
  const long long gid = cfg.env_id_base + e;
  T reward = B.reward[e];
  const bool have_act = A.actions != nullptr;
  const bool po = cfg.env_kind == FS_ENV_WAVE_ATTENUATION_PO;
  const bool accel_env = cfg.env_kind == FS_ENV_ACCEL;
  const T target = cfg.target_velocity;
  const T inv_nrl = N::div(T(1), T(cfg.n_rl > 0 ? cfg.n_rl : 1));

  int rl_veh = 0;
  for (int i = 0; i < V; ++i) if (cfg.rl_index[i] == 0) rl_veh = i;
  const int ld_veh = (rl_veh + 1 < V) ? rl_veh + 1 : 0;
  const int rl_lane = (rl_veh / VPT) * EPW + es, ld_lane = (ld_veh / VPT) * EPW + es;
  const int rl_slot = rl_veh % VPT, ld_slot = ld_veh % VPT;
  const bool rl_is_last = kExact && rl_slot == VPT - 1 && ld_slot == 0;
  const T obs_len = cfg.obs_max_length > T(0) ? cfg.obs_max_length : C;

  T max_cost = T(0);  // ||[target]*V||, rewards.py:52-53
  {
    T mc2 = T(0);
    const T t2 = mul(target, target);
    for (int i = 0; i < V; ++i) mc2 = add(mc2, t2);
    max_cost = N::sqrt(mc2);
  }
  double st_rew = 0, st_v = 0, st_v2 = 0, st_vs = 0, st_es = 0, st_cr = 0;

  auto write_obs = [&](T* o, T* ol) {
    if (po) {  // WaveAttenuationPOEnv.get_state, wave_attenuation.py:248-269
      T xr, vr, xl, vl;
      if (rl_is_last) { xr = x[VPT - 1]; vr = v[VPT - 1]; xl = x[0]; vl = v[0]; }
      else { xr = pick(x, rl_slot); vr = pick(v, rl_slot); xl = pick(x, ld_slot); vl = pick(v, ld_slot); }
      if (G > 1) {
        xr = shfl(xr, rl_lane, gmask); vr = shfl(vr, rl_lane, gmask);
        xl = shfl(xl, ld_lane, gmask); vl = shfl(vl, ld_lane, gmask);
      }
      if (g == 0 && valid) {
        const T o0 = N::div(vr, T(15)), o1 = N::div(vl - vr, T(15));
        const T o2 = N::div(fwd_dist(xl, xr, C), obs_len);
        if (o) { o[0] = o0; o[1] = o1; o[2] = o2; }
        if (ol) { ol[0] = o0; ol[1] = o1; ol[2] = o2; }
      }
    } else if (valid) {  // AccelEnv / WaveAttenuationEnv.get_state, accel.py:116-123
      FS_SLOTS(j) if (kExact || j < n_mine) {
        const T a0 = N::div(v[j], cfg.speed_limit), a1 = N::div(x[j], C);
        if (o) { o[base + j] = a0; o[V + base + j] = a1; }
        if (ol) { ol[base + j] = a0; ol[V + base + j] = a1; }
      }
    }
  };
  if (A.k_steps == 0) write_obs(nullptr, B.obs + (size_t)e * cfg.obs_dim);

  for (int k = 0; k < A.k_steps; ++k) {
    if (!done) {
      const T v_nf = (G == 1) ? v[0] : shfl(v[0], lane_next, gmask);
      T zn[NQ][4];
      if (kNoise) {
#pragma unroll
        for (int q = 0; q < NQ; ++q)
          normal4<T>(cfg.seed, gid, (base >> 2) + q, tc + episode * 1000003, zn[q]);
      }
      T act_abs_sum = T(0);
      FS_SLOTS(j) {
        const bool ok = kExact || j < n_mine;
        const T vi = v[j], hi = h[j];
        T vl;
        if (j + 1 < VPT) vl = (kExact || j + 1 < n_mine) ? v[j + 1 < VPT ? j + 1 : 0] : v_nf;
        else vl = v_nf;
        T a;
        if ((my_rl >> j) & 1u) {
          const int i = base + j;
          if (have_act) {
            // Env.apply_rl_actions: Box clip, then apply_acceleration (base.py:599-615)
            const T raw = A.actions[((size_t)k * cfg.act_dim + cfg.rl_index[i]) * E + e];
            a = N::min(N::max(raw, cfg.act_lo), cfg.act_hi);
            act_abs_sum = add(act_abs_sum, N::abs(cfg.clip_actions ? a : raw));
          } else {
            IdmRegs<T> pb;  // restated simulator default model for the un-commanded RL vehicle
            pb.load(cfg.types[type_b]);
            a = idm_fast<T, true>(vi, vl, hi, pb);
          }
        } else {
          a = idm_fast<T, true>(vi, vl, hi, pa);
          if (kNoise) {  // base_controller.py:109-110
            const int off = (G == 1) ? j : (base & 3) + j;
            T z;
            if (G == 1) z = zn[j >> 2][j & 3];
            else {
              const int qq = off >> 2, cc = off & 3;
              z = T(0);
#pragma unroll
              for (int q = 0; q < NQ; ++q)
#pragma unroll
                for (int c = 0; c < 4; ++c) z = (q == qq && c == cc) ? zn[q][c] : z;
            }
            a = add(a, mul(noise_a, z));
          }
          if (kFS) {  // base_controller.py:113-116
            if (fs_a == FS_FAILSAFE_INSTANTANEOUS)
              a = failsafe_instantaneous<T>(a, vi, hi, true, V, dt);
            else if (fs_a == FS_FAILSAFE_SAFE_VELOCITY)
              a = failsafe_safe_velocity<T>(a, vi, vl, hi, V, dt, delay_a);
          }
        }
        const T vn = N::max(add(vi, mul(a, dt)), T(0));   // vehicle/traci.py:961-962
        const T xn = wrap_fwd(add(x[j], mul(vn, dt)), C);  // simulator Euler move
        if (ok) { v[j] = vn; x[j] = xn; if (kViews) acc[kViews ? j : 0] = a; }
      }
      update_headways();
      bool cr = false;
      T sv = T(0), sq = T(0), s2 = T(0);
      FS_SLOTS(j) if (kExact || j < n_mine) {
        cr |= cfg.collision_min_gap ? (h[j] < cfg.types[type_a].sim_min_gap) : (h[j] < T(0));
        sv = add(sv, v[j]);
        if (accel_env) { const T d = v[j] - target; sq = add(sq, mul(d, d)); }
        if (A.accumulate_stats) s2 += v[j] * v[j];
      }
      if (G > 1) {
        sv = group_sum(sv, gmask, EPW);
        if (accel_env) sq = group_sum(sq, gmask, EPW);
        act_abs_sum = group_sum(act_abs_sum, gmask, EPW);
        cr = group_sum((int)cr, gmask, EPW) != 0;
        if (A.accumulate_stats) s2 = group_sum(s2, gmask, EPW);
      }
      crash = cr;
      tc += 1;
      done = crash || tc >= t_end;
      if (accel_env) {  // rewards.desired_velocity, flow/core/rewards.py:6-59
        if (crash) reward = T(0);
        else {
          const T cost = N::sqrt(sq);
          const T eps = T(1.1920928955078125e-07);
          reward = N::div(N::max(max_cost - cost, T(0)), add(max_cost, eps));
        }
      } else {  // WaveAttenuationEnv.compute_reward, wave_attenuation.py:113-139
        if (!have_act || crash) reward = T(0);
        else {
          reward = N::div(mul(T(4), N::div(sv, T(V))), T(20));
          const T mean_act = mul(act_abs_sum, inv_nrl);
          if (mean_act > T(0)) reward = add(reward, mul(T(4), T(0) - mean_act));
        }
      }
      if (A.accumulate_stats && valid && g == 0) {
        st_rew += (double)reward; st_v += (double)sv; st_v2 += (double)s2;
        st_vs += V; st_es += 1; st_cr += crash ? 1 : 0;
      }
    } else {
      reward = T(0);
    }
    const bool last = (k == A.k_steps - 1);
    if (A.obs_seq != nullptr || last)
      write_obs(A.obs_seq ? A.obs_seq + ((size_t)k * E + e) * cfg.obs_dim : nullptr,
                last ? B.obs + (size_t)e * cfg.obs_dim : nullptr);
    if (g == 0 && valid) {
      if (A.reward_seq) A.reward_seq[(size_t)k * E + e] = reward;
      if (A.done_seq) A.done_seq[(size_t)k * E + e] = done ? 1 : 0;
    }
  }

  if (valid) {
    FS_SLOTS(j) if (kExact || j < n_mine) {
      B.pos[(size_t)(base + j) * E + e] = x[j];
      B.vel[(size_t)(base + j) * E + e] = v[j];
    }
    if (g == 0) {
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
        if (A.k_steps > 0) B.v_accel[o] = acc[kViews ? j : 0];
      }
    }
  }
  if (A.accumulate_stats) {
    double vals[6] = {st_rew, st_v, st_v2, st_vs, st_es, st_cr};
#pragma unroll
    for (int q = 0; q < 6; ++q) {
      double s = vals[q];
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (lane == 0 && s != 0.0) atomicAdd(&B.stats[q], s);
    }
  }
}


// ---- host-side dispatch (one translation unit per dtype instantiates these)
template <typename T, int VPT, int G, bool kExact>
static cudaError_t launch_fast_shape(int variant, int grid, int block, cudaStream_t st,
                                     const Cfg<T>& c, const RingBufs<T>& b, const StepArgs<T>& a) {
  switch (variant) {
    case 0: ring1_fast_kernel<T, VPT, G, kExact, false, false, false><<<grid, block, 0, st>>>(c, b, a); break;
    case 1: ring1_fast_kernel<T, VPT, G, kExact, true, false, false><<<grid, block, 0, st>>>(c, b, a); break;
    case 2: ring1_fast_kernel<T, VPT, G, kExact, true, true, false><<<grid, block, 0, st>>>(c, b, a); break;
    default: ring1_fast_kernel<T, VPT, G, kExact, true, true, true><<<grid, block, 0, st>>>(c, b, a); break;
  }
  return cudaGetLastError();
}

// shape = threads per env (1, 2, 4, 8) for V = 22; variant: 0 plain, 1 noise,
// 2 noise + fail-safe, 3 = 2 + views (accel / headway / leader / follower)
template <typename T>
static cudaError_t launch_fast22(int shape, int variant, int n_envs, cudaStream_t st,
                                 const Cfg<T>& c, const RingBufs<T>& b, const StepArgs<T>& a) {
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
    default: return launch_fast_shape<T, 3, 8, false>(variant, grid, block, st, c, b, a);
  }
}

}  // namespace fs
