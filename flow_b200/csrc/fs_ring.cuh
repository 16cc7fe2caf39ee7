// fs_ring.cuh — single-lane ring step kernel (configs c1, c2, c5).
//
// Mapping: G THREADS OWN ONE ENVIRONMENT (G = 1, 2, 4, ... 32, a power of two),
// each thread keeping VPT consecutive vehicles (x, v, headway) in REGISTERS for
// all K fused steps of a launch. Lane = g * (32/G) + env-in-warp, so for a fixed
// vehicle slot the 32/G lanes of one g read 32/G consecutive envs of the
// structure-of-arrays state pos[V][E] / vel[V][E] (env index fastest): every
// sector that is fetched is fully used. G = 1 (one env per thread, no
// shuffles at all) is used when there are enough envs to fill the GPU; G > 1
// spreads small batches over more warps and keeps the unrolled body short.
//
// On a single-lane ring vehicles cannot overtake, so the leader of vehicle i
// is vehicle (i+1) mod V for ever: the only
// cross-thread traffic is one warp shuffle of the first vehicle's speed (old)
// and position (new) per step, plus a butterfly for the reward sums.
//
// One step = Env.step of the reference (flow/envs/base.py:294-412):
//   controllers + noise + fail-safes  (flow/controllers/base_controller.py:70-118)
//   apply_acceleration v' = max(v+a*dt,0) (flow/core/kernel/vehicle/traci.py:952-963)
//   simulator Euler move x' = (x+v'*dt) mod C        (SUMO, restated)
//   k.update: headway = front-to-front - leader length (vehicle/traci.py:218-250)
//   collision flag, get_state, compute_reward, done rule (base.py:378-410)
//
// kGeneric = false: every vehicle follows the IDM law (IDMController, or the
// restated simulator default for SimCarFollowing / un-commanded RL vehicles) —
// the unrolled body is ~50 instructions per vehicle. kGeneric = true: full
// controller switch (OVM, LinearOVM, BCM, CFM, LAC, Gipps, FollowerStopper,
// NonLocalFollowerStopper, PISaturation); launched with small VPT so the code
// stays inside the instruction cache.
#pragma once
#include "fs_common.cuh"

namespace fs {

template <typename T>
struct RingBufs {
  T* pos; T* vel; T* circ;          // [V][E], [V][E], [E]
  T* st0;                           // [V][E] controller state (LAC a, PI v_cmd) or null
  T* hist; double* hist_sum; int* hist_cnt;  // PISaturation history ring [H][P][E]
  int hist_len, n_pi;
  int* time_counter; int* episode;  // [E]
  uint8_t* done; uint8_t* crash;    // [E]
  T* obs; T* reward;                // [E][obs_dim], [E]
  T* v_accel; T* v_headway; int* v_leader; int* v_follower;  // views [V][E] or null
  double* stats;                    // [8]
};

template <typename T>
struct StepArgs {
  const T* actions;   // [K][A][E] or null
  T* obs_seq; T* reward_seq; uint8_t* done_seq;  // per-step outputs or null
  const uint8_t* mask;  // [E] only envs with mask!=0 step (null = all)
  int k_steps;
  int accumulate_stats;
  int ignore_done;   // warm-up inside reset
  int write_views;
  int pdl = 0;       // host-side hint: launch with programmatic stream serialization (fs_step_unfused)
};

template <typename T>
__device__ __forceinline__ T shfl(T v, int src, unsigned mask) {
  return __shfl_sync(mask, v, src);
}
template <typename T>
__device__ __forceinline__ T group_sum(T v, unsigned mask, int epw) {
  for (int o = epw; o < 32; o <<= 1) v += __shfl_xor_sync(mask, v, o);
  return v;
}
// select slot `idx` (runtime) from a register array without local memory
template <typename T, int N>
__device__ __forceinline__ T pick(const T (&a)[N], int idx) {
  T r = a[0];
#pragma unroll
  for (int j = 1; j < N; ++j) r = (j == idx) ? a[j] : r;
  return r;
}

// PISaturation (velocity_controllers.py:208-240): history ring in global memory
template <typename T>
__device__ __forceinline__ T pi_saturation(const RingBufs<T>& B, int slot, int E, int e, int tc,
                                           T v, T v_l, T dx, T dt, T max_accel, T& v_cmd) {
  typedef Num<T> N;
  const size_t so = (size_t)slot * E + e;
  int cnt = B.hist_cnt[so];
  double sum = B.hist_sum[so];
  const int H = B.hist_len;
  const int wp = tc % H;
  B.hist[((size_t)wp * B.n_pi + slot) * E + e] = v;
  sum += (double)v; cnt += 1;
  if (cnt == H) {  // len == int(38/dt): drop the oldest sample
    const int old = (wp + 1) % H;
    sum -= (double)B.hist[((size_t)old * B.n_pi + slot) * E + e];
    cnt -= 1;
  }
  B.hist_cnt[so] = cnt; B.hist_sum[so] = sum;
  const T v_des = (T)(sum / (double)cnt);
  const T dv = v_l - v;
  const T dx_s = N::max(mul(T(2), dv), T(4));
  const T v_target = add(v_des, N::min(N::max(N::div(dx - T(7), T(30) - T(7)), T(0)), T(1)));
  const T alpha = N::min(N::max(N::div(dx - dx_s, T(2)), T(0)), T(1));
  const T beta = T(1) - mul(T(0.5), alpha);
  v_cmd = add(mul(beta, add(mul(alpha, v_target), mul(T(1) - alpha, v_l))),
              mul(T(1) - beta, v_cmd));
  return N::min(N::div(v_cmd - v, dt), max_accel);
}

#define FS_SLOTS(j) _Pragma("unroll") for (int j = 0; j < VPT; ++j)

template <typename T, int VPT, int G, bool kGeneric, bool kNoise>
__global__ void __launch_bounds__(128)
ring1_kernel(const __grid_constant__ Cfg<T> cfg, const RingBufs<T> B, const StepArgs<T> A) {
  typedef Num<T> N;
  constexpr int EPW = 32 / G;  // envs per warp
  const int lane = threadIdx.x & 31;
  const int g = (G == 1) ? 0 : lane / EPW;
  const int es = (G == 1) ? lane : lane % EPW;
  const int E = cfg.n_envs, V = cfg.n_veh;
  const int e_raw = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * EPW + es;
  const bool valid = e_raw < E && (A.mask == nullptr || A.mask[e_raw < E ? e_raw : 0] != 0);
  if (!__any_sync(0xffffffffu, valid)) return;  // whole warp idle (masked reset / tail)
  const int e = e_raw < E ? e_raw : E - 1;
  unsigned gmask = 0;
#pragma unroll
  for (int q = 0; q < G; ++q) gmask |= 1u << (es + q * EPW);

  // vehicles [base, base + n_mine) live in this thread's slots 0..n_mine-1
  const int base = g * VPT;
  const int n_mine = max(0, min(VPT, V - base));
  const bool has_lead = V > 1;
  // thread holding (as its slot 0) the leader of my last vehicle / as its last
  // slot the follower of my first vehicle
  const int g_next = (base + n_mine >= V) ? 0 : g + 1;
  const int g_prev = (base == 0) ? (V - 1) / VPT : g - 1;
  const int lane_next = g_next * EPW + es, lane_prev = g_prev * EPW + es;

  T x[VPT], v[VPT], h[VPT], acc[VPT];
  double st_rew = 0, st_v = 0, st_v2 = 0, st_vs = 0, st_es = 0, st_cr = 0;

  const T C = B.circ[e];
  const T dt = cfg.dt;
  FS_SLOTS(j) {
    const bool ok = j < n_mine;
    x[j] = ok ? B.pos[(size_t)(base + j) * E + e] : T(0);
    v[j] = ok ? B.vel[(size_t)(base + j) * E + e] : T(0);
    acc[j] = T(0);
  }
  // leader length of each of my slots (type of vehicle (i+1) mod V)
  auto lead_len = [&](int j) -> T {
    int i1 = base + j + 1;
    if (i1 >= V) i1 = 0;
    return cfg.types[cfg.type_of[i1]].length;
  };
  auto update_headways = [&]() {  // k.update: vehicle/traci.py:218-250
    const T x_nf = (G == 1) ? x[0] : shfl(x[0], lane_next, gmask);
    FS_SLOTS(j) {
      const T xl = (j + 1 < VPT && j + 1 < n_mine) ? x[j + 1 < VPT ? j + 1 : 0] : x_nf;
      h[j] = has_lead ? fwd_dist(xl, x[j], C) - lead_len(j) : T(1000);
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
  const bool ma = cfg.env_kind == FS_ENV_MA_WAVE_ATTENUATION_PO;
  const bool mac = cfg.env_kind == FS_ENV_MA_ACCEL_PO;

  // WaveAttenuationPO: who holds the RL vehicle and its leader
  int rl_veh = 0;
  for (int i = 0; i < V; ++i) if (cfg.rl_index[i] == 0) rl_veh = i;
  const int ld_veh = has_lead ? ((rl_veh + 1 < V) ? rl_veh + 1 : 0) : rl_veh;
  const int rl_lane = (rl_veh / VPT) * EPW + es, ld_lane = (ld_veh / VPT) * EPW + es;
  const int rl_slot = rl_veh % VPT, ld_slot = ld_veh % VPT;

  T max_cost = T(0);  // ||[target]*V||, rewards.py:52-53
  {
    T mc2 = T(0);
    const T t2 = mul(cfg.target_velocity, cfg.target_velocity);
    for (int i = 0; i < V; ++i) mc2 = add(mc2, t2);
    max_cost = N::sqrt(mc2);
  }

  auto write_obs = [&](T* o, T* ol) {
    if (po) {
      // WaveAttenuationPOEnv.get_state, wave_attenuation.py:248-269
      T xr = pick(x, rl_slot), vr = pick(v, rl_slot), xl = pick(x, ld_slot), vl = pick(v, ld_slot);
      if (G > 1) {
        xr = shfl(xr, rl_lane, gmask); vr = shfl(vr, rl_lane, gmask);
        xl = shfl(xl, ld_lane, gmask); vl = shfl(vl, ld_lane, gmask);
      }
      if (g == 0 && valid) {
        const T o0 = N::div(vr, T(15)), o1 = N::div(vl - vr, T(15));
        const T o2 = N::div(fwd_dist(xl, xr, C),
                            cfg.obs_max_length > T(0) ? cfg.obs_max_length : C);
        if (o) { o[0] = o0; o[1] = o1; o[2] = o2; }
        if (ol) { ol[0] = o0; ol[1] = o1; ol[2] = o2; }
      }
    } else if (ma) {
      // MultiAgentWaveAttenuationPOEnv.get_state, multiagent/ring/wave_attenuation.py:195-213
      const T v_nf = (G == 1) ? v[0] : shfl(v[0], lane_next, gmask);
      const T ml = cfg.obs_max_length > T(0) ? cfg.obs_max_length : C;
      FS_SLOTS(j) if (valid && j < n_mine && cfg.rl_index[base + j] >= 0) {
        const int r = cfg.rl_index[base + j];
        const T vl = (j + 1 < VPT && j + 1 < n_mine) ? v[j + 1 < VPT ? j + 1 : 0] : v_nf;
        const T o0 = N::div(v[j], T(15)), o1 = N::div(vl - v[j], T(15)), o2 = N::div(h[j], ml);
        if (o) { o[3 * r] = o0; o[3 * r + 1] = o1; o[3 * r + 2] = o2; }
        if (ol) { ol[3 * r] = o0; ol[3 * r + 1] = o1; ol[3 * r + 2] = o2; }
      }
    } else if (mac) {
      // MultiAgentAccelPOEnv.get_state, multiagent/ring/accel.py:163-206: per RL vehicle
      // (x/L, v/vmax, (v_lead - v)/vmax, (x_lead - x - own length)/L  [not wrapped, own
      // length: as the reference computes it], (v - v_follow)/vmax, headway(follower)/L);
      // nobody ahead / behind: speeds vmax / 0, distances L
      const T v_nf = (G == 1) ? v[0] : shfl(v[0], lane_next, gmask);
      const T x_nf = (G == 1) ? x[0] : shfl(x[0], lane_next, gmask);
      const int lastj = n_mine > 0 ? n_mine - 1 : 0;
      const T vlast = pick(v, lastj), hlast = pick(h, lastj);
      const T v_pf = (G == 1) ? vlast : shfl(vlast, lane_prev, gmask);
      const T h_pf = (G == 1) ? hlast : shfl(hlast, lane_prev, gmask);
      FS_SLOTS(j) if (valid && j < n_mine && cfg.rl_index[base + j] >= 0) {
        const int r = cfg.rl_index[base + j];
        const bool in_next = !(j + 1 < VPT && j + 1 < n_mine);
        const T vl = in_next ? v_nf : v[j + 1 < VPT ? j + 1 : 0];
        const T xl = in_next ? x_nf : x[j + 1 < VPT ? j + 1 : 0];
        const T vfo = j > 0 ? v[j > 0 ? j - 1 : 0] : v_pf;
        const T hfo = j > 0 ? h[j > 0 ? j - 1 : 0] : h_pf;
        const T len = cfg.types[cfg.type_of[base + j]].length;
        T q[6];
        q[0] = N::div(x[j], C);
        q[1] = N::div(v[j], cfg.speed_limit);
        q[2] = N::div((has_lead ? vl : cfg.speed_limit) - v[j], cfg.speed_limit);
        q[3] = N::div(has_lead ? (xl - x[j]) - len : C, C);
        q[4] = N::div(v[j] - (has_lead ? vfo : T(0)), cfg.speed_limit);
        q[5] = N::div(has_lead ? hfo : C, C);
#pragma unroll
        for (int c = 0; c < 6; ++c) {
          if (o) o[6 * r + c] = q[c];
          if (ol) ol[6 * r + c] = q[c];
        }
      }
    } else if (valid) {
      // AccelEnv / WaveAttenuationEnv.get_state, accel.py:116-123
      FS_SLOTS(j) if (j < n_mine) {
        const T a0 = N::div(v[j], cfg.speed_limit), a1 = N::div(x[j], C);
        if (o) { o[base + j] = a0; o[V + base + j] = a1; }
        if (ol) { ol[base + j] = a0; ol[V + base + j] = a1; }
      }
    }
  };
  if (A.k_steps == 0)  // update-only launch (fs_set_state / reset without warm-up)
    write_obs(nullptr, B.obs + (size_t)e * cfg.obs_dim);

  for (int k = 0; k < A.k_steps; ++k) {
    if (!done) {
      // ---- pre-step exchange (old state): leader speed of my last vehicle,
      //      follower speed/headway of my first one, env mean speed
      const T v_nf = (G == 1) ? v[0] : shfl(v[0], lane_next, gmask);
      T vf = T(-1001), hf = T(-1001), mean_v = T(0);
      if (kGeneric) {
        if (cfg.any_follower & 1) {  // BCM reads the follower
          const int last = n_mine > 0 ? n_mine - 1 : 0;
          const T vlast = pick(v, last), hlast = pick(h, last);
          vf = (G == 1) ? vlast : shfl(vlast, lane_prev, gmask);
          hf = (G == 1) ? hlast : shfl(hlast, lane_prev, gmask);
        }
        if (cfg.any_follower & 2) {  // NonLocalFollowerStopper: np.mean of all speeds
          T s = T(0);
          FS_SLOTS(j) if (j < n_mine) s = add(s, v[j]);
          if (G > 1) s = group_sum(s, gmask, EPW);
          mean_v = N::div(s, T(V));
        }
      }
      T act_abs_sum = T(0);
      // ---- controllers -> apply_acceleration -> Euler move (Jacobi: all read old state)
      FS_SLOTS(j) {
        const int i = base + j;
        const bool ok = j < n_mine;
        const T vi = v[j], hi = h[j];
        const T vl = has_lead ? ((j + 1 < VPT && j + 1 < n_mine) ? v[j + 1 < VPT ? j + 1 : 0] : v_nf)
                              : T(-1001);
        const TypeP<T>& tp = cfg.types[cfg.type_of[ok ? i : 0]];
        const int ctrl = tp.ctrl;
        T a;
        if (ctrl == FS_CTRL_RL && have_act) {
          // Env.apply_rl_actions: Box clip, then apply_acceleration (base.py:599-615)
          const T raw = ok ? A.actions[((size_t)k * cfg.act_dim + cfg.rl_index[i]) * E + e] : T(0);
          a = N::min(N::max(raw, cfg.act_lo), cfg.act_hi);
          if (ok) act_abs_sum = add(act_abs_sum, N::abs(cfg.clip_actions ? a : raw));
        } else if (!kGeneric || ctrl == FS_CTRL_IDM || ctrl == FS_CTRL_SIM || ctrl == FS_CTRL_RL) {
          a = idm_accel<T>(vi, vl, hi, has_lead, tp.e_v0, tp.e_inv_v0, tp.e_T, tp.e_a, tp.e_tsab,
                           tp.e_inv_tsab, tp.e_delta, tp.e_s0);
        } else if (ctrl == FS_CTRL_PI_SATURATION) {
          T v_cmd = ok ? B.st0[(size_t)i * E + e] : T(0);
          a = (ok && valid) ? pi_saturation<T>(B, cfg.pi_index[i], E, e, tc, vi, vl, hi, dt,
                                               tp.max_accel, v_cmd)
                            : T(0);
          if (ok && valid) B.st0[(size_t)i * E + e] = v_cmd;
        } else {
          T st0 = T(0);
          const bool stateful = (ctrl == FS_CTRL_LAC);
          if (stateful && ok) st0 = B.st0[(size_t)i * E + e];
          a = ctrl_accel<T>(tp, ctrl, vi, has_lead, vl, hi, has_lead, vf, hf, mean_v, dt,
                            cfg.inv_dt, st0);
          if (stateful && ok && valid) B.st0[(size_t)i * E + e] = st0;
        }
        if (ctrl != FS_CTRL_SIM && ctrl != FS_CTRL_RL) {
          // base_controller.py:109-116: noise first, then the fail-safe
          if (kNoise) {
            if (tp.noise > T(0))
              a = add(a, mul(tp.noise, normal_sample<T>(cfg.seed, gid, i, tc + episode * 1000003)));
          }
          if (tp.fail_safe == FS_FAILSAFE_INSTANTANEOUS)
            a = failsafe_instantaneous<T>(a, vi, hi, has_lead, V, dt);
          else if (tp.fail_safe == FS_FAILSAFE_SAFE_VELOCITY)
            a = failsafe_safe_velocity<T>(a, vi, vl, hi, V, dt, tp.delay);
        }
        // a speed Flow commands passes SUMO's speed mode (restated); vehicles the simulator
        // drives itself (SimCarFollowing, un-commanded RL) do not
        T a_eff = a;
        if (tp.speed_mode != 0 && ctrl != FS_CTRL_SIM && !(ctrl == FS_CTRL_RL && !have_act))
          a_eff = speed_mode_accel<T>(tp, a, vi, vl, hi, has_lead);
        // apply_acceleration (vehicle/traci.py:961-962) + Euler move
        const T vn = N::max(add(vi, mul(a_eff, dt)), T(0));
        const T xn = wrap_fwd(add(x[j], move_dist<T>(vi, vn, dt, cfg.ballistic)), C);
        vf = vi; hf = hi;  // old values: follower data of the next slot
        if (ok) { v[j] = vn; x[j] = xn; acc[j] = a; }
      }
      // ---- k.update + collision + reward sums on the post-move state
      T h_old[VPT];
      FS_SLOTS(j) h_old[j] = h[j];
      update_headways();
      bool cr = false;
      T sv = T(0), sq = T(0);
      const T half_lap = mul(T(0.5), C);
      FS_SLOTS(j) if (j < n_mine) {
        // a vehicle that went THROUGH its leader inside the step (forward distance wraps by
        // a lap) collided on a single lane even if the two no longer overlap afterwards
        cr |= has_lead && (h[j] - h_old[j]) > half_lap;
        if (cfg.collision_min_gap) {
          int i1 = base + j + 1; if (i1 >= V) i1 = 0;
          cr |= has_lead && h[j] < cfg.types[cfg.type_of[i1]].sim_min_gap;
        } else {
          cr |= h[j] < T(0);
        }
        sv = add(sv, v[j]);
        const T d = v[j] - cfg.target_velocity;
        sq = add(sq, mul(d, d));
      }
      if (G > 1) {
        sv = group_sum(sv, gmask, EPW);
        sq = group_sum(sq, gmask, EPW);
        act_abs_sum = group_sum(act_abs_sum, gmask, EPW);
        cr = group_sum((int)cr, gmask, EPW) != 0;
      }
      crash = cr;
      tc += 1;
      done = crash || tc >= t_end;
      if (cfg.env_kind == FS_ENV_ACCEL || mac) {
        // MultiAgentAccelPOEnv.compute_reward (multiagent/ring/accel.py:154-161): the same
        // desired_velocity, shared by the agents. AccelEnv.compute_reward (accel.py:109-114): mean speed when env_params.evaluate,
        // else rewards.desired_velocity, flow/core/rewards.py:6-59
        if (cfg.evaluate && !mac) reward = N::div(sv, T(V));
        else if (crash) reward = T(0);
        else {
          const T cost = N::sqrt(sq);
          const T eps = T(1.1920928955078125e-07);  // np.finfo(np.float32).eps
          reward = N::div(N::max(max_cost - cost, T(0)), add(max_cost, eps));
        }
      } else {
        // WaveAttenuationEnv.compute_reward, wave_attenuation.py:113-139
        if (!have_act || crash) reward = T(0);
        else {
          reward = N::div(mul(T(4), N::div(sv, T(V))), T(20));
          const T mean_act = N::div(act_abs_sum, T(cfg.n_rl > 0 ? cfg.n_rl : 1));
          if (mean_act > T(0)) reward = add(reward, mul(T(4), T(0) - mean_act));
        }
      }
      if (A.accumulate_stats && valid) {
        double q = 0, s1 = 0;
        FS_SLOTS(j) if (j < n_mine) { q += (double)v[j] * (double)v[j]; s1 += (double)v[j]; }
        st_v += s1; st_v2 += q; st_vs += n_mine;
        if (g == 0) { st_rew += (double)reward; st_es += 1; st_cr += crash ? 1 : 0; }
      }
    } else {
      reward = T(0);
    }
    // ---- per-step outputs (the last step always lands in B.obs / B.reward)
    const bool last = (k == A.k_steps - 1);
    if (A.obs_seq != nullptr || last)
      write_obs(A.obs_seq ? A.obs_seq + ((size_t)k * E + e) * cfg.obs_dim : nullptr,
                last ? B.obs + (size_t)e * cfg.obs_dim : nullptr);
    if (g == 0 && valid) {
      if (A.reward_seq) A.reward_seq[(size_t)k * E + e] = reward;
      if (A.done_seq) A.done_seq[(size_t)k * E + e] = done ? 1 : 0;
    }
  }

  // ---- write back
  if (valid) {
    FS_SLOTS(j) if (j < n_mine) {
      B.pos[(size_t)(base + j) * E + e] = x[j];
      B.vel[(size_t)(base + j) * E + e] = v[j];
    }
    if (g == 0) {
      B.time_counter[e] = tc;
      B.done[e] = done ? 1 : 0;
      B.crash[e] = crash ? 1 : 0;
      B.reward[e] = reward;
    }
    if (A.write_views) {
      FS_SLOTS(j) if (j < n_mine) {
        const int i = base + j;
        const size_t o = (size_t)i * E + e;
        B.v_headway[o] = h[j];
        B.v_leader[o] = has_lead ? ((i + 1 < V) ? i + 1 : 0) : -1;
        B.v_follower[o] = has_lead ? ((i > 0) ? i - 1 : V - 1) : -1;
        if (A.k_steps > 0) B.v_accel[o] = acc[j];
      }
    }
  }
  if (A.accumulate_stats) {
    double vals[6] = {st_rew, st_v, st_v2, st_vs, st_es, st_cr};
    substep_stats(A.accumulate_stats, vals);
#pragma unroll
    for (int q = 0; q < 6; ++q) {
      double s = vals[q];
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (lane == 0 && s != 0.0) atomicAdd(&stat_stripe(B.stats)[q], s);
    }
  }
}

// host-side dispatch of the generic ring kernel (instantiated in fs_ring_f32.cu / fs_ring_f64.cu)
template <typename T>
static cudaError_t launch_ring1_impl(int V, int E, cudaStream_t st, const Cfg<T>& c,
                                     const RingBufs<T>& b, const StepArgs<T>& a) {
  const int G = V <= 8 ? 1 : (V <= 32 ? 4 : 32);
  const int epw = 32 / G;
  const int warps = (E + epw - 1) / epw;
  const int wpb = warps <= 148 * 8 ? 1 : (warps <= 148 * 32 ? 2 : 4);
  const int grid = (warps + wpb - 1) / wpb;
  if (G == 1) ring1_kernel<T, 8, 1, true, true><<<grid, wpb * 32, 0, st>>>(c, b, a);
  else if (G == 4) ring1_kernel<T, 8, 4, true, true><<<grid, wpb * 32, 0, st>>>(c, b, a);
  else ring1_kernel<T, 8, 32, true, true><<<grid, wpb * 32, 0, st>>>(c, b, a);
  return cudaGetLastError();
}

}  // namespace fs
