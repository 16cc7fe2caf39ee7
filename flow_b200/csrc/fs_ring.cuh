// fs_ring.cuh — single-lane ring step kernels (configs c1, c2, c5).
//
// Mapping: ONE THREAD OWNS ONE ENVIRONMENT. State is structure-of-arrays with
// the env index fastest (pos[V][E], vel[V][E]) so that every load/store of a
// warp is one fully used 128 B line; the thread keeps its env's x[V], v[V] and
// headway h[V] in registers (VS = V known at compile time) or in shared
// memory ([V][blockDim], bank-conflict free) for other V, for all K fused
// steps of a launch. On a single-lane ring vehicles cannot overtake, so the
// leader of vehicle i is vehicle (i+1) mod V for ever: no sort, no search,
// no shuffles.
//
// One step = Env.step of the reference (flow/envs/base.py:294-412):
//   controllers + noise + fail-safes  (flow/controllers/base_controller.py:70-118)
//   apply_acceleration v' = max(v+a*dt,0) (flow/core/kernel/vehicle/traci.py:952-963)
//   simulator Euler move x' = (x+v'*dt) mod C        (SUMO, restated)
//   k.update: headway = front-to-front - leader length (vehicle/traci.py:218-250)
//   collision flag, get_state, compute_reward, done rule (base.py:378-410)
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
};

// ---- per-thread state stores ------------------------------------------------
template <typename T, int VS>
struct RegStore {
  T x_[VS], v_[VS], h_[VS];
  __device__ __forceinline__ RegStore(void*, int, int) {}
  __device__ __forceinline__ T& x(int i) { return x_[i]; }
  __device__ __forceinline__ T& v(int i) { return v_[i]; }
  __device__ __forceinline__ T& h(int i) { return h_[i]; }
};
template <typename T>
struct SmemStore {
  T* b; int s, n;
  __device__ __forceinline__ SmemStore(void* base, int V, int nthreads)
      : b(reinterpret_cast<T*>(base) + threadIdx.x), s(nthreads), n(V) {}
  __device__ __forceinline__ T& x(int i) { return b[i * s]; }
  __device__ __forceinline__ T& v(int i) { return b[(n + i) * s]; }
  __device__ __forceinline__ T& h(int i) { return b[(2 * n + i) * s]; }
};

#define FS_VLOOP(i) _Pragma("unroll") for (int i = 0; i < (VS > 0 ? VS : V); ++i)

// get_state of the env front (flow/envs/ring/*.py), on the post-move state
template <typename T, int VS, typename Store>
__device__ __forceinline__ void write_obs(const Cfg<T>& cfg, Store& S, int V, T C, int rl,
                                          T* o, T* ol) {
  typedef Num<T> N;
  if (cfg.env_kind == FS_ENV_WAVE_ATTENUATION_PO) {
    // WaveAttenuationPOEnv.get_state, wave_attenuation.py:248-269
    const int ld = (V > 1) ? ((rl + 1 < V) ? rl + 1 : 0) : rl;
    T xr = T(0), vr = T(0), xl = T(0), vl = T(0);
    FS_VLOOP(i) {
      if (i == rl) { xr = S.x(i); vr = S.v(i); }
      if (i == ld) { xl = S.x(i); vl = S.v(i); }
    }
    const T o0 = N::div(vr, T(15)), o1 = N::div(vl - vr, T(15));
    const T o2 = N::div(fwd_dist(xl, xr, C),
                        cfg.obs_max_length > T(0) ? cfg.obs_max_length : C);
    if (o) { o[0] = o0; o[1] = o1; o[2] = o2; }
    if (ol) { ol[0] = o0; ol[1] = o1; ol[2] = o2; }
  } else {
    // AccelEnv / WaveAttenuationEnv.get_state, accel.py:116-123
    FS_VLOOP(i) {
      const T a0 = N::div(S.v(i), cfg.speed_limit), a1 = N::div(S.x(i), C);
      if (o) { o[i] = a0; o[V + i] = a1; }
      if (ol) { ol[i] = a0; ol[V + i] = a1; }
    }
  }
}

template <typename T, int VS, bool kViews>
__global__ void __launch_bounds__(128)
ring1_step_kernel(const __grid_constant__ Cfg<T> cfg, const RingBufs<T> B,
                  const StepArgs<T> A) {
  typedef Num<T> N;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int E = cfg.n_envs;
  const int V = VS > 0 ? VS : cfg.n_veh;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = e < E && (A.mask == nullptr || A.mask[e] != 0);
  typename std::conditional<(VS > 0), RegStore<T, (VS > 0 ? VS : 1)>, SmemStore<T>>::type
      S(smem_raw, V, blockDim.x);

  double st_rew = 0, st_v = 0, st_v2 = 0, st_vs = 0, st_es = 0, st_cr = 0;
  if (valid) {
    const T C = B.circ[e];
    const T dt = cfg.dt;
    FS_VLOOP(i) { S.x(i) = B.pos[(size_t)i * E + e]; S.v(i) = B.vel[(size_t)i * E + e]; }
    // k.update on the loaded state
    FS_VLOOP(i) {
      const int j = (i + 1 < V) ? i + 1 : 0;
      S.h(i) = (V > 1) ? fwd_dist(S.x(j), S.x(i), C) - cfg.types[cfg.type_of[j]].length
                       : T(1000);
    }
    int tc = B.time_counter[e];
    const int episode = B.episode[e];
    bool done = A.ignore_done ? false : (B.done[e] != 0);
    bool crash = B.crash[e] != 0;
    const int t_end = cfg.sims_per_step * (cfg.warmup_steps + cfg.horizon);
    const long long gid = cfg.env_id_base + e;
    T reward = T(0);
    T acc_last[VS > 0 ? VS : 1];
    int rl_veh = 0;
    for (int i = 0; i < V; ++i) if (cfg.rl_index[i] == 0) rl_veh = i;
    T max_cost = T(0);  // ||[target]*V||, rewards.py:52-53
    {
      T mc2 = T(0);
      const T t2 = mul(cfg.target_velocity, cfg.target_velocity);
      for (int i = 0; i < V; ++i) mc2 = add(mc2, t2);
      max_cost = N::sqrt(mc2);
    }
    if (A.k_steps == 0)  // update-only launch (fs_set_state / reset without warm-up)
      write_obs<T, VS>(cfg, S, V, C, rl_veh, nullptr, B.obs + (size_t)e * cfg.obs_dim);

    for (int k = 0; k < A.k_steps; ++k) {
      if (!done) {
        T mean_v = T(0);
        if (cfg.any_follower & 2) {  // NonLocalFollowerStopper present
          T s = T(0);
          FS_VLOOP(i) s = add(s, S.v(i));
          mean_v = N::div(s, T(V));
        }
        const T v0_old = S.v(0);
        T vf = S.v(V - 1), hf = S.h(V - 1);
        T act_abs_sum = T(0);
        const bool have_act = A.actions != nullptr;
        FS_VLOOP(i) {
          const T vi = S.v(i), hi = S.h(i);
          const bool has_lead = V > 1;
          const T vl = has_lead ? ((i + 1 < V) ? S.v(i + 1 < V ? i + 1 : 0) : v0_old) : T(-1001);
          const TypeP<T>& tp = cfg.types[cfg.type_of[i]];
          const int ctrl = tp.ctrl;
          T a;
          if (ctrl == FS_CTRL_RL && have_act) {
            const T raw = A.actions[((size_t)k * cfg.act_dim + cfg.rl_index[i]) * E + e];
            a = N::min(N::max(raw, cfg.act_lo), cfg.act_hi);
            act_abs_sum = add(act_abs_sum, N::abs(cfg.clip_actions ? a : raw));
          } else {
            T st0 = T(0);
            const bool stateful = (ctrl == FS_CTRL_LAC);
            if (stateful) st0 = B.st0[(size_t)i * E + e];
            a = ctrl_accel<T>(tp, ctrl, vi, has_lead, vl, hi, has_lead,
                              has_lead ? vf : T(-1001), has_lead ? hf : T(-1001), mean_v, dt,
                              cfg.inv_dt, st0);
            if (stateful) B.st0[(size_t)i * E + e] = st0;
            if (ctrl != FS_CTRL_SIM && ctrl != FS_CTRL_RL) {
              if (tp.noise > T(0))
                a = add(a, mul(tp.noise, normal_sample<T>(cfg.seed, gid, i,
                                                          tc + episode * 1000003)));
              if (tp.fail_safe == FS_FAILSAFE_INSTANTANEOUS)
                a = failsafe_instantaneous<T>(a, vi, hi, has_lead, V, dt);
              else if (tp.fail_safe == FS_FAILSAFE_SAFE_VELOCITY)
                a = failsafe_safe_velocity<T>(a, vi, vl, hi, V, dt, tp.delay);
            }
          }
          // apply_acceleration + Euler move
          const T vn = N::max(add(vi, mul(a, dt)), T(0));
          const T xn = wrap_fwd(add(S.x(i), mul(vn, dt)), C);
          vf = vi; hf = hi;
          S.v(i) = vn; S.x(i) = xn;
          if (kViews && VS > 0) acc_last[VS > 0 ? i : 0] = a;
          if (kViews && VS == 0) B.v_accel[(size_t)i * E + e] = a;
        }
        // k.update + collision + reward sums
        bool cr = false;
        T sv = T(0), sq = T(0);
        FS_VLOOP(i) {
          const int j = (i + 1 < V) ? i + 1 : 0;
          const TypeP<T>& tl = cfg.types[cfg.type_of[j]];
          const T hh = (V > 1) ? fwd_dist(S.x(j), S.x(i), C) - tl.length : T(1000);
          S.h(i) = hh;
          cr |= cfg.collision_min_gap ? (hh < tl.sim_min_gap && V > 1) : (hh < T(0));
          const T vi = S.v(i);
          sv = add(sv, vi);
          const T d = vi - cfg.target_velocity;
          sq = add(sq, mul(d, d));
        }
        crash = cr;
        tc += 1;
        done = crash || tc >= t_end;
        // compute_reward
        if (cfg.env_kind == FS_ENV_ACCEL) {
          // rewards.desired_velocity, flow/core/rewards.py:6-59
          if (crash) reward = T(0);
          else {
            const T cost = N::sqrt(sq);
            const T eps = T(1.1920928955078125e-07);
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
        if (A.accumulate_stats) {
          st_rew += (double)reward; st_v += (double)sv; st_vs += V; st_es += 1;
          double q = 0;
          FS_VLOOP(i) q += (double)S.v(i) * (double)S.v(i);
          st_v2 += q;
          st_cr += crash ? 1 : 0;
        }
      } else {
        reward = T(0);
      }
      // per-step outputs (last step always lands in B.obs / B.reward)
      const bool last = (k == A.k_steps - 1);
      if (A.obs_seq != nullptr || last)
        write_obs<T, VS>(cfg, S, V, C, rl_veh,
                         A.obs_seq ? A.obs_seq + ((size_t)k * E + e) * cfg.obs_dim : nullptr,
                         last ? B.obs + (size_t)e * cfg.obs_dim : nullptr);
      if (A.reward_seq) A.reward_seq[(size_t)k * E + e] = reward;
      if (A.done_seq) A.done_seq[(size_t)k * E + e] = done ? 1 : 0;
    }
    // write back
    FS_VLOOP(i) { B.pos[(size_t)i * E + e] = S.x(i); B.vel[(size_t)i * E + e] = S.v(i); }
    B.time_counter[e] = tc;
    B.done[e] = done ? 1 : 0;
    B.crash[e] = crash ? 1 : 0;
    B.reward[e] = reward;
    if (kViews) {
      FS_VLOOP(i) {
        const size_t o = (size_t)i * E + e;
        B.v_headway[o] = S.h(i);
        B.v_leader[o] = (V > 1) ? ((i + 1 < V) ? i + 1 : 0) : -1;
        B.v_follower[o] = (V > 1) ? ((i > 0) ? i - 1 : V - 1) : -1;
        if (VS > 0) B.v_accel[o] = acc_last[VS > 0 ? i : 0];
      }
    }
  }
  if (A.accumulate_stats) {
    double vals[6] = {st_rew, st_v, st_v2, st_vs, st_es, st_cr};
#pragma unroll
    for (int q = 0; q < 6; ++q) {
      double x = vals[q];
      for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
      if ((threadIdx.x & 31) == 0 && x != 0.0) atomicAdd(&B.stats[q], x);
    }
  }
}

}  // namespace fs
