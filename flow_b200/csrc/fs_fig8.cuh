// fs_fig8.cuh — FigureEightNetwork step kernel: the closed loop of
// flow/networks/figure_eight.py that passes its centre junction twice, stepped under AccelEnv
// (examples/exp_configs/rl/singleagent/singleagent_figure_eight.py).
//
// Reference arithmetic followed line by line:
//   AccelEnv.get_state / compute_reward / _apply_rl_actions   flow/envs/ring/accel.py:100-123
//   rewards.desired_velocity                                  flow/core/rewards.py:6-59
//   get_x_by_id on the network's edge starts                  flow/networks/figure_eight.py:217-263
//     through TraCIKernelNetwork.get_x                        flow/core/kernel/network/traci.py:273-289
//     (the four 0.28 m node slots return their START, the two junction passes start + position)
//   BaseController.get_action: no command on an internal edge flow/controllers/base_controller.py:96-99
//   controllers / noise / fail-safes                          fs_common.cuh
// RESTATED SUMO ("parity unpinned"; oracle/figure8.py is the same statement on the
// CPU): along the loop a single-lane ring (leader = nearest vehicle ahead mod L, a vehicle that
// went through its leader inside a step collided); the crossing is a merge followed by a
// diverge: with delta = c - s the distance of the front bumper to the junction box of the
// vehicle's own pass, a vehicle with -J < delta <= zone also regards the vehicles of the OTHER
// pass that have not cleared the box (-(J + length) < delta_j <= zone) and are closer to /
// further through it as ahead of it, at a gap (delta_i - delta_j) - length_j; its leader is the
// nearer of the two. Collision <=> negative gap towards the loop leader, or towards a crossing
// leader while both bodies are inside the box; a negative gap towards a crossing leader outside
// the box reads as 1 mm for the controllers.
//
// Mapping: one warp per env, vehicle s = lane + 32 r, state env-major, (s, v, delta, flags)
// published to the warp's shared-memory slice once per step. Loop leader: vehicles are placed in
// loop order and nobody overtakes, so it is index s + 1 whenever the published positions are a
// cyclic shift of an ascending sequence (one ballot count), else an all-pairs scan. Crossing
// leader: the vehicles with a role at the junction are compacted into two short lists.
#pragma once
#include "fs_common.cuh"
#include "fs_ring.cuh"

namespace fs {

template <typename T>
struct F8Bufs {
  T* pos; T* vel;                                   // [E][V]
  int* time_counter; int* episode;
  uint8_t* done; uint8_t* crash;
  T* obs; T* reward;
  T* v_accel; T* v_headway; int* v_leader; int* v_follower;   // views [E][V] or null
  double* stats;
};

template <typename T>
__host__ __device__ inline size_t f8_smem_per_warp(int V) {
  return (size_t)V * (3 * sizeof(T) + 3 * sizeof(int));
}

// index of the segment that holds loop coordinate s (the last start <= s)
template <typename T>
__device__ __forceinline__ int f8_segment(const Cfg<T>& c, T s) {
  int k = 0;
#pragma unroll
  for (int q = 1; q < 12; ++q) k = s >= c.f8_start[q] ? q : k;
  return k;
}
// get_x_by_id: the node slots (segments 0, 4, 6, 10) give their start, else start + position
template <typename T>
__device__ __forceinline__ T f8_x_abs(const Cfg<T>& c, T s) {
  const int k = f8_segment(c, s);
  const T st = c.f8_start[k];
  return (k == 0 || k == 4 || k == 6 || k == 10) ? st : add(st, s - st);
}

template <typename T, int SLOTS>
__global__ void __launch_bounds__(128)
fig8_kernel(const __grid_constant__ Cfg<T> cfg, const __grid_constant__ F8Bufs<T> B,
            const StepArgs<T> A) {
  typedef Num<T> N;
  constexpr unsigned FULL = 0xffffffffu;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane_id = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int E = cfg.n_envs, V = cfg.n_veh;
  const int e = blockIdx.x * (blockDim.x >> 5) + wib;
  if (e >= E) return;
  if (A.mask != nullptr && A.mask[e] == 0) return;
  unsigned char* mine = smem_raw + f8_smem_per_warp<T>(V) * wib;
  T* xs = reinterpret_cast<T*>(mine);      // loop coordinate
  T* vs = xs + V;
  T* dl = vs + V;                          // delta: distance to the box of the own pass
  int* fl = reinterpret_cast<int*>(dl + V);   // pass | regarded << 1
  int* zl = fl + V;                        // role lists: pass 0 from the front, pass 1 from the back
  int* ld = zl + V;                        // leader (views / follower search)

  T x[SLOTS], v[SLOTS], h[SLOTS], acc[SLOTS], aeff[SLOTS], len[SLOTS];
  int lead[SLOTS];
  bool cross[SLOTS];
  const size_t eo = (size_t)e * V;
#pragma unroll
  for (int r = 0; r < SLOTS; ++r) {
    const int s = lane_id + 32 * r;
    const bool ok = s < V;
    x[r] = ok ? B.pos[eo + s] : T(0);
    v[r] = ok ? B.vel[eo + s] : T(0);
    len[r] = ok ? cfg.types[cfg.type_of[s]].length : T(0);
    acc[r] = T(0); aeff[r] = T(0); lead[r] = -1; h[r] = T(1000); cross[r] = false;
  }
  const T L = cfg.f8_L, cA = cfg.f8_cA, cB = cfg.f8_cB, J = cfg.f8_J, zone = cfg.f8_zone;
  const T qL = mul(T(0.25), L), half_lap = mul(T(0.5), L);
  const T dt = cfg.dt;
  int tc = B.time_counter[e];
  const int episode = B.episode[e];
  bool done = A.ignore_done ? false : (B.done[e] != 0);
  bool crash = B.crash[e] != 0;
  const int t_end = cfg.sims_per_step * (cfg.warmup_steps + cfg.horizon);
  const long long gid = cfg.env_id_base + e;
  T reward = B.reward[e];
  double st_rew = 0, st_v = 0, st_v2 = 0, st_vs = 0, st_es = 0, st_cr = 0;

  auto in_box = [&](T d, T ln) -> bool { return (T(0) - add(J, ln)) < d && d < T(0); };

  // k.update on the loop + the crossing; returns whether somebody went through its loop leader
  // since the previous update (checked against the leader / headway still held in registers)
  auto update = [&](bool first) -> bool {
    __syncwarp();
#pragma unroll
    for (int r = 0; r < SLOTS; ++r) {
      const int s = lane_id + 32 * r;
      if (s < V) {
        const T dA = cA - x[r];
        const int pass = dA > (T(0) - qL) ? 0 : 1;
        const T d = pass ? cB - x[r] : dA;
        const bool regarded = (T(0) - add(J, len[r])) < d && d <= zone;
        xs[s] = x[r]; vs[s] = v[r]; dl[s] = d; fl[s] = pass | (regarded ? 2 : 0);
      }
    }
    __syncwarp();
    bool thr = false;
    int ndesc = 0;
    bool tie = false;
    int zn0 = 0, zn1 = 0;
#pragma unroll
    for (int r = 0; r < SLOTS; ++r) {
      const int s = lane_id + 32 * r;
      const bool on = s < V;
      if (on && !first && lead[r] >= 0 && !cross[r]) {
        const int j = lead[r];
        const T d_new = fwd_dist(xs[j], x[r], L) - cfg.types[cfg.type_of[j]].length;
        thr |= (d_new - h[r]) > half_lap;
      }
      const int nx = s + 1 == V ? 0 : s + 1;
      const T xn = on ? xs[nx] : T(0);
      ndesc += __popc(__ballot_sync(FULL, on && xn < x[r]));
      tie |= __any_sync(FULL, on && xn == x[r]) != 0;
      const bool rg = on && (fl[s] & 2);
      const unsigned m1 = __ballot_sync(FULL, rg && (fl[s] & 1));
      const unsigned m0 = __ballot_sync(FULL, rg && !(fl[s] & 1));
      if (rg) {
        const unsigned lower = (1u << lane_id) - 1u;
        if (fl[s] & 1) zl[V - 1 - (zn1 + __popc(m1 & lower))] = s;
        else zl[zn0 + __popc(m0 & lower)] = s;
      }
      zn0 += __popc(m0); zn1 += __popc(m1);
    }
    const bool ordered = V >= 2 && ndesc == 1 && !tie;
    __syncwarp();
#pragma unroll
    for (int r = 0; r < SLOTS; ++r) {
      const int s = lane_id + 32 * r;
      int bj = -1; T bd = T(0); bool cr = false;
      if (s < V) {
        if (ordered) {
          bj = s + 1 == V ? 0 : s + 1;
          bd = fwd_dist(xs[bj], x[r], L);
        } else {
          for (int j = 0; j < V; ++j) {
            if (j == s) continue;
            const T d = fwd_dist(xs[j], x[r], L);
            if (bj < 0 || d < bd) { bj = j; bd = d; }
          }
        }
        const T di = dl[s];
        if ((T(0) - J) < di && di <= zone) {          // regards the other pass
          const int other = (fl[s] & 1) ^ 1, zn = other ? zn1 : zn0;
          for (int q = 0; q < zn; ++q) {
            const int j = other ? zl[V - 1 - q] : zl[q];
            const T dj = dl[j];
            if (!(dj < di || (dj == di && j < s))) continue;
            const T d = di - dj;
            if (bj < 0 || d < bd) { bj = j; bd = d; cr = true; }
          }
        }
      }
      lead[r] = bj; cross[r] = cr;
      h[r] = bj >= 0 ? bd - cfg.types[cfg.type_of[bj]].length : T(1000);
    }
    __syncwarp();
#pragma unroll
    for (int r = 0; r < SLOTS; ++r) {
      const int s = lane_id + 32 * r;
      if (s < V) ld[s] = lead[r];
    }
    return __any_sync(FULL, thr) != 0;
  };
  update(true);

  // AccelEnv.get_state (accel.py:116-123), get_ids order: speeds / max_speed, then x / length
  auto write_obs = [&](T* o, T* ol) {
#pragma unroll
    for (int r = 0; r < SLOTS; ++r) {
      const int s = lane_id + 32 * r;
      if (s >= V) continue;
      const T q0 = N::div(v[r], cfg.speed_limit), q1 = N::div(f8_x_abs<T>(cfg, x[r]), L);
      if (o) { o[s] = q0; o[V + s] = q1; }
      if (ol) { ol[s] = q0; ol[V + s] = q1; }
    }
  };
  if (A.k_steps == 0) write_obs(nullptr, B.obs + (size_t)e * cfg.obs_dim);

  const T max_cost = N::sqrt(mul(mul(cfg.target_velocity, cfg.target_velocity), T(V)));
  for (int k = 0; k < A.k_steps; ++k) {
    bool stepped = false;
#pragma unroll 1
    for (int q = 0; q < cfg.sims_per_step; ++q) {
      if (done) break;
      stepped = true;
      T mean_v = T(0);
      if (cfg.any_follower & 2) {
        T sm = T(0);
#pragma unroll
        for (int r = 0; r < SLOTS; ++r) if (lane_id + 32 * r < V) sm = add(sm, v[r]);
        for (int o = 16; o > 0; o >>= 1) sm += __shfl_xor_sync(FULL, sm, o);
        mean_v = N::div(sm, T(V));
      }
#pragma unroll
      for (int r = 0; r < SLOTS; ++r) {
        const int s = lane_id + 32 * r;
        if (s >= V) continue;
        const TypeP<T>& tp = cfg.types[cfg.type_of[s]];
        const bool has_lead = lead[r] >= 0;
        const T vi = v[r];
        const T vl = has_lead ? vs[lead[r]] : T(-1001);
        // an overlapping crossing leader outside the box reads as bumper to bumper
        T hi = h[r];
        if (has_lead && cross[r] && hi < T(0) && !(in_box(dl[s], len[r]) &&
                                                   in_box(dl[lead[r]], cfg.types[cfg.type_of[lead[r]]].length)))
          hi = T(1e-3);
        const int rl = cfg.rl_index[s];
        T a;
        bool commanded;
        if (rl >= 0 && A.actions != nullptr) {
          // _apply_rl_actions -> apply_acceleration, whatever edge the vehicle is on
          a = A.actions[((size_t)k * cfg.act_dim + rl) * E + e];
          if (cfg.clip_actions) a = N::min(N::max(a, cfg.act_lo), cfg.act_hi);
          commanded = true;
        } else {
          // base_controller.py:96-99: on an internal edge get_action returns None and the
          // simulator's model drives (even segment index = internal)
          const bool internal = (f8_segment<T>(cfg, x[r]) & 1) == 0;
          const int ctrl = (internal || tp.ctrl == FS_CTRL_RL) ? FS_CTRL_SIM : tp.ctrl;
          T st0 = T(0);
          a = ctrl_accel<T>(tp, ctrl, vi, has_lead, vl, hi, false, T(-1001), T(-1001), mean_v, dt,
                            cfg.inv_dt, st0);
          commanded = ctrl != FS_CTRL_SIM;
          if (commanded) {   // base_controller.py:109-116: noise, then the fail-safe
            if (tp.noise > T(0))
              a = add(a, mul(tp.noise, normal_sample<T>(cfg.seed, gid, s, tc + episode * 1000003)));
            if (tp.fail_safe == FS_FAILSAFE_INSTANTANEOUS)
              a = failsafe_instantaneous<T>(a, vi, hi, has_lead, V, dt);
            else if (tp.fail_safe == FS_FAILSAFE_SAFE_VELOCITY)
              a = failsafe_safe_velocity<T>(a, vi, vl, hi, V, dt, tp.delay);
          }
        }
        acc[r] = a;
        aeff[r] = (commanded && tp.speed_mode != 0) ? speed_mode_accel<T>(tp, a, vi, vl, hi, has_lead) : a;
      }
      // apply_acceleration (vehicle/traci.py:961-962) + the restated Euler move round the loop
#pragma unroll
      for (int r = 0; r < SLOTS; ++r) {
        if (lane_id + 32 * r >= V) continue;
        const T vn = N::max(add(v[r], mul(aeff[r], dt)), T(0));
        x[r] = wrap_fwd(add(x[r], mul(vn, dt)), L);
        v[r] = vn;
      }
      tc += 1;
      const bool through = update(false);
      bool cr = false;
      T sv = T(0), sqd = T(0), s2 = T(0);
#pragma unroll
      for (int r = 0; r < SLOTS; ++r) {
        const int s = lane_id + 32 * r;
        if (s >= V) continue;
        if (lead[r] >= 0 && h[r] < T(0))
          cr |= !cross[r] || (in_box(dl[s], len[r]) &&
                              in_box(dl[lead[r]], cfg.types[cfg.type_of[lead[r]]].length));
        sv = add(sv, v[r]);
        const T d = v[r] - cfg.target_velocity;
        sqd = add(sqd, mul(d, d));
        s2 += v[r] * v[r];
      }
      for (int o = 16; o > 0; o >>= 1) {
        sv += __shfl_xor_sync(FULL, sv, o);
        sqd += __shfl_xor_sync(FULL, sqd, o);
        s2 += __shfl_xor_sync(FULL, s2, o);
      }
      crash = through || __any_sync(FULL, cr);
      done = crash || tc >= t_end;
      // AccelEnv.compute_reward (accel.py:108-114)
      if (cfg.evaluate) reward = N::div(sv, T(V));
      else if (crash) reward = T(0);
      else reward = N::div(N::max(max_cost - N::sqrt(sqd), T(0)), add(max_cost, T(1.1920928955078125e-07)));
      if (A.accumulate_stats && lane_id == 0) {
        st_v += (double)sv; st_v2 += (double)s2; st_vs += V;
        if (done || q == cfg.sims_per_step - 1) { st_rew += (double)reward; st_es += 1; st_cr += crash ? 1 : 0; }
      }
    }
    if (!stepped) reward = T(0);
    const bool last = (k == A.k_steps - 1);
    if (A.obs_seq != nullptr || last)
      write_obs(A.obs_seq ? A.obs_seq + ((size_t)k * E + e) * cfg.obs_dim : nullptr,
                last ? B.obs + (size_t)e * cfg.obs_dim : nullptr);
    if (lane_id == 0) {
      if (A.reward_seq) A.reward_seq[(size_t)k * E + e] = reward;
      if (A.done_seq) A.done_seq[(size_t)k * E + e] = done ? 1 : 0;
    }
  }

  if (A.write_views) {       // follower search reads every vehicle's headway: `dl` is free now
    __syncwarp();
#pragma unroll
    for (int r = 0; r < SLOTS; ++r) if (lane_id + 32 * r < V) dl[lane_id + 32 * r] = h[r];
    __syncwarp();
  }
#pragma unroll
  for (int r = 0; r < SLOTS; ++r) {
    const int s = lane_id + 32 * r;
    if (s >= V) continue;
    B.pos[eo + s] = x[r]; B.vel[eo + s] = v[r];
    if (A.write_views) {
      B.v_headway[eo + s] = h[r];
      B.v_leader[eo + s] = lead[r];
      int fj = -1; T fh = T(0);      // the closest vehicle that calls this one its leader
      for (int j = 0; j < V; ++j)
        if (ld[j] == s && (fj < 0 || dl[j] < fh)) { fj = j; fh = dl[j]; }
      B.v_follower[eo + s] = fj;
      if (A.k_steps > 0) B.v_accel[eo + s] = acc[r];
    }
  }
  if (lane_id == 0) {
    B.time_counter[e] = tc;
    B.done[e] = done ? 1 : 0;
    B.crash[e] = crash ? 1 : 0;
    B.reward[e] = reward;
    if (A.accumulate_stats) {
      double vals[6] = {st_rew, st_v, st_v2, st_vs, st_es, st_cr};
      for (int q = 0; q < 6; ++q) if (vals[q] != 0.0) atomicAdd(&stat_stripe(B.stats)[q], vals[q]);
    }
  }
}

// Env.reset: gen_even_start_pos (flow/core/kernel/network/base.py:256-392, edges_distribution
// 'all', one lane) walked sequentially by one thread per env in double, on the loop's segment
// table: x advances by increment + 5 + min_gap modulo the NON-internal length, is read as an
// edge through the edge-start table (get_edge), and a position inside a junction moves to the
// start of the next edge. The network's edge-start table also carries two entries "for aimsun"
// ('right_to_left' at 6.2 + 3 * 0.28 = 7.04 m, 'bottom_to_top' at the first junction pass) that
// get_edge reads like internal edges: every x from 7.04 m up to the start of 'top' is moved to
// the start of 'top', as on the reference. The perturbed / random placements come from the host
// (fs_reset_stage + fs_set_state, flow_b200/placement.py).
template <typename T>
__global__ void fig8_reset_kernel(const __grid_constant__ Cfg<T> cfg, const F8Bufs<T> B,
                                  const F8Geo g, const uint8_t* mask, T* length) {
  const int E = cfg.n_envs, V = cfg.n_veh;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  if (mask != nullptr && mask[e] == 0) return;
  const size_t eo = (size_t)e * V;
  const double min_gap = fmax(0.0, (double)cfg.init_min_gap), efs = add(min_gap, 5.0);
  PySum ps;             // usable real edges in get_edge_list order: bottom top right left upper lower
  const int order[6] = {1, 3, 7, 9, 5, 11};
  for (int q = 0; q < 6; ++q) if (g.len[order[q]] > efs) ps.push(g.len[order[q]]);
  // explicit roundings (no FMA contraction): Python rounds every product and every sum
  const double available = add(add(ps.value(), -(double)cfg.bunching), -mul((double)V, efs));
  const double increment = available / V;
  double xx = (double)cfg.x0;
  for (int i = 0; i < V; ++i) {
    int k = 0;
    for (int q = 1; q < 12; ++q) k = xx >= g.start[q] ? q : k;
    double p = xx - g.start[k];
    if (xx >= g.aimsun && xx < g.start[3]) { k = 3; xx = g.start[3]; p = 0.0; }
    for (int guard = 0; guard < 24; ++guard) {
      if ((k & 1) == 0) {            // internal: the start of the next edge in the table
        k = k == 11 ? 0 : k + 1;
        xx = g.start[k]; p = 0.0;
        continue;
      }
      if (!(g.len[k] > efs)) {       // an edge too short to hold a vehicle
        xx = fmod(add(xx, g.len[k]), g.non_internal);
        k = 0;
        for (int q = 1; q < 12; ++q) k = xx >= g.start[q] ? q : k;
        p = xx - g.start[k];
        continue;
      }
      break;
    }
    B.pos[eo + i] = (T)add(g.start[k], p);
    B.vel[eo + i] = T(0);
    xx = fmod(add(add(add(xx, increment), 5.0), min_gap), g.non_internal);
  }
  length[e] = (T)g.L;
  B.time_counter[e] = 0;
  B.episode[e] = B.episode[e] + 1;
  B.done[e] = 0; B.crash[e] = 0;
}

// host-side dispatch (instantiated in fs_f8_f32.cu / fs_f8_f64.cu)
template <typename K>
static cudaError_t f8_smem_attr(K kernel, size_t smem) {
  return smem > 48 * 1024 ? cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 (int)smem) : cudaSuccess;
}
template <typename T>
static cudaError_t launch_f8_impl(int V, int grid, int block, size_t smem, cudaStream_t st,
                                  const Cfg<T>& c, const F8Bufs<T>& b, const StepArgs<T>& a) {
#define FS_F8(S)                                                                   \
  do {                                                                             \
    cudaError_t rc = f8_smem_attr(fig8_kernel<T, S>, smem);                        \
    if (rc != cudaSuccess) return rc;                                              \
    fig8_kernel<T, S><<<grid, block, smem, st>>>(c, b, a);                         \
  } while (0)
  if (V <= 32) FS_F8(1);
  else if (V <= 64) FS_F8(2);
  else if (V <= 128) FS_F8(4);
  else FS_F8(8);
#undef FS_F8
  return cudaGetLastError();
}
template <typename T>
static cudaError_t launch_f8_reset_impl(int E, cudaStream_t st, const Cfg<T>& c, const F8Bufs<T>& b,
                                        const F8Geo& g, const uint8_t* mask, T* length) {
  fig8_reset_kernel<T><<<(E + 127) / 128, 128, 0, st>>>(c, b, g, mask, length);
  return cudaGetLastError();
}

}  // namespace fs
