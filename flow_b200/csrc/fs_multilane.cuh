// fs_multilane.cuh — multi-lane ring step kernel (config c3: LaneChangeAccelEnv,
// 3 lanes x 22 columns = 66 vehicles, IDM humans + RL vehicles that change lane).
//
// Mapping: ONE WARP OWNS ONE ENVIRONMENT; vehicle i lives in lane (i & 31),
// register slot (i >> 5). State is env-major here (pos[E][V], vehicle index
// fastest) so a warp's loads/stores are contiguous. Every step the warp
// publishes (x, v, lane) of its vehicles to its slice of shared memory and each
// thread finds, for each of its vehicles, the nearest vehicle ahead IN THE SAME
// LANE (what SUMO's getLeader returns through
// flow/core/kernel/vehicle/traci.py:218-250) by scanning that slice with
// broadcast reads; lane sets are re-derived every step, so lane changes need
// no list maintenance. Reductions (collision flag, reward norm) are warp
// shuffles / ballots.
//
// Restated from the reference:
//   LaneChangeAccelEnv._apply_rl_actions   flow/envs/ring/lane_change_accel.py:132-154
//   apply_lane_change (clip to lanes)      flow/core/kernel/vehicle/traci.py:965-997
//   get_last_lc returns the headway (bug)  flow/core/kernel/vehicle/traci.py:604-614
//   LaneChangeAccelEnv.get_state / reward  flow/envs/ring/lane_change_accel.py:100-130
// Restated SUMO part ("parity unpinned"): a requested change
// happens in the same simulation step unless the vehicle would overlap its new
// leader or follower in the target lane; humans (SimLaneChangeController /
// lane-change mode 512) never change lane.
#pragma once
#include "fs_common.cuh"
#include "fs_ring.cuh"

namespace fs {

template <typename T>
struct MlBufs {
  T* pos; T* vel; uint8_t* lane; T* circ;     // [E][V] (vehicle fastest), [E]
  T* st0;                                     // [V][E] controller state or null
  T* hist; double* hist_sum; int* hist_cnt; int hist_len, n_pi;
  int* last_lc;                               // [E][V] step of the last lane change
  int* time_counter; int* episode;
  uint8_t* done; uint8_t* crash;
  T* obs; T* reward;
  T* v_accel; T* v_headway; int* v_leader; int* v_follower;  // views [E][V] or null
  double* stats;
  int lc_bug_compat;
};

// shared bytes of ONE environment's slice (W = warps that share the env)
template <typename T>
__host__ __device__ inline size_t ml_smem_per_env(int V, int W) {
  const size_t b = (size_t)V * (5 * sizeof(T) + 5 * sizeof(int)) + 16 * sizeof(T) +
                   (16 + 8 + 8 * 8 * (size_t)W) * sizeof(int);
  return (b + 15) & ~(size_t)15;   // slices of several envs in one block stay aligned for T
}

// W = 1 (default): one warp owns an env, vehicle i lives in lane (i & 31), slot (i >> 5);
//        several envs per block.
// W > 1 (FS_ML_SPLIT=1, experiment kept under test): ONE BLOCK of W warps owns an env,
//        vehicle i = tid + 32 W s; warp collectives become block collectives (__syncthreads,
//        shared-memory partials). Same results; slower on c3 (see launch_ml in flowstep.cu).
// kGeneric = false: every type follows the IDM law (IDM, or the simulator model of SIM / RL
// vehicles): no controller switch, no follower search, no controller state
template <typename T, int SLOTS, int W, bool kGeneric = true>
__global__ void __launch_bounds__(128)
ring_ml_kernel(const __grid_constant__ Cfg<T> cfg, const MlBufs<T> B, const StepArgs<T> A) {
  typedef Num<T> N;
  constexpr int NT = 32 * W;       // threads per env
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane_id = threadIdx.x & 31;
  const int wid = W == 1 ? 0 : (int)(threadIdx.x >> 5);     // warp inside the env's group
  const int tid = W == 1 ? lane_id : (int)threadIdx.x;      // thread inside the env's group
  const int wib = W == 1 ? (int)(threadIdx.x >> 5) : 0;     // env slice inside the block
  const int E = cfg.n_envs, V = cfg.n_veh, L = cfg.n_lanes;
  const int e = W == 1 ? blockIdx.x * (blockDim.x >> 5) + wib : (int)blockIdx.x;
  if (e >= E) return;                                  // block-uniform when W > 1
  if (A.mask != nullptr && A.mask[e] == 0) return;
  // per-env shared slice: xs vs hs cx lens [V] redT[16] (T) | ls ld cidx rq_i rq_t [V]
  //                       loff[16] rq_n[8] wcnt[8][8][W] (int)
  const size_t per_env = ml_smem_per_env<T>(V, W);
  unsigned char* mine = smem_raw + per_env * wib;
  T* xs = reinterpret_cast<T*>(mine);
  T* vs = xs + V;
  T* hs = vs + V;
  T* cx = hs + V;                  // positions grouped by lane (index order inside a lane)
  T* lens = cx + V;                // vehicle lengths (static; one LDS instead of two indexed LDCs)
  T* redT = lens + V;              // W > 1: per-warp partial sums
  int* ls = reinterpret_cast<int*>(redT + 16);
  int* ld = ls + V;
  int* cidx = ld + V;              // vehicle index of every entry of cx
  int* rq_i = cidx + V;            // W > 1: lane-change requests of this step (vehicle, target)
  int* rq_t = rq_i + V;
  int* loff = rq_t + V;            // lane l occupies entries [loff[l], loff[l+1])
  int* rq_n = loff + 16;
  int* wcnt = rq_n + 8;            // W > 1: members of lane l in slot s of warp w

  auto grp_sync = [&]() { if (W == 1) __syncwarp(); else __syncthreads(); };
  // sums over the env's threads (W > 1: warp tree, then the W partials in warp order)
  auto grp_sum4 = [&](T& a0, T& a1, T& a2, T& a3) {
    for (int o = 16; o > 0; o >>= 1) {
      a0 += __shfl_xor_sync(0xffffffffu, a0, o);
      a1 += __shfl_xor_sync(0xffffffffu, a1, o);
      a2 += __shfl_xor_sync(0xffffffffu, a2, o);
      a3 += __shfl_xor_sync(0xffffffffu, a3, o);
    }
    if (W > 1) {
      if (lane_id == 0) { redT[wid] = a0; redT[4 + wid] = a1; redT[8 + wid] = a2; redT[12 + wid] = a3; }
      __syncthreads();
      a0 = redT[0]; a1 = redT[4]; a2 = redT[8]; a3 = redT[12];
      for (int w = 1; w < W; ++w) { a0 += redT[w]; a1 += redT[4 + w]; a2 += redT[8 + w]; a3 += redT[12 + w]; }
      __syncthreads();
    }
  };
  auto grp_any = [&](bool p) -> bool {
    if (W == 1) return __any_sync(0xffffffffu, p);
    return __syncthreads_or(p ? 1 : 0) != 0;
  };
  auto grp_count = [&](int n) -> int {   // sum of small per-thread counts
    for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    if (W > 1) {
      if (lane_id == 0) rq_n[1 + wid] = n;
      __syncthreads();
      n = 0;
      for (int w = 0; w < W; ++w) n += rq_n[1 + w];
      __syncthreads();
    }
    return n;
  };

  T x[SLOTS], v[SLOTS], h[SLOTS], acc[SLOTS];
  int ln[SLOTS], lead[SLOTS], foll[SLOTS];
  const T C = B.circ[e];
  const T dt = cfg.dt;
  const size_t eo = (size_t)e * V;
#pragma unroll
  for (int s = 0; s < SLOTS; ++s) {
    const int i = tid + NT * s;
    const bool ok = i < V;
    if (ok) lens[i] = cfg.types[cfg.type_of[i]].length;
    x[s] = ok ? B.pos[eo + i] : T(0);
    v[s] = ok ? B.vel[eo + i] : T(0);
    ln[s] = ok ? (int)B.lane[eo + i] : -1;
    acc[s] = T(0); foll[s] = -1;
  }

  if (W > 1 && tid == 0) rq_n[0] = 0;
  auto publish = [&]() {
    grp_sync();
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) {
      const int i = tid + NT * s;
      if (i < V) { xs[i] = x[s]; vs[i] = v[s]; ls[i] = ln[s]; }
    }
    grp_sync();
  };
  // k.update: nearest same-lane vehicle ahead; ties -> lowest index (oracle order).
  // The warp first groups the vehicles by lane (ballot compaction, index order kept), so a
  // vehicle scans the members of its own lane only: V/L entries instead of V, no lane test.
  const unsigned lt_mask = (1u << lane_id) - 1u;
  auto update = [&](bool need_followers) {
    publish();
    if (W == 1) {
      int base = 0;
      for (int l = 0; l < L; ++l) {
        if (lane_id == 0) loff[l] = base;
#pragma unroll
        for (int s = 0; s < SLOTS; ++s) {
          const int i = tid + NT * s;
          const bool m = i < V && ln[s] == l;
          const unsigned b = __ballot_sync(0xffffffffu, m);
          if (m) { const int p = base + __popc(b & lt_mask); cidx[p] = i; cx[p] = x[s]; }
          base += __popc(b);
        }
      }
      if (lane_id == 0) loff[L] = base;
      __syncwarp();
    } else {
      // members of lane l in (slot s, warp w), then exclusive offsets in (l, s, w) order,
      // which is index order inside a lane (i = lane + 32 w + 32 W s)
      int rk[SLOTS];
      for (int l = 0; l < L; ++l) {
#pragma unroll
        for (int s = 0; s < SLOTS; ++s) {
          const bool m = tid + NT * s < V && ln[s] == l;
          const unsigned b = __ballot_sync(0xffffffffu, m);
          if (m) rk[s] = __popc(b & lt_mask);
          if (lane_id == 0) wcnt[(l * SLOTS + s) * W + wid] = __popc(b);
        }
      }
      __syncthreads();
      if (tid <= L) {
        int base = 0;
        for (int q = 0; q < tid * SLOTS * W; ++q) base += wcnt[q];
        loff[tid] = base;
      }
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) {
        const int i = tid + NT * s;
        if (i < V) {
          const int mine_q = (ln[s] * SLOTS + s) * W + wid;
          int p = rk[s];
          for (int q = 0; q < mine_q; ++q) p += wcnt[q];
          cidx[p] = i; cx[p] = x[s];
        }
      }
      __syncthreads();
    }
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) {
      const int i = tid + NT * s;
      T best = T(0); int bj = -1;
      if (i < V) {
        const int p1 = loff[ln[s] + 1];
        const T xi = x[s];
#pragma unroll 4
        for (int p = loff[ln[s]]; p < p1; ++p) {
          const int j = cidx[p];
          const T d = fwd_dist(cx[p], xi, C);
          if (j != i && (bj < 0 || d < best)) { best = d; bj = j; }
        }
      }
      lead[s] = bj;
      h[s] = (bj >= 0) ? best - lens[bj] : T(1000);
    }
    if (need_followers) {  // vehicle/traci.py:243-250: closest vehicle that calls me leader
      grp_sync();
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) {
        const int i = tid + NT * s;
        if (i < V) { ld[i] = lead[s]; hs[i] = h[s]; }
      }
      grp_sync();
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) {
        const int i = tid + NT * s;
        int f = -1; T fh = T(0);
        if (i < V) {   // only a vehicle of my lane can call me leader
          const int p1 = loff[ln[s] + 1];
          for (int p = loff[ln[s]]; p < p1; ++p) {
            const int j = cidx[p];
            if (ld[j] == i && (f < 0 || hs[j] < fh)) { f = j; fh = hs[j]; }
          }
        }
        foll[s] = f;
      }
    }
  };
  const bool want_foll = (kGeneric && (cfg.any_follower & 1)) || A.write_views;
  update(want_foll);

  int tc = B.time_counter[e];
  const int episode = B.episode[e];
  bool done = A.ignore_done ? false : (B.done[e] != 0);
  bool crash = B.crash[e] != 0;
  const int t_end = cfg.sims_per_step * (cfg.warmup_steps + cfg.horizon);
  const long long gid = cfg.env_id_base + e;
  T reward = B.reward[e];
  const bool have_act = A.actions != nullptr;
  const bool lc_env = cfg.env_kind == FS_ENV_LANE_CHANGE_ACCEL ||
                      cfg.env_kind == FS_ENV_LANE_CHANGE_ACCEL_PO;
  int last_lc[SLOTS];
#pragma unroll
  for (int s = 0; s < SLOTS; ++s) {
    const int i = tid + NT * s;
    last_lc[s] = (i < V) ? B.last_lc[eo + i] : 0;
  }
  T max_cost;
  {
    T mc2 = T(0);
    const T t2 = mul(cfg.target_velocity, cfg.target_velocity);
    for (int i = 0; i < V; ++i) mc2 = add(mc2, t2);
    max_cost = N::sqrt(mc2);
  }
  double st_rew = 0, st_v = 0, st_v2 = 0, st_vs = 0, st_es = 0, st_cr = 0;

  const bool po_env = cfg.env_kind == FS_ENV_LANE_CHANGE_ACCEL_PO;
  auto rl_vehicle = [&](int r) {  // r-th RL vehicle in get_rl_ids order
    for (int i = 0; i < V; ++i) if (cfg.rl_index[i] == r) return i;
    return 0;
  };

  auto write_obs = [&](T* o, T* ol) {
    if (po_env) {
      // LaneChangeAccelPOEnv.get_state (lane_change_accel.py:215-263). Per RL vehicle a
      // block [headway_l | tailway_l | v_leader_l | v_follower_l] (l = lane), then ego speed.
      // As the reference BEHAVES (default): the `return` sits inside the loop over RL
      // vehicles, so only the FIRST block is filled, the others stay 0 and ONE raw ego
      // speed follows (4*n_rl*L + 1 values); headways / tailways are copied before the
      // in-place normalisation, i.e. they stay in metres (1000 = empty lane).
      // lc_corrected: every block, distances / network length (1 when the lane is empty),
      // n_rl ego speeds / max_speed (4*n_rl*L + n_rl values, the declared observation_space).
      // Lane neighbours = TraCIVehicle._multi_lane_headways on a closed loop
      // (vehicle/traci.py:699-937), scanned in the slice published by update().
      const bool fixed = !B.lc_bug_compat;
      const int nb = 4 * cfg.n_rl * L;
      const int n_blk = fixed ? cfg.n_rl : 1;
      for (int q = 4 * L * n_blk + tid; q < nb; q += NT) {
        if (o) o[q] = T(0);
        if (ol) ol[q] = T(0);
      }
      for (int r = tid; r < n_blk; r += NT) {
        const T sp = vs[rl_vehicle(r)];
        const T ego = fixed ? N::div(sp, cfg.speed_limit) : sp;
        if (o) o[nb + r] = ego;
        if (ol) ol[nb + r] = ego;
      }
      for (int q = tid; q < n_blk * L; q += NT) {
        const int r = q / L, lq = q - r * L;
        const int me = rl_vehicle(r);
        const T p = xs[me];
        int jl = -1, jf = -1;
        T bd = T(0), bb = T(0);
        for (int j = 0; j < V; ++j) {
          if (ls[j] != lq || j == me) continue;
          T d = xs[j] - p;             // a vehicle level with us in another lane is ahead (d = 0)
          if (d < T(0)) d = d + C;
          if (jl < 0 || d < bd) { jl = j; bd = d; }          // ties: lowest index
          T b = p - xs[j];
          if (!(xs[j] < p)) b = b + C;
          if (jf < 0 || b <= bb) { jf = j; bb = b; }         // ties: highest index
        }
        if (jl < 0 && ls[me] == lq) { jl = jf = me; bd = bb = C; }  // alone: itself, one lap away
        T hw = jl >= 0 ? bd - lens[jl] : T(1000);
        T tw = jf >= 0 ? bb - lens[me] : T(1000);
        if (fixed) {
          hw = jl >= 0 ? N::div(hw, C) : T(1);
          tw = jf >= 0 ? N::div(tw, C) : T(1);
        }
        const T vl = jl >= 0 ? N::div(vs[jl], cfg.speed_limit) : T(0);
        const T vf = jf >= 0 ? N::div(vs[jf], cfg.speed_limit) : T(0);
        const int at = 4 * L * r + lq;
        if (o) { o[at] = hw; o[at + L] = tw; o[at + 2 * L] = vl; o[at + 3 * L] = vf; }
        if (ol) { ol[at] = hw; ol[at + L] = tw; ol[at + 2 * L] = vl; ol[at + 3 * L] = vf; }
      }
      return;
    }
    // LaneChangeAccelEnv.get_state (lane_change_accel.py:114-130) / AccelEnv.get_state
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) {
      const int i = tid + NT * s;
      if (i >= V) continue;
      const T a0 = N::div(v[s], cfg.speed_limit), a1 = N::div(x[s], C);
      const T a2 = N::div(T(ln[s]), T(L));
      if (o) { o[i] = a0; o[V + i] = a1; if (lc_env) o[2 * V + i] = a2; }
      if (ol) { ol[i] = a0; ol[V + i] = a1; if (lc_env) ol[2 * V + i] = a2; }
    }
  };
  if (A.k_steps == 0) write_obs(nullptr, B.obs + (size_t)e * cfg.obs_dim);

  // action rows of this thread's RL vehicles, fetched one step ahead of their use
  T nxt_a[SLOTS], nxt_d[SLOTS];
  auto fetch_actions = [&](int k) {
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) {
      const int i = tid + NT * s;
      nxt_a[s] = T(0); nxt_d[s] = T(0);
      if (have_act && i < V && cfg.rl_index[i] >= 0) {
        const int slot = cfg.rl_index[i];
        const size_t arow = (size_t)k * cfg.act_dim + (lc_env ? 2 * slot : slot);
        nxt_a[s] = A.actions[arow * E + e];
        if (lc_env) nxt_d[s] = A.actions[(arow + 1) * E + e];
      }
    }
  };
  if (A.k_steps > 0) fetch_actions(0);

  for (int k = 0; k < A.k_steps; ++k) {
    T act_a[SLOTS], act_d[SLOTS];
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) { act_a[s] = nxt_a[s]; act_d[s] = nxt_d[s]; }
    if (k + 1 < A.k_steps) fetch_actions(k + 1);
    if (!done) {
      T mean_v = T(0);
      if (kGeneric && (cfg.any_follower & 2)) {
        T sm = T(0);
#pragma unroll
        for (int s = 0; s < SLOTS; ++s) if (tid + NT * s < V) sm = add(sm, v[s]);
        T z1 = T(0), z2 = T(0), z3 = T(0);
        grp_sum4(sm, z1, z2, z3);
        mean_v = N::div(sm, T(V));
      }
      int new_ln[SLOTS];
      T act_abs_sum = T(0);
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) new_ln[s] = ln[s];
      if (lc_env && have_act) {
        // LaneChangeAccelEnv._apply_rl_actions + apply_lane_change + the restated execution.
        // Requests are checked by the WHOLE warp, one requesting vehicle at a time: every
        // thread tests its share of the target lane's members (pre-move lane lists).
#pragma unroll
        for (int s = 0; s < SLOTS; ++s) {
          const int i = tid + NT * s;
          int target = -1;
          T len_i = T(0);
          if (i < V && cfg.rl_index[i] >= 0) {
            // direction zeroed while time_counter <= lane_change_duration + get_last_lc
            // (== headway under the reference bug)
            T dirf = N::min(N::max(act_d[s], T(-1)), T(1));
            const T last = B.lc_bug_compat ? h[s] : T(last_lc[s]);
            if (T(tc + 1) <= add(T(cfg.lane_change_duration), last)) dirf = T(0);
            const int dir = dirf > T(0.5) ? 1 : (dirf < T(-0.5) ? -1 : 0);
            int t = ln[s] + dir;
            t = t < 0 ? 0 : (t > L - 1 ? L - 1 : t);
            if (t != ln[s]) target = t;
            len_i = lens[i];
          }
          if (W > 1) {   // block-wide: queue the request, checked below by all W warps
            if (target >= 0) { const int q = atomicAdd(rq_n, 1); rq_i[q] = i; rq_t[q] = target; }
            continue;
          }
          unsigned req = __ballot_sync(0xffffffffu, target >= 0);
          while (req) {
            const int src = __ffs(req) - 1;
            req &= req - 1;
            const int t = __shfl_sync(0xffffffffu, target, src);
            const T xi = __shfl_sync(0xffffffffu, x[s], src);
            const T li = __shfl_sync(0xffffffffu, len_i, src);
            const int p1 = loff[t + 1];
            bool bad = false;   // no overlap with any vehicle of the target lane
            for (int p = loff[t] + lane_id; p < p1; p += 32) {
              const T fwd = fwd_dist(cx[p], xi, C), back = fwd_dist(xi, cx[p], C);
              bad |= fwd - lens[cidx[p]] < T(0) || back - li < T(0);
            }
            const bool ok = !__any_sync(0xffffffffu, bad);
            if (lane_id == src && ok) { new_ln[s] = t; last_lc[s] = tc + 1; }
          }
        }
        if (W > 1) {
          __syncthreads();
          const int n_req = rq_n[0];
          for (int r = 0; r < n_req; ++r) {   // requests are independent: all read pre-move lists
            const int i = rq_i[r], t = rq_t[r];
            const T xi = xs[i], li = lens[i];
            const int p1 = loff[t + 1];
            bool bad = false;
            for (int p = loff[t] + tid; p < p1; p += NT) {
              const T fwd = fwd_dist(cx[p], xi, C), back = fwd_dist(xi, cx[p], C);
              bad |= fwd - lens[cidx[p]] < T(0) || back - li < T(0);
            }
            const bool ok = !grp_any(bad);
#pragma unroll
            for (int s = 0; s < SLOTS; ++s)
              if (ok && tid + NT * s == i) { new_ln[s] = t; last_lc[s] = tc + 1; }
          }
          __syncthreads();
          if (tid == 0) rq_n[0] = 0;
        }
      }
      T zn[SLOTS > 4 ? 8 : 4];
      if (cfg.any_noise) {   // one Philox call = four normals = the thread's first four slots
        if (W == 1) {
          normal4<T>(cfg.seed, gid, lane_id, tc + episode * 1000003, zn);
          if (SLOTS > 4) normal4<T>(cfg.seed, gid, lane_id + 32, tc + episode * 1000003, zn + 4);
        } else {
          // the same stream whatever W: vehicle i draws output (i >> 5) & 3 of call (i & 31) + 32 (i >> 7)
#pragma unroll
          for (int s = 0; s < SLOTS; ++s) {
            const int i = tid + NT * s;
            T z4[4];
            normal4<T>(cfg.seed, gid, (i & 31) + 32 * (i >> 7), tc + episode * 1000003, z4);
            zn[s] = z4[(i >> 5) & 3];
          }
        }
      }
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) {
        const int i = tid + NT * s;
        if (i >= V) continue;
        const TypeP<T>& tp = cfg.types[cfg.type_of[i]];
        const int ctrl = tp.ctrl;
        const bool has_lead = lead[s] >= 0;
        const T vi = v[s], hi = h[s];
        const T vl = has_lead ? vs[lead[s]] : T(-1001);
        T a;
        if (ctrl == FS_CTRL_RL && have_act) {
          const T raw = act_a[s];
          a = N::min(N::max(raw, cfg.act_lo), cfg.act_hi);
          act_abs_sum = add(act_abs_sum, N::abs(cfg.clip_actions ? a : raw));
        } else if (!kGeneric) {   // same arithmetic as ctrl_accel's IDM / SIM cases
          a = idm_accel<T>(vi, vl, hi, has_lead, tp.e_v0, tp.e_inv_v0, tp.e_T, tp.e_a, tp.e_tsab,
                           tp.e_inv_tsab, tp.e_delta, tp.e_s0);
        } else if (ctrl == FS_CTRL_PI_SATURATION) {
          T v_cmd = B.st0[(size_t)i * E + e];
          RingBufs<T> rb;
          rb.hist = B.hist; rb.hist_sum = B.hist_sum; rb.hist_cnt = B.hist_cnt;
          rb.hist_len = B.hist_len; rb.n_pi = B.n_pi;
          a = pi_saturation<T>(rb, cfg.pi_index[i], E, e, tc, vi, vl, hi, dt, tp.max_accel, v_cmd);
          B.st0[(size_t)i * E + e] = v_cmd;
        } else {
          T st0 = T(0);
          const bool stateful = (ctrl == FS_CTRL_LAC);
          if (stateful) st0 = B.st0[(size_t)i * E + e];
          const bool has_f = foll[s] >= 0;
          a = ctrl_accel<T>(tp, ctrl, vi, has_lead, vl, hi, has_f, has_f ? vs[foll[s]] : T(-1001),
                            has_f ? hs[foll[s]] : T(-1001), mean_v, dt, cfg.inv_dt, st0);
          if (stateful) B.st0[(size_t)i * E + e] = st0;
        }
        if (ctrl != FS_CTRL_SIM && ctrl != FS_CTRL_RL) {
          if (tp.noise > T(0))
            a = add(a, mul(tp.noise, zn[s]));
          if (tp.fail_safe == FS_FAILSAFE_INSTANTANEOUS)
            a = failsafe_instantaneous<T>(a, vi, hi, has_lead, V, dt);
          else if (tp.fail_safe == FS_FAILSAFE_SAFE_VELOCITY)
            a = failsafe_safe_velocity<T>(a, vi, vl, hi, V, dt, tp.delay);
        }
        acc[s] = a;
      }
      // all reads of the published pre-move state are done; move
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) {
        if (tid + NT * s >= V) continue;
        const T vn = N::max(add(v[s], mul(acc[s], dt)), T(0));
        x[s] = wrap_fwd(add(x[s], move_dist<T>(v[s], vn, dt, cfg.ballistic)), C);
        v[s] = vn;
        ln[s] = new_ln[s];
      }
      update(want_foll);
      bool cr = false;
      T sv = T(0), sq = T(0), s2 = T(0);
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) {
        const int i = tid + NT * s;
        if (i >= V) continue;
        if (cfg.collision_min_gap)
          cr |= lead[s] >= 0 && h[s] < cfg.types[cfg.type_of[lead[s]]].sim_min_gap;
        else
          cr |= h[s] < T(0);
        sv = add(sv, v[s]);
        const T d = v[s] - cfg.target_velocity;
        sq = add(sq, mul(d, d));
        s2 += v[s] * v[s];
      }
      grp_sum4(sv, sq, s2, act_abs_sum);
      crash = grp_any(cr);
      tc += 1;
      done = crash || tc >= t_end;
      if (cfg.env_kind == FS_ENV_WAVE_ATTENUATION || cfg.env_kind == FS_ENV_WAVE_ATTENUATION_PO) {
        if (!have_act || crash) reward = T(0);
        else {
          reward = N::div(mul(T(4), N::div(sv, T(V))), T(20));
          const T mean_act = N::div(act_abs_sum, T(cfg.n_rl > 0 ? cfg.n_rl : 1));
          if (mean_act > T(0)) reward = add(reward, mul(T(4), T(0) - mean_act));
        }
      } else {
        // rewards.desired_velocity (+ LaneChangeAccelEnv lane-change penalty);
        // AccelEnv with env_params.evaluate: mean speed (accel.py:111-112)
        if (cfg.evaluate && !lc_env) reward = N::div(sv, T(V));
        else if (crash) reward = T(0);
        else {
          const T cost = N::sqrt(sq);
          const T eps = T(1.1920928955078125e-07);
          reward = N::div(N::max(max_cost - cost, T(0)), add(max_cost, eps));
        }
        if (lc_env) {
          int pen = 0;
#pragma unroll
          for (int s = 0; s < SLOTS; ++s) {
            const int i = tid + NT * s;
            if (i < V && cfg.rl_index[i] >= 0) {
              const T last = B.lc_bug_compat ? h[s] : T(last_lc[s]);
              pen += (last == T(tc)) ? 1 : 0;
            }
          }
          pen = grp_count(pen);
          for (int q = 0; q < pen; ++q) reward = reward - T(0.1);
        }
      }
      if (A.accumulate_stats && tid == 0) {
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
    if (tid == 0) {
      if (A.reward_seq) A.reward_seq[(size_t)k * E + e] = reward;
      if (A.done_seq) A.done_seq[(size_t)k * E + e] = done ? 1 : 0;
    }
  }

#pragma unroll
  for (int s = 0; s < SLOTS; ++s) {
    const int i = tid + NT * s;
    if (i >= V) continue;
    B.pos[eo + i] = x[s]; B.vel[eo + i] = v[s]; B.lane[eo + i] = (uint8_t)ln[s];
    B.last_lc[eo + i] = last_lc[s];
    if (A.write_views) {
      B.v_headway[eo + i] = h[s]; B.v_leader[eo + i] = lead[s]; B.v_follower[eo + i] = foll[s];
      if (A.k_steps > 0) B.v_accel[eo + i] = acc[s];
    }
  }
  if (tid == 0) {
    B.time_counter[e] = tc;
    B.done[e] = done ? 1 : 0;
    B.crash[e] = crash ? 1 : 0;
    B.reward[e] = reward;
    if (A.accumulate_stats) {
      const double vals[6] = {st_rew, st_v, st_v2, st_vs, st_es, st_cr};
      for (int q = 0; q < 6; ++q) if (vals[q] != 0.0) atomicAdd(&stat_stripe(B.stats)[q], vals[q]);
    }
  }
}

}  // namespace fs
