// fs_multilane.cuh — multi-lane ring step kernel (config c3: LaneChangeAccelEnv,
// 3 lanes x 22 columns = 66 vehicles, IDM humans + RL vehicles that change lane).
//
// Mapping: ONE WARP OWNS ONE ENVIRONMENT; vehicle i lives in lane (i & 31),
// register slot (i >> 5). State is env-major here (pos[E][V], vehicle index
// fastest) so a warp's loads/stores are contiguous.
//
// Per-lane ordering (north_star "leaders are found by per-lane ordering"): inside
// a lane vehicles cannot overtake, so the cyclic order of a lane's members only
// changes when a vehicle CHANGES LANE (RL vehicles only: humans run
// SimLaneChangeController under lane-change mode 512 and never do). Every vehicle
// therefore keeps the index of its leader in a REGISTER across the steps of a
// launch. A step publishes the moved (x, v) to the warp's shared-memory slice and
// reads the leader's entry back — two stores and two loads per vehicle, no search.
// The ordered lane lists are (re)built — ballot + popc compaction by lane, index
// order kept inside a lane, then a scan of the own lane for the nearest vehicle
// ahead, ties to the lowest index like the reference's sort + bisect_left
// (flow/core/kernel/vehicle/traci.py:723-738, 826-841) — only
//   * once at the top of a launch (state may have been set from outside),
//   * in a step in which a lane change was executed,
//   * if some headway jumped by more than half a lap (a vehicle got past its
//     leader inside one step; the oracle's per-step search would re-order), and
//   * every step when the views (leader / follower / headway) are recorded.
//
// Restated from the reference:
//   LaneChangeAccelEnv._apply_rl_actions   flow/envs/ring/lane_change_accel.py:132-154
//   apply_lane_change (clip to lanes)      flow/core/kernel/vehicle/traci.py:965-997
//   get_last_lc returns the headway (bug)  flow/core/kernel/vehicle/traci.py:604-614
//   LaneChangeAccelEnv.get_state / reward  flow/envs/ring/lane_change_accel.py:100-130
// Restated SUMO part ("parity unpinned"): a requested change
// happens in the same simulation step unless the vehicle would overlap its new
// leader or follower in the target lane; humans never change lane.
#pragma once
#include "fs_common.cuh"
#include "fs_ring_fast.cuh"

namespace fs {

// kGeneric = false: host-precomputed constants of a type table that holds ONE IDM-law human
// type (IDMController, or SimCarFollowingController = the restated simulator model) and at most
// one RL type, all of the same length (what every shipped multi-lane config builds)
template <typename T>
struct MlFast {
  IdmRegs<T> pa, pb;     // human type / un-commanded RL type (simulator default model)
  IdmRegs<T> ps;         // simulator model of the HUMAN type (its speed mode's safe speed)
  T len, noise, delay, min_gap;
  T cap_a, floor_a, cap_b, floor_b;   // speed-mode bits 1 / 2 of the two types (+-inf when off)
  int safe_a, safe_b;                 // speed-mode bit 0
  T inv_speed_limit, inv_lanes, rew_inv;   // fp32 observation / reward reciprocals
  int fail_safe, human_is_sim;
};

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
  MlFast<T> fp;
};

template <typename T>
static inline MlFast<T> make_mlfast(const Cfg<T>& c) {
  MlFast<T> p;
  memset(&p, 0, sizeof(p));
  int ta = -1, tb = -1;
  for (int i = 0; i < c.n_veh; ++i) {
    if (c.rl_index[i] >= 0) tb = c.type_of[i]; else ta = c.type_of[i];
  }
  if (ta < 0) ta = tb < 0 ? 0 : tb;
  if (tb < 0) tb = ta;
  p.pa.load(c.types[ta]); p.pb.load(c.types[tb]);
  p.len = c.types[ta].length; p.noise = c.types[ta].noise; p.delay = c.types[ta].delay;
  p.min_gap = c.types[ta].sim_min_gap; p.fail_safe = c.types[ta].fail_safe;
  p.human_is_sim = c.types[ta].ctrl != FS_CTRL_IDM;
  load_sim_block(p.ps, c.types[ta]);
  speed_mode_consts(c.types[ta], p.safe_a, p.cap_a, p.floor_a);
  speed_mode_consts(c.types[tb], p.safe_b, p.cap_b, p.floor_b);
  p.inv_speed_limit = T(1) / c.speed_limit; p.inv_lanes = T(1) / T(c.n_lanes);
  T mc2 = T(0);
  for (int i = 0; i < c.n_veh; ++i) mc2 = mc2 + c.target_velocity * c.target_velocity;
  p.rew_inv = T(1) / ((T)sqrt((double)mc2) + T(1.1920928955078125e-07));
  return p;
}

// shared bytes of ONE environment's slice
template <typename T>
__host__ __device__ inline size_t ml_smem_per_env(int V) {
  const size_t b = (size_t)V * (5 * sizeof(T) + 3 * sizeof(int)) + 16 * sizeof(int);
  return (b + 15) & ~(size_t)15;   // slices of several envs in one block stay aligned for T
}

template <typename T, int SLOTS, bool kGeneric = true>
__global__ void __launch_bounds__(128)
ring_ml_kernel(const __grid_constant__ Cfg<T> cfg, const __grid_constant__ MlBufs<T> B,
               const StepArgs<T> A) {
  typedef Num<T> N;
  constexpr unsigned FULL = 0xffffffffu;
  constexpr bool kF32 = sizeof(T) == 4;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane_id = threadIdx.x & 31;
  const int wib = (int)(threadIdx.x >> 5);     // env slice inside the block
  const int E = cfg.n_envs, V = cfg.n_veh, L = cfg.n_lanes;
  const int e = blockIdx.x * (blockDim.x >> 5) + wib;
  if (e >= E) return;
  if (A.mask != nullptr && A.mask[e] == 0) return;
  // per-env shared slice: xs vs hs cx lens [V] (T) | ls ld cidx [V] loff[16] (int)
  unsigned char* mine = smem_raw + ml_smem_per_env<T>(V) * wib;
  T* xs = reinterpret_cast<T*>(mine);
  T* vs = xs + V;
  T* hs = vs + V;
  T* cx = hs + V;                  // positions grouped by lane (index order inside a lane)
  T* lens = cx + V;                // vehicle lengths (static)
  int* ls = reinterpret_cast<int*>(lens + V);
  int* ld = ls + V;
  int* cidx = ld + V;              // vehicle index of every entry of cx
  int* loff = cidx + V;            // lane l occupies entries [loff[l], loff[l+1])
  const MlFast<T>& P = B.fp;

  T x[SLOTS], v[SLOTS], h[SLOTS], vl[SLOTS], acc[SLOTS], aeff[SLOTS];
  int ln[SLOTS], lead[SLOTS], foll[SLOTS];
  int tt[SLOTS], rls[SLOTS];       // type index / action slot (-1: not RL) of my vehicles
  const T C = B.circ[e];
  const T half_lap = mul(T(0.5), C);
  const T dt = cfg.dt;
  const size_t eo = (size_t)e * V;
  unsigned rl_slots = 0;           // bit s: some thread of the warp has an RL vehicle in slot s
#pragma unroll
  for (int s = 0; s < SLOTS; ++s) {
    const int i = lane_id + 32 * s;
    const bool ok = i < V;
    tt[s] = ok ? (int)cfg.type_of[i] : 0;
    rls[s] = ok ? (int)cfg.rl_index[i] : -1;
    if (ok) lens[i] = cfg.types[tt[s]].length;
    x[s] = ok ? B.pos[eo + i] : T(0);
    v[s] = ok ? B.vel[eo + i] : T(0);
    ln[s] = ok ? (int)B.lane[eo + i] : -1;
    acc[s] = T(0); foll[s] = -1; lead[s] = -1; h[s] = T(1000); vl[s] = T(-1001);
    if (__any_sync(FULL, rls[s] >= 0)) rl_slots |= 1u << s;
  }
  auto len_of = [&](int j) -> T { return kGeneric ? lens[j] : P.len; };

  const unsigned lt_mask = (1u << lane_id) - 1u;
  // (re)build the ordered lane lists and every vehicle's leader (and follower on request)
  auto rebuild = [&](bool need_followers) {
    __syncwarp();
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) {
      const int i = lane_id + 32 * s;
      if (i < V) { xs[i] = x[s]; vs[i] = v[s]; ls[i] = ln[s]; }
    }
    int base = 0;
    for (int l = 0; l < L; ++l) {
      if (lane_id == 0) loff[l] = base;
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) {
        const int i = lane_id + 32 * s;
        const bool m = i < V && ln[s] == l;
        const unsigned b = __ballot_sync(FULL, m);
        if (m) { const int p = base + __popc(b & lt_mask); cidx[p] = i; cx[p] = x[s]; }
        base += __popc(b);
      }
    }
    if (lane_id == 0) loff[L] = base;
    __syncwarp();
    // nearest same-lane vehicle ahead; ties -> lowest index (oracle order)
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) {
      const int i = lane_id + 32 * s;
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
      h[s] = (bj >= 0) ? best - len_of(bj) : T(1000);
      vl[s] = (bj >= 0) ? vs[bj] : T(-1001);
    }
    if (need_followers) {  // vehicle/traci.py:243-250: closest vehicle that calls me leader
      __syncwarp();
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) {
        const int i = lane_id + 32 * s;
        if (i < V) { ld[i] = lead[s]; hs[i] = h[s]; }
      }
      __syncwarp();
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) {
        const int i = lane_id + 32 * s;
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
  // the step's k.update when no lane list changed: publish the moved state, read the kept
  // leader's entry back. Returns true if some headway wrapped by a lap (order broken).
  auto refresh = [&](bool need_followers) -> bool {
    __syncwarp();          // every read of the previous publish is done
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) {
      const int i = lane_id + 32 * s;
      if (i < V) { xs[i] = x[s]; vs[i] = v[s]; }
    }
    __syncwarp();
    bool odd = false;
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) {
      const int j = lead[s];
      if (j >= 0) {        // lead stays -1 for the idle slots
        const T hn = fwd_dist(xs[j], x[s], C) - len_of(j);
        vl[s] = vs[j];
        odd |= hn - h[s] > half_lap;
        h[s] = hn;
      }
    }
    if (need_followers) {  // BCM reads its follower's headway
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) {
        const int i = lane_id + 32 * s;
        if (i < V) hs[i] = h[s];
      }
      __syncwarp();
    }
    return __any_sync(FULL, odd);
  };
  const bool want_foll = (kGeneric && (cfg.any_follower & 1)) || A.write_views;
  rebuild(want_foll);

  int tc = B.time_counter[e];
  const int tc0 = tc;
  const int episode = B.episode[e];
  bool done = A.ignore_done ? false : (B.done[e] != 0);
  bool crash = B.crash[e] != 0;
  const int t_end = cfg.sims_per_step * (cfg.warmup_steps + cfg.horizon);
  const long long gid = cfg.env_id_base + e;
  T reward = B.reward[e];
  const bool have_act = A.actions != nullptr;
  const bool lc_env = cfg.env_kind == FS_ENV_LANE_CHANGE_ACCEL ||
                      cfg.env_kind == FS_ENV_LANE_CHANGE_ACCEL_PO;
  const bool wave_env = cfg.env_kind == FS_ENV_WAVE_ATTENUATION ||
                        cfg.env_kind == FS_ENV_WAVE_ATTENUATION_PO;
  int last_lc[SLOTS];
#pragma unroll
  for (int s = 0; s < SLOTS; ++s) {
    const int i = lane_id + 32 * s;
    last_lc[s] = (i < V) ? B.last_lc[eo + i] : 0;
  }
  T max_cost;
  {
    T mc2 = T(0);
    const T t2 = mul(cfg.target_velocity, cfg.target_velocity);
    for (int i = 0; i < V; ++i) mc2 = add(mc2, t2);
    max_cost = N::sqrt(mc2);
  }
  const T crash_thr = cfg.collision_min_gap ? P.min_gap : T(0);   // kGeneric = false only
  // rollout statistics: speeds as per-thread partial sums (reduced once at the end of the
  // launch), reward / crashes on lane 0
  double st_rew = 0, st_v = 0, st_v2 = 0, st_cr = 0;

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
      // (vehicle/traci.py:699-937), scanned in the published slice (xs / vs are refreshed by
      // every step, ls whenever a lane changed).
      const bool fixed = !B.lc_bug_compat;
      const int nb = 4 * cfg.n_rl * L;
      const int n_blk = fixed ? cfg.n_rl : 1;
      for (int q = 4 * L * n_blk + lane_id; q < nb; q += 32) {
        if (o) o[q] = T(0);
        if (ol) ol[q] = T(0);
      }
      for (int r = lane_id; r < n_blk; r += 32) {
        const T sp = vs[rl_vehicle(r)];
        const T ego = fixed ? N::div(sp, cfg.speed_limit) : sp;
        if (o) o[nb + r] = ego;
        if (ol) ol[nb + r] = ego;
      }
      for (int q = lane_id; q < n_blk * L; q += 32) {
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
        const T vlq = jl >= 0 ? N::div(vs[jl], cfg.speed_limit) : T(0);
        const T vfq = jf >= 0 ? N::div(vs[jf], cfg.speed_limit) : T(0);
        const int at = 4 * L * r + lq;
        if (o) { o[at] = hw; o[at + L] = tw; o[at + 2 * L] = vlq; o[at + 3 * L] = vfq; }
        if (ol) { ol[at] = hw; ol[at + L] = tw; ol[at + 2 * L] = vlq; ol[at + 3 * L] = vfq; }
      }
      return;
    }
    // LaneChangeAccelEnv.get_state (lane_change_accel.py:114-130) / AccelEnv.get_state
    const T inv_C = kF32 ? T(1) / C : T(0);
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) {
      const int i = lane_id + 32 * s;
      if (i >= V) continue;
      T a0, a1, a2;
      if (kF32 && !kGeneric) {   // reciprocal multiplies (fp32 tolerance 1e-4)
        a0 = v[s] * P.inv_speed_limit; a1 = x[s] * inv_C; a2 = T(ln[s]) * P.inv_lanes;
      } else {
        a0 = N::div(v[s], cfg.speed_limit); a1 = N::div(x[s], C); a2 = N::div(T(ln[s]), T(L));
      }
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
      nxt_a[s] = T(0); nxt_d[s] = T(0);
      if (!((rl_slots >> s) & 1u)) continue;   // warp-uniform
      if (have_act && rls[s] >= 0) {
        const size_t arow = (size_t)k * cfg.act_dim + (lc_env ? 2 * rls[s] : rls[s]);
        nxt_a[s] = __ldg(A.actions + arow * E + e);
        if (lc_env) nxt_d[s] = __ldg(A.actions + (arow + 1) * E + e);
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
        for (int s = 0; s < SLOTS; ++s) if (lane_id + 32 * s < V) sm = add(sm, v[s]);
        for (int o = 16; o > 0; o >>= 1) sm += __shfl_xor_sync(FULL, sm, o);
        mean_v = N::div(sm, T(V));
      }
      int new_ln[SLOTS];
      T act_abs_sum = T(0);
      bool changed = false;        // warp-uniform: a lane change is executed in this step
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) new_ln[s] = ln[s];
      if (lc_env && have_act) {
        // LaneChangeAccelEnv._apply_rl_actions + apply_lane_change + the restated execution.
        // Requests are checked by the WHOLE warp, one requesting vehicle at a time: every
        // thread tests its own vehicles of the target lane (pre-move registers).
#pragma unroll
        for (int s = 0; s < SLOTS; ++s) {
          if (!((rl_slots >> s) & 1u)) continue;   // warp-uniform
          int target = -1;
          if (rls[s] >= 0) {
            // direction zeroed while time_counter <= lane_change_duration + get_last_lc
            // (== headway under the reference bug)
            T dirf = N::min(N::max(act_d[s], T(-1)), T(1));
            const T last = B.lc_bug_compat ? h[s] : T(last_lc[s]);
            if (T(tc + 1) <= add(T(cfg.lane_change_duration), last)) dirf = T(0);
            const int dir = dirf > T(0.5) ? 1 : (dirf < T(-0.5) ? -1 : 0);
            int t = ln[s] + dir;
            t = t < 0 ? 0 : (t > L - 1 ? L - 1 : t);
            if (t != ln[s]) target = t;
          }
          unsigned req = __ballot_sync(FULL, target >= 0);
          while (req) {
            const int src = __ffs(req) - 1;
            req &= req - 1;
            const int t = __shfl_sync(FULL, target, src);
            const T xi = __shfl_sync(FULL, x[s], src);
            const T li = kGeneric ? lens[src + 32 * s] : P.len;
            bool bad = false;   // no overlap with any vehicle of the target lane
#pragma unroll
            for (int s2 = 0; s2 < SLOTS; ++s2) {
              const int i2 = lane_id + 32 * s2;
              if (i2 < V && ln[s2] == t) {
                const T fwd = fwd_dist(x[s2], xi, C), back = fwd_dist(xi, x[s2], C);
                bad |= fwd - len_of(i2) < T(0) || back - li < T(0);
              }
            }
            const bool ok = !__any_sync(FULL, bad);
            changed |= ok;
            if (lane_id == src && ok) { new_ln[s] = t; last_lc[s] = tc + 1; }
          }
        }
      }
      const bool noisy = kGeneric ? cfg.any_noise != 0 : P.noise > T(0);
      T zn[SLOTS > 4 ? 8 : 4];
      if (noisy) {   // one Philox call = four normals = the thread's first four slots
        normal4<T>(cfg.seed, gid, lane_id, tc + episode * 1000003, zn);
        if (SLOTS > 4) normal4<T>(cfg.seed, gid, lane_id + 32, tc + episode * 1000003, zn + 4);
      }
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) {
        const int i = lane_id + 32 * s;
        if (i >= V) continue;
        const bool has_lead = lead[s] >= 0;
        const T vi = v[s], hi = h[s];
        const T vlead = vl[s];       // -1001 without a leader
        T a, ae;             // commanded / what moves the vehicle (after SUMO's speed mode)
        if (!kGeneric) {
          // ONE IDM-law human type (+ RL): parameters are uniform kernel-argument constants
          if (rls[s] >= 0) {
            if (have_act) {
              const T raw = act_a[s];
              a = N::min(N::max(raw, cfg.act_lo), cfg.act_hi);
              act_abs_sum = add(act_abs_sum, N::abs(cfg.clip_actions ? a : raw));
              ae = speed_mode_fast<T>(a, vi, vlead, hi, has_lead, P.pb, P.safe_b, P.cap_b, P.floor_b);
            } else {   // un-commanded RL vehicle: restated simulator default model
              a = idm_accel<T>(vi, vlead, hi, has_lead, P.pb.v0, P.pb.inv_v0, P.pb.Tt, P.pb.a,
                               P.pb.tsab, P.pb.inv_tsab, P.pb.delta, P.pb.s0);
              ae = a;
            }
          } else {
            T a_sim = T(0);
            const bool pair = kF32 && P.safe_a && !P.human_is_sim && P.pa.delta == T(4);   // uniform
            if (pair) {   // controller law + the simulator model behind the speed mode, fused
              float af, asf;
              idm_pair_f32((float)vi, (float)vlead, (float)hi,
                           reinterpret_cast<const IdmRegs<float>&>(P.pa),
                           reinterpret_cast<const IdmRegs<float>&>(P.ps), af, asf, has_lead);
              a = (T)af; a_sim = (T)asf;
            } else {
              a = idm_accel<T>(vi, vlead, hi, has_lead, P.pa.v0, P.pa.inv_v0, P.pa.Tt, P.pa.a,
                               P.pa.tsab, P.pa.inv_tsab, P.pa.delta, P.pa.s0);
            }
            ae = a;
            if (!P.human_is_sim) {   // base_controller.py:109-116: noise, then fail-safe
              if (noisy) a = add(a, mul(P.noise, zn[s]));
              if (P.fail_safe == FS_FAILSAFE_INSTANTANEOUS)
                a = failsafe_instantaneous<T>(a, vi, hi, has_lead, V, dt);
              else if (P.fail_safe == FS_FAILSAFE_SAFE_VELOCITY)
                a = failsafe_safe_velocity<T>(a, vi, vlead, hi, V, dt, P.delay);
              if (pair) ae = N::max(N::min(N::min(a, a_sim), P.cap_a), P.floor_a);
              else ae = speed_mode_fast<T>(a, vi, vlead, hi, has_lead, P.ps, P.safe_a, P.cap_a, P.floor_a);
            }
          }
        } else {
          const TypeP<T>& tp = cfg.types[tt[s]];
          const int ctrl = tp.ctrl;
          if (ctrl == FS_CTRL_RL && have_act) {
            const T raw = act_a[s];
            a = N::min(N::max(raw, cfg.act_lo), cfg.act_hi);
            act_abs_sum = add(act_abs_sum, N::abs(cfg.clip_actions ? a : raw));
          } else if (ctrl == FS_CTRL_PI_SATURATION) {
            T v_cmd = B.st0[(size_t)i * E + e];
            RingBufs<T> rb;
            rb.hist = B.hist; rb.hist_sum = B.hist_sum; rb.hist_cnt = B.hist_cnt;
            rb.hist_len = B.hist_len; rb.n_pi = B.n_pi;
            a = pi_saturation<T>(rb, cfg.pi_index[i], E, e, tc, vi, vlead, hi, dt, tp.max_accel, v_cmd);
            B.st0[(size_t)i * E + e] = v_cmd;
          } else {
            T st0 = T(0);
            const bool stateful = (ctrl == FS_CTRL_LAC);
            if (stateful) st0 = B.st0[(size_t)i * E + e];
            const bool has_f = foll[s] >= 0;
            a = ctrl_accel<T>(tp, ctrl, vi, has_lead, vlead, hi, has_f, has_f ? vs[foll[s]] : T(-1001),
                              has_f ? hs[foll[s]] : T(-1001), mean_v, dt, cfg.inv_dt, st0);
            if (stateful) B.st0[(size_t)i * E + e] = st0;
          }
          if (ctrl != FS_CTRL_SIM && ctrl != FS_CTRL_RL) {
            if (tp.noise > T(0))
              a = add(a, mul(tp.noise, zn[s]));
            if (tp.fail_safe == FS_FAILSAFE_INSTANTANEOUS)
              a = failsafe_instantaneous<T>(a, vi, hi, has_lead, V, dt);
            else if (tp.fail_safe == FS_FAILSAFE_SAFE_VELOCITY)
              a = failsafe_safe_velocity<T>(a, vi, vlead, hi, V, dt, tp.delay);
          }
          ae = a;
          if (tp.speed_mode != 0 && ctrl != FS_CTRL_SIM && !(ctrl == FS_CTRL_RL && !have_act))
            ae = speed_mode_accel<T>(tp, a, vi, vlead, hi, has_lead);
        }
        acc[s] = a;
        aeff[s] = ae;
      }
      // all reads of the published pre-move state are done (refresh / rebuild sync first); move
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) {
        if (lane_id + 32 * s >= V) continue;
        const T vn = N::max(add(v[s], mul(aeff[s], dt)), T(0));
        x[s] = wrap_fwd(add(x[s], move_dist<T>(v[s], vn, dt, cfg.ballistic)), C);
        v[s] = vn;
        ln[s] = new_ln[s];
      }
      if (changed || A.write_views) rebuild(want_foll);
      else if (refresh(want_foll)) rebuild(want_foll);
      bool cr = false;
      T sv = T(0), sq = T(0), s2 = T(0);
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) {
        const int i = lane_id + 32 * s;
        if (i >= V) continue;
        if (!kGeneric)
          cr |= h[s] < crash_thr;          // h = 1000 without a leader
        else if (cfg.collision_min_gap)
          cr |= lead[s] >= 0 && h[s] < cfg.types[cfg.type_of[lead[s]]].sim_min_gap;
        else
          cr |= h[s] < T(0);
        sv = add(sv, v[s]);
        const T d = v[s] - cfg.target_velocity;
        sq = add(sq, mul(d, d));
        s2 += v[s] * v[s];
      }
      if (A.accumulate_stats) { st_v += (double)sv; st_v2 += (double)s2; }
      crash = __any_sync(FULL, cr);
      tc += 1;
      done = crash || tc >= t_end;
      if (wave_env) {
        for (int o = 16; o > 0; o >>= 1) {
          sv += __shfl_xor_sync(FULL, sv, o);
          act_abs_sum += __shfl_xor_sync(FULL, act_abs_sum, o);
        }
        if (!have_act || crash) reward = T(0);
        else {
          reward = N::div(mul(T(4), N::div(sv, T(V))), T(20));
          const T mean_act = N::div(act_abs_sum, T(cfg.n_rl > 0 ? cfg.n_rl : 1));
          if (mean_act > T(0)) reward = add(reward, mul(T(4), T(0) - mean_act));
        }
      } else {
        // rewards.desired_velocity (+ LaneChangeAccelEnv lane-change penalty);
        // AccelEnv with env_params.evaluate: mean speed (accel.py:111-112)
        if (cfg.evaluate && !lc_env) {
          for (int o = 16; o > 0; o >>= 1) sv += __shfl_xor_sync(FULL, sv, o);
          reward = N::div(sv, T(V));
        } else if (crash) reward = T(0);
        else {
          for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(FULL, sq, o);
          const T cost = N::sqrt(sq);
          if (kF32 && !kGeneric) reward = N::max(max_cost - cost, T(0)) * P.rew_inv;
          else reward = N::div(N::max(max_cost - cost, T(0)), add(max_cost, T(1.1920928955078125e-07)));
        }
        if (lc_env && rl_slots) {
          int pen = 0;
#pragma unroll
          for (int s = 0; s < SLOTS; ++s) {
            if (rls[s] >= 0) {
              const T last = B.lc_bug_compat ? h[s] : T(last_lc[s]);
              pen += (last == T(tc)) ? 1 : 0;
            }
          }
          pen = __reduce_add_sync(FULL, pen);
          for (int q = 0; q < pen; ++q) reward = reward - T(0.1);
        }
      }
      if (A.accumulate_stats && lane_id == 0) { st_rew += (double)reward; st_cr += crash ? 1 : 0; }
    } else {
      reward = T(0);
    }
    const bool last = (k == A.k_steps - 1);
    if (A.obs_seq != nullptr || last)
      write_obs(A.obs_seq ? A.obs_seq + ((size_t)k * E + e) * cfg.obs_dim : nullptr,
                last ? B.obs + (size_t)e * cfg.obs_dim : nullptr);
    if (lane_id == 0) {
      if (A.reward_seq) A.reward_seq[(size_t)k * E + e] = reward;
      if (A.done_seq) A.done_seq[(size_t)k * E + e] = done ? 1 : 0;
    }
  }

#pragma unroll
  for (int s = 0; s < SLOTS; ++s) {
    const int i = lane_id + 32 * s;
    if (i >= V) continue;
    B.pos[eo + i] = x[s]; B.vel[eo + i] = v[s]; B.lane[eo + i] = (uint8_t)ln[s];
    B.last_lc[eo + i] = last_lc[s];
    if (A.write_views) {
      B.v_headway[eo + i] = h[s]; B.v_leader[eo + i] = lead[s]; B.v_follower[eo + i] = foll[s];
      if (A.k_steps > 0) B.v_accel[eo + i] = acc[s];
    }
  }
  if (lane_id == 0) {
    B.time_counter[e] = tc;
    B.done[e] = done ? 1 : 0;
    B.crash[e] = crash ? 1 : 0;
    B.reward[e] = reward;
  }
  if (A.accumulate_stats) {
    // stats[]: sum reward, sum v, sum v^2, vehicle-steps, env-steps, crashes
    for (int o = 16; o > 0; o >>= 1) {
      st_v += __shfl_xor_sync(FULL, st_v, o);
      st_v2 += __shfl_xor_sync(FULL, st_v2, o);
    }
    if (lane_id == 0) {
      const double adv = (double)(tc - tc0);
      double vals[6] = {st_rew, st_v, st_v2, adv * V, adv, st_cr};
      substep_stats(A.accumulate_stats, vals);
      for (int q = 0; q < 6; ++q) if (vals[q] != 0.0) atomicAdd(&stat_stripe(B.stats)[q], vals[q]);
    }
  }
}

// host-side dispatch (instantiated in fs_ml_f32.cu / fs_ml_f64.cu)
template <typename T>
static cudaError_t launch_ml_impl(int V, bool uniform, int grid, int block, size_t smem,
                                  cudaStream_t st, const Cfg<T>& c, const MlBufs<T>& b,
                                  const StepArgs<T>& a) {
#define FS_ML(SL) \
  do { if (uniform) ring_ml_kernel<T, SL, false><<<grid, block, smem, st>>>(c, b, a); \
       else ring_ml_kernel<T, SL, true><<<grid, block, smem, st>>>(c, b, a); } while (0)
  if (V <= 32) FS_ML(1); else if (V <= 96) FS_ML(3); else FS_ML(8);
#undef FS_ML
  return cudaGetLastError();
}

}  // namespace fs
