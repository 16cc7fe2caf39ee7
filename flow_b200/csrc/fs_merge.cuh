// fs_merge.cuh — MergeNetwork step kernel: single-lane highway with a
// single-lane on-ramp, stepped under MergePOEnv.
//
// Reference arithmetic followed line by line:
//   MergePOEnv.get_state / compute_reward / _apply_rl_actions / additional_command
//                                                   flow/envs/merge.py:106-220
//   get_x_by_id on MergeNetwork's edge starts       flow/networks/merge.py:196-221 with
//   TraCIKernelNetwork.get_x's junction quirk       flow/core/kernel/network/traci.py:273-289
//   rewards.desired_velocity                        flow/core/rewards.py:6-59
//   controllers / noise / fail-safes                fs_common.cuh
// RESTATED SUMO ("parity unpinned"; oracle/merge.py is the same statement on the
// CPU): both arms lie on one axis u = distance travelled on the arm + (LH - L_arm), so the merge
// point (start of edge 'center') is u = LH for both. A vehicle's leader is the nearest vehicle
// ahead on that axis among the vehicles of its own arm, the vehicles past the merge point and -
// once both are inside the look-ahead zone before the merge point - the vehicles of the other
// arm (a zipper by position stands in for SUMO's junction right-of-way); ties in u: the lower
// slot is ahead. Collision <=> negative headway towards a leader that shares the road (same
// arm, or both inside the junction / on 'center'). Vehicles leave at the end of 'center'
// (warp-ballot compaction, order kept); every <flow> emits at most one vehicle per step at the
// start of its arm, earliest scheduled departure first, once the gap to the arm's last vehicle
// reaches the simulator model's desired gap s* for the depart speed.
//
// Mapping: one warp per env, slot s = lane + 32 r in insertion order (= departure order, which
// is what the reference's rl_queue follows), state env-major. Leader search: previous slot of
// the own arm + a bisection of the other arm's in-zone list (see update()); the all-pairs scan
// only where the slot order is not the road order.
#pragma once
#include "fs_common.cuh"
#include "fs_ring.cuh"

namespace fs {

template <typename T>
struct MgBufs {
  T* pos; T* vel; uint8_t* arm; uint8_t* vtype; uint8_t* active;   // [E][CAP]
  int* vid; int* seq;                                               // [E][CAP] id, departure order
  int* n_active; int* departed; int* arrived;                       // [E]
  int* spawned;                                                     // [FS_MAX_INFLOWS][E]
  int* n_ctrl; int* hi_seq; int* next_seq;                          // [E] rl_veh bookkeeping
  int* ctrl_id;                                                     // [E][A] the rl_veh list
  int* time_counter; int* episode;
  uint8_t* done; uint8_t* crash;
  T* obs; T* reward;
  T* v_accel; T* v_headway; int* v_leader; int* v_follower;         // views [E][CAP] or null
  double* stats;
  int n_flows, max_agents;
  double period[FS_MAX_INFLOWS], begin[FS_MAX_INFLOWS];
  T depart_speed[FS_MAX_INFLOWS];
  int inflow_type[FS_MAX_INFLOWS], inflow_arm[FS_MAX_INFLOWS];
};

template <typename T>
__host__ __device__ inline size_t mg_smem_per_warp(int cap) {
  return (size_t)cap * (5 * sizeof(T) + 3 * sizeof(int));
}

// get_x_by_id of a vehicle p metres into its arm (oracle/merge.py MergeGeometry.x_abs)
template <typename T>
__device__ __forceinline__ T mg_x_abs(const Cfg<T>& c, int arm, T p) {
  const T E0 = T(100), JS = T(0.1);
  const T j_center = add(add(E0, c.mg_pre), JS);              // start of ':center'
  if (arm == 0) {
    if (p < E0) return p;
    if (p < add(E0, JS)) return E0;                            // ':left': its start, no position
    if (p < add(add(E0, JS), c.mg_pre)) return p;
    if (p < c.mg_LH) return j_center;
    return p;
  }
  const T x0 = add(add(add(add(E0, c.mg_pre), c.mg_post), JS), T(22.5));   // 'inflow_merge'
  if (p < E0) return add(x0, p);
  if (p < add(E0, JS)) return add(x0, E0);                     // ':bottom'
  if (p < add(add(E0, JS), c.mg_ramp)) return add(x0, p);
  if (p < c.mg_LM) return j_center;
  return add(c.mg_LH, p - c.mg_LM);                            // on 'center'
}

template <typename T, int SLOTS>
__global__ void __launch_bounds__(128)
merge_kernel(const __grid_constant__ Cfg<T> cfg, const __grid_constant__ MgBufs<T> B,
             const StepArgs<T> A) {
  typedef Num<T> N;
  constexpr unsigned FULL = 0xffffffffu;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane_id = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int E = cfg.n_envs, CAP = cfg.n_veh, AG = B.max_agents;
  const int e = blockIdx.x * (blockDim.x >> 5) + wib;
  if (e >= E) return;
  if (A.mask != nullptr && A.mask[e] == 0) return;
  unsigned char* mine = smem_raw + mg_smem_per_warp<T>(CAP) * wib;
  T* us = reinterpret_cast<T*>(mine);      // axis coordinate u of every slot
  T* vs = us + CAP;
  T* hs = vs + CAP;                        // headway
  T* cmd = hs + CAP;                       // commanded RL acceleration of the step (NaN: none)
  T* xp = cmd + CAP;                       // distance travelled on the own arm
  int* at = reinterpret_cast<int*>(xp + CAP);    // arm | type << 8
  int* ld = at + CAP;                      // leader slot
  int* is = ld + CAP;                      // vehicle id

  T x[SLOTS], v[SLOTS], h[SLOTS], acc[SLOTS], aeff[SLOTS];
  int ty[SLOTS], arm[SLOTS], id[SLOTS], sq[SLOTS], lead[SLOTS];
  const size_t eo = (size_t)e * CAP;
  int n = B.n_active[e];
#pragma unroll
  for (int r = 0; r < SLOTS; ++r) {
    const int s = lane_id + 32 * r;
    const bool ok = s < n;
    x[r] = ok ? B.pos[eo + s] : T(0);
    v[r] = ok ? B.vel[eo + s] : T(0);
    ty[r] = ok ? (int)B.vtype[eo + s] : 0;
    arm[r] = ok ? (int)B.arm[eo + s] : 0;
    id[r] = ok ? B.vid[eo + s] : -1;
    sq[r] = ok ? B.seq[eo + s] : 0;
    acc[r] = T(0); lead[r] = -1; h[r] = T(1000);
  }
  const T LH = cfg.mg_LH, Ltot = add(cfg.mg_LH, cfg.mg_post);
  const T off1 = LH - cfg.mg_LM;           // ramp arm -> axis
  const T z0 = LH - cfg.mg_zone, zc = LH - T(22.5);
  const T dt = cfg.dt;
  int tc = B.time_counter[e];
  const int episode = B.episode[e];
  bool done = A.ignore_done ? false : (B.done[e] != 0);
  bool crash = B.crash[e] != 0;
  const int t_end = cfg.sims_per_step * (cfg.warmup_steps + cfg.horizon);
  const long long gid = cfg.env_id_base + e;
  T reward = B.reward[e];
  int departed = B.departed[e], arrived = B.arrived[e];
  int spawned[FS_MAX_INFLOWS];
#pragma unroll
  for (int f = 0; f < FS_MAX_INFLOWS; ++f) spawned[f] = f < B.n_flows ? B.spawned[(size_t)f * E + e] : 0;
  // rl_veh: lane k < n_ctrl holds the id of controlled vehicle k and the slot it sits in now
  int n_ctrl = B.n_ctrl[e], hi_seq = B.hi_seq[e], next_seq = B.next_seq[e];
  int my_ctrl = (lane_id < n_ctrl && lane_id < AG) ? B.ctrl_id[(size_t)e * AG + lane_id] : -1;
  int my_slot = -1;
  unsigned rl_types = 0;
  for (int t = 0; t < cfg.n_types; ++t) rl_types |= (cfg.types[t].ctrl == FS_CTRL_RL ? 1u : 0u) << t;
  double st_rew = 0, st_vs = 0, st_es = 0, st_cr = 0;
  double pl_v = 0, pl_v2 = 0;      // per-lane sums of v and v^2, reduced once at the end

  auto u_of = [&](int r) -> T { return arm[r] ? add(x[r], off1) : x[r]; };
  // k.update on the merged axis + slot of every controlled vehicle
  // (every per-row loop below stops at the last populated slot row: n is warp-uniform, so the
  //  bound is a uniform branch and the ballots inside stay full-warp)
  auto update = [&]() {
    const int nrows = (n + 31) >> 5;
    __syncwarp();
#pragma unroll
    for (int r = 0; r < SLOTS && r < nrows; ++r) {
      const int s = lane_id + 32 * r;
      if (s < n) {
        us[s] = u_of(r); xp[s] = x[r]; vs[s] = v[r]; at[s] = arm[r] | (ty[r] << 8); is[s] = id[r];
      }
    }
    __syncwarp();
    // Slots are in departure order and nobody overtakes on its own arm, so an arm's vehicles
    // normally sit in the slot array by descending u: the own-arm leader is the previous slot of
    // the arm (ballot + clz), and the other arm only matters through its vehicles inside the
    // zone, kept as two short sorted lists (ballot compaction into `ld`, arm 0 from the front,
    // arm 1 from the back) that a vehicle inside the zone bisects. Anything else (initial
    // vehicles placed in another order, states after a collision) takes the full scan below.
    int pred[SLOTS];
    int last0 = -1, last1 = -1, zn0 = 0, zn1 = 0, np0 = 0, np1 = 0;
    bool sorted = true;
#pragma unroll
    for (int r = 0; r < SLOTS && r < nrows; ++r) {
      const int s = lane_id + 32 * r;
      const bool on = s < n;
      const T ui = on ? us[s] : T(0);
      const unsigned mv = __ballot_sync(FULL, on);
      const unsigned m1 = __ballot_sync(FULL, on && arm[r] != 0), m0 = mv & ~m1;
      const unsigned mine_m = arm[r] ? m1 : m0, below = mine_m & ((1u << lane_id) - 1u);
      pred[r] = below ? 32 * r + 31 - __clz(below) : (arm[r] ? last1 : last0);
      if (on && pred[r] >= 0) sorted &= us[pred[r]] > ui;
      if (m0) last0 = 32 * r + 31 - __clz(m0);
      if (m1) last1 = 32 * r + 31 - __clz(m1);
      const bool zin = on && ui >= z0;
      const unsigned z1 = __ballot_sync(FULL, zin && arm[r] != 0);
      const unsigned z0m = __ballot_sync(FULL, zin && arm[r] == 0);
      const unsigned p1 = __ballot_sync(FULL, on && arm[r] != 0 && ui >= LH);
      const unsigned p0 = __ballot_sync(FULL, on && arm[r] == 0 && ui >= LH);
      if (zin) {
        const unsigned lower = (1u << lane_id) - 1u;
        if (arm[r]) ld[CAP - 1 - (zn1 + __popc(z1 & lower))] = s;
        else ld[zn0 + __popc(z0m & lower)] = s;
      }
      zn0 += __popc(z0m); zn1 += __popc(z1); np0 += __popc(p0); np1 += __popc(p1);
    }
    sorted = __all_sync(FULL, sorted);
    __syncwarp();
    if (sorted) {
#pragma unroll
      for (int r = 0; r < SLOTS && r < nrows; ++r) {
        const int s = lane_id + 32 * r;
        int bj = -1; T bd = T(0);
        if (s < n) {
          const T ui = us[s];
          bj = pred[r];
          if (bj >= 0) bd = us[bj] - ui;
          // other arm: nearest vehicle ahead inside the zone when this one is inside it too,
          // else the last vehicle past the merge point
          const int zn = arm[r] ? zn0 : zn1, np = arm[r] ? np0 : np1;
          int c = np;
          if (ui >= z0) {
            int lo = 0, hi = zn;              // entries [0, lo) are ahead of this vehicle
            while (lo < hi) {
              const int mid = (lo + hi) >> 1;
              const int j = arm[r] ? ld[mid] : ld[CAP - 1 - mid];
              const T uj = us[j];
              if (uj > ui || (uj == ui && j < s)) lo = mid + 1; else hi = mid;
            }
            c = lo;
          }
          if (c > 0) {
            const int j = arm[r] ? ld[c - 1] : ld[CAP - 1 - (c - 1)];
            const T d = us[j] - ui;
            if (bj < 0 || d < bd || (d == bd && j < bj)) { bj = j; bd = d; }
          }
        }
        lead[r] = bj;
        h[r] = bj >= 0 ? bd - cfg.types[at[bj] >> 8].length : T(1000);
      }
    } else {
#pragma unroll
      for (int r = 0; r < SLOTS && r < nrows; ++r) {
        const int s = lane_id + 32 * r;
        int bj = -1; T bd = T(0);
        if (s < n) {
          const T ui = us[s];
          const bool in_zone = ui >= z0;
          for (int j = 0; j < n; ++j) {
            const T uj = us[j];
            if (j == s || !(uj > ui || (uj == ui && j < s))) continue;
            const bool ok = (at[j] & 0xff) == arm[r] || uj >= LH || (in_zone && uj >= z0);
            if (!ok) continue;
            const T d = uj - ui;
            if (bj < 0 || d < bd) { bj = j; bd = d; }
          }
        }
        lead[r] = bj;
        h[r] = bj >= 0 ? bd - cfg.types[at[bj] >> 8].length : T(1000);
      }
    }
    __syncwarp();
#pragma unroll
    for (int r = 0; r < SLOTS && r < nrows; ++r) {
      const int s = lane_id + 32 * r;
      if (s < n) { ld[s] = lead[r]; hs[s] = h[r]; }
    }
    __syncwarp();
    my_slot = -1;          // row a of rl_veh: which slot carries its vehicle now
    for (int a = 0; a < n_ctrl; ++a) {
      const int cid = __shfl_sync(FULL, my_ctrl, a);
#pragma unroll
      for (int r = 0; r < SLOTS && r < nrows; ++r) {
        const unsigned hit = __ballot_sync(FULL, lane_id + 32 * r < n && id[r] == cid);
        if (hit && lane_id == a) my_slot = 32 * r + __ffs(hit) - 1;
      }
    }
  };
  update();

  // MergePOEnv.get_state (merge.py:114-159): row k <- controlled vehicle k; a vehicle that has
  // left reads as the kernel's error values (speed -1001, nobody ahead / behind)
  auto write_obs = [&](T* o, T* ol) {
    const T ms = cfg.speed_limit;
    const T ml = add(add(add(add(add(T(200), cfg.mg_pre), cfg.mg_ramp), cfg.mg_post), T(0.2)), T(22.5));
    if (lane_id < AG) {
      T q[5] = {T(0), T(0), T(0), T(0), T(0)};
      if (lane_id < n_ctrl) {
        const int s = my_slot;
        const T sp = s >= 0 ? vs[s] : T(-1001);
        const int lj = s >= 0 ? ld[s] : -1;
        int fj = -1; T fh = T(0);
        if (s >= 0)     // follower: the closest vehicle that calls me leader (vehicle/traci.py:243-250)
          for (int j = 0; j < n; ++j) if (ld[j] == s && (fj < 0 || hs[j] < fh)) { fj = j; fh = hs[j]; }
        T lsp = ms, lhd = ml, fsp = T(0), fhd = ml;
        if (lj >= 0) {
          lsp = vs[lj];
          lhd = (mg_x_abs<T>(cfg, at[lj] & 0xff, xp[lj]) - mg_x_abs<T>(cfg, at[s] & 0xff, xp[s])) -
                cfg.types[at[s] >> 8].length;
        }
        if (fj >= 0) { fsp = vs[fj]; fhd = fh; }
        q[0] = N::div(sp, ms); q[1] = N::div(lsp - sp, ms); q[2] = N::div(lhd, ml);
        q[3] = N::div(sp - fsp, ms); q[4] = N::div(fhd, ml);
      }
#pragma unroll
      for (int c = 0; c < 5; ++c) {
        if (o) o[lane_id * 5 + c] = q[c];
        if (ol) ol[lane_id * 5 + c] = q[c];
      }
    }
  };
  if (A.k_steps == 0) write_obs(nullptr, B.obs + (size_t)e * cfg.obs_dim);

  // k_steps counts Env.steps: each runs sims_per_step simulation steps on the same action row
  // (flow/envs/base.py:325-383; every shipped merge config uses 5) and observes once; a collision
  // ends the env, which skips its remaining sub-steps like the reference's `break`
  for (int k = 0; k < A.k_steps; ++k) {
    bool stepped = false;
#pragma unroll 1
    for (int q = 0; q < cfg.sims_per_step; ++q) {
    if (!done) {
      stepped = true;
      const int nrows = (n + 31) >> 5, mrows = nrows;
      // ---- _apply_rl_actions (merge.py:106-112): action row k -> controlled vehicle k, if it is
      //      still in the network; NaN = no command (the simulator model drives)
      __syncwarp();
#pragma unroll
      for (int r = 0; r < SLOTS && r < nrows; ++r) { const int s = lane_id + 32 * r; if (s < n) cmd[s] = N::nan(); }
      __syncwarp();
      if (A.actions != nullptr && lane_id < n_ctrl && my_slot >= 0)
        cmd[my_slot] = A.actions[((size_t)k * cfg.act_dim + lane_id) * E + e];
      __syncwarp();
      T mean_v = T(0);
      if (cfg.any_follower & 2) {
        T sm = T(0);
#pragma unroll
        for (int r = 0; r < SLOTS && r < nrows; ++r) if (lane_id + 32 * r < n) sm = add(sm, v[r]);
        for (int o = 16; o > 0; o >>= 1) sm += __shfl_xor_sync(FULL, sm, o);
        mean_v = N::div(sm, T(n > 0 ? n : 1));
      }
#pragma unroll
      for (int r = 0; r < SLOTS && r < nrows; ++r) {
        const int s = lane_id + 32 * r;
        if (s >= n) continue;
        const TypeP<T>& tp = cfg.types[ty[r]];
        const int ctrl = tp.ctrl;
        const bool has_lead = lead[r] >= 0;
        const T vi = v[r];
        const T vl = has_lead ? vs[lead[r]] : T(-1001);
        // an overlapping leader on the OTHER approach road reads as bumper to bumper: with a
        // negative gap an IDM accelerates again once |gap| > s* (include/flowstep.h FS_TOPO_MERGE)
        T hi = h[r];
        if (has_lead && hi < T(0) && (at[lead[r]] & 0xff) != arm[r] && !(us[s] >= zc && us[lead[r]] >= zc))
          hi = T(1e-3);
        T st0 = T(0);
        T a = ctrl_accel<T>(tp, ctrl, vi, has_lead, vl, hi, false, T(-1001), T(-1001), mean_v, dt,
                            cfg.inv_dt, st0);
        bool commanded = ctrl != FS_CTRL_SIM && ctrl != FS_CTRL_RL;
        if (commanded) {   // base_controller.py:109-116: noise, then the fail-safe
          if (tp.noise > T(0))
            a = add(a, mul(tp.noise, normal_sample<T>(cfg.seed, gid, s, tc + episode * 1000003)));
          if (tp.fail_safe == FS_FAILSAFE_INSTANTANEOUS)
            a = failsafe_instantaneous<T>(a, vi, hi, has_lead, n, dt);
          else if (tp.fail_safe == FS_FAILSAFE_SAFE_VELOCITY)
            a = failsafe_safe_velocity<T>(a, vi, vl, hi, n, dt, tp.delay);
        }
        const T c = cmd[s];
        if (c == c) { a = c; commanded = true; }
        acc[r] = a;
        aeff[r] = (commanded && tp.speed_mode != 0) ? speed_mode_accel<T>(tp, a, vi, vl, hi, has_lead) : a;
      }
      // ---- additional_command (merge.py:191-220), before the move: rows of vehicles that left
      //      close up, waiting RL vehicles (departure order = slot order) fill the free rows
      {
        const bool keep = lane_id < n_ctrl && my_slot >= 0;
        const unsigned km = __ballot_sync(FULL, keep);
        if (km != (n_ctrl >= 32 ? FULL : (1u << n_ctrl) - 1u)) {   // a controlled vehicle left
          const int dst = __popc(km & ((1u << lane_id) - 1u));
          int moved = -1;
#pragma unroll 1
          for (unsigned m = km; m; m &= m - 1) {    // lane dst takes the id of the dst-th kept row
            const int src = __ffs(m) - 1;
            const int idv = __shfl_sync(FULL, my_ctrl, src);
            const int d = __shfl_sync(FULL, dst, src);
            if (d == lane_id) moved = idv;
          }
          n_ctrl = __popc(km);
          my_ctrl = lane_id < n_ctrl ? moved : -1;
        }
        if (n_ctrl < AG) {                          // a row is free: the longest-waiting RL vehicles
          int base = n_ctrl;
#pragma unroll
          for (int r = 0; r < SLOTS && r < nrows; ++r) {
            const int s = lane_id + 32 * r;
            const bool w = s < n && ((rl_types >> ty[r]) & 1u) && sq[r] > hi_seq;
            const unsigned wm = __ballot_sync(FULL, w);
            const int row = w ? base + __popc(wm & ((1u << lane_id) - 1u)) : -1;
            // hand (id, seq) of waiting vehicle -> row `row` while rows are left
#pragma unroll 1
            for (unsigned m = wm; m; m &= m - 1) {
              const int src = __ffs(m) - 1;
              const int rw = __shfl_sync(FULL, row, src);
              if (rw >= AG) break;
              const int idv = __shfl_sync(FULL, id[r], src);
              const int sv = __shfl_sync(FULL, sq[r], src);
              if (lane_id == rw) my_ctrl = idv;
              hi_seq = sv;
            }
            base += __popc(wm);
          }
          n_ctrl = base < AG ? base : AG;
        }
      }
      tc += 1;
      // ---- move, exits (u >= end of 'center'), order-preserving ballot compaction
      int base = 0;
      int dst[SLOTS];
#pragma unroll
      for (int r = 0; r < SLOTS && r < mrows; ++r) {
        const int s = lane_id + 32 * r;
        bool keep = false;
        if (s < n) {
          const T vn = N::max(add(v[r], mul(aeff[r], dt)), T(0));
          x[r] = add(x[r], mul(vn, dt));
          v[r] = vn;
          keep = u_of(r) < Ltot;
        }
        const unsigned m = __ballot_sync(FULL, keep);
        dst[r] = keep ? base + __popc(m & ((1u << lane_id) - 1u)) : -1;
        base += __popc(m);
      }
      const int n_keep = base;
      arrived += n - n_keep;
      if (n_keep != n) {
        __syncwarp();
#pragma unroll
        for (int r = 0; r < SLOTS && r < mrows; ++r)
          if (dst[r] >= 0) {
            xp[dst[r]] = x[r]; vs[dst[r]] = v[r]; hs[dst[r]] = acc[r];
            at[dst[r]] = arm[r] | (ty[r] << 8); is[dst[r]] = id[r]; ld[dst[r]] = sq[r];
          }
        __syncwarp();
#pragma unroll
        for (int r = 0; r < SLOTS && r < mrows; ++r) {
          const int s = lane_id + 32 * r;
          if (s < n_keep) {
            x[r] = xp[s]; v[r] = vs[s]; acc[r] = hs[s]; arm[r] = at[s] & 0xff; ty[r] = at[s] >> 8;
            id[r] = is[s]; sq[r] = ld[s];
          } else { x[r] = T(0); v[r] = T(0); acc[r] = T(0); arm[r] = 0; ty[r] = 0; id[r] = -1; sq[r] = 0; }
        }
        n = n_keep;
      }
      // ---- inflows: earliest scheduled departure first (ties: definition order). Everything is
      //      indexed with compile-time constants (selects over the four flows): no local memory
      {
        const double t = (double)tc * (double)cfg.dt;
        double due[FS_MAX_INFLOWS];
        unsigned open = 0;
#pragma unroll
        for (int f = 0; f < FS_MAX_INFLOWS; ++f) {
          const bool on = f < B.n_flows && B.period[f] > 0.0;
          due[f] = on ? B.begin[f] + (double)spawned[f] * B.period[f] : 0.0;
          // a flow that is not due yet (or finds the slots full) can be dropped right away:
          // serving order only matters among the flows that do emit
          if (on && t >= due[f]) open |= 1u << f;
        }
#pragma unroll 1
        while (open != 0 && n < CAP) {
          int f = -1; double bd = 0.0;
#pragma unroll
          for (int q = 0; q < FS_MAX_INFLOWS; ++q)
            if (((open >> q) & 1u) && (f < 0 || due[q] < bd)) { f = q; bd = due[q]; }
          open &= ~(1u << f);
          const int fa = B.inflow_arm[f];
          const TypeP<T>& nt = cfg.types[B.inflow_type[f]];
          // last vehicle of the arm: smallest distance travelled
          T bx = T(0), bv = T(0); int bt = -1;
#pragma unroll
          for (int r = 0; r < SLOTS && r < ((n + 31) >> 5); ++r)
            if (lane_id + 32 * r < n && arm[r] == fa && (bt < 0 || x[r] < bx)) { bx = x[r]; bv = v[r]; bt = ty[r]; }
          if (sizeof(T) == 4) {
            // x >= 0: the float's bits order like the float; one REDUX instead of a shuffle tree
            const unsigned key = bt >= 0 ? __float_as_uint((float)bx) : 0xffffffffu;
            const unsigned kmin = __reduce_min_sync(FULL, key);
            const unsigned owners = __ballot_sync(FULL, key == kmin && bt >= 0);
            if (owners) {
              const int src = __ffs(owners) - 1;
              bx = __shfl_sync(FULL, bx, src); bv = __shfl_sync(FULL, bv, src);
              bt = __shfl_sync(FULL, bt, src);
            } else bt = -1;
          } else {
            for (int o = 16; o > 0; o >>= 1) {
              const T ox = __shfl_xor_sync(FULL, bx, o), ov = __shfl_xor_sync(FULL, bv, o);
              const int ot = __shfl_xor_sync(FULL, bt, o);
              if (ot >= 0 && (bt < 0 || ox < bx)) { bx = ox; bv = ov; bt = ot; }
            }
          }
          bool spawn = true;
          if (bt >= 0) {   // restated SUMO insertion: gap >= the simulator model's s* at depart speed
            const T vd = B.depart_speed[f];
            const T q = N::div(mul(vd, vd - bv), mul(T(2), N::sqrt(mul(nt.max_accel, nt.max_decel))));
            const T s_star = add(nt.sim_min_gap, N::max(T(0), add(mul(vd, nt.sim_tau), q)));
            spawn = (bx - cfg.types[bt].length) >= s_star;
          }
          if (spawn) {
            int sp = 0;
#pragma unroll
            for (int q = 0; q < FS_MAX_INFLOWS; ++q) if (q == f) { sp = spawned[q]; spawned[q] += 1; }
#pragma unroll
            for (int r = 0; r < SLOTS && r < ((n + 32) >> 5); ++r)
              if (lane_id + 32 * r == n) {
                x[r] = T(0); v[r] = B.depart_speed[f]; ty[r] = B.inflow_type[f]; arm[r] = fa;
                acc[r] = T(0); id[r] = ((f + 1) << 24) | sp; sq[r] = next_seq;
              }
            n += 1; departed += 1; next_seq += 1;
          }
        }
      }
      update();
      // ---- collision (shared road only), desired_velocity, time-headway penalty
      bool cr = false;
      T sv = T(0), sqd = T(0);
#pragma unroll
      for (int r = 0; r < SLOTS && r < ((n + 31) >> 5); ++r) {
        const int s = lane_id + 32 * r;
        if (s >= n) continue;
        if (lead[r] >= 0 && h[r] < T(0)) {
          const int lj = lead[r];
          cr |= (at[lj] & 0xff) == arm[r] || (us[s] >= zc && us[lj] >= zc);
        }
        sv = add(sv, v[r]);
        const T d = v[r] - cfg.target_velocity;
        sqd = add(sqd, mul(d, d));
        if (A.accumulate_stats) pl_v2 += (double)v[r] * (double)v[r];
      }
      if (A.accumulate_stats) { pl_v += (double)sv; if (lane_id == 0) st_vs += n; }
      crash = __any_sync(FULL, cr);
      done = crash || tc >= t_end;
      // the reward is read once per Env.step: after the last sub-step or the one that collided
      if (done || q == cfg.sims_per_step - 1) {
        T c2 = T(0);     // merge.py:176-184, t_min = 1
        if (lane_id < n_ctrl && my_slot >= 0 && ld[my_slot] >= 0 && vs[my_slot] > T(0))
          c2 = N::min(N::max(N::div(hs[my_slot], vs[my_slot]), T(0)) - T(1), T(0));
        for (int o = 16; o > 0; o >>= 1) {
          sv += __shfl_xor_sync(FULL, sv, o);
          sqd += __shfl_xor_sync(FULL, sqd, o);
          c2 += __shfl_xor_sync(FULL, c2, o);
        }
        if (cfg.evaluate) reward = N::div(sv, T(n));
        else if (crash) reward = T(0);
        else {
          T cost1 = T(0);
          if (n > 0) {
            const T mc2 = mul(mul(cfg.target_velocity, cfg.target_velocity), T(n));
            const T max_cost = N::sqrt(mc2), cost = N::sqrt(sqd);
            cost1 = N::div(N::max(max_cost - cost, T(0)), add(max_cost, T(1.1920928955078125e-07)));
          }
          reward = N::max(add(cost1, mul(T(0.1), c2)), T(0));
        }
        if (A.accumulate_stats && lane_id == 0) { st_rew += (double)reward; st_es += 1; st_cr += crash ? 1 : 0; }
      }
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

#pragma unroll
  for (int r = 0; r < SLOTS; ++r) {
    const int s = lane_id + 32 * r;
    if (s >= CAP) continue;
    B.pos[eo + s] = x[r]; B.vel[eo + s] = v[r]; B.vtype[eo + s] = (uint8_t)ty[r];
    B.arm[eo + s] = (uint8_t)arm[r];
    B.active[eo + s] = s < n ? 1 : 0;
    B.vid[eo + s] = s < n ? id[r] : -1;
    B.seq[eo + s] = sq[r];
    if (A.write_views) {
      B.v_headway[eo + s] = s < n ? h[r] : T(0);
      B.v_leader[eo + s] = s < n ? lead[r] : -1;
      int fj = -1; T fh = T(0);      // the closest vehicle that calls this one its leader
      if (s < n)
        for (int j = 0; j < n; ++j) if (ld[j] == s && (fj < 0 || hs[j] < fh)) { fj = j; fh = hs[j]; }
      B.v_follower[eo + s] = fj;
      if (A.k_steps > 0) B.v_accel[eo + s] = s < n ? acc[r] : T(0);
    }
  }
  if (A.accumulate_stats)
    for (int o = 16; o > 0; o >>= 1) {
      pl_v += __shfl_xor_sync(FULL, pl_v, o);
      pl_v2 += __shfl_xor_sync(FULL, pl_v2, o);
    }
  if (lane_id < AG) B.ctrl_id[(size_t)e * AG + lane_id] = lane_id < n_ctrl ? my_ctrl : -1;
  if (lane_id == 0) {
    B.n_active[e] = n; B.departed[e] = departed; B.arrived[e] = arrived;
#pragma unroll
    for (int f = 0; f < FS_MAX_INFLOWS; ++f) if (f < B.n_flows) B.spawned[(size_t)f * E + e] = spawned[f];
    B.n_ctrl[e] = n_ctrl; B.hi_seq[e] = hi_seq; B.next_seq[e] = next_seq;
    B.time_counter[e] = tc;
    B.done[e] = done ? 1 : 0;
    B.crash[e] = crash ? 1 : 0;
    B.reward[e] = reward;
    if (A.accumulate_stats) {
      double vals[6] = {st_rew, pl_v, pl_v2, st_vs, st_es, st_cr};
      for (int q = 0; q < 6; ++q) if (vals[q] != 0.0) atomicAdd(&stat_stripe(B.stats)[q], vals[q]);
    }
  }
}

// Env.reset on the merge network: n_initial vehicles (types from the table) in the first slots,
// standing at the start of the highway arm until the caller places them (fs_reset_stage
// FS_RESET_PLACE + fs_set_state: gen_even_start_pos over the five edges is drawn on the host
// side, flow_b200/placement.py); empty bookkeeping.
template <typename T>
__global__ void merge_reset_kernel(const __grid_constant__ Cfg<T> cfg, const MgBufs<T> B,
                                   const uint8_t* mask, int n_initial, T* length) {
  const int E = cfg.n_envs, CAP = cfg.n_veh;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  if (mask != nullptr && mask[e] == 0) return;
  const size_t eo = (size_t)e * CAP;
  for (int s = 0; s < CAP; ++s) {
    const bool on = s < n_initial;
    B.pos[eo + s] = T(0); B.vel[eo + s] = T(0); B.arm[eo + s] = 0;
    B.vtype[eo + s] = on ? cfg.type_of[s] : 0; B.active[eo + s] = on ? 1 : 0;
    B.vid[eo + s] = on ? s : -1; B.seq[eo + s] = on ? s : 0;
  }
  for (int a = 0; a < B.max_agents; ++a) B.ctrl_id[(size_t)e * B.max_agents + a] = -1;
  length[e] = add(add(add(add(add(T(200), cfg.mg_pre), cfg.mg_ramp), cfg.mg_post), T(0.2)), T(22.5));
  B.n_active[e] = n_initial; B.departed[e] = n_initial; B.arrived[e] = 0;
  for (int f = 0; f < FS_MAX_INFLOWS; ++f) B.spawned[(size_t)f * E + e] = 0;
  B.n_ctrl[e] = 0; B.hi_seq[e] = -1; B.next_seq[e] = n_initial;
  B.time_counter[e] = 0;
  B.episode[e] = B.episode[e] + 1;
  B.done[e] = 0; B.crash[e] = 0;
}

// host-side dispatch (instantiated in fs_mg_f32.cu / fs_mg_f64.cu)
template <typename T>
static cudaError_t launch_mg_impl(int cap, int grid, int block, size_t smem, cudaStream_t st,
                                  const Cfg<T>& c, const MgBufs<T>& b, const StepArgs<T>& a) {
  // 4 warps x 256 slots in fp64 need more than the 48 KB a kernel gets without opting in
#define FS_MG(S)                                                                          \
  do {                                                                                    \
    if (smem > 48 * 1024) {                                                               \
      cudaError_t rc = cudaFuncSetAttribute(merge_kernel<T, S>,                           \
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
      if (rc != cudaSuccess) return rc;                                                   \
    }                                                                                     \
    merge_kernel<T, S><<<grid, block, smem, st>>>(c, b, a);                               \
  } while (0)
  if (cap <= 64) FS_MG(2);
  else if (cap <= 128) FS_MG(4);
  else FS_MG(8);
#undef FS_MG
  return cudaGetLastError();
}
template <typename T>
static cudaError_t launch_mg_reset_impl(int E, cudaStream_t st, const Cfg<T>& c, const MgBufs<T>& b,
                                        const uint8_t* mask, int n_initial, T* length) {
  merge_reset_kernel<T><<<(E + 127) / 128, 128, 0, st>>>(c, b, mask, n_initial, length);
  return cudaGetLastError();
}

}  // namespace fs
