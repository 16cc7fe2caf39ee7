// fs_highway.cuh — single-lane open highway with an inflow (config c4).
//
// Mapping: ONE WARP OWNS ONE ENVIRONMENT; vehicles live head-first in a
// fixed-capacity slot array (slot 0 = furthest downstream), slot s in lane
// (s & 31), register (s >> 5); state is env-major ([E][CAP]) so the warp's
// loads/stores are contiguous. The leader of slot s is slot s-1 (no overtaking
// on one lane), read through a per-warp shared-memory copy. Vehicles that
// reach the end of the edge are removed by WARP-BALLOT STREAM COMPACTION
// (__ballot_sync + popc prefix -> new slot), and the inflow appends at most
// one vehicle per step at the tail.
//
// Restated from the reference / SUMO (see oracle/highway.py for the rules):
//   arrivals + departures bookkeeping   flow/core/kernel/vehicle/traci.py:145-174,212-216
//   <flow vehsPerHour begin departSpeed> flow/core/kernel/network/traci.py:801-833,
//                                        flow/core/params.py:1080-1213
//   no-leader headway 1000               flow/core/kernel/vehicle/traci.py:234-238
//   desired_velocity over present vehicles  flow/core/rewards.py:6-59
//   MultiAgentHighwayPOEnv (kMA)        flow/envs/multiagent/highway.py:97-196,
//                                       flow/envs/multiagent/base.py:116-225
// SUMO's insertion / arrival rules have no reference fixture: parity unpinned.
//
// kMA = true (FS_ENV_MA_HIGHWAY_PO): RL vehicles may enter through an inflow, so
// every vehicle carries a persistent id through the compaction, the RL vehicles
// present are ranked from the downstream end by ballot + popc each step
// (rank a <-> action row a <-> observation row a), and each agent gets its own
// observation row and reward.
#pragma once
#include "fs_common.cuh"
#include "fs_ring.cuh"
#include "fs_ring_fast.cuh"

namespace fs {

template <typename T>
struct HwBufs {
  T* pos; T* vel; uint8_t* vtype; uint8_t* active;   // [E][CAP]
  T* length;                                         // [E] edge length
  int* n_active; int* departed; int* arrived;  // [E]
  int* spawned;                                // [FS_MAX_INFLOWS][E] vehicles emitted per inflow
  int* pending;                                // [FS_MAX_INFLOWS][E] probabilistic flows: due, not yet in
  int* time_counter; int* episode;
  uint8_t* done; uint8_t* crash;
  T* obs; T* reward;
  T* v_accel; T* v_headway; int* v_leader; int* v_follower;  // views [E][CAP] or null
  double* stats;
  int n_flows;
  double period[FS_MAX_INFLOWS], begin[FS_MAX_INFLOWS], end[FS_MAX_INFLOWS];   // s (end <= 0: none)
  double prob[FS_MAX_INFLOWS];                 // per second; > 0: probabilistic flow
  T depart_speed[FS_MAX_INFLOWS];
  int inflow_type[FS_MAX_INFLOWS];
  // multi-agent front (kMA)
  int* vid;            // [E][CAP] persistent vehicle id
  T* agent_reward;     // [E][A]
  int* agent_id;       // [E][A]
  int max_agents;
};

template <typename T>
__host__ __device__ inline size_t hw_smem_per_warp(int cap) {
  return (size_t)cap * (3 * sizeof(T) + 2 * sizeof(int));
}

// NF = inflows compiled in (1: the single <flow> of c4; FS_MAX_INFLOWS otherwise);
// kProb = probabilistic flows and flow ends compiled in;
// kGeneric = false: every type follows the IDM law (IDM, or the simulator model of SIM / RL
// vehicles), so the controller switch, follower and mean-speed inputs are compiled out
template <typename T, int SLOTS, bool kMA, int NF, bool kProb, bool kGeneric>
__global__ void __launch_bounds__(128)
highway_kernel(const __grid_constant__ Cfg<T> cfg, const HwBufs<T> B, const StepArgs<T> A) {
  typedef Num<T> N;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane_id = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int E = cfg.n_envs, CAP = cfg.n_veh;
  const int e = blockIdx.x * (blockDim.x >> 5) + wib;
  if (e >= E) return;
  if (A.mask != nullptr && A.mask[e] == 0) return;
  unsigned char* mine = smem_raw + hw_smem_per_warp<T>(CAP) * wib;
  T* xs = reinterpret_cast<T*>(mine);
  T* vs = xs + CAP;
  T* hs = vs + CAP;
  int* ts = reinterpret_cast<int*>(hs + CAP);
  int* is = ts + CAP;   // vehicle ids during the compaction (kMA)

  T x[SLOTS], v[SLOTS], h[SLOTS], acc[SLOTS], aeff[SLOTS];
  int ty[SLOTS];
  int id[kMA ? SLOTS : 1], rank[kMA ? SLOTS : 1];
  const size_t eo = (size_t)e * CAP;
  int n = B.n_active[e];
#pragma unroll
  for (int r = 0; r < SLOTS; ++r) {
    const int s = lane_id + 32 * r;
    const bool ok = s < n;
    x[r] = ok ? B.pos[eo + s] : T(0);
    v[r] = ok ? B.vel[eo + s] : T(0);
    ty[r] = ok ? (int)B.vtype[eo + s] : 0;
    acc[r] = T(0);
    if (kMA) id[r] = ok ? B.vid[eo + s] : -1;
  }
  const int AG = kMA ? B.max_agents : 0;
  int n_rl = 0;
  const T Lr = B.length[e];
  const T dt = cfg.dt;
  int tc = B.time_counter[e];
  const int episode = B.episode[e];
  bool done = A.ignore_done ? false : (B.done[e] != 0);
  bool crash = B.crash[e] != 0;
  const int t_end = cfg.sims_per_step * (cfg.warmup_steps + cfg.horizon);
  const long long gid = cfg.env_id_base + e;
  T reward = B.reward[e];
  int departed = B.departed[e], arrived = B.arrived[e];
  int spawned[NF];
#pragma unroll
  for (int f = 0; f < NF; ++f) spawned[f] = f < B.n_flows ? B.spawned[(size_t)f * E + e] : 0;
  int pending[NF];
#pragma unroll
  for (int f = 0; f < NF; ++f)
    pending[f] = (kProb && f < B.n_flows && B.prob[f] > 0.0) ? B.pending[(size_t)f * E + e] : 0;
  double st_rew = 0, st_v = 0, st_v2 = 0, st_vs = 0, st_es = 0, st_cr = 0;

  // Slot rows (32 slots each) that hold a vehicle: at 2000 veh/h about 14 of c4's 160 slots are
  // occupied, so every per-row loop below stops at the populated rows (warp-uniform bound);
  // rows beyond keep x = v = 0, type 0 and are neither published nor read.
  auto rows_of = [&](int m) { const int q = (m + 31) >> 5; return q < SLOTS ? q : SLOTS; };
  auto publish = [&]() {
    __syncwarp();
    const int rows = rows_of(n);
#pragma unroll
    for (int r = 0; r < SLOTS; ++r) {
      if (r >= rows) break;
      const int s = lane_id + 32 * r;
      if (s < CAP) { xs[s] = x[r]; vs[s] = v[r]; ts[s] = ty[r]; }
    }
    __syncwarp();
  };
  auto update = [&]() {  // headway to slot s-1; the head sees 1000 (no leader)
    publish();
    const int rows = rows_of(n);
#pragma unroll
    for (int r = 0; r < SLOTS; ++r) {
      if (r >= rows) { h[r] = T(1000); continue; }
      const int s = lane_id + 32 * r;
      h[r] = (s > 0 && s < n) ? (xs[s - 1] - x[r]) - cfg.types[ts[s - 1]].length : T(1000);
    }
    __syncwarp();
#pragma unroll
    for (int r = 0; r < SLOTS; ++r) {
      if (r >= rows) break;
      const int s = lane_id + 32 * r;
      if (s < CAP) hs[s] = h[r];
    }
    __syncwarp();
  };
  update();

  // RL vehicles present, ranked head first: rank a <-> action row a <-> observation row a
  unsigned rl_types = 0;   // bit t: type t is driven by RL actions (no table lookup per slot and step)
  if (kMA)
    for (int t = 0; t < cfg.n_types; ++t) rl_types |= (cfg.types[t].ctrl == FS_CTRL_RL ? 1u : 0u) << t;
  auto rank_agents = [&]() {
    if (!kMA) return;
    n_rl = 0;
    const int rows = rows_of(n);
#pragma unroll
    for (int r = 0; r < SLOTS; ++r) {
      if (r >= rows) { rank[kMA ? r : 0] = -1; continue; }
      const int s = lane_id + 32 * r;
      const bool rl = s < n && ((rl_types >> ty[r]) & 1u);
      const unsigned m = __ballot_sync(0xffffffffu, rl);
      rank[kMA ? r : 0] = rl ? n_rl + __popc(m & ((1u << lane_id) - 1u)) : -1;
      n_rl += __popc(m);
    }
  };
  rank_agents();
  // MultiAgentHighwayPOEnv.get_state / compute_reward (multiagent/highway.py:97-172) for the
  // agents present after the step; cost1 = rewards.desired_velocity of the whole network
  auto write_agents = [&](T* o, T* ol, bool with_reward, T cost1, bool fail) {
    if (!kMA) return;
    T rsum = T(0);
    const int rows = rows_of(n);
#pragma unroll
    for (int r = 0; r < SLOTS; ++r) {
      if (r >= rows) break;
      const int s = lane_id + 32 * r;
      const int a = rank[kMA ? r : 0];
      if (a < 0 || a >= AG) continue;
      const T vi = v[r];
      const bool has_lead = s > 0, has_f = s + 1 < n;
      // "lead_head" is the LEADER's own headway (highway.py:118), 1000 for the head vehicle
      const T lead_speed = has_lead ? vs[s - 1] : cfg.speed_limit;
      const T lead_head = has_lead ? hs[s - 1] : Lr;
      const T follow_speed = has_f ? vs[s + 1] : T(0);
      const T follow_head = has_f ? hs[s + 1] : Lr;
      const T q[5] = {N::div(vi, cfg.speed_limit), N::div(lead_speed - vi, cfg.speed_limit),
                      N::div(lead_head, Lr), N::div(vi - follow_speed, cfg.speed_limit),
                      N::div(follow_head, Lr)};
#pragma unroll
      for (int c = 0; c < 5; ++c) {
        if (o) o[a * 5 + c] = q[c];
        if (ol) ol[a * 5 + c] = q[c];
      }
      if (ol) B.agent_id[(size_t)e * AG + a] = id[kMA ? r : 0];
      if (with_reward) {
        T rw;
        if (cfg.evaluate) rw = vi;
        else if (fail) rw = T(0);
        else {
          T cost2 = T(0);
          if (has_lead && vi > T(0)) {
            const T th = N::max(N::div(h[r], vi), T(0));   // time headway, t_min = 1
            cost2 = N::min(N::div(th - T(1), T(1)), T(0));
          }
          rw = N::max(add(mul(T(1), cost1), mul(T(0.1), cost2)), T(0));
        }
        rsum = add(rsum, rw);
        if (ol) B.agent_reward[(size_t)e * AG + a] = rw;
      }
    }
    for (int a = lane_id; a < AG; a += 32) {
      if (a < n_rl) continue;
#pragma unroll
      for (int c = 0; c < 5; ++c) {
        if (o) o[a * 5 + c] = T(0);
        if (ol) ol[a * 5 + c] = T(0);
      }
      if (ol) { B.agent_id[(size_t)e * AG + a] = -1; B.agent_reward[(size_t)e * AG + a] = T(0); }
    }
    if (with_reward) {   // env-level reward = mean over the agents present (statistics only)
      for (int of = 16; of > 0; of >>= 1) rsum += __shfl_xor_sync(0xffffffffu, rsum, of);
      const int na = n_rl < AG ? n_rl : AG;
      reward = na > 0 ? N::div(rsum, T(na)) : T(0);
    }
  };

  auto write_obs = [&](T* o, T* ol) {
    if (kMA) { write_agents(o, ol, false, T(0), false); return; }
#pragma unroll
    for (int r = 0; r < SLOTS; ++r) {
      const int s = lane_id + 32 * r;
      if (s >= CAP) continue;
      const bool live_row = 32 * r < n;     // warp-uniform
      const T a0 = (live_row && s < n) ? N::div(v[r], cfg.speed_limit) : T(0);
      const T a1 = (live_row && s < n) ? N::div(x[r], Lr) : T(0);
      if (o) { o[s] = a0; o[CAP + s] = a1; }
      if (ol) { ol[s] = a0; ol[CAP + s] = a1; }
    }
  };
  if (A.k_steps == 0) write_obs(nullptr, B.obs + (size_t)e * cfg.obs_dim);

  for (int k = 0; k < A.k_steps; ++k) {
    const bool done_before = done;
    if (!done) {
      T mean_v = T(0);
      if (kGeneric && (cfg.any_follower & 2)) {
        T sm = T(0);
#pragma unroll
        for (int r = 0; r < SLOTS; ++r) if (lane_id + 32 * r < n) sm = add(sm, v[r]);
        for (int o = 16; o > 0; o >>= 1) sm += __shfl_xor_sync(0xffffffffu, sm, o);
        mean_v = N::div(sm, T(n > 0 ? n : 1));
      }
      // controllers -> apply_acceleration -> Euler move (all read the published state)
      const int rows0 = rows_of(n);
#pragma unroll
      for (int r = 0; r < SLOTS; ++r) {
        if (r >= rows0) break;
        const int s = lane_id + 32 * r;
        if (s >= n) continue;
        const TypeP<T>& tp = cfg.types[ty[r]];
        const int ctrl = tp.ctrl;
        const bool has_lead = s > 0, has_f = s + 1 < n;
        const T vi = v[r], hi = h[r];
        const T vl = has_lead ? vs[s - 1] : T(-1001);
        T st0 = T(0);
        T a, a_sim = T(0);
        bool have_sim = false;
        if (kGeneric)
          a = ctrl_accel<T>(tp, ctrl, vi, has_lead, vl, hi, has_f, has_f ? vs[s + 1] : T(-1001),
                            has_f ? hs[s + 1] : T(-1001), mean_v, dt, cfg.inv_dt, st0);
        else if (sizeof(T) == 4 && (tp.speed_mode & 1) && tp.e_delta == T(4)) {
          // IDM-law vehicle under a speed mode: its law and the simulator model behind the
          // mode's safe speed evaluated together (idm_pair_f32, fs_ring_fast.cuh)
          IdmRegs<float> pp, pq;
          pp.s0 = (float)tp.e_s0; pp.Tt = (float)tp.e_T; pp.inv_tsab = (float)tp.e_inv_tsab;
          pp.inv_v0 = (float)tp.e_inv_v0; pp.a = (float)tp.e_a;
          pq.s0 = (float)tp.sim_min_gap; pq.Tt = (float)tp.sim_tau; pq.inv_tsab = (float)tp.sim_inv_two_sqrt_ab;
          pq.inv_v0 = (float)tp.sim_inv_v0; pq.a = (float)tp.max_accel;
          float af, asf;
          idm_pair_f32((float)vi, (float)vl, (float)hi, pp, pq, af, asf, has_lead);
          a = (T)af; a_sim = (T)asf; have_sim = true;
        } else   // the same arithmetic as ctrl_accel's IDM / SIM cases (fs_ring.cuh does the same)
          a = idm_accel<T>(vi, vl, hi, has_lead, tp.e_v0, tp.e_inv_v0, tp.e_T, tp.e_a, tp.e_tsab,
                           tp.e_inv_tsab, tp.e_delta, tp.e_s0);
        bool commanded = ctrl != FS_CTRL_SIM && ctrl != FS_CTRL_RL;
        if (kMA && ctrl == FS_CTRL_RL && A.actions != nullptr && rank[kMA ? r : 0] < AG) {
          // MultiAgentHighwayPOEnv._apply_rl_actions: apply_acceleration(rl_id, actions[0]),
          // not clipped; NaN = this agent got no command (stays with the simulator model)
          const T raw = A.actions[((size_t)k * cfg.act_dim + rank[kMA ? r : 0]) * E + e];
          if (raw == raw) { a = raw; commanded = true; }
        }
        if (ctrl != FS_CTRL_SIM && ctrl != FS_CTRL_RL) {
          if (tp.noise > T(0))
            a = add(a, mul(tp.noise, normal_sample<T>(cfg.seed, gid, s, tc + episode * 1000003)));
          if (tp.fail_safe == FS_FAILSAFE_INSTANTANEOUS)
            a = failsafe_instantaneous<T>(a, vi, hi, has_lead, n, dt);
          else if (tp.fail_safe == FS_FAILSAFE_SAFE_VELOCITY)
            a = failsafe_safe_velocity<T>(a, vi, vl, hi, n, dt, tp.delay);
        }
        acc[r] = a;      // commanded (the `accel` view)
        // a speed Flow commands passes SUMO's speed mode (restated, fs_common.cuh)
        T ae = a;
        if (commanded && tp.speed_mode != 0) {
          if (have_sim) {
            ae = N::min(a, a_sim);
            if (tp.speed_mode & 2) ae = N::min(ae, tp.max_accel);
            if (tp.speed_mode & 4) ae = N::max(ae, T(0) - tp.max_decel);
          } else {
            ae = speed_mode_accel<T>(tp, a, vi, vl, hi, has_lead);
          }
        }
        aeff[r] = ae;
      }
      __syncwarp();
      tc += 1;
      // move, flag arrivals, compact survivors with ballots (order preserved)
      int base = 0;
      int dst[SLOTS];
#pragma unroll
      for (int r = 0; r < SLOTS; ++r) {
        if (r >= rows0) { dst[r] = -1; continue; }
        const int s = lane_id + 32 * r;
        bool keep = false;
        if (s < n) {
          const T vn = N::max(add(v[r], mul(aeff[r], dt)), T(0));
          x[r] = add(x[r], move_dist<T>(v[r], vn, dt, cfg.ballistic));
          v[r] = vn;
          keep = x[r] < Lr;
        }
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        dst[r] = keep ? base + __popc(m & ((1u << lane_id) - 1u)) : -1;
        base += __popc(m);
      }
      const int n_keep = base;
      arrived += n - n_keep;
      if (n_keep != n) {  // scatter survivors to their new slots through shared memory
        __syncwarp();
#pragma unroll
        for (int r = 0; r < SLOTS; ++r)
          if (dst[r] >= 0) {
            xs[dst[r]] = x[r]; vs[dst[r]] = v[r]; ts[dst[r]] = ty[r]; hs[dst[r]] = acc[r];
            if (kMA) is[dst[r]] = id[kMA ? r : 0];
          }
        __syncwarp();
#pragma unroll
        for (int r = 0; r < SLOTS; ++r) {
          if (r >= rows0) break;
          const int s = lane_id + 32 * r;
          if (s < n_keep) { x[r] = xs[s]; v[r] = vs[s]; ty[r] = ts[s]; acc[r] = hs[s]; }
          else { x[r] = T(0); v[r] = T(0); ty[r] = 0; acc[r] = T(0); }
          if (kMA) id[kMA ? r : 0] = s < n_keep ? is[s] : -1;
        }
        n = n_keep;
      }
      // inflows, in definition order: each emits at most one due vehicle per step at the
      // tail, if the entry gap to the last vehicle allows it
#pragma unroll
      for (int f = 0; f < NF; ++f) {
        if (f >= B.n_flows) continue;
        const double t = (double)tc * (double)cfg.dt;
        const bool probabilistic = kProb && B.prob[f] > 0.0;
        bool want;
        if (probabilistic) {   // one Bernoulli(probability * dt) draw per step inside [begin, end)
          if (t >= B.begin[f] && (B.end[f] <= 0.0 || t < B.end[f])) {
            uint32_t rr[4];
            Philox::gen((uint32_t)tc, 0x40000000u | (uint32_t)f, (uint32_t)gid, (uint32_t)(gid >> 32),
                        (uint32_t)cfg.seed ^ ((uint32_t)episode * 0x9E3779B9u),
                        (uint32_t)(cfg.seed >> 32), rr);
            if ((double)rr[0] * (1.0 / 4294967296.0) < B.prob[f] * (double)cfg.dt) pending[f] += 1;
          }
          want = pending[f] > 0;
        } else {
          if (!(B.period[f] > 0.0)) continue;
          const double due = B.begin[f] + (double)spawned[f] * B.period[f];
          want = t >= due && (!kProb || B.end[f] <= 0.0 || due < B.end[f]);
        }
        bool spawn = want && n < CAP;
        if (spawn && n > 0) {
          const int sl = n - 1;  // last vehicle
          T xl = T(0); int tl = 0;
#pragma unroll
          for (int r = 0; r < SLOTS; ++r) if (lane_id + 32 * r == sl) { xl = x[r]; tl = ty[r]; }
          xl = __shfl_sync(0xffffffffu, xl, sl & 31);
          tl = __shfl_sync(0xffffffffu, tl, sl & 31);
          spawn = (xl - cfg.types[tl].length) >= cfg.types[B.inflow_type[f]].sim_min_gap;
        }
        if (spawn) {
#pragma unroll
          for (int r = 0; r < SLOTS; ++r)
            if (lane_id + 32 * r == n) {
              x[r] = T(0); v[r] = B.depart_speed[f]; ty[r] = B.inflow_type[f]; acc[r] = T(0);
              if (kMA) id[kMA ? r : 0] = ((f + 1) << 24) | spawned[f];
            }
          n += 1; spawned[f] += 1; departed += 1;
          if (probabilistic) pending[f] -= 1;
        }
      }
      update();
      bool cr = false;
      T sv = T(0), sq = T(0), s2 = T(0);
      const int rows1 = rows_of(n);
#pragma unroll
      for (int r = 0; r < SLOTS; ++r) {
        if (r >= rows1) break;
        const int s = lane_id + 32 * r;
        if (s >= n) continue;
        cr |= h[r] < T(0);
        sv = add(sv, v[r]);
        const T d = v[r] - cfg.target_velocity;
        sq = add(sq, mul(d, d));
        s2 += v[r] * v[r];
      }
      for (int o = 16; o > 0; o >>= 1) {
        sv += __shfl_xor_sync(0xffffffffu, sv, o);
        sq += __shfl_xor_sync(0xffffffffu, sq, o);
        s2 += __shfl_xor_sync(0xffffffffu, s2, o);
      }
      crash = __any_sync(0xffffffffu, cr);
      done = crash || tc >= t_end;
      if (cfg.evaluate) reward = N::div(sv, T(n));   // accel.py:111-112 (np.mean: NaN when empty)
      else if (crash || n == 0) reward = T(0);
      else {  // rewards.desired_velocity over the vehicles present
        // || [v_t] * n ||: n equal squares; their sum is exact (and so independent of numpy's
        // summation order) whenever v_t^2 * n is representable, which one multiply gives as well
        const T t2 = mul(cfg.target_velocity, cfg.target_velocity);
        const T mc2 = mul(t2, T(n));
        const T max_cost = N::sqrt(mc2), cost = N::sqrt(sq);
        reward = N::div(N::max(max_cost - cost, T(0)), add(max_cost, T(1.1920928955078125e-07)));
      }
      if (kMA) {
        rank_agents();
        const bool lastk = (k == A.k_steps - 1);
        // fail = the engine's collision flag (the reference's MultiEnv zeroes it, base.py:174)
        write_agents(A.obs_seq ? A.obs_seq + ((size_t)k * E + e) * cfg.obs_dim : nullptr,
                     lastk ? B.obs + (size_t)e * cfg.obs_dim : nullptr, true,
                     (crash || n == 0) ? T(0) : reward, crash);
      }
      if (A.accumulate_stats && lane_id == 0) {
        st_rew += (double)reward; st_v += (double)sv; st_v2 += (double)s2;
        st_vs += n; st_es += 1; st_cr += crash ? 1 : 0;
      }
    } else {
      reward = T(0);
    }
    const bool last = (k == A.k_steps - 1);
    if (kMA) {
      if (done_before) write_obs(A.obs_seq ? A.obs_seq + ((size_t)k * E + e) * cfg.obs_dim : nullptr,
                                 last ? B.obs + (size_t)e * cfg.obs_dim : nullptr);
    } else if (A.obs_seq != nullptr || last)
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
    B.active[eo + s] = s < n ? 1 : 0;
    if (kMA) B.vid[eo + s] = s < n ? id[kMA ? r : 0] : -1;
    if (A.write_views) {
      B.v_headway[eo + s] = s < n ? h[r] : T(0);
      B.v_leader[eo + s] = (s > 0 && s < n) ? s - 1 : -1;
      B.v_follower[eo + s] = (s + 1 < n) ? s + 1 : -1;
      if (A.k_steps > 0) B.v_accel[eo + s] = s < n ? acc[r] : T(0);
    }
  }
  if (lane_id == 0) {
    B.n_active[e] = n; B.departed[e] = departed; B.arrived[e] = arrived;
#pragma unroll
    for (int f = 0; f < NF; ++f) if (f < B.n_flows) {
      B.spawned[(size_t)f * E + e] = spawned[f];
      if (kProb && B.prob[f] > 0.0) B.pending[(size_t)f * E + e] = pending[f];
    }
    B.time_counter[e] = tc;
    B.done[e] = done ? 1 : 0;
    B.crash[e] = crash ? 1 : 0;
    B.reward[e] = reward;
    if (A.accumulate_stats) {
      double vals[6] = {st_rew, st_v, st_v2, st_vs, st_es, st_cr};
      substep_stats(A.accumulate_stats, vals);
      for (int q = 0; q < 6; ++q) if (vals[q] != 0.0) atomicAdd(&stat_stripe(B.stats)[q], vals[q]);
    }
  }
}

// Env.reset on the highway: initial vehicles by gen_even_start_pos on the single
// edge (flow/core/kernel/network/base.py:263-391), stored head first.
template <typename T>
__global__ void highway_reset_kernel(const __grid_constant__ Cfg<T> cfg, const HwBufs<T> B,
                                     const uint8_t* mask, int n_initial) {
  const int E = cfg.n_envs, CAP = cfg.n_veh;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  if (mask != nullptr && mask[e] == 0) return;
  const double L = (double)cfg.length_lo;
  const double min_gap = fmax(0.0, (double)cfg.init_min_gap);
  const double bunching = fmax(0.0, (double)cfg.bunching);
  const size_t eo = (size_t)e * CAP;
  if (n_initial > 0) {
    const double available = (L - bunching) - n_initial * (min_gap + 5.0);
    const double increment = available / n_initial;
    // positions grow with the index until the `% length` of gen_even_start_pos wraps them to
    // the start of the edge (x0 > 0; at most once, the vehicles span less than the edge):
    // indices [0, w) lie downstream of indices [w, n). Head first: w-1 .. 0, n-1 .. w.
    int w = n_initial;
    {
      double x = (double)cfg.x0;
      for (int i = 1; i < n_initial; ++i) {
        const double xn = fmod(__dadd_rn(__dadd_rn(__dadd_rn(x, increment), 5.0), min_gap), L);
        if (xn < x) { w = i; break; }
        x = xn;
      }
    }
    double x = (double)cfg.x0;
    for (int i = 0; i < n_initial; ++i) {
      const int s = i < w ? w - 1 - i : n_initial - 1 - (i - w);
      B.pos[eo + s] = (T)x; B.vel[eo + s] = T(0);
      B.vtype[eo + s] = cfg.type_of[i]; B.active[eo + s] = 1;
      if (B.vid) B.vid[eo + s] = i;
      x = fmod(__dadd_rn(__dadd_rn(__dadd_rn(x, increment), 5.0), min_gap), L);
    }
  }
  for (int s = n_initial; s < CAP; ++s) {
    B.pos[eo + s] = T(0); B.vel[eo + s] = T(0); B.vtype[eo + s] = 0; B.active[eo + s] = 0;
    if (B.vid) B.vid[eo + s] = -1;
  }
  B.length[e] = (T)L;
  B.n_active[e] = n_initial; B.departed[e] = n_initial; B.arrived[e] = 0;
  for (int f = 0; f < FS_MAX_INFLOWS; ++f) { B.spawned[(size_t)f * E + e] = 0; B.pending[(size_t)f * E + e] = 0; }
  B.time_counter[e] = 0;
  B.episode[e] = B.episode[e] + 1;
  B.done[e] = 0; B.crash[e] = 0;
}

// host-side dispatch (instantiated in fs_hw_f32.cu / fs_hw_f64.cu). multi_agent:
// FS_ENV_MA_HIGHWAY_PO; single_periodic: one periodic <flow> without end (c4)
template <typename T>
static cudaError_t launch_hw_impl(int cap, bool multi_agent, bool single_periodic, bool generic,
                                  int grid, int block, size_t smem, cudaStream_t st,
                                  const Cfg<T>& cfg, const HwBufs<T>& b, const StepArgs<T>& a) {
  constexpr int NF = FS_MAX_INFLOWS;
#define FS_HW2(SL, MA, F, PR, GEN) \
  highway_kernel<T, SL, MA, F, PR, GEN><<<grid, block, smem, st>>>(cfg, b, a)
#define FS_HW(SL, MA, F, PR) \
  do { if (generic) FS_HW2(SL, MA, F, PR, true); else FS_HW2(SL, MA, F, PR, false); } while (0)
  if (multi_agent) {
    if (cap <= 64) FS_HW(2, true, NF, true); else if (cap <= 160) FS_HW(5, true, NF, true); else FS_HW(8, true, NF, true);
  } else if (single_periodic) {
    if (cap <= 64) FS_HW(2, false, 1, false); else if (cap <= 160) FS_HW(5, false, 1, false); else FS_HW(8, false, 1, false);
  } else {
    if (cap <= 64) FS_HW(2, false, NF, true); else if (cap <= 160) FS_HW(5, false, NF, true); else FS_HW(8, false, NF, true);
  }
#undef FS_HW
#undef FS_HW2
  return cudaGetLastError();
}

}  // namespace fs
