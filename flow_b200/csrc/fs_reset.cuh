// fs_reset.cuh — device-side Env.reset for ring environments.
//
// Restates (always in double, then cast to T so fp64 engines match the
// reference bit for bit):
//   WaveAttenuationEnv.reset ring-length draw   flow/envs/ring/wave_attenuation.py:171-174
//   gen_even_start_pos / _get_start_pos_util    flow/core/kernel/network/base.py:263-391,515-608
//   RingNetwork edge-start tables               flow/networks/ring.py:194-216
//   Env.reset counters                          flow/envs/base.py:431
// The reference draws the length with Python's global MT19937; here it is a
// Philox draw keyed by (seed, global env id, episode) unless the host passes
// explicit lengths (the single-env gym view does, to reproduce random.randint).
#pragma once
#include "fs_common.cuh"

namespace fs {

template <typename T>
struct ResetBufs {
  T* pos; T* vel; T* circ; T* st0;
  double* hist_sum; int* hist_cnt; int n_pi;
  uint8_t* lane; uint8_t* done; uint8_t* crash;
  int* time_counter; int* episode;
  int* last_lc;             // [E][V] or null (multi-lane)
  int vs, es;               // element strides of pos/vel/lane: vehicle, env
  const uint8_t* mask;      // [E] or null
  const double* lengths;    // [E] or null ; <= 0 means "draw"
};

template <typename T>
__global__ void ring_reset_kernel(const __grid_constant__ Cfg<T> cfg, const ResetBufs<T> B) {
  const int E = cfg.n_envs, V = cfg.n_veh;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  if (B.mask != nullptr && B.mask[e] == 0) return;
  const long long gid = cfg.env_id_base + e;
  const int episode = B.episode[e] + 1;

  // ring length: integer in [lo, hi]
  double L;
  if (B.lengths != nullptr && B.lengths[e] > 0.0) {
    L = B.lengths[e];
  } else {
    const long long lo = (long long)cfg.length_lo, hi = (long long)cfg.length_hi;
    if (hi > lo) {
      uint32_t r[4];
      Philox::gen((uint32_t)episode, 0x52534554u /* 'RSET' */, (uint32_t)gid,
                  (uint32_t)(gid >> 32), (uint32_t)cfg.seed, (uint32_t)(cfg.seed >> 32), r);
      const unsigned long long n = (unsigned long long)(hi - lo + 1);
      L = (double)(lo + (long long)(((unsigned long long)r[0] * n) >> 32));
    } else {
      L = (double)cfg.length_lo;
    }
  }
  const double j = (double)cfg.junction_len;
  const double C = L + 4 * j;

  // _get_start_pos_util
  const double min_gap = fmax(0.0, (double)cfg.init_min_gap);
  const double bunching = fmax(0.0, (double)cfg.bunching);
  int lanes_distr = cfg.lanes_distribution;
  if (lanes_distr > cfg.n_lanes) lanes_distr = cfg.n_lanes;
  if (lanes_distr < 1) lanes_distr = 1;
  const double edge_len = L / 4.0;
  double distribution_length = 0.0;
  if (edge_len > min_gap + 5.0) {
    const double a = edge_len * (double)min(cfg.n_lanes, lanes_distr);
    for (int q = 0; q < 4; ++q) distribution_length = __dadd_rn(distribution_length, a);
  }
  const double available = (distribution_length - lanes_distr * bunching) - V * (min_gap + 5.0);
  // (available < 0 is rejected on the host before launch: FatalFlowError analogue)
  const double increment = available / V;

  // total_edgestarts, ordered: bottom, :right_0, right, :top_0, top, :left_0, left, :bottom_0
  const double starts[8] = {0.0,           0.25 * L,          0.25 * L + j,     0.5 * L + j,
                            0.5 * L + 2 * j, 0.75 * L + 2 * j, 0.75 * L + 3 * j, L + 3 * j};
  double x = (double)cfg.x0;
  int count = 0;
  const int per_col = min(cfg.n_lanes, lanes_distr);
  while (count < V) {
    int idx = 0;
    for (int q = 0; q < 8; ++q) if (x >= starts[q]) idx = q;
    if (idx & 1) {  // inside a junction slot -> start of the next edge
      idx = (idx == 7) ? 0 : idx + 1;
      x = starts[idx];
    }
    for (int lane = 0; lane < per_col && count < V; ++lane) {
      const size_t so = (size_t)count * B.vs + (size_t)e * B.es;
      B.pos[so] = (T)x;
      B.vel[so] = T(0);
      B.lane[so] = (uint8_t)lane;
      if (B.last_lc) B.last_lc[so] = -(1 << 30);  // -inf: vehicle/traci.py:178-180
      if (B.st0) {
        const TypeP<T>& tp = cfg.types[cfg.type_of[count]];
        B.st0[(size_t)count * E + e] = (tp.ctrl == FS_CTRL_LAC) ? tp.p[4] : T(0);
      }
      ++count;
    }
    x = fmod(__dadd_rn(__dadd_rn(__dadd_rn(x, increment), 5.0), min_gap), L);
  }
  for (int p = 0; p < B.n_pi; ++p) {
    B.hist_sum[(size_t)p * E + e] = 0.0;
    B.hist_cnt[(size_t)p * E + e] = 0;
  }
  B.circ[e] = (T)C;
  B.time_counter[e] = 0;
  B.episode[e] = episode;
  B.done[e] = 0;
  B.crash[e] = 0;
}

}  // namespace fs
