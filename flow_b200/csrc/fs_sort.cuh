// fs_sort.cuh — additional_params['sort_vehicles'] (FS_ENVF_SORT_VEHICLES).
//
// Restated from the reference:
//   AccelEnv.sorted_ids / _get_abs_position     flow/envs/ring/accel.py:150-169
//   AccelEnv.additional_command (bookkeeping)   flow/envs/ring/accel.py:125-148
//   AccelEnv.reset (keys := positions)          flow/envs/ring/accel.py:170-183
//   AccelEnv._apply_rl_actions (sorted RL ids)  flow/envs/ring/accel.py:100-106
//   LaneChangeAccelEnv.get_state / _apply_rl_actions use sorted_ids too
//   (lane_change_accel.py:114-154) but its additional_command (:156-161) drops the
//   bookkeeping, so its keys stay what AccelEnv.reset stored.
//
// The option is off in every shipped config, so it stays out of the register-resident
// step kernels: a step with the flag set is  sort_pre -> step kernel (k = 1) -> sort_obs.
//   sort_pre : action slot r goes to the r-th RL vehicle in key order (keys as the
//              PREVIOUS additional_command left them), then - AccelEnv only - the
//              bookkeeping on the pre-move positions:
//                abs <- (abs + (x - prev)) % length ; prev <- x
//              (a vehicle without an entry yet: abs = x % length, change = 0).
//   sort_obs : rows of the observation ([speed | position (| lane)], V each) permuted
//              by the stable rank of the keys; a vehicle without a key has -1001, so the
//              very first reset keeps id order like the reference.
// One warp per environment; keys live in [E][V] arrays of their own and persist across
// resets exactly like the reference's dicts.
#pragma once
#include "fs_common.cuh"

namespace fs {

template <typename T>
struct SortBufs {
  const T* pos; int vs, es;       // strides of pos: vehicle, env
  T* abs_pos; T* prev_pos;        // [E][V]
  uint8_t* keys_valid;            // [E]
  const T* circ;                  // [E]
  const uint8_t* done;            // [E]
  const uint8_t* mask;            // [E] or null
};

template <typename T>
__device__ __forceinline__ T py_mod(T a, T m) {  // Python float %, m > 0
  T r = sizeof(T) == 8 ? (T)fmod((double)a, (double)m) : (T)fmodf((float)a, (float)m);
  if (r < T(0)) r += m;
  return r;
}

// stable rank of vehicle i among the vehicles selected by `sel` (key order, ties by index)
template <typename T, typename Sel>
__device__ __forceinline__ int key_rank(const T* keys, bool valid, int V, int i, Sel sel) {
  if (!valid) {  // every key is -1001: id order
    int r = 0;
    for (int j = 0; j < i; ++j) r += sel(j) ? 1 : 0;
    return r;
  }
  const T ki = keys[i];
  int r = 0;
  for (int j = 0; j < V; ++j)
    if (sel(j) && (keys[j] < ki || (keys[j] == ki && j < i))) ++r;
  return r;
}

template <typename T>
__global__ void __launch_bounds__(128)
sort_pre_kernel(const __grid_constant__ Cfg<T> cfg, const SortBufs<T> B, const T* actions,
                T* act_sorted, int update_abs, int ignore_done) {
  const int lane_id = threadIdx.x & 31;
  const int e = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int E = cfg.n_envs, V = cfg.n_veh;
  if (e >= E) return;
  if (B.mask != nullptr && B.mask[e] == 0) return;
  const bool valid = B.keys_valid[e] != 0;
  T* keys = B.abs_pos + (size_t)e * V;
  T* prev = B.prev_pos + (size_t)e * V;
  if (actions != nullptr) {
    const int rows = cfg.n_rl > 0 ? cfg.act_dim / cfg.n_rl : 1;   // 2 for [accel, direction]
    for (int i = lane_id; i < V; i += 32) {
      const int slot = cfg.rl_index[i];
      if (slot < 0) continue;
      const int r = key_rank<T>(keys, valid, V, i, [&](int j) { return cfg.rl_index[j] >= 0; });
      for (int q = 0; q < rows; ++q)
        act_sorted[(size_t)(rows * slot + q) * E + e] = actions[(size_t)(rows * r + q) * E + e];
    }
  }
  __syncwarp();
  const bool frozen = !ignore_done && B.done[e] != 0;
  if (update_abs && !frozen) {
    const T C = B.circ[e];
    for (int i = lane_id; i < V; i += 32) {
      const T x = B.pos[(size_t)i * B.vs + (size_t)e * B.es];
      const T change = valid ? x - prev[i] : T(0);
      keys[i] = py_mod<T>((valid ? keys[i] : x) + change, C);
      prev[i] = x;
    }
    __syncwarp();
    if (lane_id == 0) B.keys_valid[e] = 1;
  }
}

// AccelEnv.reset tail: absolute_position = prev_pos = get_x_by_id
template <typename T>
__global__ void sort_setkeys_kernel(const __grid_constant__ Cfg<T> cfg, const SortBufs<T> B) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= cfg.n_envs) return;
  if (B.mask != nullptr && B.mask[e] == 0) return;
  const int V = cfg.n_veh;
  for (int i = 0; i < V; ++i) {
    const T x = B.pos[(size_t)i * B.vs + (size_t)e * B.es];
    B.abs_pos[(size_t)e * V + i] = x;
    B.prev_pos[(size_t)e * V + i] = x;
  }
  B.keys_valid[e] = 1;
}

template <typename T, int SLOTS>
__global__ void __launch_bounds__(128)
sort_obs_kernel(const __grid_constant__ Cfg<T> cfg, const SortBufs<T> B, T* obs, T* obs_step,
                int rows) {
  const int lane_id = threadIdx.x & 31;
  const int e = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int E = cfg.n_envs, V = cfg.n_veh;
  if (e >= E) return;
  if (B.mask != nullptr && B.mask[e] == 0) return;
  const bool valid = B.keys_valid[e] != 0;
  if (!valid) return;  // identity
  const T* keys = B.abs_pos + (size_t)e * V;
  int rank[SLOTS];
#pragma unroll
  for (int s = 0; s < SLOTS; ++s) {
    const int i = lane_id + 32 * s;
    rank[s] = i < V ? key_rank<T>(keys, true, V, i, [](int) { return true; }) : 0;
  }
  T* dst[2] = {obs ? obs + (size_t)e * cfg.obs_dim : nullptr,
               obs_step ? obs_step + (size_t)e * cfg.obs_dim : nullptr};
  for (int d = 0; d < 2; ++d) {
    T* o = dst[d];
    if (!o) continue;
    for (int q = 0; q < rows; ++q) {
      T val[SLOTS];
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) {
        const int i = lane_id + 32 * s;
        val[s] = i < V ? o[q * V + i] : T(0);
      }
      __syncwarp();
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) {
        const int i = lane_id + 32 * s;
        if (i < V) o[q * V + rank[s]] = val[s];
      }
      __syncwarp();
    }
  }
}

}  // namespace fs
