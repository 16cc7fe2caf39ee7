// flowstep.cu — engine object + C-ABI (include/flowstep.h) of libflowstep.so.
// Host side only orchestrates: all arithmetic of the step path runs in the
// CUDA kernels of fs_ring.cuh / fs_reset.cuh. There is NO CPU fallback: every
// entry point fails with FS_ECUDA when no sm_100 device is usable.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <type_traits>
#include <vector>

#include "../../include/flowstep.h"
#include "fs_common.cuh"
#include "fs_reset.cuh"
#include "fs_launch.h"
#include "fs_sort.cuh"

static thread_local std::string g_create_error;

struct fs_engine {
  fs_config cfg;
  bool f64 = false;
  bool types_set = false;
  bool generic = false;   // some controller is not IDM/SIM/RL
  bool fast_ok = false;   // fs_ring_fast.cuh applies (see fs_set_types)
  int fast_variant = 0;   // 0 plain, 1 noise, 2 noise+fail-safe+speed mode, 3 = 2 + views, 4 = noise+speed mode
  bool ml_uniform = false;  // multi-lane kernel: one IDM-law human type + at most one RL type
  int cl_mode = 0;                // FS_CL_MODE at fs_create: zero-copy bits of the closed-loop host step
  bool ml_force_generic = false;  // FS_ML_GENERIC=1 at fs_create: table-driven instantiation (A/B tests)
  bool hot_noise = false;   // the human type of the hot-kernel pattern carries noise
  bool hot_ok = false;    // c2/c5 pattern: 21 IDM + RL vehicle last, WaveAttenuation(PO), no fail-safe / views
  bool no_hot = false;    // FS_NO_HOT=1: keep ring1_fast_kernel (A/B tests)
  int ring_shape = 0;     // FS_RING_SHAPE override: threads per env (1,2,4,8), 0 = auto
  int obs_dim = 0, act_dim = 0, n_rl = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = true;
  std::string err;
  int64_t launches = 0;
  // device state
  void *pos = nullptr, *vel = nullptr, *circ = nullptr, *st0 = nullptr, *hist = nullptr;
  double* hist_sum = nullptr;
  int* hist_cnt = nullptr;
  int hist_len = 0, n_pi = 0;
  uint8_t *lane = nullptr, *active = nullptr, *done = nullptr, *crash = nullptr, *mask = nullptr;
  int *time_counter = nullptr, *episode = nullptr, *leader = nullptr, *follower = nullptr;
  int* last_lc = nullptr;
  int* counters = nullptr;   // highway: [3 + FS_MAX_INFLOWS][E] in network, departed, arrived, emitted per inflow
  uint8_t* vtype = nullptr;  // highway: [E][CAP] type of the vehicle in each slot
  // highway inflows (fs_set_inflows; flow 0 defaults to the one in fs_config)
  int n_flows = 0;
  fs_inflow flows[FS_MAX_INFLOWS] = {};
  // FS_ENV_MA_HIGHWAY_PO
  int* veh_id = nullptr;       // [E][CAP]
  int* agent_id = nullptr;     // [E][A]
  void* agent_reward = nullptr;  // T [E][A]
  bool ma_highway() const { return cfg.env_kind == FS_ENV_MA_HIGHWAY_PO; }
  bool merge() const { return cfg.topology == FS_TOPO_MERGE; }
  bool fig8() const { return cfg.topology == FS_TOPO_FIGURE8; }
  int* seq = nullptr;          // merge: [E][CAP] departure order of the vehicle in each slot
  // fs_step_unfused: the k single-step launches of the last call, kept as an instantiated graph
  cudaGraphExec_t uf_exec = nullptr;
  const void* uf_key[4] = {nullptr, nullptr, nullptr, nullptr};
  int uf_k = 0;
  int64_t uf_launches = 0;
  bool pdl_hint = false;   // set while fs_step_unfused issues its launches
  // closed-loop host step (fs_step_host with few steps per call): ONE pinned host block and ONE
  // device block laid out [actions | obs | reward | done] for cl_k steps, and the sequence
  // H2D(actions) -> step kernel(s) -> D2H(obs | reward | done) kept as an instantiated graph
  unsigned char* cl_host = nullptr;
  unsigned char* cl_hdev = nullptr;   // device-side address of cl_host (mapped pinned memory)
  unsigned char* cl_dev = nullptr;
  int cl_k = 0;
  size_t cl_off[5] = {0, 0, 0, 0, 0};   // byte offsets of actions, obs, reward, done, end
  cudaGraphExec_t cl_exec = nullptr;
  int cl_exec_k = 0;
  bool cl_exec_act = false;
  int64_t cl_launches = 0;
  int64_t epoch = 0;       // bumped by every call that changes what a captured launch would do
  bool env_major = false;  // pos/vel/lane stored [E][V] (multi-lane, highway)
  // FS_ENVF_SORT_VEHICLES (fs_sort.cuh): keys [E][V], their validity [E], permuted actions [A][E]
  void *abs_pos = nullptr, *prev_pos = nullptr, *act_sorted = nullptr;
  uint8_t* keys_valid = nullptr;
  bool sort_on() const { return (cfg.env_flags & FS_ENVF_SORT_VEHICLES) != 0; }
  void *headway = nullptr, *accel = nullptr, *obs = nullptr, *reward = nullptr;
  double *stats = nullptr, *lengths_dev = nullptr;
  // staging for fs_step_host
  void *h_act = nullptr, *h_obs = nullptr, *h_rew = nullptr;
  uint8_t* h_done = nullptr;
  size_t h_act_cap = 0, h_obs_cap = 0, h_rew_cap = 0, h_done_cap = 0;
  // fs_step_host_async: three staging slots, copy streams and events, so that the
  // H2D of call n+1, the kernel of call n and the D2H of call n-1 overlap
  struct Slot {
    void *act = nullptr, *obs = nullptr, *rew = nullptr, *done = nullptr;
    size_t act_cap = 0, obs_cap = 0, rew_cap = 0, done_cap = 0;
    cudaEvent_t up = nullptr, run = nullptr, down = nullptr;
    bool used = false;
  } slots[3];
  cudaStream_t up_stream = nullptr, down_stream = nullptr;
  int64_t async_calls = 0;
  // parameter tables (one of the two is used)
  fs::Cfg<float> cf;
  fs::Cfg<double> cd;
  std::vector<void*> allocs;
  bool env_needs_rl() const {
    return cfg.env_kind == FS_ENV_WAVE_ATTENUATION_PO || cfg.env_kind == FS_ENV_MA_WAVE_ATTENUATION_PO ||
           cfg.env_kind == FS_ENV_MA_ACCEL_PO ||
           cfg.env_kind == FS_ENV_LANE_CHANGE_ACCEL_PO;
  }
  // warp-per-env multi-lane kernel (also the only one with the per-lane neighbour observation)
  bool multilane() const { return cfg.n_lanes > 1 || cfg.env_kind == FS_ENV_LANE_CHANGE_ACCEL_PO; }
};

#define FS_CUDA(e, call)                                                        \
  do {                                                                          \
    cudaError_t _st = (call);                                                   \
    if (_st != cudaSuccess) {                                                   \
      (e)->err = std::string(#call) + ": " + cudaGetErrorString(_st);           \
      return FS_ECUDA;                                                          \
    }                                                                           \
  } while (0)

static int fail(fs_engine* e, int code, const std::string& msg) {
  if (e) e->err = msg; else g_create_error = msg;
  return code;
}

// Captured graphs hold the kernel arguments BY VALUE (type tables, inflow rows, the kernel
// instantiation picked from them) and the stream they were captured on: anything that changes
// those drops the graphs, the next call re-captures.
static void drop_graphs(fs_engine* e) {
  if (e->uf_exec) { cudaGraphExecDestroy(e->uf_exec); e->uf_exec = nullptr; }
  if (e->cl_exec) { cudaGraphExecDestroy(e->cl_exec); e->cl_exec = nullptr; }
  e->epoch++;
}

template <typename P>
static int dev_alloc(fs_engine* e, P** out, size_t bytes, bool zero = true) {
  void* p = nullptr;
  if (bytes == 0) bytes = 16;
  cudaError_t st = cudaMalloc(&p, bytes);
  if (st != cudaSuccess) {
    e->err = std::string("cudaMalloc: ") + cudaGetErrorString(st);
    return st == cudaErrorMemoryAllocation ? FS_ENOMEM : FS_ECUDA;
  }
  if (zero) cudaMemsetAsync(p, 0, bytes, e->stream);
  e->allocs.push_back(p);
  *out = reinterpret_cast<P*>(p);
  return FS_OK;
}

static size_t tsize(const fs_engine* e) { return e->f64 ? 8 : 4; }

extern "C" int fs_abi_version(void) { return FS_ABI_VERSION; }

extern "C" const char* fs_last_error(const fs_engine* e) {
  return e ? e->err.c_str() : g_create_error.c_str();
}

static int obs_dim_for(const fs_config& c) {
  switch (c.env_kind) {
    case FS_ENV_ACCEL:
    case FS_ENV_WAVE_ATTENUATION: return 2 * c.n_veh;
    case FS_ENV_WAVE_ATTENUATION_PO: return 3;
    case FS_ENV_LANE_CHANGE_ACCEL: return 3 * c.n_veh;
    case FS_ENV_LANE_CHANGE_ACCEL_PO:   // capacity; 4 * n_rl * n_lanes + (1 | n_rl) after fs_set_types
      return c.topology == FS_TOPO_RING ? 4 * c.n_veh * c.n_lanes + c.n_veh : -1;
    case FS_ENV_HIGHWAY: return 2 * c.n_veh;
    case FS_ENV_MA_HIGHWAY_PO:
      return c.topology == FS_TOPO_HIGHWAY && c.max_agents >= 1 && c.max_agents <= c.n_veh
                 ? 5 * c.max_agents : -1;
    case FS_ENV_MERGE_PO:
      return c.topology == FS_TOPO_MERGE && c.max_agents >= 1 && c.max_agents <= 32 &&
                     c.max_agents <= c.n_veh ? 5 * c.max_agents : -1;
    case FS_ENV_MA_WAVE_ATTENUATION_PO: return 3 * c.n_veh;  // capacity; 3 * n_rl after fs_set_types
    case FS_ENV_MA_ACCEL_PO:                                  // capacity; 6 * n_rl after fs_set_types
      return c.topology == FS_TOPO_RING && c.n_lanes == 1 ? 6 * c.n_veh : -1;
    default: return -1;
  }
}

extern "C" int fs_create(const fs_config* cfg, fs_engine** out) {
  if (!cfg || !out) return fail(nullptr, FS_EINVAL, "fs_create: null argument");
  if (cfg->abi_version != FS_ABI_VERSION)
    return fail(nullptr, FS_EINVAL, "fs_create: abi_version mismatch");
  if (cfg->n_envs <= 0 || cfg->n_veh <= 0 || cfg->n_veh > FS_MAX_VEH)
    return fail(nullptr, FS_EINVAL, "fs_create: n_envs/n_veh out of range (n_veh <= 256)");
  if (cfg->dtype != FS_F32 && cfg->dtype != FS_F64)
    return fail(nullptr, FS_EINVAL, "fs_create: dtype must be FS_F32 or FS_F64");
  if (!(cfg->sim_step > 0)) return fail(nullptr, FS_EINVAL, "fs_create: sim_step must be > 0");
  if (cfg->n_lanes < 1 || cfg->n_lanes > 8)
    return fail(nullptr, FS_EINVAL, "fs_create: n_lanes must be in 1..8");
  if (obs_dim_for(*cfg) < 0)
    return fail(nullptr, FS_EINVAL, "fs_create: unsupported env_kind");
  if ((cfg->topology == FS_TOPO_MERGE) != (cfg->env_kind == FS_ENV_MERGE_PO))
    return fail(nullptr, FS_EINVAL, "fs_create: FS_TOPO_MERGE goes with FS_ENV_MERGE_PO and vice versa");
  if (cfg->topology == FS_TOPO_MERGE &&
      (!(cfg->merge_pre > 0) || !(cfg->merge_ramp > 0) || !(cfg->merge_post > 0) || cfg->n_lanes != 1 ||
       cfg->merge_zone < 0 || (cfg->env_flags & ~FS_ENVF_EVALUATE)))
    return fail(nullptr, FS_EINVAL, "fs_create: merge network needs merge_pre / merge_ramp / merge_post "
                                    "> 0, one lane per road, merge_zone >= 0, no env flag but evaluate");
  if (cfg->topology == FS_TOPO_FIGURE8 &&
      (cfg->env_kind != FS_ENV_ACCEL || cfg->n_lanes != 1 || !(cfg->fig8_radius > 0) ||
       cfg->fig8_zone < 0 || cfg->fig8_zone > cfg->fig8_radius ||
       !(cfg->fig8_radius > fmax(0.0, cfg->init_min_gap) + 5.0) || !(cfg->fig8_radius > 7.04) ||
       (cfg->env_flags & ~FS_ENVF_EVALUATE)))
    return fail(nullptr, FS_EINVAL, "fs_create: the figure-eight network runs AccelEnv on one lane, needs "
                                    "fig8_radius > max(min_gap + 5, 7.04), 0 <= fig8_zone <= fig8_radius and no env "
                                    "flag but evaluate");
  if ((cfg->env_flags & FS_ENVF_SORT_VEHICLES) &&
      (cfg->topology != FS_TOPO_RING ||
       (cfg->env_kind != FS_ENV_ACCEL && cfg->env_kind != FS_ENV_LANE_CHANGE_ACCEL &&
        cfg->env_kind != FS_ENV_LANE_CHANGE_ACCEL_PO)))
    return fail(nullptr, FS_EINVAL, "fs_create: sort_vehicles exists for AccelEnv / "
                                    "LaneChangeAccel(PO)Env on the ring only");
  if (cfg->env_flags & ~(FS_ENVF_EVALUATE | FS_ENVF_SORT_VEHICLES | FS_ENVF_BALLISTIC))
    return fail(nullptr, FS_EINVAL, "fs_create: unknown env_flags bit");
  int ndev = 0;
  cudaError_t st = cudaGetDeviceCount(&ndev);
  if (st != cudaSuccess || ndev == 0)
    return fail(nullptr, FS_ECUDA,
                std::string("fs_create: no CUDA device (") + cudaGetErrorString(st) +
                    "); libflowstep has no CPU path");
  if (cfg->device < 0 || cfg->device >= ndev)
    return fail(nullptr, FS_EINVAL, "fs_create: bad device ordinal");
  st = cudaSetDevice(cfg->device);
  if (st != cudaSuccess) return fail(nullptr, FS_ECUDA, cudaGetErrorString(st));

  fs_engine* e = new fs_engine();
  e->cfg = *cfg;
  e->f64 = cfg->dtype == FS_F64;
  e->obs_dim = obs_dim_for(*cfg);
  if (const char* rs = getenv("FS_RING_SHAPE")) e->ring_shape = atoi(rs);
  if (const char* nh = getenv("FS_NO_HOT")) e->no_hot = atoi(nh) != 0;
  if (const char* mg = getenv("FS_ML_GENERIC")) e->ml_force_generic = atoi(mg) != 0;
  if (const char* zc = getenv("FS_CL_MODE")) e->cl_mode = atoi(zc) & 3;
  st = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking);
  if (st != cudaSuccess) {
    g_create_error = cudaGetErrorString(st);
    delete e;
    return FS_ECUDA;
  }
  const size_t E = cfg->n_envs, V = cfg->n_veh, ts = tsize(e);
  int rc = FS_OK;
  auto A = [&](auto** p, size_t bytes) { if (rc == FS_OK) rc = dev_alloc(e, p, bytes); };
  A(&e->pos, V * E * ts); A(&e->vel, V * E * ts); A(&e->circ, E * ts);
  A(&e->lane, V * E); A(&e->active, V * E);
  A(&e->done, E); A(&e->crash, E); A(&e->mask, E);
  A(&e->time_counter, E * 4); A(&e->episode, E * 4);
  A(&e->obs, E * e->obs_dim * ts); A(&e->reward, E * ts);
  A(&e->stats, FS_STAT_STRIPES * 8 * sizeof(double)); A(&e->lengths_dev, E * sizeof(double));
  e->env_major = e->multilane() || cfg->topology == FS_TOPO_HIGHWAY || e->merge() || e->fig8();
  if (e->env_major) A(&e->last_lc, V * E * 4);
  if (e->sort_on()) { A(&e->abs_pos, V * E * ts); A(&e->prev_pos, V * E * ts); A(&e->keys_valid, E); }
  if (e->merge()) {
    A(&e->counters, (3 + 2 * FS_MAX_INFLOWS) * E * 4); A(&e->vtype, V * E);
    A(&e->veh_id, V * E * 4); A(&e->seq, V * E * 4);
    A(&e->agent_id, E * (size_t)cfg->max_agents * 4);
  }
  if (cfg->topology == FS_TOPO_HIGHWAY) {
    A(&e->counters, (3 + 2 * FS_MAX_INFLOWS) * E * 4); A(&e->vtype, V * E);
    if (cfg->inflow_vehs_per_hour > 0) {
      e->n_flows = 1;
      e->flows[0].vtype = cfg->inflow_type; e->flows[0].vehs_per_hour = cfg->inflow_vehs_per_hour;
      e->flows[0].depart_speed = cfg->inflow_depart_speed; e->flows[0].begin = cfg->inflow_begin;
    }
  }
  if (e->ma_highway()) {
    A(&e->veh_id, V * E * 4); A(&e->agent_id, E * (size_t)cfg->max_agents * 4);
    A(&e->agent_reward, E * (size_t)cfg->max_agents * ts);
  }
  if (cfg->record_views) {
    A(&e->headway, V * E * ts); A(&e->accel, V * E * ts);
    A(&e->leader, V * E * 4); A(&e->follower, V * E * 4);
  }
  if (rc != FS_OK) {
    g_create_error = e->err;
    fs_destroy(e);
    return rc;
  }
  cudaMemsetAsync(e->active, 1, V * E, e->stream);
  *out = e;
  return FS_OK;
}

extern "C" void fs_destroy(fs_engine* e) {
  if (!e) return;
  cudaSetDevice(e->cfg.device);
  if (e->stream) cudaStreamSynchronize(e->stream);
  if (e->down_stream) { cudaStreamSynchronize(e->down_stream); cudaStreamDestroy(e->down_stream); }
  if (e->up_stream) { cudaStreamSynchronize(e->up_stream); cudaStreamDestroy(e->up_stream); }
  for (auto& sl : e->slots) {
    if (sl.up) cudaEventDestroy(sl.up);
    if (sl.run) cudaEventDestroy(sl.run);
    if (sl.down) cudaEventDestroy(sl.down);
  }
  if (e->uf_exec) cudaGraphExecDestroy(e->uf_exec);
  if (e->cl_exec) cudaGraphExecDestroy(e->cl_exec);
  if (e->cl_host) cudaFreeHost(e->cl_host);
  for (void* p : e->allocs) if (p) cudaFree(p);
  if (e->stream && e->own_stream) cudaStreamDestroy(e->stream);
  delete e;
}

template <typename T>
static void fill_cfg(fs_engine* e, fs::Cfg<T>& c, const fs_vtype* types, int n_types,
                     const uint8_t* type_of) {
  const fs_config& g = e->cfg;
  memset(&c, 0, sizeof(c));
  c.n_envs = g.n_envs; c.n_veh = g.n_veh; c.n_lanes = g.n_lanes; c.env_kind = g.env_kind;
  c.n_types = n_types; c.warmup_steps = g.warmup_steps; c.horizon = g.horizon;
  c.sims_per_step = g.sims_per_step > 0 ? g.sims_per_step : 1;
  c.clip_actions = g.clip_actions; c.lane_change_duration = g.lane_change_duration;
  c.obs_dim = e->obs_dim; c.collision_min_gap = g.collision_min_gap;
  c.env_id_base = g.env_id_base; c.seed = g.seed;
  c.dt = (T)g.sim_step; c.inv_dt = (T)(1.0 / g.sim_step);
  c.junction_len = (T)g.junction_len; c.speed_limit = (T)g.speed_limit;
  c.inv_speed_limit = (T)(1.0 / g.speed_limit);
  c.x0 = (T)g.x0; c.bunching = (T)g.bunching; c.init_min_gap = (T)g.init_min_gap;
  c.target_velocity = (T)g.target_velocity; c.act_lo = (T)g.act_lo; c.act_hi = (T)g.act_hi;
  c.obs_max_length = (T)g.obs_max_length;
  c.length_lo = (T)g.length_lo; c.length_hi = (T)g.length_hi;
  c.lanes_distribution = g.lanes_distribution;
  {  // merge geometry, summed in the order oracle/merge.py MergeGeometry sums it
    const double LH = ((100.0 + 0.1) + g.merge_pre) + 22.5, LM = ((100.0 + 0.1) + g.merge_ramp) + 22.5;
    c.mg_LH = (T)LH; c.mg_LM = (T)LM; c.mg_post = (T)g.merge_post;
    c.mg_zone = (T)(g.merge_zone > 0 ? g.merge_zone : g.merge_ramp + 22.5);
    c.mg_pre = (T)g.merge_pre; c.mg_ramp = (T)g.merge_ramp;
  }
  if (g.topology == FS_TOPO_FIGURE8) {
    const fs::F8Geo q = fs::f8_geometry(g.fig8_radius);
    c.f8_L = (T)q.L; c.f8_cA = (T)q.cA; c.f8_cB = (T)q.cB; c.f8_J = (T)q.J;
    c.f8_zone = (T)(g.fig8_zone > 0 ? g.fig8_zone : g.fig8_radius);
    for (int k = 0; k < 12; ++k) c.f8_start[k] = (T)q.start[k];
  }
  c.evaluate = (g.env_flags & FS_ENVF_EVALUATE) ? 1 : 0;
  c.ballistic = (g.env_flags & FS_ENVF_BALLISTIC) ? 1 : 0;
  for (int t = 0; t < n_types; ++t) {
    const fs_vtype& s = types[t];
    fs::TypeP<T>& d = c.types[t];
    d.ctrl = s.controller; d.fail_safe = s.fail_safe; d.lc = s.lc_controller;
    d.speed_mode = s.speed_mode & 7;
    for (int q = 0; q < FS_NPARAM; ++q) d.p[q] = (T)s.p[q];
    const bool sim_driven = s.controller == FS_CTRL_SIM || s.controller == FS_CTRL_RL;
    d.noise = sim_driven ? (T)0 : (T)s.noise; d.delay = (T)s.delay; d.max_accel = (T)s.max_accel;
    if (sim_driven) d.fail_safe = FS_FAILSAFE_NONE;  // get_accel -> None: no Flow noise / fail-safe
    d.max_decel = (T)fabs(s.max_decel); d.length = (T)s.length;
    d.sim_tau = (T)s.sim_tau; d.sim_min_gap = (T)s.sim_min_gap;
    const double sim_v0 = fmin(s.sim_max_speed, g.speed_limit);
    d.sim_v0 = (T)sim_v0;
    const double tsab = 2.0 * sqrt(s.p[2] * s.p[3]);
    const double stsab = 2.0 * sqrt(s.max_accel * fabs(s.max_decel));
    d.two_sqrt_ab = (T)tsab; d.sim_two_sqrt_ab = (T)stsab;
    d.inv_v0 = (T)(s.p[0] != 0 ? 1.0 / s.p[0] : 0.0); d.sim_inv_v0 = (T)(1.0 / sim_v0);
    d.inv_two_sqrt_ab = (T)(tsab > 0 ? 1.0 / tsab : 0.0);
    d.sim_inv_two_sqrt_ab = (T)(stsab > 0 ? 1.0 / stsab : 0.0);
    const bool own = s.controller == FS_CTRL_IDM;
    d.e_v0 = own ? d.p[0] : d.sim_v0; d.e_inv_v0 = own ? d.inv_v0 : d.sim_inv_v0;
    d.e_T = own ? d.p[1] : d.sim_tau; d.e_a = own ? d.p[2] : d.max_accel;
    d.e_tsab = own ? d.two_sqrt_ab : d.sim_two_sqrt_ab;
    d.e_inv_tsab = own ? d.inv_two_sqrt_ab : d.sim_inv_two_sqrt_ab;
    d.e_delta = own ? d.p[4] : (T)4; d.e_s0 = own ? d.p[5] : d.sim_min_gap;
    if (s.noise > 0 && !sim_driven) c.any_noise = 1;
    if (s.controller == FS_CTRL_BCM) c.any_follower |= 1;
    if (s.controller == FS_CTRL_NONLOCAL_FOLLOWER_STOPPER) c.any_follower |= 2;
  }
  int n_rl = 0, n_pi = 0;
  for (int i = 0; i < g.n_veh; ++i) {
    c.type_of[i] = type_of[i];
    c.rl_index[i] = -1;
    c.pi_index[i] = -1;
    if (types[type_of[i]].controller == FS_CTRL_PI_SATURATION) c.pi_index[i] = (int16_t)n_pi++;
    if (types[type_of[i]].controller == FS_CTRL_RL) c.rl_index[i] = (int16_t)n_rl++;
  }
  c.n_rl = n_rl;
  const bool lc_env = g.env_kind == FS_ENV_LANE_CHANGE_ACCEL ||
                      g.env_kind == FS_ENV_LANE_CHANGE_ACCEL_PO;
  c.act_dim = lc_env ? 2 * n_rl : n_rl;
  if (g.env_kind == FS_ENV_MA_WAVE_ATTENUATION_PO) { e->obs_dim = 3 * n_rl; c.obs_dim = e->obs_dim; }
  if (g.env_kind == FS_ENV_MA_ACCEL_PO) { e->obs_dim = 6 * n_rl; c.obs_dim = e->obs_dim; }
  // LaneChangeAccelPOEnv.get_state returns inside its loop over RL vehicles
  // (lane_change_accel.py:215-263): 4 * n_rl * lanes zeros-or-first-block + ONE ego speed
  // (with lc_corrected: every block and n_rl ego speeds, the declared observation_space)
  if (g.env_kind == FS_ENV_LANE_CHANGE_ACCEL_PO) {
    e->obs_dim = 4 * n_rl * g.n_lanes + (g.lc_corrected ? n_rl : 1); c.obs_dim = e->obs_dim;
  }
  if (g.env_kind == FS_ENV_MA_HIGHWAY_PO || g.env_kind == FS_ENV_MERGE_PO)
    c.act_dim = g.max_agents;  // one action row per agent / controlled-vehicle row
  e->n_rl = n_rl;
  e->act_dim = c.act_dim;
}

extern "C" int fs_set_types(fs_engine* e, const fs_vtype* types, int n_types,
                            const uint8_t* type_of_vehicle) {
  if (!e) return FS_EINVAL;
  if (!types || !type_of_vehicle || n_types < 1 || n_types > FS_MAX_TYPES)
    return fail(e, FS_EINVAL, "fs_set_types: need 1..16 types and a type_of_vehicle table");
  cudaSetDevice(e->cfg.device);
  drop_graphs(e);
  e->generic = e->fast_ok = e->hot_ok = e->ml_uniform = false;   // a second call starts clean
  e->fast_variant = 0;
  int n_pi = 0, n_state = 0;
  for (int i = 0; i < e->cfg.n_veh; ++i) {
    if (type_of_vehicle[i] >= n_types)
      return fail(e, FS_EINVAL, "fs_set_types: type_of_vehicle entry out of range");
    const int c = types[type_of_vehicle[i]].controller;
    if (c < 0 || c > FS_CTRL_RL) return fail(e, FS_EINVAL, "fs_set_types: unknown controller");
    if (c != FS_CTRL_IDM && c != FS_CTRL_SIM && c != FS_CTRL_RL) e->generic = true;
    if (c == FS_CTRL_PI_SATURATION) ++n_pi;
    if (c == FS_CTRL_PI_SATURATION || c == FS_CTRL_LAC) ++n_state;
  }
  {
    // hot-kernel preconditions: V = 22, IDM-law vehicles only, ONE human type
    // and at most one RL type, equal lengths, delta = 4
    int ta = -1, tb = -1;
    bool ok = e->cfg.n_veh == 22 && !e->generic && e->cfg.topology == FS_TOPO_RING &&
              e->cfg.n_lanes == 1 && !(e->cfg.env_flags & (FS_ENVF_EVALUATE | FS_ENVF_BALLISTIC)) &&
              e->cfg.env_kind != FS_ENV_MA_ACCEL_PO;   // follower view: generic kernel only
    for (int i = 0; ok && i < e->cfg.n_veh; ++i) {
      const int t = type_of_vehicle[i];
      const fs_vtype& s = types[t];
      // several type rows are fine as long as they are byte-identical (e.g. the
      // "human_0" / "human_1" groups of examples/exp_configs/rl/multiagent/multiagent_ring.py)
      if (s.controller == FS_CTRL_RL) {
        if (tb >= 0 && tb != t && memcmp(&types[tb], &s, sizeof(s)) != 0) ok = false;
        tb = t;
      } else {
        if (ta >= 0 && ta != t && memcmp(&types[ta], &s, sizeof(s)) != 0) ok = false;
        ta = t;
      }
      if (s.length != types[type_of_vehicle[0]].length) ok = false;
      if (s.controller == FS_CTRL_IDM && s.p[4] != 4.0) ok = false;
    }
    if (ta < 0) ok = false;
    // the IDM fast paths keep ONE collision threshold (the human type's minGap): with
    // collision_min_gap set, types whose minGap differ need the generic kernel's per-leader test
    if (ok && e->cfg.collision_min_gap && tb >= 0 && types[ta].sim_min_gap != types[tb].sim_min_gap)
      ok = false;
    e->fast_ok = ok;
    {
      // hot-kernel pattern: exactly one RL vehicle and it is the last one
      int n_rl = 0;
      for (int i = 0; i < e->cfg.n_veh; ++i) n_rl += types[type_of_vehicle[i]].controller == FS_CTRL_RL;
      e->hot_ok = ok && n_rl == 1 &&
                  types[type_of_vehicle[e->cfg.n_veh - 1]].controller == FS_CTRL_RL &&
                  e->cfg.env_kind == FS_ENV_WAVE_ATTENUATION_PO && !e->cfg.record_views &&
                  // the hot kernel omits the went-through-its-leader test of the other ring
                  // kernels: with dt <= 0.1 that needs a closing speed above 50 m/s
                  e->cfg.sim_step <= 0.1;
    }
    if (ok) {
      const fs_vtype& s = types[ta];
      const bool sim = s.controller == FS_CTRL_SIM;
      e->fast_variant = e->cfg.record_views ? 3
                        : (!sim && s.fail_safe != FS_FAILSAFE_NONE) ? 2
                        : (!sim && s.noise > 0) ? 1 : 0;
      if (e->fast_variant > 1) e->hot_ok = false;   // (before the speed mode picks variant 4)
      // a SUMO speed mode on some type: the fast kernel applies it in its fail-safe variants,
      // the hot kernel in its kSafe instantiations
      const bool any_mode = (!sim && (s.speed_mode & 7) != 0) ||
                            (tb >= 0 && (types[tb].speed_mode & 7) != 0);
      if (any_mode && e->fast_variant < 2) e->fast_variant = 4;   // speed mode, no fail-safe code
      e->hot_noise = !sim && s.noise > 0;
    }
  }
  {
    // multi-lane kernel, kGeneric = false: every non-RL vehicle of one (byte-identical) IDM-law
    // type, every RL vehicle of one type, equal lengths, one collision threshold
    int ta = -1, tb = -1;
    bool ok = !e->generic;
    for (int i = 0; ok && i < e->cfg.n_veh; ++i) {
      const int t = type_of_vehicle[i];
      const fs_vtype& s = types[t];
      int& slot = s.controller == FS_CTRL_RL ? tb : ta;
      if (slot >= 0 && slot != t && memcmp(&types[slot], &s, sizeof(s)) != 0) ok = false;
      slot = t;
      if (s.length != types[type_of_vehicle[0]].length) ok = false;
    }
    if (ok && ta >= 0 && tb >= 0 && e->cfg.collision_min_gap &&
        types[ta].sim_min_gap != types[tb].sim_min_gap) ok = false;
    e->ml_uniform = ok;
  }
  if (e->fig8())
    for (int t = 0; t < n_types; ++t) {
      const int c = types[t].controller;
      if (c == FS_CTRL_PI_SATURATION || c == FS_CTRL_LAC || c == FS_CTRL_BCM)
        return fail(e, FS_EINVAL, "fs_set_types: PISaturation / LAC / BCM vehicles are not supported on "
                                  "the figure-eight network (controller state / follower input)");
    }
  if (e->merge()) {
    if (e->cfg.n_initial < 0 || e->cfg.n_initial > e->cfg.n_veh)
      return fail(e, FS_EINVAL, "fs_set_types: n_initial must be within the slot capacity");
    for (int t = 0; t < n_types; ++t) {
      const int c = types[t].controller;
      if (c == FS_CTRL_PI_SATURATION || c == FS_CTRL_LAC || c == FS_CTRL_BCM)
        return fail(e, FS_EINVAL, "fs_set_types: PISaturation / LAC / BCM vehicles are not supported on "
                                  "the merge network (per-slot controller state / follower input)");
    }
  }
  if (e->cfg.topology == FS_TOPO_HIGHWAY) {
    // inflow vehicles may be of a type no initial vehicle has: look at every type
    for (int t = 0; t < n_types; ++t) {
      const int c = types[t].controller;
      if (c != FS_CTRL_IDM && c != FS_CTRL_SIM && c != FS_CTRL_RL) e->generic = true;
    }
    if (e->cfg.n_initial < 0 || e->cfg.n_initial > e->cfg.n_veh)
      return fail(e, FS_EINVAL, "fs_set_types: n_initial must be within the slot capacity");
    if (e->cfg.inflow_type < 0 || e->cfg.inflow_type >= n_types)
      return fail(e, FS_EINVAL, "fs_set_types: inflow_type out of range");
    for (int t = 0; t < n_types; ++t) {
      const int c = types[t].controller;
      if (c == FS_CTRL_PI_SATURATION || c == FS_CTRL_LAC || (c == FS_CTRL_RL && !e->ma_highway()))
        return fail(e, FS_EINVAL, "fs_set_types: PISaturation / LAC vehicles are not supported on the "
                                  "highway topology (slots are re-packed); RL vehicles need "
                                  "FS_ENV_MA_HIGHWAY_PO");
    }
  }
  if (e->f64) fill_cfg<double>(e, e->cd, types, n_types, type_of_vehicle);
  else fill_cfg<float>(e, e->cf, types, n_types, type_of_vehicle);
  if (e->env_needs_rl() && e->n_rl < 1)
    return fail(e, FS_EINVAL, "fs_set_types: this env_kind needs at least one RL vehicle");
  const size_t E = e->cfg.n_envs, V = e->cfg.n_veh, ts = tsize(e);
  int rc = FS_OK;
  if (n_state > 0 && !e->st0) rc = dev_alloc(e, &e->st0, V * E * ts);
  if (rc == FS_OK && n_pi > 0 && !e->hist) {
    e->n_pi = n_pi;
    e->hist_len = (int)(38.0 / e->cfg.sim_step);  // int(38 / sim_step), velocity_controllers.py:221
    rc = dev_alloc(e, &e->hist, (size_t)e->hist_len * n_pi * E * ts);
    if (rc == FS_OK) rc = dev_alloc(e, &e->hist_sum, (size_t)n_pi * E * sizeof(double));
    if (rc == FS_OK) rc = dev_alloc(e, &e->hist_cnt, (size_t)n_pi * E * sizeof(int));
  }
  if (rc == FS_OK && e->sort_on() && e->act_dim > 0 && !e->act_sorted)
    rc = dev_alloc(e, &e->act_sorted, (size_t)e->act_dim * E * ts);
  if (rc != FS_OK) return rc;
  e->types_set = true;
  return FS_OK;
}

// ---------------------------------------------------------------- launches
template <typename T>
static fs::RingBufs<T> ring_bufs(fs_engine* e) {
  fs::RingBufs<T> b;
  b.pos = (T*)e->pos; b.vel = (T*)e->vel; b.circ = (T*)e->circ; b.st0 = (T*)e->st0;
  b.hist = (T*)e->hist; b.hist_sum = e->hist_sum; b.hist_cnt = e->hist_cnt;
  b.hist_len = e->hist_len; b.n_pi = e->n_pi;
  b.time_counter = e->time_counter; b.episode = e->episode;
  b.done = e->done; b.crash = e->crash;
  b.obs = (T*)e->obs; b.reward = (T*)e->reward;
  b.v_accel = (T*)e->accel; b.v_headway = (T*)e->headway;
  b.v_leader = e->leader; b.v_follower = e->follower;
  b.stats = e->stats;
  return b;
}

template <typename T>
static const fs::Cfg<T>& cfg_of(fs_engine* e);
template <> const fs::Cfg<float>& cfg_of<float>(fs_engine* e) { return e->cf; }
template <> const fs::Cfg<double>& cfg_of<double>(fs_engine* e) { return e->cd; }

template <typename T>
static int launch_ring1(fs_engine* e, const fs::StepArgs<T>& a) {
  FS_CUDA(e, fs_launch_ring1<T>(e->cfg.n_veh, e->cfg.n_envs, e->stream, cfg_of<T>(e), ring_bufs<T>(e), a));
  e->launches++;
  return FS_OK;
}

template <typename T>
static int launch_ml(fs_engine* e, const fs::StepArgs<T>& a) {
  fs::MlBufs<T> b;
  b.pos = (T*)e->pos; b.vel = (T*)e->vel; b.lane = e->lane; b.circ = (T*)e->circ;
  b.st0 = (T*)e->st0; b.hist = (T*)e->hist; b.hist_sum = e->hist_sum; b.hist_cnt = e->hist_cnt;
  b.hist_len = e->hist_len; b.n_pi = e->n_pi; b.last_lc = e->last_lc;
  b.time_counter = e->time_counter; b.episode = e->episode; b.done = e->done; b.crash = e->crash;
  b.obs = (T*)e->obs; b.reward = (T*)e->reward;
  b.v_accel = (T*)e->accel; b.v_headway = (T*)e->headway;
  b.v_leader = e->leader; b.v_follower = e->follower; b.stats = e->stats;
  b.lc_bug_compat = e->cfg.lc_corrected ? 0 : 1;
  b.fp = fs::make_mlfast(cfg_of<T>(e));
  const int E = e->cfg.n_envs, V = e->cfg.n_veh;
  // one warp per env; one warp per block balances small batches over the 148 SMs (+4 % on c3
  // at 2048 envs), four per block once occupancy matters
  const int wpb = E <= 148 * 32 ? 1 : 4;
  const int grid = (E + wpb - 1) / wpb;
  const size_t smem = (size_t)wpb * fs::ml_smem_per_env<T>(V);
  // kGeneric = false: ONE IDM-law human type + at most one RL type (fs_set_types: ml_uniform);
  // FS_ML_GENERIC=1 keeps the table-driven instantiation for A/B runs
  const bool uni = e->ml_uniform && !e->ml_force_generic;
  FS_CUDA(e, fs_launch_ml<T>(V, uni, grid, wpb * 32, smem, e->stream, cfg_of<T>(e), b, a));
  e->launches++;
  FS_CUDA(e, cudaGetLastError());
  return FS_OK;
}

template <typename T>
static fs::HwBufs<T> hw_bufs(fs_engine* e) {
  fs::HwBufs<T> b;
  const size_t E = e->cfg.n_envs;
  b.pos = (T*)e->pos; b.vel = (T*)e->vel; b.vtype = e->vtype; b.active = e->active;
  b.length = (T*)e->circ;
  b.n_active = e->counters; b.departed = e->counters + E; b.arrived = e->counters + 2 * E;
  b.spawned = e->counters + 3 * E;
  b.pending = e->counters + (3 + FS_MAX_INFLOWS) * E;
  b.time_counter = e->time_counter; b.episode = e->episode; b.done = e->done; b.crash = e->crash;
  b.obs = (T*)e->obs; b.reward = (T*)e->reward;
  b.v_accel = (T*)e->accel; b.v_headway = (T*)e->headway;
  b.v_leader = e->leader; b.v_follower = e->follower; b.stats = e->stats;
  b.n_flows = e->n_flows;
  for (int f = 0; f < FS_MAX_INFLOWS; ++f) {
    const fs_inflow& q = e->flows[f];
    const bool on = f < e->n_flows && (q.vehs_per_hour > 0 || q.probability > 0);
    b.period[f] = on && q.vehs_per_hour > 0 ? 3600.0 / q.vehs_per_hour : 0.0;
    b.prob[f] = on && !(q.vehs_per_hour > 0) ? q.probability : 0.0;
    b.begin[f] = on ? q.begin : 0.0;
    b.end[f] = on ? q.end : 0.0;
    b.depart_speed[f] = on ? (T)q.depart_speed : T(0);
    b.inflow_type[f] = on ? q.vtype : 0;
  }
  b.vid = e->veh_id; b.agent_reward = (T*)e->agent_reward; b.agent_id = e->agent_id;
  b.max_agents = e->ma_highway() ? e->cfg.max_agents : 0;
  return b;
}

template <typename T>
static fs::MgBufs<T> mg_bufs(fs_engine* e) {
  fs::MgBufs<T> b;
  const size_t E = e->cfg.n_envs;
  b.pos = (T*)e->pos; b.vel = (T*)e->vel; b.arm = e->lane; b.vtype = e->vtype; b.active = e->active;
  b.vid = e->veh_id; b.seq = e->seq;
  b.n_active = e->counters; b.departed = e->counters + E; b.arrived = e->counters + 2 * E;
  b.spawned = e->counters + 3 * E;
  int* extra = e->counters + (3 + FS_MAX_INFLOWS) * E;     // the highway's `pending` rows
  b.n_ctrl = extra; b.hi_seq = extra + E; b.next_seq = extra + 2 * E;
  b.ctrl_id = e->agent_id;
  b.time_counter = e->time_counter; b.episode = e->episode; b.done = e->done; b.crash = e->crash;
  b.obs = (T*)e->obs; b.reward = (T*)e->reward;
  b.v_accel = (T*)e->accel; b.v_headway = (T*)e->headway;
  b.v_leader = e->leader; b.v_follower = e->follower; b.stats = e->stats;
  b.n_flows = e->n_flows; b.max_agents = e->cfg.max_agents;
  for (int f = 0; f < FS_MAX_INFLOWS; ++f) {
    const fs_inflow& q = e->flows[f];
    const bool on = f < e->n_flows && q.vehs_per_hour > 0;
    b.period[f] = on ? 3600.0 / q.vehs_per_hour : 0.0;
    b.begin[f] = on ? q.begin : 0.0;
    b.depart_speed[f] = on ? (T)q.depart_speed : T(0);
    b.inflow_type[f] = on ? q.vtype : 0;
    b.inflow_arm[f] = on ? q.arm : 0;
  }
  return b;
}

template <typename T>
static int launch_mg(fs_engine* e, const fs::StepArgs<T>& a) {
  const int E = e->cfg.n_envs, CAP = e->cfg.n_veh;
  const int wpb = E <= 148 * 8 ? 1 : 4;
  const int grid = (E + wpb - 1) / wpb;
  const size_t smem = (size_t)wpb * fs::mg_smem_per_warp<T>(CAP);
  FS_CUDA(e, fs_launch_mg<T>(CAP, grid, wpb * 32, smem, e->stream, cfg_of<T>(e), mg_bufs<T>(e), a));
  e->launches++;
  return FS_OK;
}

template <typename T>
static fs::F8Bufs<T> f8_bufs(fs_engine* e) {
  fs::F8Bufs<T> b;
  b.pos = (T*)e->pos; b.vel = (T*)e->vel;
  b.time_counter = e->time_counter; b.episode = e->episode; b.done = e->done; b.crash = e->crash;
  b.obs = (T*)e->obs; b.reward = (T*)e->reward;
  b.v_accel = (T*)e->accel; b.v_headway = (T*)e->headway;
  b.v_leader = e->leader; b.v_follower = e->follower; b.stats = e->stats;
  return b;
}

template <typename T>
static int launch_f8(fs_engine* e, const fs::StepArgs<T>& a) {
  const int E = e->cfg.n_envs, V = e->cfg.n_veh;
  const int wpb = E <= 148 * 8 ? 1 : 4;
  const int grid = (E + wpb - 1) / wpb;
  const size_t smem = (size_t)wpb * fs::f8_smem_per_warp<T>(V);
  FS_CUDA(e, fs_launch_f8<T>(V, grid, wpb * 32, smem, e->stream, cfg_of<T>(e), f8_bufs<T>(e), a));
  e->launches++;
  return FS_OK;
}

template <typename T>
static int launch_hw(fs_engine* e, const fs::StepArgs<T>& a) {
  const int E = e->cfg.n_envs, CAP = e->cfg.n_veh;
  const int wpb = E <= 148 * 8 ? 1 : 4;
  const int grid = (E + wpb - 1) / wpb;
  const size_t smem = (size_t)wpb * fs::hw_smem_per_warp<T>(CAP);
  const fs::HwBufs<T> b = hw_bufs<T>(e);
  bool any_prob = false;
  for (int f = 0; f < e->n_flows; ++f) any_prob |= b.prob[f] > 0.0 || b.end[f] > 0.0;
  FS_CUDA(e, fs_launch_hw<T>(CAP, e->ma_highway(), e->n_flows <= 1 && !any_prob, e->generic, grid,
                             wpb * 32, smem, e->stream, cfg_of<T>(e), b, a));
  e->launches++;
  FS_CUDA(e, cudaGetLastError());
  return FS_OK;
}

static cudaError_t launch_fast(int shape, int variant, int E, cudaStream_t st,
                               const fs::FastP<float>& c, const fs::RingBufs<float>& b,
                               const fs::StepArgs<float>& a) {
  return fs_launch_fast22_f32(shape, variant, E, st, c, b, a);
}
static cudaError_t launch_fast(int shape, int variant, int E, cudaStream_t st,
                               const fs::FastP<double>& c, const fs::RingBufs<double>& b,
                               const fs::StepArgs<double>& a) {
  return fs_launch_fast22_f64(shape, variant, E, st, c, b, a);
}

static cudaError_t launch_hot(int shape, bool noise, int E, cudaStream_t st,
                              const fs::FastP<float>& c, const fs::RingBufs<float>& b,
                              const fs::StepArgs<float>& a) {
  return fs_launch_hot22_f32(shape, noise, E, st, c, b, a);
}
static cudaError_t launch_hot(int shape, bool noise, int E, cudaStream_t st,
                              const fs::FastP<double>& c, const fs::RingBufs<double>& b,
                              const fs::StepArgs<double>& a) {
  return fs_launch_hot22_f64(shape, noise, E, st, c, b, a);
}

template <typename T>
static int run_steps(fs_engine* e, const fs::StepArgs<T>& a) {
  if (e->cfg.topology == FS_TOPO_HIGHWAY) return launch_hw<T>(e, a);
  if (e->merge()) return launch_mg<T>(e, a);
  if (e->fig8()) return launch_f8<T>(e, a);
  const int V = e->cfg.n_veh, E = e->cfg.n_envs;
  if (e->multilane()) return launch_ml<T>(e, a);
  if (e->fast_ok) {
    // IDM-law hot kernel; threads per env chosen from the batch size
    int shape = e->ring_shape;
    if (shape != 1 && shape != 2 && shape != 4 && shape != 8 && shape != 16 && shape != 32)
    {
      // about 1024 warps (1.7 per SM sub-partition) whatever the batch: the fewest
      // threads per env (most vehicles per thread, least per-env overhead) that
      // still spread the batch over every scheduler. Measured on c2 / c5.
      shape = 1;
      while (shape < 32 && (long long)E * shape < 32768) shape <<= 1;
    }
    // fp32 keeps one thread per env for the largest batches although two threads x 11 vehicles
    // fuse 3.6 % faster on c5 (4.09 against 3.95 x 10^11): the one-step-per-launch variant of the
    // same engine is HBM-bound and loses 11 % with it (0.43 -> 0.38 of the measured HBM peak),
    // and the noise stream's counter layout is per shape, so both variants must share one
    if (sizeof(T) == 8 && shape == 1 && e->ring_shape == 0) shape = 2;  // f64 x 22 slots spills
    if (e->hot_ok && !e->no_hot) {   // c2 / c5 configuration: fully specialised kernel
      FS_CUDA(e, launch_hot(shape, e->hot_noise, E, e->stream,
                            fs::make_fastp(cfg_of<T>(e)), ring_bufs<T>(e), a));
      e->launches++;
      return FS_OK;
    }
    FS_CUDA(e, launch_fast(shape, e->fast_variant, E, e->stream, fs::make_fastp(cfg_of<T>(e)),
                           ring_bufs<T>(e), a));
    e->launches++;
    return FS_OK;
  }
  return launch_ring1<T>(e, a);
}

template <typename T>
static fs::SortBufs<T> sort_bufs(fs_engine* e, const uint8_t* mask) {
  fs::SortBufs<T> b;
  b.pos = (const T*)e->pos;
  b.vs = e->env_major ? 1 : e->cfg.n_envs; b.es = e->env_major ? e->cfg.n_veh : 1;
  b.abs_pos = (T*)e->abs_pos; b.prev_pos = (T*)e->prev_pos; b.keys_valid = e->keys_valid;
  b.circ = (const T*)e->circ; b.done = e->done; b.mask = mask;
  return b;
}

// sort_vehicles: every step is  sort_pre -> step kernel (k = 1) -> sort_obs  (fs_sort.cuh)
template <typename T>
static int do_step_sorted(fs_engine* e, const void* actions, int k, void* obs_seq,
                          void* reward_seq, uint8_t* done_seq, const uint8_t* mask, int stats,
                          int ignore_done) {
  const int E = e->cfg.n_envs, V = e->cfg.n_veh, wpb = 4, grid = (E + wpb - 1) / wpb;
  const fs::SortBufs<T> sb = sort_bufs<T>(e, mask);
  const bool accel_env = e->cfg.env_kind == FS_ENV_ACCEL;
  const int rows = e->cfg.env_kind == FS_ENV_LANE_CHANGE_ACCEL ? 3 : 2;
  const bool permute_obs = e->cfg.env_kind != FS_ENV_LANE_CHANGE_ACCEL_PO;
  auto post = [&](T* obs_step) {
    if (!permute_obs) return;
    if (V <= 32) fs::sort_obs_kernel<T, 1><<<grid, wpb * 32, 0, e->stream>>>(cfg_of<T>(e), sb, (T*)e->obs, obs_step, rows);
    else if (V <= 96) fs::sort_obs_kernel<T, 3><<<grid, wpb * 32, 0, e->stream>>>(cfg_of<T>(e), sb, (T*)e->obs, obs_step, rows);
    else fs::sort_obs_kernel<T, 8><<<grid, wpb * 32, 0, e->stream>>>(cfg_of<T>(e), sb, (T*)e->obs, obs_step, rows);
    e->launches++;
  };
  fs::StepArgs<T> a;
  a.mask = mask; a.accumulate_stats = stats; a.ignore_done = ignore_done;
  a.write_views = e->cfg.record_views != 0;
  a.pdl = 0;
  if (k == 0) {  // observation of the current state (reset without warm-up)
    a.actions = nullptr; a.obs_seq = nullptr; a.reward_seq = nullptr; a.done_seq = nullptr;
    a.k_steps = 0;
    int rc = run_steps<T>(e, a);
    if (rc != FS_OK) return rc;
    post(nullptr);
    FS_CUDA(e, cudaGetLastError());
    return FS_OK;
  }
  const size_t act_row = (size_t)e->act_dim * E, obs_row = (size_t)E * e->obs_dim;
  for (int s = 0; s < k; ++s) {
    const T* act = actions ? (const T*)actions + s * act_row : nullptr;
    const bool remap = act != nullptr && e->n_rl > 1;
    if (remap || accel_env) {
      fs::sort_pre_kernel<T><<<grid, wpb * 32, 0, e->stream>>>(
          cfg_of<T>(e), sb, remap ? act : nullptr, (T*)e->act_sorted, accel_env ? 1 : 0, ignore_done);
      e->launches++;
    }
    a.actions = remap ? (const T*)e->act_sorted : act;
    a.obs_seq = obs_seq ? (T*)obs_seq + s * obs_row : nullptr;
    a.reward_seq = reward_seq ? (T*)reward_seq + (size_t)s * E : nullptr;
    a.done_seq = done_seq ? done_seq + (size_t)s * E : nullptr;
    a.k_steps = 1;
    int rc = run_steps<T>(e, a);
    if (rc != FS_OK) return rc;
    post(a.obs_seq);
  }
  FS_CUDA(e, cudaGetLastError());
  return FS_OK;
}

template <typename T>
static int do_step1(fs_engine* e, const void* actions, int k, void* obs_seq, void* reward_seq,
                    uint8_t* done_seq, const uint8_t* mask, int stats, int ignore_done);

// Env.step runs sims_per_step simulation steps on the same actions and observes once
// (flow/envs/base.py:325-383). Every shipped ring / highway config uses 1; other values take
// one launch per simulation step, outputs from the last one (a collision freezes the env, so
// the remaining sub-steps of that env are no-ops, like the reference's `break`).
template <typename T>
static int do_step(fs_engine* e, const void* actions, int k, void* obs_seq, void* reward_seq,
                   uint8_t* done_seq, const uint8_t* mask, int stats, int ignore_done) {
  const int S = e->cfg.sims_per_step;
  if (S <= 1 || k == 0 || e->merge() || e->fig8())   // these kernels loop over the sub-steps themselves
    return do_step1<T>(e, actions, k, obs_seq, reward_seq, done_seq, mask, stats, ignore_done);
  const size_t E = e->cfg.n_envs, act_row = (size_t)e->act_dim * E, obs_row = E * e->obs_dim;
  for (int s = 0; s < k; ++s) {
    const T* act = actions ? (const T*)actions + s * act_row : nullptr;
    for (int q = 0; q < S; ++q) {
      const bool last = q == S - 1;
      int rc = do_step1<T>(e, act, 1, last && obs_seq ? (T*)obs_seq + s * obs_row : nullptr,
                           last && reward_seq ? (T*)reward_seq + s * E : nullptr,
                           last && done_seq ? done_seq + s * E : nullptr, mask,
                           stats ? (last ? 1 : 2) : 0, ignore_done);   // 2: see substep_stats
      if (rc != FS_OK) return rc;
    }
  }
  return FS_OK;
}

template <typename T>
static int do_step1(fs_engine* e, const void* actions, int k, void* obs_seq, void* reward_seq,
                    uint8_t* done_seq, const uint8_t* mask, int stats, int ignore_done) {
  if (e->sort_on())
    return do_step_sorted<T>(e, actions, k, obs_seq, reward_seq, done_seq, mask, stats, ignore_done);
  fs::StepArgs<T> a;
  a.actions = (const T*)actions; a.obs_seq = (T*)obs_seq; a.reward_seq = (T*)reward_seq;
  a.done_seq = done_seq; a.mask = mask; a.k_steps = k; a.accumulate_stats = stats;
  a.ignore_done = ignore_done;
  a.write_views = e->cfg.record_views != 0;
  a.pdl = e->pdl_hint ? 1 : 0;
  return run_steps<T>(e, a);
}

extern "C" int fs_step(fs_engine* e, const void* actions, int k, void* obs_seq,
                       void* reward_seq, uint8_t* done_seq) {
  if (!e) return FS_EINVAL;
  if (!e->types_set) return fail(e, FS_ESTATE, "fs_step: call fs_set_types first");
  if (k < 1) return fail(e, FS_EINVAL, "fs_step: k must be >= 1");
  cudaSetDevice(e->cfg.device);
  return e->f64 ? do_step<double>(e, actions, k, obs_seq, reward_seq, done_seq, nullptr, 1, 0)
                : do_step<float>(e, actions, k, obs_seq, reward_seq, done_seq, nullptr, 1, 0);
}

// k steps as k SINGLE-STEP launches (state round-trips through memory every step: the
// HBM-bound variant of SURVEY 8d). The launch sequence is issued from C and, when the engine's
// stream can be captured, recorded into a CUDA graph that later calls with the same buffers
// replay with one cudaGraphLaunch (the pointers are baked into the nodes).
template <typename T>
static int unfused_loop(fs_engine* e, const void* actions, int k, void* obs_seq, void* reward_seq,
                        uint8_t* done_seq) {
  const size_t E = e->cfg.n_envs, act_row = (size_t)e->act_dim * E, obs_row = E * e->obs_dim;
  static const bool no_pdl = getenv("FS_NO_PDL") && atoi(getenv("FS_NO_PDL")) != 0;
  // measured (c5, 65 536 envs: 8.9 -> 8.0 us per step; c2, 4096 envs: 6.2 -> 6.9 us): only
  // launches with a long per-thread run gain from starting the next grid early
  e->pdl_hint = !no_pdl && E >= 32768;
  int rc = FS_OK;
  for (int s = 0; s < k && rc == FS_OK; ++s)
    rc = do_step<T>(e, actions ? (const T*)actions + s * act_row : nullptr, 1,
                    obs_seq ? (T*)obs_seq + s * obs_row : nullptr,
                    reward_seq ? (T*)reward_seq + s * E : nullptr,
                    done_seq ? done_seq + s * E : nullptr, nullptr, 1, 0);
  e->pdl_hint = false;
  return rc;
}

extern "C" int fs_step_unfused(fs_engine* e, const void* actions, int k, void* obs_seq,
                               void* reward_seq, uint8_t* done_seq) {
  if (!e) return FS_EINVAL;
  if (!e->types_set) return fail(e, FS_ESTATE, "fs_step_unfused: call fs_set_types first");
  if (k < 1) return fail(e, FS_EINVAL, "fs_step_unfused: k must be >= 1");
  cudaSetDevice(e->cfg.device);
  auto loop = [&]() {
    return e->f64 ? unfused_loop<double>(e, actions, k, obs_seq, reward_seq, done_seq)
                  : unfused_loop<float>(e, actions, k, obs_seq, reward_seq, done_seq);
  };
  const bool same = e->uf_exec && e->uf_k == k && e->uf_key[0] == actions &&
                    e->uf_key[1] == obs_seq && e->uf_key[2] == reward_seq && e->uf_key[3] == done_seq;
  if (same) {
    FS_CUDA(e, cudaGraphLaunch(e->uf_exec, e->stream));
    e->launches += e->uf_launches;
    return FS_OK;
  }
  if (e->uf_exec) { cudaGraphExecDestroy(e->uf_exec); e->uf_exec = nullptr; }
  static const bool no_graph = getenv("FS_NO_GRAPH") && atoi(getenv("FS_NO_GRAPH")) != 0;
  // the legacy default stream cannot be captured; k == 1 gains nothing from a graph
  if (no_graph || k == 1 || e->stream == nullptr || e->stream == cudaStreamLegacy) return loop();
  if (cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
    cudaGetLastError();
    return loop();
  }
  const int64_t l0 = e->launches;
  const int rc = loop();
  cudaGraph_t g = nullptr;
  cudaError_t st = cudaStreamEndCapture(e->stream, &g);
  if (rc != FS_OK) { if (g) cudaGraphDestroy(g); return rc; }
  if (st != cudaSuccess || !g) {
    cudaGetLastError();
    e->launches = l0;
    return loop();   // capture refused (e.g. another capture in flight): plain launches
  }
  e->uf_launches = e->launches - l0;
  st = cudaGraphInstantiate(&e->uf_exec, g, 0);
  cudaGraphDestroy(g);
  if (st != cudaSuccess) {
    e->uf_exec = nullptr;
    e->launches = l0;
    cudaGetLastError();
    return loop();
  }
  e->uf_k = k; e->uf_key[0] = actions; e->uf_key[1] = obs_seq; e->uf_key[2] = reward_seq;
  e->uf_key[3] = done_seq;
  FS_CUDA(e, cudaGraphLaunch(e->uf_exec, e->stream));
  return FS_OK;
}

template <typename T>
static int do_reset(fs_engine* e, const uint8_t* mask_dev, const double* lengths_dev, int stages) {
  fs::ResetBufs<T> b;
  b.pos = (T*)e->pos; b.vel = (T*)e->vel; b.circ = (T*)e->circ; b.st0 = (T*)e->st0;
  b.hist_sum = e->hist_sum; b.hist_cnt = e->hist_cnt; b.n_pi = e->n_pi;
  b.lane = e->lane; b.done = e->done; b.crash = e->crash;
  b.time_counter = e->time_counter; b.episode = e->episode;
  b.mask = mask_dev; b.lengths = lengths_dev;
  b.last_lc = e->last_lc;
  b.vs = e->env_major ? 1 : e->cfg.n_envs; b.es = e->env_major ? e->cfg.n_veh : 1;
  const int E = e->cfg.n_envs, block = 128;
  if (stages & FS_RESET_PLACE) {
    if (e->merge())
      FS_CUDA(e, fs_launch_mg_reset<T>(E, e->stream, cfg_of<T>(e), mg_bufs<T>(e), mask_dev,
                                       e->cfg.n_initial, (T*)e->circ));
    else if (e->fig8())
      FS_CUDA(e, fs_launch_f8_reset<T>(E, e->stream, cfg_of<T>(e), f8_bufs<T>(e),
                                       fs::f8_geometry(e->cfg.fig8_radius), mask_dev, (T*)e->circ));
    else if (e->cfg.topology == FS_TOPO_HIGHWAY)
      fs::highway_reset_kernel<T><<<(E + block - 1) / block, block, 0, e->stream>>>(
          cfg_of<T>(e), hw_bufs<T>(e), mask_dev, e->cfg.n_initial);
    else
      fs::ring_reset_kernel<T><<<(E + block - 1) / block, block, 0, e->stream>>>(cfg_of<T>(e), b);
    e->launches++;
    FS_CUDA(e, cudaGetLastError());
  }
  if (!(stages & FS_RESET_WARMUP)) return FS_OK;
  // warm-up: Env.reset steps with rl_actions=None (flow/envs/base.py:554-555)
  int rc = do_step<T>(e, nullptr, e->cfg.warmup_steps, nullptr, nullptr, nullptr, mask_dev, 0, 1);
  if (rc == FS_OK && e->sort_on()) {  // AccelEnv.reset tail (accel.py:178-181), after the observation
    fs::sort_setkeys_kernel<T><<<(E + block - 1) / block, block, 0, e->stream>>>(
        cfg_of<T>(e), sort_bufs<T>(e, mask_dev));
    e->launches++;
    FS_CUDA(e, cudaGetLastError());
  }
  return rc;
}

static int check_room(fs_engine* e, double L) {
  // _get_start_pos_util raises FatalFlowError when vehicles do not fit
  const fs_config& g = e->cfg;
  if (g.topology == FS_TOPO_MERGE) return FS_OK;   // initial vehicles are placed by the caller
  if (g.topology == FS_TOPO_FIGURE8) {
    const fs::F8Geo q = fs::f8_geometry(g.fig8_radius);
    if (q.non_internal - fmax(0.0, g.bunching) - g.n_veh * (fmax(0.0, g.init_min_gap) + 5.0) < 0)
      return fail(e, FS_EINVAL, "There is not enough space to place all vehicles in the network.");
    return FS_OK;
  }
  const double min_gap = fmax(0.0, g.init_min_gap), bunching = fmax(0.0, g.bunching);
  int ld = g.lanes_distribution;
  if (ld > g.n_lanes) ld = g.n_lanes;
  if (ld < 1) ld = 1;
  const bool hw = g.topology == FS_TOPO_HIGHWAY;
  const double edge_len = hw ? L : L / 4.0;
  const double dist = edge_len > min_gap + 5.0 ? (hw ? 1 : 4) * edge_len * ld : 0.0;
  if (dist - ld * bunching - (hw ? g.n_initial : g.n_veh) * (min_gap + 5.0) < 0)
    return fail(e, FS_EINVAL, "There is not enough space to place all vehicles in the network.");
  return FS_OK;
}

extern "C" int fs_set_inflows(fs_engine* e, const fs_inflow* flows, int n_flows) {
  if (!e) return FS_EINVAL;
  if (e->cfg.topology != FS_TOPO_HIGHWAY && !e->merge())
    return fail(e, FS_EINVAL, "fs_set_inflows: inflows exist on the highway and merge topologies only");
  if (!e->types_set) return fail(e, FS_ESTATE, "fs_set_inflows: call fs_set_types first");
  if (n_flows < 0 || n_flows > FS_MAX_INFLOWS || (n_flows > 0 && !flows))
    return fail(e, FS_EINVAL, "fs_set_inflows: 0..4 flows");
  const int n_types = e->f64 ? e->cd.n_types : e->cf.n_types;
  for (int f = 0; f < n_flows; ++f) {
    if (flows[f].vtype < 0 || flows[f].vtype >= n_types)
      return fail(e, FS_EINVAL, "fs_set_inflows: vtype out of range");
    const bool periodic = flows[f].vehs_per_hour > 0;
    const bool prob = flows[f].probability > 0 && flows[f].probability <= 1;
    if (periodic == prob || flows[f].probability < 0 || flows[f].vehs_per_hour < 0 ||
        !(flows[f].depart_speed >= 0) || !(flows[f].begin >= 0))
      return fail(e, FS_EINVAL, "fs_set_inflows: exactly one of vehs_per_hour > 0 / probability in "
                                "(0, 1]; depart_speed >= 0, begin >= 0");
    const fs::TypeP<double>* td = e->f64 ? &e->cd.types[flows[f].vtype] : nullptr;
    const int ctrl = td ? td->ctrl : e->cf.types[flows[f].vtype].ctrl;
    if (ctrl == FS_CTRL_RL && !e->ma_highway() && !e->merge())
      return fail(e, FS_EINVAL, "fs_set_inflows: an RL inflow needs FS_ENV_MA_HIGHWAY_PO / FS_ENV_MERGE_PO");
    if (e->merge() ? (flows[f].arm != 0 && flows[f].arm != 1 || !periodic) : flows[f].arm != 0)
      return fail(e, FS_EINVAL, "fs_set_inflows: arm is 0 (1 = the ramp, merge network only); merge "
                                "flows are periodic (vehs_per_hour)");
  }
  drop_graphs(e);
  e->n_flows = n_flows;
  for (int f = 0; f < FS_MAX_INFLOWS; ++f) e->flows[f] = f < n_flows ? flows[f] : fs_inflow{};
  return FS_OK;
}

extern "C" int fs_agents(fs_engine* e, fs_agent_views* out) {
  if (!e || !out) return FS_EINVAL;
  if (!e->ma_highway() && !e->merge())
    return fail(e, FS_EINVAL, "fs_agents: FS_ENV_MA_HIGHWAY_PO / FS_ENV_MERGE_PO only");
  out->agent_obs = e->obs; out->agent_reward = e->agent_reward; out->agent_id = e->agent_id;
  out->veh_id = e->veh_id; out->veh_type = e->vtype;
  out->max_agents = e->cfg.max_agents; out->obs_per_agent = 5;
  return FS_OK;
}

extern "C" int fs_reset(fs_engine* e, const uint8_t* env_mask, const double* lengths) {
  return fs_reset_stage(e, env_mask, lengths, FS_RESET_PLACE | FS_RESET_WARMUP);
}

extern "C" int fs_reset_stage(fs_engine* e, const uint8_t* env_mask, const double* lengths,
                              int stages) {
  if (!e) return FS_EINVAL;
  if (!e->types_set) return fail(e, FS_ESTATE, "fs_reset: call fs_set_types first");
  if (stages < 1 || stages > (FS_RESET_PLACE | FS_RESET_WARMUP))
    return fail(e, FS_EINVAL, "fs_reset_stage: stages must be FS_RESET_PLACE, FS_RESET_WARMUP or both");
  cudaSetDevice(e->cfg.device);
  const size_t E = e->cfg.n_envs;
  int rc = check_room(e, e->cfg.length_lo);
  if (rc != FS_OK) return rc;
  const uint8_t* mask_dev = nullptr;
  const double* len_dev = nullptr;
  if (env_mask) {
    FS_CUDA(e, cudaMemcpyAsync(e->mask, env_mask, E, cudaMemcpyHostToDevice, e->stream));
    mask_dev = e->mask;
  }
  if (lengths) {
    for (size_t i = 0; i < E; ++i)
      if ((!env_mask || env_mask[i]) && lengths[i] > 0 && (rc = check_room(e, lengths[i])) != FS_OK)
        return rc;
    FS_CUDA(e, cudaMemcpyAsync(e->lengths_dev, lengths, E * sizeof(double),
                               cudaMemcpyHostToDevice, e->stream));
    len_dev = e->lengths_dev;
  }
  rc = e->f64 ? do_reset<double>(e, mask_dev, len_dev, stages)
              : do_reset<float>(e, mask_dev, len_dev, stages);
  if (rc != FS_OK) return rc;
  // host buffers may be reused by the caller right away
  if (env_mask || lengths) FS_CUDA(e, cudaStreamSynchronize(e->stream));
  return FS_OK;
}

static int grow(fs_engine* e, void** p, size_t* cap, size_t need) {
  if (need <= *cap) return FS_OK;
  if (*p) {
    cudaStreamSynchronize(e->stream);
    cudaFree(*p);
    for (auto& q : e->allocs) if (q == *p) q = nullptr;
  }
  void* n = nullptr;
  cudaError_t st = cudaMalloc(&n, need);
  if (st != cudaSuccess) { *p = nullptr; *cap = 0; return fail(e, FS_ENOMEM, cudaGetErrorString(st)); }
  e->allocs.push_back(n);
  *p = n; *cap = need;
  return FS_OK;
}

// ---- closed-loop host step -------------------------------------------------------------
static const int FS_CL_MAX_K = 8;   // calls of up to this many steps take the staged / graph path

static size_t align256(size_t n) { return (n + 255) & ~(size_t)255; }

static int cl_ensure(fs_engine* e, int k) {
  if (k <= e->cl_k) return FS_OK;
  const size_t E = e->cfg.n_envs, ts = tsize(e);
  size_t off[5];
  off[0] = 0;
  off[1] = align256((size_t)k * e->act_dim * E * ts);
  off[2] = off[1] + align256((size_t)k * E * e->obs_dim * ts);
  off[3] = off[2] + align256((size_t)k * E * ts);
  off[4] = off[3] + align256((size_t)k * E);
  FS_CUDA(e, cudaStreamSynchronize(e->stream));
  if (e->cl_exec) { cudaGraphExecDestroy(e->cl_exec); e->cl_exec = nullptr; }
  if (e->cl_host) { cudaFreeHost(e->cl_host); e->cl_host = nullptr; }
  if (e->cl_dev) {
    cudaFree(e->cl_dev);
    for (auto& q : e->allocs) if (q == e->cl_dev) q = nullptr;
    e->cl_dev = nullptr;
  }
  e->cl_k = 0;
  FS_CUDA(e, cudaHostAlloc((void**)&e->cl_host, off[4], cudaHostAllocMapped));
  e->cl_hdev = nullptr;
  if (cudaHostGetDevicePointer((void**)&e->cl_hdev, e->cl_host, 0) != cudaSuccess) {
    cudaGetLastError();
    e->cl_hdev = nullptr;     // no mapped access: the copy-engine path below still works
  }
  void* d = nullptr;
  cudaError_t st = cudaMalloc(&d, off[4]);
  if (st != cudaSuccess) return fail(e, FS_ENOMEM, cudaGetErrorString(st));
  e->allocs.push_back(d);
  e->cl_dev = (unsigned char*)d;
  memcpy(e->cl_off, off, sizeof(off));
  e->cl_k = k;
  return FS_OK;
}

extern "C" int fs_host_staging(fs_engine* e, int k, fs_host_views* out) {
  if (!e || !out) return FS_EINVAL;
  if (!e->types_set) return fail(e, FS_ESTATE, "fs_host_staging: call fs_set_types first");
  if (k < 1 || k > FS_CL_MAX_K) return fail(e, FS_EINVAL, "fs_host_staging: k must be in 1..8");
  cudaSetDevice(e->cfg.device);
  int rc = cl_ensure(e, k);
  if (rc != FS_OK) return rc;
  out->actions = e->cl_host + e->cl_off[0]; out->obs = e->cl_host + e->cl_off[1];
  out->reward = e->cl_host + e->cl_off[2]; out->done = e->cl_host + e->cl_off[3];
  out->k = e->cl_k; out->n_envs = e->cfg.n_envs; out->obs_dim = e->obs_dim; out->act_dim = e->act_dim;
  out->dtype = e->cfg.dtype;
  return FS_OK;
}

// enqueue [H2D(actions)] -> k steps -> [D2H(obs | reward | done)] on the engine's stream;
// zc bit 0: the kernel reads the actions from the mapped pinned block itself, bit 1: it writes
// obs / reward / done straight into it (no copy-engine operation for that direction)
static int cl_enqueue(fs_engine* e, bool with_act, int k, int zc) {
  const size_t E = e->cfg.n_envs, ts = tsize(e);
  unsigned char* in = (zc & 1) ? e->cl_hdev : e->cl_dev;
  unsigned char* out = (zc & 2) ? e->cl_hdev : e->cl_dev;
  if (with_act && !(zc & 1))
    FS_CUDA(e, cudaMemcpyAsync(e->cl_dev + e->cl_off[0], e->cl_host + e->cl_off[0],
                               (size_t)k * e->act_dim * E * ts, cudaMemcpyHostToDevice, e->stream));
  int rc = fs_step(e, with_act ? in + e->cl_off[0] : nullptr, k, out + e->cl_off[1],
                   out + e->cl_off[2], out + e->cl_off[3]);
  if (rc != FS_OK) return rc;
  if (!(zc & 2)) {
    // ONE copy brings obs, reward and done back (the three regions are contiguous)
    const size_t nb_out = e->cl_off[3] + (size_t)k * E - e->cl_off[1];
    FS_CUDA(e, cudaMemcpyAsync(e->cl_host + e->cl_off[1], e->cl_dev + e->cl_off[1], nb_out,
                               cudaMemcpyDeviceToHost, e->stream));
  }
  return FS_OK;
}

static int step_host_closed_loop(fs_engine* e, const void* actions, int k, void* obs_seq,
                                 void* reward_seq, uint8_t* done_seq) {
  int rc = cl_ensure(e, k);
  if (rc != FS_OK) return rc;
  const size_t E = e->cfg.n_envs, ts = tsize(e);
  const size_t nb_act = (size_t)k * e->act_dim * E * ts, nb_obs = (size_t)k * E * e->obs_dim * ts,
               nb_rew = (size_t)k * E * ts, nb_done = (size_t)k * E;
  unsigned char* h = e->cl_host;
  const bool with_act = actions && e->act_dim > 0;
  // the staging block is handed out by fs_host_staging: a caller that works in it directly
  // pays no host copy; any other host buffer (pageable or pinned) is copied through it
  if (with_act && actions != h + e->cl_off[0]) memcpy(h + e->cl_off[0], actions, nb_act);
  // cl_mode (FS_CL_MODE at fs_create, default 0 = upload node -> kernel -> download node, one
  // graph launch): bit 0 = the kernel reads the actions from the mapped pinned block, bit 1 = it
  // writes obs / reward / done straight into it. Measured per call on c2 (4096 envs, 16 KB up,
  // 70 KB down; this box: launch + sync of an EMPTY kernel 12.4 us, fs_step(k=1) + sync on
  // device buffers 14.3 us): mode 0 23.8 us, 1 27.9, 2 25.5, 3 23.7 - SM-issued PCIe traffic is
  // no cheaper than the two copy-engine operations (~4.7 us each), so the default keeps them.
  int zc = e->cl_hdev ? e->cl_mode : 0;
  if (nb_obs + nb_rew + nb_done > (512u << 10)) zc &= ~2;
  if (nb_act > (512u << 10)) zc &= ~1;
  if (zc == 3 || (zc == 2 && !with_act)) {   // no copy at all: plain launch + sync
    rc = cl_enqueue(e, with_act, k, 3);
    if (rc != FS_OK) return rc;
    FS_CUDA(e, cudaStreamSynchronize(e->stream));
    if (obs_seq && obs_seq != h + e->cl_off[1]) memcpy(obs_seq, h + e->cl_off[1], nb_obs);
    if (reward_seq && reward_seq != h + e->cl_off[2]) memcpy(reward_seq, h + e->cl_off[2], nb_rew);
    if (done_seq && done_seq != h + e->cl_off[3]) memcpy(done_seq, h + e->cl_off[3], nb_done);
    return FS_OK;
  }
  static const bool no_graph = getenv("FS_NO_GRAPH") && atoi(getenv("FS_NO_GRAPH")) != 0;
  const bool capturable = !no_graph && e->stream != nullptr && e->stream != cudaStreamLegacy;
  if (e->cl_exec && (e->cl_exec_k != k || e->cl_exec_act != with_act)) {
    cudaGraphExecDestroy(e->cl_exec); e->cl_exec = nullptr;
  }
  if (e->cl_exec) {
    FS_CUDA(e, cudaGraphLaunch(e->cl_exec, e->stream));
    e->launches += e->cl_launches;
  } else {
    bool launched = false;
    if (capturable && cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal) == cudaSuccess) {
      const int64_t l0 = e->launches;
      rc = cl_enqueue(e, with_act, k, zc);
      cudaGraph_t g = nullptr;
      cudaError_t st = cudaStreamEndCapture(e->stream, &g);
      if (rc != FS_OK) { if (g) cudaGraphDestroy(g); return rc; }
      if (st == cudaSuccess && g && cudaGraphInstantiate(&e->cl_exec, g, 0) == cudaSuccess) {
        e->cl_launches = e->launches - l0;
        e->cl_exec_k = k; e->cl_exec_act = with_act;
        FS_CUDA(e, cudaGraphLaunch(e->cl_exec, e->stream));
        launched = true;
      } else {
        e->cl_exec = nullptr;
        e->launches = l0;
      }
      if (g) cudaGraphDestroy(g);
      cudaGetLastError();
    } else {
      cudaGetLastError();
    }
    if (!launched && (rc = cl_enqueue(e, with_act, k, zc)) != FS_OK) return rc;
  }
  FS_CUDA(e, cudaStreamSynchronize(e->stream));
  if (obs_seq && obs_seq != h + e->cl_off[1]) memcpy(obs_seq, h + e->cl_off[1], nb_obs);
  if (reward_seq && reward_seq != h + e->cl_off[2]) memcpy(reward_seq, h + e->cl_off[2], nb_rew);
  if (done_seq && done_seq != h + e->cl_off[3]) memcpy(done_seq, h + e->cl_off[3], nb_done);
  return FS_OK;
}

extern "C" int fs_step_host(fs_engine* e, const void* actions, int k, void* obs_seq,
                            void* reward_seq, uint8_t* done_seq) {
  if (!e) return FS_EINVAL;
  if (!e->types_set) return fail(e, FS_ESTATE, "fs_step_host: call fs_set_types first");
  if (k < 1) return fail(e, FS_EINVAL, "fs_step_host: k must be >= 1");
  cudaSetDevice(e->cfg.device);
  if (k <= FS_CL_MAX_K) return step_host_closed_loop(e, actions, k, obs_seq, reward_seq, done_seq);
  const size_t E = e->cfg.n_envs, ts = tsize(e);
  const size_t nb_act = (size_t)k * e->act_dim * E * ts, nb_obs = (size_t)k * E * e->obs_dim * ts,
               nb_rew = (size_t)k * E * ts, nb_done = (size_t)k * E;
  int rc;
  void* d_act = nullptr;
  if (actions && e->act_dim > 0) {
    if ((rc = grow(e, &e->h_act, &e->h_act_cap, nb_act)) != FS_OK) return rc;
    FS_CUDA(e, cudaMemcpyAsync(e->h_act, actions, nb_act, cudaMemcpyHostToDevice, e->stream));
    d_act = e->h_act;
  }
  if (obs_seq && (rc = grow(e, &e->h_obs, &e->h_obs_cap, nb_obs)) != FS_OK) return rc;
  if (reward_seq && (rc = grow(e, &e->h_rew, &e->h_rew_cap, nb_rew)) != FS_OK) return rc;
  if (done_seq && (rc = grow(e, (void**)&e->h_done, &e->h_done_cap, nb_done)) != FS_OK) return rc;
  rc = fs_step(e, d_act, k, obs_seq ? e->h_obs : nullptr, reward_seq ? e->h_rew : nullptr,
               done_seq ? e->h_done : nullptr);
  if (rc != FS_OK) return rc;
  if (obs_seq)
    FS_CUDA(e, cudaMemcpyAsync(obs_seq, e->h_obs, nb_obs, cudaMemcpyDeviceToHost, e->stream));
  if (reward_seq)
    FS_CUDA(e, cudaMemcpyAsync(reward_seq, e->h_rew, nb_rew, cudaMemcpyDeviceToHost, e->stream));
  if (done_seq)
    FS_CUDA(e, cudaMemcpyAsync(done_seq, e->h_done, nb_done, cudaMemcpyDeviceToHost, e->stream));
  FS_CUDA(e, cudaStreamSynchronize(e->stream));
  return FS_OK;
}

// ---- fs_step_host_noalloc -----------------------------------------------------------------
// No counterpart in the reference: added for callers
// that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
// that forbid it). Same contract as fs_step_host, but the device staging the step kernels read
// the actions from and write every step's obs / reward / done into does NOT come from
// cudaMalloc (fs_step_host grows its staging with cudaMalloc / cudaFree whenever a call is
// larger than the last): it is this file-scope device array of FS_NOALLOC_SCRATCH_BYTES.
// Every GPU that loads the library holds these bytes until the process exits.
//
// In HOST code the name `fs_noalloc_scratch` is only a stand-in for the symbol, not a device
// address: the address is looked up with cudaGetSymbolAddress once per device, with that
// device current (scratch_base below), and that pointer is what the kernels get.
//
// All calls on one GPU share the array (also calls on different engines): it keeps the last
// call's contents, so the output regions a call uses are cleared first, and calls that overlap
// on the GPU (two host threads, two engines on different streams) race - serialise them.
static __device__ __align__(256) unsigned char fs_noalloc_scratch[FS_NOALLOC_SCRATCH_BYTES];

// never launched: names the array in device code so that it is kept in the device image even
// though only host code (cudaGetSymbolAddress) refers to it
__global__ void fs_noalloc_keepalive(unsigned char** out) { *out = fs_noalloc_scratch; }

static const int FS_NOALLOC_MAX_DEVICES = 64;
static unsigned char* g_noalloc_base[FS_NOALLOC_MAX_DEVICES] = {};

// device address of the array on the engine's device (which the caller made current)
static int scratch_base(fs_engine* e, unsigned char** out) {
  const int dev = e->cfg.device;
  if (dev < 0 || dev >= FS_NOALLOC_MAX_DEVICES)
    return fail(e, FS_EINVAL, "fs_step_host_noalloc: device ordinal beyond 63");
  if (!g_noalloc_base[dev]) {
    void* ptr = nullptr;
    cudaError_t st = cudaGetSymbolAddress(&ptr, fs_noalloc_scratch);
    if (st != cudaSuccess || !ptr) {
      e->err = std::string("fs_step_host_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(st);
      return FS_ECUDA;
    }
    g_noalloc_base[dev] = (unsigned char*)ptr;
  }
  *out = g_noalloc_base[dev];
  return FS_OK;
}

extern "C" int fs_step_host_noalloc(fs_engine* e, const void* actions, int k, void* obs_seq,
                                    void* reward_seq, uint8_t* done_seq) {
  if (!e) return FS_EINVAL;
  if (!e->types_set) return fail(e, FS_ESTATE, "fs_step_host_noalloc: call fs_set_types first");
  if (k < 0) return fail(e, FS_EINVAL, "fs_step_host_noalloc: k must be >= 0");
  cudaSetDevice(e->cfg.device);
  unsigned char* base = nullptr;
  int rc = scratch_base(e, &base);
  if (rc != FS_OK || k == 0) return rc;     // k == 0: set-up only (resolve the address, e.g. before a capture)
  const size_t E = e->cfg.n_envs, ts = tsize(e);
  const bool with_act = actions && e->act_dim > 0;
  const size_t nb_act = with_act ? (size_t)k * e->act_dim * E * ts : 0,
               nb_obs = obs_seq ? (size_t)k * E * e->obs_dim * ts : 0,
               nb_rew = reward_seq ? (size_t)k * E * ts : 0, nb_done = done_seq ? (size_t)k * E : 0;
  const size_t o_obs = align256(nb_act), o_rew = o_obs + align256(nb_obs),
               o_done = o_rew + align256(nb_rew), total = o_done + align256(nb_done);
  if (total > (size_t)FS_NOALLOC_SCRATCH_BYTES) {
    char msg[256];
    snprintf(msg, sizeof(msg), "fs_step_host_noalloc: this call needs %zu bytes of device staging "
             "(k = %d steps x %zu envs), the static scratch holds FS_NOALLOC_SCRATCH_BYTES = %d; "
             "use fewer steps per call or fs_step_host", total, k, E, FS_NOALLOC_SCRATCH_BYTES);
    return fail(e, FS_EINVAL, msg);
  }
  if (with_act)
    FS_CUDA(e, cudaMemcpyAsync(base, actions, nb_act, cudaMemcpyHostToDevice, e->stream));
  // the array keeps the previous call's results: clear the regions this call hands back
  if (total > o_obs) FS_CUDA(e, cudaMemsetAsync(base + o_obs, 0, total - o_obs, e->stream));
  rc = fs_step(e, with_act ? base : nullptr, k, obs_seq ? base + o_obs : nullptr,
               reward_seq ? base + o_rew : nullptr, done_seq ? base + o_done : nullptr);
  if (rc != FS_OK) return rc;
  if (obs_seq)
    FS_CUDA(e, cudaMemcpyAsync(obs_seq, base + o_obs, nb_obs, cudaMemcpyDeviceToHost, e->stream));
  if (reward_seq)
    FS_CUDA(e, cudaMemcpyAsync(reward_seq, base + o_rew, nb_rew, cudaMemcpyDeviceToHost, e->stream));
  if (done_seq)
    FS_CUDA(e, cudaMemcpyAsync(done_seq, base + o_done, nb_done, cudaMemcpyDeviceToHost, e->stream));
  FS_CUDA(e, cudaStreamSynchronize(e->stream));
  return FS_OK;
}

extern "C" int fs_step_host_async(fs_engine* e, const void* actions, int k, void* obs_seq,
                                  void* reward_seq, uint8_t* done_seq) {
  if (!e) return FS_EINVAL;
  if (!e->types_set) return fail(e, FS_ESTATE, "fs_step_host_async: call fs_set_types first");
  if (k < 1) return fail(e, FS_EINVAL, "fs_step_host_async: k must be >= 1");
  cudaSetDevice(e->cfg.device);
  if (!e->up_stream) {
    FS_CUDA(e, cudaStreamCreateWithFlags(&e->up_stream, cudaStreamNonBlocking));
    FS_CUDA(e, cudaStreamCreateWithFlags(&e->down_stream, cudaStreamNonBlocking));
    for (auto& sl : e->slots) {
      FS_CUDA(e, cudaEventCreateWithFlags(&sl.up, cudaEventDisableTiming));
      FS_CUDA(e, cudaEventCreateWithFlags(&sl.run, cudaEventDisableTiming));
      FS_CUDA(e, cudaEventCreateWithFlags(&sl.down, cudaEventDisableTiming));
    }
  }
  fs_engine::Slot& sl = e->slots[e->async_calls % 3];
  const size_t E = e->cfg.n_envs, ts = tsize(e);
  const size_t nb_act = (size_t)k * e->act_dim * E * ts, nb_obs = (size_t)k * E * e->obs_dim * ts,
               nb_rew = (size_t)k * E * ts, nb_done = (size_t)k * E;
  int rc;
  const bool want_act = actions && e->act_dim > 0;
  if ((want_act && nb_act > sl.act_cap) || (obs_seq && nb_obs > sl.obs_cap) ||
      (reward_seq && nb_rew > sl.rew_cap) || (done_seq && nb_done > sl.done_cap)) {
    // (re)size ALL slots at once (cudaMalloc synchronises the device: do it on the
    // first call of a size, not once per slot in the middle of a pipelined run)
    FS_CUDA(e, cudaDeviceSynchronize());
    for (auto& q : e->slots) {
      if (want_act && (rc = grow(e, &q.act, &q.act_cap, nb_act)) != FS_OK) return rc;
      if (obs_seq && (rc = grow(e, &q.obs, &q.obs_cap, nb_obs)) != FS_OK) return rc;
      if (reward_seq && (rc = grow(e, &q.rew, &q.rew_cap, nb_rew)) != FS_OK) return rc;
      if (done_seq && (rc = grow(e, &q.done, &q.done_cap, nb_done)) != FS_OK) return rc;
    }
  }
  // the slot's previous results must have left the device before it is reused
  if (sl.used) FS_CUDA(e, cudaStreamWaitEvent(e->up_stream, sl.down, 0));
  void* d_act = nullptr;
  if (want_act) {
    FS_CUDA(e, cudaMemcpyAsync(sl.act, actions, nb_act, cudaMemcpyHostToDevice, e->up_stream));
    d_act = sl.act;
  }
  FS_CUDA(e, cudaEventRecord(sl.up, e->up_stream));
  FS_CUDA(e, cudaStreamWaitEvent(e->stream, sl.up, 0));
  if (sl.used) FS_CUDA(e, cudaStreamWaitEvent(e->stream, sl.down, 0));  // outputs of the slot's last use
  rc = fs_step(e, d_act, k, obs_seq ? sl.obs : nullptr, reward_seq ? sl.rew : nullptr,
               done_seq ? (uint8_t*)sl.done : nullptr);
  if (rc != FS_OK) return rc;
  FS_CUDA(e, cudaEventRecord(sl.run, e->stream));
  FS_CUDA(e, cudaStreamWaitEvent(e->down_stream, sl.run, 0));
  if (obs_seq)
    FS_CUDA(e, cudaMemcpyAsync(obs_seq, sl.obs, nb_obs, cudaMemcpyDeviceToHost, e->down_stream));
  if (reward_seq)
    FS_CUDA(e, cudaMemcpyAsync(reward_seq, sl.rew, nb_rew, cudaMemcpyDeviceToHost, e->down_stream));
  if (done_seq)
    FS_CUDA(e, cudaMemcpyAsync(done_seq, sl.done, nb_done, cudaMemcpyDeviceToHost, e->down_stream));
  FS_CUDA(e, cudaEventRecord(sl.down, e->down_stream));
  sl.used = true;
  e->async_calls++;
  return FS_OK;
}

// ---------------------------------------------------------------- DLPack
// Minimal DLPack v0 ABI (dlpack.h: DLDevice, DLDataType, DLTensor, DLManagedTensor),
// declared here so the library has no build dependency. The managed tensor does
// NOT own the device memory (the engine does); its deleter only frees the
// descriptor block, in C, so consumers may drop tensors at any time — including
// interpreter shutdown — without calling back into Python.
namespace {
struct DLDev { int32_t device_type; int32_t device_id; };
struct DLType { uint8_t code; uint8_t bits; uint16_t lanes; };
struct DLTens { void* data; DLDev device; int32_t ndim; DLType dtype; int64_t* shape;
                int64_t* strides; uint64_t byte_offset; };
struct DLManaged { DLTens dl_tensor; void* manager_ctx; void (*deleter)(DLManaged*); };
void dl_free(DLManaged* m) { free(m); }
}  // namespace

extern "C" void* fs_dlpack_wrap(void* dev_ptr, int device, int type_code, int bits, int ndim,
                                const int64_t* shape, const int64_t* strides) {
  if (!dev_ptr || ndim < 1 || ndim > 8 || !shape) return nullptr;
  const size_t bytes = sizeof(DLManaged) + 2 * ndim * sizeof(int64_t);
  DLManaged* m = (DLManaged*)malloc(bytes);
  if (!m) return nullptr;
  int64_t* tail = (int64_t*)(m + 1);
  m->dl_tensor.data = dev_ptr;
  m->dl_tensor.device = DLDev{2 /* kDLCUDA */, device};
  m->dl_tensor.ndim = ndim;
  m->dl_tensor.dtype = DLType{(uint8_t)type_code, (uint8_t)bits, 1};
  m->dl_tensor.shape = tail;
  m->dl_tensor.strides = strides ? tail + ndim : nullptr;
  m->dl_tensor.byte_offset = 0;
  for (int i = 0; i < ndim; ++i) { tail[i] = shape[i]; if (strides) tail[ndim + i] = strides[i]; }
  m->manager_ctx = nullptr;
  m->deleter = dl_free;
  return m;
}

extern "C" int fs_buffers(fs_engine* e, fs_views* out) {
  if (!e || !out) return FS_EINVAL;
  memset(out, 0, sizeof(*out));
  out->pos = e->pos; out->vel = e->vel; out->lane = e->lane; out->active = e->active;
  out->headway = e->headway; out->leader = e->leader; out->follower = e->follower;
  out->accel = e->accel; out->length = e->circ; out->obs = e->obs; out->reward = e->reward;
  out->done = e->done; out->crash = e->crash; out->time_counter = e->time_counter;
  out->counters = e->counters;
  out->n_envs = e->cfg.n_envs; out->n_veh = e->cfg.n_veh; out->n_lanes = e->cfg.n_lanes;
  out->obs_dim = e->obs_dim; out->act_dim = e->act_dim; out->dtype = e->cfg.dtype;
  out->veh_stride = e->env_major ? 1 : e->cfg.n_envs;
  out->env_stride = e->env_major ? e->cfg.n_veh : 1;
  return FS_OK;
}

extern "C" int fs_set_state(fs_engine* e, const void* pos, const void* vel, const uint8_t* lane,
                            const double* lengths) {
  if (!e) return FS_EINVAL;
  if (!e->types_set) return fail(e, FS_ESTATE, "fs_set_state: call fs_set_types first");
  cudaSetDevice(e->cfg.device);
  const size_t E = e->cfg.n_envs, V = e->cfg.n_veh, ts = tsize(e);
  // host arrays are [V][E]; env-major engines store [E][V]
  std::vector<unsigned char> tp, tv, tl;
  auto transposed = [&](const void* src, size_t es, std::vector<unsigned char>& buf) -> const void* {
    if (!e->env_major) return src;
    buf.resize(V * E * es);
    const unsigned char* s8 = (const unsigned char*)src;
    for (size_t i = 0; i < V; ++i)
      for (size_t q = 0; q < E; ++q) memcpy(&buf[(q * V + i) * es], &s8[(i * E + q) * es], es);
    return buf.data();
  };
  if (pos) FS_CUDA(e, cudaMemcpyAsync(e->pos, transposed(pos, ts, tp), V * E * ts, cudaMemcpyHostToDevice, e->stream));
  if (vel) FS_CUDA(e, cudaMemcpyAsync(e->vel, transposed(vel, ts, tv), V * E * ts, cudaMemcpyHostToDevice, e->stream));
  if (lane) FS_CUDA(e, cudaMemcpyAsync(e->lane, transposed(lane, 1, tl), V * E, cudaMemcpyHostToDevice, e->stream));
  if (e->env_major) FS_CUDA(e, cudaStreamSynchronize(e->stream));
  if (lengths) {
    std::vector<double> c64(E);
    std::vector<float> c32(E);
    for (size_t i = 0; i < E; ++i) {
      c64[i] = lengths[i] + 4 * e->cfg.junction_len;
      c32[i] = (float)c64[i];
    }
    FS_CUDA(e, cudaMemcpyAsync(e->circ, e->f64 ? (void*)c64.data() : (void*)c32.data(), E * ts,
                               cudaMemcpyHostToDevice, e->stream));
    FS_CUDA(e, cudaStreamSynchronize(e->stream));
  }
  // k.update on the injected state (refreshes views + obs); k = 0 moves nothing
  int rc = e->f64 ? do_step<double>(e, nullptr, 0, nullptr, nullptr, nullptr, nullptr, 0, 1)
                  : do_step<float>(e, nullptr, 0, nullptr, nullptr, nullptr, nullptr, 0, 1);
  if (rc != FS_OK) return rc;
  FS_CUDA(e, cudaStreamSynchronize(e->stream));
  return FS_OK;
}

extern "C" int fs_stats(fs_engine* e, double* host_out, int n, int reset) {
  if (!e || !host_out || n < 1) return FS_EINVAL;
  cudaSetDevice(e->cfg.device);
  double tmp[FS_STAT_STRIPES * 8];   // striped accumulators, see fs_common.cuh
  FS_CUDA(e, cudaMemcpyAsync(tmp, e->stats, sizeof(tmp), cudaMemcpyDeviceToHost, e->stream));
  if (reset) FS_CUDA(e, cudaMemsetAsync(e->stats, 0, sizeof(tmp), e->stream));
  FS_CUDA(e, cudaStreamSynchronize(e->stream));
  for (int i = 0; i < n && i < 8; ++i) {
    double s = 0.0;
    for (int q = 0; q < FS_STAT_STRIPES; ++q) s += tmp[q * 8 + i];
    host_out[i] = s;
  }
  return FS_OK;
}

extern "C" int fs_sync(fs_engine* e) {
  if (!e) return FS_EINVAL;
  cudaSetDevice(e->cfg.device);
  if (e->up_stream) FS_CUDA(e, cudaStreamSynchronize(e->up_stream));
  FS_CUDA(e, cudaStreamSynchronize(e->stream));
  if (e->down_stream) FS_CUDA(e, cudaStreamSynchronize(e->down_stream));
  return FS_OK;
}

extern "C" void* fs_stream(fs_engine* e) { return e ? (void*)e->stream : nullptr; }
extern "C" int64_t fs_launch_count(const fs_engine* e) { return e ? e->launches : 0; }

extern "C" int fs_set_stream(fs_engine* e, void* cuda_stream) {
  if (!e) return FS_EINVAL;
  cudaSetDevice(e->cfg.device);
  FS_CUDA(e, cudaStreamSynchronize(e->stream));
  drop_graphs(e);
  if (e->stream && e->own_stream) cudaStreamDestroy(e->stream);
  e->stream = (cudaStream_t)cuda_stream;
  e->own_stream = false;
  return FS_OK;
}
