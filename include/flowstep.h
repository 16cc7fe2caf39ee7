/*
 * flowstep.h — C-ABI of libflowstep.so, the B200 batched traffic-microsimulation
 * step engine that stands in for Flow's per-vehicle Python/SUMO loop.
 *
 * Every entry point replaces a piece of the reference interface (paths are
 * relative to the reference checkout, parthjaggi/flow):
 *
 *   fs_create / fs_destroy   Kernel(simulator, sim_params) + start_simulation
 *                            flow/core/kernel/kernel.py:48-78,
 *                            flow/core/kernel/simulation/traci.py:70-174
 *   fs_set_types             VehicleParams.add(...) controller/param tables
 *                            flow/core/params.py:236-351,
 *                            flow/core/kernel/vehicle/traci.py:261-372
 *   fs_reset                 Env.reset + WaveAttenuationEnv.reset + warm-up
 *                            flow/envs/base.py:414-560,
 *                            flow/envs/ring/wave_attenuation.py:157-210,
 *                            flow/core/kernel/network/base.py:263-391
 *   fs_step / fs_step_unfused / fs_step_host (+ fs_host_staging) / fs_step_host_async
 *                            Env.step (controllers, fail-safes, apply_acceleration,
 *                            simulation_step, k.update, get_state, compute_reward)
 *                            flow/envs/base.py:294-412
 *   fs_buffers               KernelVehicle getters (get_speed, get_position,
 *                            get_lane, get_headway, get_leader, get_follower)
 *                            flow/core/kernel/vehicle/traci.py:541-602
 *   fs_set_state             TraCIVehicle.test_set_speed / set_headway hooks
 *                            flow/core/kernel/vehicle/traci.py:411-425,639-689
 *   fs_stats                 Experiment.run return/velocity bookkeeping
 *                            flow/core/experiment.py:123-176
 *   fs_set_inflows           NetParams(inflows=InFlows()) with several <flow> rows
 *                            flow/core/params.py:1080-1213,
 *                            flow/core/kernel/network/traci.py:801-833
 *   fs_agents                MultiEnv's dict-keyed observations / rewards and the
 *                            vehicle ids behind them (get_rl_ids, get_arrived_rl_ids)
 *                            flow/envs/multiagent/base.py:116-225,
 *                            flow/envs/multiagent/highway.py:97-196
 *
 * Conventions: every call returns 0 on success or a negative FS_E* code;
 * fs_last_error() gives the message. The engine owns all device memory;
 * pointers handed out by fs_buffers stay valid until fs_destroy. All work is
 * enqueued on the engine's stream; only fs_stats, fs_step_host and the *_sync
 * helpers block. One engine per device; a handle is not thread-safe.
 * No torch types cross this boundary.
 */
#ifndef FLOWSTEP_H_
#define FLOWSTEP_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FS_ABI_VERSION 5

/* error codes */
#define FS_OK 0
#define FS_EINVAL (-1)  /* bad argument / unsupported configuration */
#define FS_ECUDA (-2)   /* CUDA runtime error (message has the detail) */
#define FS_ENOMEM (-3)
#define FS_ESTATE (-4)  /* call order violated (e.g. step before set_types) */

/* arithmetic type of the state and of every buffer except done/lane/index */
#define FS_F32 0
#define FS_F64 1

/* topology */
#define FS_TOPO_RING 0
#define FS_TOPO_HIGHWAY 1
#define FS_TOPO_MERGE 2   /* MergeNetwork (flow/networks/merge.py), single-lane highway + single-lane
                           * on-ramp: edges inflow_highway (100 m) - left (pre_merge_length) -
                           * center (post_merge_length), ramp inflow_merge (100 m) - bottom
                           * (merge_length) joining at 'center'; junction lengths 0.1 / 22.5 m as
                           * specify_edge_starts implies. n_veh = slot capacity, vehicles enter
                           * through fs_set_inflows (fs_inflow.arm) and leave at the end of 'center';
                           * `lane` holds the ARM of a slot's vehicle (0 highway, 1 ramp), `pos` the
                           * distance travelled on that arm. SUMO's junction is RESTATED:
                           * both arms on one axis aligned at the merge point, cross-arm car following
                           * inside merge_zone, collision only on a shared road. */

#define FS_TOPO_FIGURE8 3 /* FigureEightNetwork (flow/networks/figure_eight.py), one lane: ONE closed loop
                           * of length 4 r + 3 pi r + 2 * 6.2 + 4 * 0.28 that passes the centre
                           * junction twice (northbound bottom -> top, westbound right -> left);
                           * `pos` is the loop coordinate (0 = start of ':bottom'). Runs FS_ENV_ACCEL
                           * (singleagent_figure_eight.py); fs_reset places with gen_even_start_pos
                           * over the loop's edges. Controllers give no command on internal edges
                           * (base_controller.py:96-99), the simulator model drives there. SUMO's
                           * junction is RESTATED: the crossing is a merge followed by a
                           * diverge - inside fig8_zone before the junction box a vehicle also follows
                           * the vehicles of the other pass that are closer to / inside the box;
                           * collision at the crossing = overlapping while both bodies are in the box. */

/* environment front: selects observation + reward */
#define FS_ENV_ACCEL 0             /* AccelEnv: obs 2V, desired_velocity            */
#define FS_ENV_WAVE_ATTENUATION 1  /* WaveAttenuationEnv: obs 2V, WA reward         */
#define FS_ENV_WAVE_ATTENUATION_PO 2 /* WaveAttenuationPOEnv: obs 3, WA reward      */
#define FS_ENV_LANE_CHANGE_ACCEL 3 /* LaneChangeAccelEnv: obs 3V, dv - lc penalty   */
#define FS_ENV_LANE_CHANGE_ACCEL_PO 4 /* LaneChangeAccelPOEnv (lane_change_accel.py:164-269), ring only:
                                      per-lane leader / follower of the RL vehicles
                                      (TraCIVehicle._multi_lane_headways, vehicle/traci.py:699-937).
                                      obs as the reference BEHAVES: block of the FIRST RL vehicle
                                      [headway_l | tailway_l (m, 1000 = empty lane) | v_lead_l/vmax |
                                      v_follow_l/vmax], zeros for the other RL vehicles, ONE raw ego
                                      speed -> 4*n_rl*lanes + 1. lc_corrected=1: every block,
                                      distances / ring length, n_rl ego speeds / vmax. */
#define FS_ENV_HIGHWAY 5           /* open highway + inflow: obs 2*slots, dv reward */
#define FS_ENV_MA_WAVE_ATTENUATION_PO 6 /* MultiAgentWaveAttenuationPOEnv
                                      (flow/envs/multiagent/ring/wave_attenuation.py:143-312):
                                      obs [n_rl][3] per env = (v/15, (v_lead-v)/15, headway/ring_length[1])
                                      of every RL vehicle in action-slot order, shared WA reward */

/* acceleration controllers (flow/controllers/*.py) */
#define FS_CTRL_SIM 0          /* SimCarFollowingController: simulator default model */
#define FS_CTRL_IDM 1          /* p: v0,T,a,b,delta,s0                               */
#define FS_CTRL_OVM 2          /* p: alpha,beta,h_st,h_go,v_max                      */
#define FS_CTRL_LINEAR_OVM 3   /* p: v_max,adaptation,h_st                           */
#define FS_CTRL_BCM 4          /* p: k_d,k_v,k_c,d_des,v_des                         */
#define FS_CTRL_CFM 5          /* p: k_d,k_v,k_c,d_des,v_des                         */
#define FS_CTRL_LAC 6          /* p: k_1,k_2,h,tau,a0          (stateful: a)         */
#define FS_CTRL_GIPPS 7        /* p: v0,acc,b,b_l,s0,tau                             */
#define FS_CTRL_FOLLOWER_STOPPER 8        /* p: v_des                                */
#define FS_CTRL_NONLOCAL_FOLLOWER_STOPPER 9
#define FS_CTRL_PI_SATURATION 10          /* stateful: v_history, v_cmd              */
#define FS_CTRL_RL 11          /* RLController: accel = clipped action, else SIM     */

#define FS_FAILSAFE_NONE 0
#define FS_FAILSAFE_INSTANTANEOUS 1
#define FS_FAILSAFE_SAFE_VELOCITY 2

/* lane-change controllers */
#define FS_LC_SIM 0     /* SimLaneChangeController: never changes (mode 512)   */
#define FS_LC_STATIC 1  /* StaticLaneChanger: direction 0                      */
#define FS_LC_RL 2      /* direction comes from the action vector              */

#define FS_ENV_MA_ACCEL_PO 7 /* MultiAgentAccelPOEnv (flow/envs/multiagent/ring/accel.py:87-226), single-lane
                                ring: obs [n_rl][6] per env = (x/L, v/vmax, (v_lead-v)/vmax,
                                (x_lead - x - len)/L  -- NOT wrapped, like the reference --,
                                (v-v_follow)/vmax, headway(follower)/L); shared desired_velocity reward */

#define FS_ENV_MERGE_PO 9 /* MergePOEnv (flow/envs/merge.py:28-231) on FS_TOPO_MERGE: max_agents = num_rl
                           * controlled vehicles (the reference's `rl_veh` list: FIFO from the queue of
                           * RL vehicles in the network, a row closes up when its vehicle left), obs
                           * 5 * num_rl, one action row per controlled vehicle, reward desired_velocity +
                           * 0.1 * time-headway penalty; fs_agents gives the ids behind the rows */
#define FS_ENV_MA_HIGHWAY_PO 8 /* MultiAgentHighwayPOEnv (flow/envs/multiagent/highway.py:97-196) on the
                                single-lane open highway: every RL vehicle IN the network is an agent.
                                obs [max_agents][5] per env, row a = a-th RL vehicle counted from the
                                downstream end: (v/vmax, (v_lead-v)/vmax, headway(LEADER)/L,
                                (v-v_follow)/vmax, headway(follower)/L), defaults vmax / L / 0 / L
                                when there is no leader / follower; action row a drives the same
                                vehicle, UNCLIPPED (MultiEnv.clip_actions is a no-op,
                                multiagent/base.py:372-395), NaN = no command (simulator model);
                                per-agent reward max(desired_velocity + 0.1 min(h/v - 1, 0), 0) */

/* fs_config.env_flags */
#define FS_ENVF_EVALUATE 1      /* EnvParams.evaluate: AccelEnv reward = mean speed
                                   (flow/envs/ring/accel.py:109-114)                    */
#define FS_ENVF_SORT_VEHICLES 2 /* additional_params['sort_vehicles']: observation rows and RL
                                   action slots in order of AccelEnv.absolute_position
                                   (flow/envs/ring/accel.py:100-169); ring topology only  */

#define FS_ENVF_BALLISTIC 4     /* SumoParams.use_ballistic (flow/core/kernel/simulation/traci.py:95-96):
                                   the restated move becomes x += (v_old + v_new)/2 dt           */

#define FS_MAX_TYPES 16
#define FS_NPARAM 8

typedef struct fs_engine fs_engine; /* opaque */

typedef struct fs_vtype {
  int32_t controller;    /* FS_CTRL_*                                           */
  int32_t fail_safe;     /* FS_FAILSAFE_*                                       */
  int32_t lc_controller; /* FS_LC_*                                             */
  int32_t speed_mode;    /* SumoCarFollowingParams.speed_mode (flow/core/params.py:12-18, default
                          * 25 there; 0 = 'aggressive' = Flow's command is taken as is). Applies to
                          * speeds Flow commands (controller or RL action), RESTATED SUMO: bit 0 caps
                          * the commanded speed by the simulator model's own next speed (IDM with
                          * the sim_* / max_accel / max_decel fields below), bit 1 caps the
                          * acceleration at max_accel, bit 2 floors it at -max_decel; bits 3-4
                          * (right of way, red lights) have nothing to act on here                 */
  double p[FS_NPARAM];   /* controller parameters, order as listed above        */
  double noise;          /* std of the Gaussian added to accel (0 = none)       */
  double delay;          /* used by the safe_velocity fail-safe                 */
  double max_accel;      /* car_following_params.controller_params['accel']     */
  double max_decel;      /* abs(controller_params['decel'])                     */
  double length;         /* vehicle length (SUMO default 5 m)                   */
  /* SumoCarFollowingParams used when the simulator drives the vehicle          */
  double sim_tau, sim_min_gap, sim_max_speed;
} fs_vtype;

typedef struct fs_config {
  int32_t abi_version;   /* FS_ABI_VERSION                                      */
  int32_t device;        /* CUDA device ordinal                                 */
  int32_t dtype;         /* FS_F32 / FS_F64                                     */
  int32_t topology;      /* FS_TOPO_*                                           */
  int32_t env_kind;      /* FS_ENV_*                                            */
  int32_t n_envs;        /* E on this device                                    */
  int32_t n_veh;         /* V: vehicles per env (ring) or slot capacity (hwy)   */
  int32_t n_lanes;
  int64_t env_id_base;   /* global index of env 0 (keeps RNG invariant to #GPUs)*/
  uint64_t seed;
  double sim_step;       /* dt                                                  */
  /* network */
  double length_lo, length_hi; /* ring: integer draw U{lo..hi} per reset (lo==hi fixed);
                                  highway: length = length_lo                    */
  double junction_len;   /* ring: 4 junctions of this length are part of the loop */
  double speed_limit;
  /* initial placement (InitialConfig) */
  double x0, bunching, init_min_gap;
  int32_t lanes_distribution; /* clipped to n_lanes                            */
  /* env params */
  int32_t warmup_steps, horizon, sims_per_step;
  int32_t clip_actions;  /* reward sees clipped actions iff 1                  */
  int32_t lane_change_duration;
  double target_velocity;
  double act_lo, act_hi; /* accel bounds of the action space (-|max_decel|, max_accel) */
  double obs_max_length; /* WaveAttenuationPO: ring_length[1] normaliser        */
  /* highway inflow */
  double inflow_vehs_per_hour, inflow_depart_speed;
  int32_t inflow_type;   /* vehicle type index of inflow vehicles               */
  int32_t collision_min_gap; /* 0: crash iff gap<0 ; 1: crash iff gap<minGap    */
  int32_t record_views;  /* 1: keep views.accel/headway/leader/follower current  */
  int32_t lc_corrected;  /* 0: reference behaviour, bugs included: get_last_lc returns the
                            headway (vehicle/traci.py:614), LaneChangeAccelPOEnv.get_state
                            returns inside its loop (lane_change_accel.py:263);
                            1: the step of the last change / the full observation */
  /* highway: vehicles present at reset (types = type_of_vehicle[0..n_initial)) */
  int32_t n_initial;
  int32_t env_flags;     /* FS_ENVF_* bits (0 = reference defaults)              */
  double inflow_begin;   /* s; InFlows.add(begin=1) */
  int32_t max_agents;    /* FS_ENV_MA_HIGHWAY_PO: observation / action rows per env (<= n_veh);
                          * FS_ENV_MERGE_PO: num_rl (<= 32) */
  int32_t reserved0;
  /* FS_TOPO_MERGE (net_params.additional_params of MergeNetwork); merge_zone = metres before the
   * merge point in which vehicles of the two arms follow one another (0: merge_length + 22.5) */
  double merge_pre, merge_ramp, merge_post, merge_zone;
  /* FS_TOPO_FIGURE8: radius_ring of the network; fig8_zone = metres before the junction box in
   * which a vehicle regards the other pass (0: the approach edge = fig8_radius; <= fig8_radius) */
  double fig8_radius, fig8_zone;
} fs_config;

/* one <flow> of the open highway (InFlows.add: veh_type, vehs_per_hour | period | probability,
 * depart_speed, begin, end). Periodic flow (vehs_per_hour > 0): vehicle k is due at
 * begin + k * 3600 / vehs_per_hour as long as that is before `end`. Probabilistic flow
 * (probability > 0, per second): in every simulation step inside [begin, end) one more vehicle
 * becomes due with probability `probability * sim_step` (counter-based Philox keyed by seed,
 * global env id, episode, flow and step, so the draw does not depend on the GPU count). Due
 * vehicles that find the entry blocked wait. Vehicle k of flow f gets the id
 * ((f + 1) << 24) | k   ("flow_<f>.<k>" in the reference). */
#define FS_MAX_INFLOWS 4
typedef struct fs_inflow {
  int32_t vtype;         /* vehicle type index                                  */
  int32_t arm;           /* FS_TOPO_MERGE: 0 = edge 'inflow_highway', 1 = 'inflow_merge' (else 0) */
  double vehs_per_hour, depart_speed, begin;
  double end;            /* s; <= 0: never ends (InFlows.add default 86400)      */
  double probability;    /* per second, in (0, 1]; 0 = periodic flow             */
} fs_inflow;

/* device views of the multi-agent front (FS_ENV_MA_HIGHWAY_PO); A = max_agents */
typedef struct fs_agent_views {
  void* agent_obs;        /* T  [E][A][5] (= views.obs)                          */
  void* agent_reward;     /* T  [E][A]   reward of the agent in row a (0 if none) */
  int32_t* agent_id;      /* i32[E][A]   vehicle id of row a, -1 = empty row      */
  int32_t* veh_id;        /* i32[E][V]   vehicle id in each highway slot: initial vehicle i -> i,
                                         k-th vehicle of inflow f -> ((f + 1) << 24) | k */
  uint8_t* veh_type;      /* u8 [E][V]   vehicle type in each highway slot         */
  int32_t max_agents, obs_per_agent;
} fs_agent_views;

/* device views; T = float or double per cfg.dtype; [a][b] = a slow, b fast */
typedef struct fs_views {
  void* pos;         /* T  [V][E]  continuous route position                    */
  void* vel;         /* T  [V][E]                                               */
  uint8_t* lane;     /* u8 [V][E]                                               */
  uint8_t* active;   /* u8 [V][E]  highway slots; ring: all 1                   */
  void* headway;     /* T  [V][E]  bumper-to-bumper gap to leader (1000 if none)*/
  int32_t* leader;   /* i32[V][E]  vehicle index of leader, -1 if none          */
  int32_t* follower; /* i32[V][E]                                               */
  void* accel;       /* T  [V][E]  acceleration applied in the last step        */
  void* length;      /* T  [E]     loop circumference incl. junctions           */
  void* obs;         /* T  [E][obs_dim]                                         */
  void* reward;      /* T  [E]                                                  */
  uint8_t* done;     /* u8 [E]                                                  */
  uint8_t* crash;    /* u8 [E]                                                  */
  int32_t* time_counter; /* i32 [E]                                             */
  int32_t* counters;     /* i32 [3 + 2 FS_MAX_INFLOWS][E] highway: vehicles in network, departed,
                            arrived, vehicles emitted by inflow 0..3, vehicles of the
                            probabilistic inflows 0..3 still waiting to enter */
  int32_t n_envs, n_veh, n_lanes, obs_dim, act_dim, dtype;
  /* element (vehicle i, env e) of pos/vel/lane/headway/leader/follower/accel is at
   * i*veh_stride + e*env_stride: single-lane rings are [V][E] (E, 1), multi-lane
   * rings and highways are [E][V] (1, V) */
  int32_t veh_stride, env_stride;
} fs_views;

int fs_abi_version(void);
int fs_create(const fs_config* cfg, fs_engine** out);
void fs_destroy(fs_engine* e);
const char* fs_last_error(const fs_engine* e); /* e may be NULL: last create error */

/* types: n_types entries; type_of_vehicle: host u8 [V] */
int fs_set_types(fs_engine* e, const fs_vtype* types, int n_types,
                 const uint8_t* type_of_vehicle);

/* Reset the envs whose mask byte is non-zero (host u8 [E]; NULL = all):
 * draw ring length (or take lengths[E] host doubles if non-NULL), place
 * vehicles, refresh leader/headway, then run cfg.warmup_steps steps with no
 * RL action. Observation of the reset envs is left in views.obs. */
int fs_reset(fs_engine* e, const uint8_t* env_mask, const double* lengths);

/* The two halves of fs_reset, for placements the engine does not generate itself
 * (InitialConfig(perturbation > 0 / spacing = 'random' / shuffle = True),
 * flow/core/kernel/network/base.py:384-389, 393-487, flow/envs/base.py:275-277: the reference
 * draws them from Python's `random` and numpy's global generator ON THE HOST, and so does the
 * host side here, with the same streams): FS_RESET_PLACE = everything fs_reset does before the
 * warm-up (ring length, even placement, counters, flags, episode + 1); the caller then replaces
 * positions / lanes with fs_set_state; FS_RESET_WARMUP = the warm-up steps (and the
 * sort_vehicles key snapshot) on the state as it is now. Both bits = fs_reset. */
#define FS_RESET_PLACE  1
#define FS_RESET_WARMUP 2
int fs_reset_stage(fs_engine* e, const uint8_t* env_mask, const double* lengths, int stages);

/* k fused steps. actions: DEVICE pointer, T [k][act_dim][E], or NULL (no RL
 * action: RL vehicles are driven by the simulator model). If obs_seq /
 * reward_seq / done_seq are non-NULL (device, T [k][E][obs_dim], T [k][E],
 * u8 [k][E]) the per-step results are stored there; the last step's results
 * are always left in views.obs / reward / done. */
int fs_step(fs_engine* e, const void* actions, int k, void* obs_seq,
            void* reward_seq, uint8_t* done_seq);

/* The same k steps as k SINGLE-STEP launches - the variant whose vehicle state makes the round
 * trip through device memory every step (SURVEY 8d "single-step variant"; fs_step fuses the k
 * steps in one launch and keeps the state in registers). Results are identical to fs_step's.
 * The launches are issued from C and recorded into a CUDA graph; a later call with the same
 * k and buffer addresses replays it with one cudaGraphLaunch (new action values are picked up:
 * only the addresses are baked in). FS_NO_GRAPH=1 keeps plain launches. */
int fs_step_unfused(fs_engine* e, const void* actions, int k, void* obs_seq,
                    void* reward_seq, uint8_t* done_seq);

/* Same with HOST buffers (pageable or pinned): copies actions up, runs the k
 * steps, copies obs/reward/done of every step back, and synchronises. */
int fs_step_host(fs_engine* e, const void* actions, int k, void* obs_seq,
                 void* reward_seq, uint8_t* done_seq);

/* Closed-loop cadence (Env.step(policy(obs)), flow/envs/base.py:294-412 called once per
 * policy forward): calls of k <= 8 steps run through ONE pinned host block and ONE device block
 * laid out [actions | obs | reward | done]; upload -> step kernel(s) -> ONE download are
 * captured into a CUDA graph on the first call and replayed by one cudaGraphLaunch afterwards.
 * The host block is mapped; FS_CL_MODE (read at fs_create) lets the kernel read the actions
 * (bit 0) and / or write the results (bit 1) in place instead of using the copy engine - kept
 * for A/B runs, not faster on the measured box (flowstep.cu, step_host_closed_loop). fs_host_staging hands out that pinned block: a caller that writes its actions
 * into views.actions and passes the four pointers of `out` to fs_step_host pays no host-side
 * copy at all; any other host buffer is copied through the block. The pointers stay valid
 * until a later fs_host_staging call asks for a larger k, or fs_destroy. */
typedef struct fs_host_views {
  void* actions;        /* T  [k][act_dim][E]   pinned host                               */
  void* obs;            /* T  [k][E][obs_dim]                                             */
  void* reward;         /* T  [k][E]                                                      */
  uint8_t* done;        /* u8 [k][E]                                                      */
  int32_t k, n_envs, obs_dim, act_dim, dtype, reserved;
} fs_host_views;
int fs_host_staging(fs_engine* e, int k, fs_host_views* out);

/* fs_step_host for callers that may not allocate device memory. NO COUNTERPART IN THE
 * REFERENCE: added (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc). Same arguments and results as fs_step_host
 * (HOST buffers in and out, blocks until the results are on the host), but the device staging
 * between the copies and the step kernels is a file-scope static __device__ array of
 * FS_NOALLOC_SCRATCH_BYTES instead of cudaMalloc'ed blocks: the call makes no cudaMalloc /
 * cudaFree and no allocating library call. Limits and rules:
 *   - a call whose actions + obs + reward + done of all k steps (each rounded up to 256 B)
 *     exceed FS_NOALLOC_SCRATCH_BYTES is rejected with FS_EINVAL and a message naming both sizes
 *     (c2, 4096 envs fp32: 86 KB per step -> k <= 48; 65 536 envs: k <= 2);
 *   - every GPU that loads the library holds the array until the process exits;
 *   - ALL calls on one GPU share it, whatever the engine: calls that overlap on the GPU race,
 *     the caller serialises them; the regions a call hands back are cleared first, nothing of an
 *     earlier call is visible;
 *   - k = 0 does the one-time set-up only (looks the array's device address up for the
 *     engine's device), e.g. ahead of a stream capture. */
#define FS_NOALLOC_SCRATCH_BYTES (4 * 1024 * 1024)
int fs_step_host_noalloc(fs_engine* e, const void* actions, int k, void* obs_seq,
                         void* reward_seq, uint8_t* done_seq);

/* Pipelined form of fs_step_host for open-loop rollouts (the action sequence
 * is known up front): enqueues H2D actions -> k fused steps -> D2H results on
 * three internal streams and returns without waiting, so the upload of call
 * n+1, the kernel of call n and the download of call n-1 overlap (three
 * staging slots). The HOST buffers must be pinned for the copies to overlap and
 * must stay valid and untouched until fs_sync(), which waits for everything. */
int fs_step_host_async(fs_engine* e, const void* actions, int k, void* obs_seq,
                       void* reward_seq, uint8_t* done_seq);

int fs_buffers(fs_engine* e, fs_views* out);

/* Highway: replace the single inflow of fs_config by n_flows (0..FS_MAX_INFLOWS) flows, served
 * in this order when several are due in one step (each needs the entry gap on its own). Call
 * after fs_set_types and before fs_reset. */
int fs_set_inflows(fs_engine* e, const fs_inflow* flows, int n_flows);

/* Multi-agent views (FS_ENV_MA_HIGHWAY_PO only; FS_EINVAL otherwise). */
int fs_agents(fs_engine* e, fs_agent_views* out);

/* DLPack export of engine-owned device memory (obs / reward / state views, or
 * any sub-range of them): returns a malloc'ed DLManagedTensor* (DLPack v0 ABI,
 * device type kDLCUDA) describing `dev_ptr` with the given element type
 * (type_code 0 int, 1 uint, 2 float; bits 8/32/64), shape and optional strides
 * (in elements; NULL = C-contiguous). The tensor does not own the memory; its
 * deleter frees only the descriptor. Wrap it in a PyCapsule named "dltensor"
 * and hand it to torch.from_dlpack / cupy / jax. NULL on bad arguments. */
void* fs_dlpack_wrap(void* dev_ptr, int device, int type_code, int bits, int ndim,
                     const int64_t* shape, const int64_t* strides);

/* Overwrite state from host arrays (T [V][E], any may be NULL), then refresh
 * leader/headway exactly as k.update does. lane: u8 [V][E]. Only positions, speeds, lanes and
 * lengths are replaced: done / crash flags, time counters, the sort_vehicles keys and the
 * controller state of PISaturation / LAC vehicles keep their values (call fs_reset first for a
 * clean episode). */
int fs_set_state(fs_engine* e, const void* pos, const void* vel,
                 const uint8_t* lane, const double* lengths);

/* host_out[0..7] = sum reward, sum speed, sum speed^2, vehicle-steps,
 * env-steps, crashes, 0, 0 accumulated since the last fs_stats(reset=1). */
int fs_stats(fs_engine* e, double* host_out, int n, int reset);

int fs_sync(fs_engine* e);
/* cudaStream_t of the engine, for callers that want to order their own work */
void* fs_stream(fs_engine* e);
/* Adopt a caller-owned cudaStream_t (e.g. torch's current stream; NULL = the
 * legacy default stream) so engine work is ordered with the caller's. */
int fs_set_stream(fs_engine* e, void* cuda_stream);
/* number of kernel launches issued by this engine since creation */
int64_t fs_launch_count(const fs_engine* e);

#ifdef __cplusplus
}
#endif
#endif /* FLOWSTEP_H_ */
