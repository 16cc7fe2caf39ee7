"""The BASELINE.json workloads as engine configurations (synthetic inputs of
SURVEY.md §8d). Each returns (types, type_of_vehicle, engine_kwargs)."""
from . import engine as EN

IDM_DEFAULT = [30, 1, 1, 1.5, 4, 2]  # flow/controllers/car_following_models.py:437-447
# SumoCarFollowingParams.speed_mode: every shipped config leaves the default 'right_of_way'
# (flow/core/params.py:12-18, 841) in place, so the workloads carry it (0 = 'aggressive')
SPEED_MODE = 25


def ring_c1(speed_mode=SPEED_MODE):
    """examples/exp_configs/non_rl/ring.py: 22 IDM, 230 m, bunching 20 (noise 0)."""
    types = [EN.make_vtype(EN.CTRL_IDM, IDM_DEFAULT, speed_mode=speed_mode)]
    kw = dict(env_kind=EN.ENV_ACCEL, length_lo=230.0, length_hi=230.0, bunching=20.0,
              target_velocity=10.0, horizon=3000, warmup_steps=0, act_lo=-3.0, act_hi=3.0)
    return types, [0] * 22, kw


def ring_c2(noise=0.0, speed_mode=SPEED_MODE):
    """examples/exp_configs/rl/singleagent/singleagent_ring.py:
    WaveAttenuationPOEnv, 21 IDM (min_gap 0) + 1 RL, length ~ U{220..270},
    bunching 50, warm-up 750, horizon 3000, actions in [-1, 1]."""
    types = [EN.make_vtype(EN.CTRL_IDM, IDM_DEFAULT, noise=noise, sim_min_gap=0.0,
                           speed_mode=speed_mode),
             EN.make_vtype(EN.CTRL_RL, speed_mode=speed_mode)]
    kw = dict(env_kind=EN.ENV_WAVE_ATTENUATION_PO, length_lo=220.0, length_hi=270.0,
              bunching=50.0, init_min_gap=0.0, warmup_steps=750, horizon=3000,
              act_lo=-1.0, act_hi=1.0, obs_max_length=270.0, clip_actions=0)
    return types, [0] * 21 + [1], kw


def ring_ma(noise=0.0, speed_mode=SPEED_MODE):
    """examples/exp_configs/rl/multiagent/multiagent_ring.py:
    MultiAgentWaveAttenuationPOEnv, (1 RL + 10 IDM) x 2 on one ring."""
    human = EN.make_vtype(EN.CTRL_IDM, IDM_DEFAULT, noise=noise, sim_min_gap=0.0,
                          speed_mode=speed_mode)
    types = [EN.make_vtype(EN.CTRL_RL, speed_mode=speed_mode), human]
    kw = dict(env_kind=EN.ENV_MA_WAVE_ATTENUATION_PO, length_lo=220.0, length_hi=270.0,
              bunching=50.0, init_min_gap=0.0, warmup_steps=750, horizon=3000,
              act_lo=-1.0, act_hi=1.0, obs_max_length=270.0, clip_actions=0)
    return types, [0] + [1] * 10 + [0] + [1] * 10, kw


def ring_c3(noise=0.0, speed_mode=SPEED_MODE):
    """LaneChangeAccelEnv on a 3-lane 230 m ring, 66 vehicles 3-abreast
    (63 IDM that keep their lane + 3 RL that change lane), actions [accel, direction] per RL."""
    types = [EN.make_vtype(EN.CTRL_IDM, IDM_DEFAULT, noise=noise, speed_mode=speed_mode),
             EN.make_vtype(EN.CTRL_RL, lc_controller=EN.LC_RL, speed_mode=speed_mode)]
    kw = dict(env_kind=EN.ENV_LANE_CHANGE_ACCEL, n_lanes=3, length_lo=230.0, length_hi=230.0,
              target_velocity=10.0, horizon=3000, warmup_steps=0, act_lo=-3.0, act_hi=3.0,
              lane_change_duration=5)
    return types, [0] * 63 + [1] * 3, kw


def highway_c4(noise=0.0, speed_mode=SPEED_MODE):
    """single-lane 1 km HighwayNetwork, IDM vehicles, inflow 2000 veh/h at
    20 m/s, 160 vehicle slots per env, departures / arrivals compacted every step."""
    types = [EN.make_vtype(EN.CTRL_IDM, IDM_DEFAULT, noise=noise, speed_mode=speed_mode)]
    kw = dict(topology=EN.TOPO_HIGHWAY, env_kind=EN.ENV_HIGHWAY, length_lo=1000.0, length_hi=1000.0,
              junction_len=0.0, target_velocity=10.0, horizon=3000, warmup_steps=0,
              inflow_vehs_per_hour=2000.0, inflow_depart_speed=20.0, inflow_type=0,
              capacity=160)
    return types, [0] * 20, kw


def highway_ma(noise=0.0, speed_mode=SPEED_MODE):
    """MultiAgentHighwayPOEnv (flow/envs/multiagent/highway.py) on the single-lane 1 km
    HighwayNetwork with the inflow split of examples/exp_configs/rl/multiagent/
    multiagent_highway.py (20 % RL penetration): an RL inflow of 400 veh/h served before an IDM
    inflow of 1600 veh/h, 160 slots, up to 32 agent rows per env."""
    types = [EN.make_vtype(EN.CTRL_IDM, IDM_DEFAULT, noise=noise, speed_mode=speed_mode),
             EN.make_vtype(EN.CTRL_RL, speed_mode=speed_mode)]
    kw = dict(topology=EN.TOPO_HIGHWAY, env_kind=EN.ENV_MA_HIGHWAY_PO, length_lo=1000.0,
              length_hi=1000.0, junction_len=0.0, target_velocity=25.0, horizon=3000,
              warmup_steps=0, capacity=160, max_agents=32,
              inflows=[(1, 400.0, 20.0, 1.0), (0, 1600.0, 20.0, 1.0)])
    return types, [0] * 20, kw


def merge_mg(noise=0.0, speed_mode=SPEED_MODE):
    """MergePOEnv on MergeNetwork with the numbers of examples/exp_configs/rl/
    singleagent/singleagent_merge.py (pre-merge 500 m, ramp and post-merge 100 m, 2000 veh/h on
    the highway of which 10 % RL, 100 veh/h on the ramp, 5 controlled rows, sim_step 0.2 with
    sims_per_step 5, horizon 600). The network starts empty and fills during 100 warm-up
    Env.steps; 128 slots."""
    types = [EN.make_vtype(EN.CTRL_IDM, IDM_DEFAULT, noise=noise, speed_mode=speed_mode),
             EN.make_vtype(EN.CTRL_RL, speed_mode=speed_mode)]
    kw = dict(topology=EN.TOPO_MERGE, env_kind=EN.ENV_MERGE_PO, merge_pre=500.0, merge_ramp=100.0,
              merge_post=100.0, max_agents=5, sim_step=0.2, sims_per_step=5, target_velocity=20.0,
              horizon=600, warmup_steps=100, capacity=128,
              inflows=[(0, 1800.0, 10.0, 1.0, 0.0, 0.0, 0), (1, 200.0, 10.0, 1.0, 0.0, 0.0, 0),
                       (0, 100.0, 7.5, 1.0, 0.0, 0.0, 1)])
    return types, [], kw


def figure8_f8(noise=0.0, speed_mode=SPEED_MODE):
    """AccelEnv on FigureEightNetwork with the numbers of examples/exp_configs/
    rl/singleagent/singleagent_figure_eight.py (radius 30, 13 IDM humans + 1 RL vehicle, decel 1.5,
    target velocity 20, horizon 1500)."""
    types = [EN.make_vtype(EN.CTRL_IDM, IDM_DEFAULT, noise=noise, speed_mode=speed_mode, max_decel=1.5),
             EN.make_vtype(EN.CTRL_RL, speed_mode=speed_mode, max_decel=1.5)]
    kw = dict(topology=EN.TOPO_FIGURE8, env_kind=EN.ENV_ACCEL, fig8_radius=30.0, sim_step=0.1,
              target_velocity=20.0, horizon=1500, warmup_steps=0, act_lo=-3.0, act_hi=3.0,
              clip_actions=1)
    return types, [0] * 13 + [1], kw


WORKLOADS = {
    'f8': dict(builder=figure8_f8, n_envs=4096, n_veh=14,
               name='AccelEnv on FigureEightNetwork (radius 30), 13 IDM + 1 RL, 4096 envs/GPU, random '
                    'RL accelerations'),
    'mg': dict(builder=merge_mg, n_envs=4096, n_veh=128,
               name='MergePOEnv on MergeNetwork (500 / 100 / 100 m), inflows 1800 + 200 RL + 100 veh/h, '
                    '5 controlled rows, 5 simulation steps per Env.step, 128 slots, 4096 envs/GPU, '
                    'random accelerations'),
    'mah': dict(builder=highway_ma, n_envs=1024, n_veh=160,
                name='MultiAgentHighwayPOEnv open 1 km highway, IDM inflow 1600 + RL inflow 400 veh/h, '
                     '160 slots, 32 agent rows, 1024 envs/GPU, agents without command (simulator model)'),
    'c3': dict(builder=ring_c3, n_envs=2048, n_veh=66,
               name='LaneChangeAccelEnv 3-lane ring, 63 IDM + 3 RL lane changers, 2048 envs/GPU, '
                    'random accelerations and directions'),
    'c4': dict(builder=highway_c4, n_envs=1024, n_veh=160,
               name='open 1 km highway, inflow 2000 veh/h, 20 initial IDM vehicles, 160 slots, '
                    '1024 envs/GPU'),
    'c2': dict(builder=ring_c2, n_envs=4096, n_veh=22,
               name='WaveAttenuationPOEnv ring 21 IDM + 1 RL, 4096 envs/GPU, random RL actions'),
    'ma': dict(builder=ring_ma, n_envs=4096, n_veh=22,
               name='MultiAgentWaveAttenuationPOEnv ring (1 RL + 10 IDM) x 2, 4096 envs/GPU, '
                    'random RL actions'),
    'c5': dict(builder=ring_c2, n_envs=65536, n_veh=22,
               name='WaveAttenuationPOEnv ring stress, 65536 envs/GPU, fused 100-step launches'),
}
