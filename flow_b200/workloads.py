"""The BASELINE.json workloads as engine configurations (synthetic inputs of
SURVEY.md §8d). Each returns (types, type_of_vehicle, engine_kwargs)."""
from . import engine as EN

IDM_DEFAULT = [30, 1, 1, 1.5, 4, 2]  # flow/controllers/car_following_models.py:437-447


def ring_c1():
    """examples/exp_configs/non_rl/ring.py: 22 IDM, 230 m, bunching 20 (noise 0)."""
    types = [EN.make_vtype(EN.CTRL_IDM, IDM_DEFAULT)]
    kw = dict(env_kind=EN.ENV_ACCEL, length_lo=230.0, length_hi=230.0, bunching=20.0,
              target_velocity=10.0, horizon=3000, warmup_steps=0, act_lo=-3.0, act_hi=3.0)
    return types, [0] * 22, kw


def ring_c2(noise=0.0):
    """examples/exp_configs/rl/singleagent/singleagent_ring.py:
    WaveAttenuationPOEnv, 21 IDM (min_gap 0) + 1 RL, length ~ U{220..270},
    bunching 50, warm-up 750, horizon 3000, actions in [-1, 1]."""
    types = [EN.make_vtype(EN.CTRL_IDM, IDM_DEFAULT, noise=noise, sim_min_gap=0.0),
             EN.make_vtype(EN.CTRL_RL)]
    kw = dict(env_kind=EN.ENV_WAVE_ATTENUATION_PO, length_lo=220.0, length_hi=270.0,
              bunching=50.0, init_min_gap=0.0, warmup_steps=750, horizon=3000,
              act_lo=-1.0, act_hi=1.0, obs_max_length=270.0, clip_actions=0)
    return types, [0] * 21 + [1], kw


def ring_ma(noise=0.0):
    """examples/exp_configs/rl/multiagent/multiagent_ring.py:
    MultiAgentWaveAttenuationPOEnv, (1 RL + 10 IDM) x 2 on one ring."""
    human = EN.make_vtype(EN.CTRL_IDM, IDM_DEFAULT, noise=noise, sim_min_gap=0.0)
    types = [EN.make_vtype(EN.CTRL_RL), human]
    kw = dict(env_kind=EN.ENV_MA_WAVE_ATTENUATION_PO, length_lo=220.0, length_hi=270.0,
              bunching=50.0, init_min_gap=0.0, warmup_steps=750, horizon=3000,
              act_lo=-1.0, act_hi=1.0, obs_max_length=270.0, clip_actions=0)
    return types, [0] + [1] * 10 + [0] + [1] * 10, kw


WORKLOADS = {
    'c2': dict(builder=ring_c2, n_envs=4096, n_veh=22,
               name='WaveAttenuationPOEnv ring 21 IDM + 1 RL, 4096 envs/GPU, random RL actions'),
    'ma': dict(builder=ring_ma, n_envs=4096, n_veh=22,
               name='MultiAgentWaveAttenuationPOEnv ring (1 RL + 10 IDM) x 2, 4096 envs/GPU, '
                    'random RL actions'),
    'c5': dict(builder=ring_c2, n_envs=65536, n_veh=22,
               name='WaveAttenuationPOEnv ring stress, 65536 envs/GPU, fused 100-step launches'),
}
