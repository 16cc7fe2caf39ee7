"""The callers either side of the step path: the
`flow_params.json` round trip RLlib experiments use, and a vector-env adapter so
that one policy forward feeds thousands of engine envs.

Reference counterparts (parthjaggi/flow):
  FlowParamsEncoder, get_flow_params     flow/utils/rllib.py:22-192
  make_create_env / register_env use     examples/train.py:110-212
  ray.rllib.env.VectorEnv contract       vector_reset / reset_at / vector_step /
                                         get_unwrapped (ray is not a dependency here:
                                         the adapter is duck-typed to that interface)

The JSON written here has the reference's layout (VehicleParams -> its
`initial` rows with controller class NAMES, env / network classes as dotted
paths, every other parameter object as its __dict__), so a file produced by the
reference's encoder is read back by `get_flow_params` below and vice versa;
classes are resolved by their last path component against `flow_b200.flow_api`.
"""
import inspect
import json
from copy import deepcopy

import numpy as np

from . import flow_api as F


class FlowParamsEncoder(json.JSONEncoder):
    """json.dumps(flow_params, cls=FlowParamsEncoder, sort_keys=True, indent=4)"""

    def default(self, obj):
        if type(obj).__name__ == 'VehicleParams':
            rows = deepcopy(obj.initial)
            for r in rows:
                for key in ('acceleration_controller', 'lane_change_controller',
                            'routing_controller'):
                    if r.get(key) is not None:
                        r[key] = (r[key][0].__name__, r[key][1])
            return rows
        if inspect.isclass(obj):
            if any(b.__name__ in ('Env', 'Network') for b in obj.__mro__):
                return '{}.{}'.format(obj.__module__, obj.__name__)
            return obj.__name__
        if hasattr(obj, '__name__'):
            return obj.__name__
        if isinstance(obj, (np.integer,)):
            return int(obj)
        if isinstance(obj, (np.floating,)):
            return float(obj)
        return obj.__dict__


def _lookup(name, table, what):
    return F._resolve(name.split('.')[-1], table, what)


_ROUTERS = {'ContinuousRouter': F.ContinuousRouter}


def get_flow_params(config):
    """flow_params dict from an RLlib config dict (`config['env_config']['flow_params']`
    holds the JSON text) or from the path of a flow_params json file."""
    if isinstance(config, dict):
        fp = json.loads(config['env_config']['flow_params'])
    else:
        with open(config, 'r') as f:
            fp = json.load(f)

    veh = F.VehicleParams()
    for row in fp['veh']:
        row = dict(row)
        acc = row.pop('acceleration_controller')
        lc = row.pop('lane_change_controller')
        rt = row.pop('routing_controller')
        cf = F.SumoCarFollowingParams()
        cf.__dict__ = row.pop('car_following_params')
        lcp = F.SumoLaneChangeParams()
        lcp.__dict__ = row.pop('lane_change_params')
        rt_ctrl = None
        if rt is not None:
            if rt[0] not in _ROUTERS:
                raise NotImplementedError('routing controller %r (rings and straight highways '
                                          'need none)' % rt[0])
            rt_ctrl = (_ROUTERS[rt[0]], rt[1])
        veh.add(acceleration_controller=(_lookup(acc[0], F._ACCEL_BY_NAME, 'acceleration controller'),
                                         acc[1]),
                lane_change_controller=(_lookup(lc[0], F._LC_BY_NAME, 'lane-change controller'),
                                        lc[1]),
                routing_controller=rt_ctrl, car_following_params=cf, lane_change_params=lcp,
                **row)

    def obj(cls, d):
        o = cls()
        o.__dict__ = dict(d)
        return o

    sim = obj(F.SumoParams, fp['sim'])
    net = obj(F.NetParams, fp['net'])
    net.inflows = F.InFlows()
    if fp['net'].get('inflows'):
        net.inflows.__dict__ = dict(fp['net']['inflows'])
    env = obj(F.EnvParams, fp['env'])
    initial = obj(F.InitialConfig, fp['initial']) if 'initial' in fp else F.InitialConfig()
    if initial.lanes_distribution is None:
        initial.lanes_distribution = float('inf')
    tls = obj(F.TrafficLightParams, fp['tls']) if 'tls' in fp else F.TrafficLightParams()
    fp['env_name'] = _lookup(fp['env_name'], F._ENV_BY_NAME, 'environment')
    fp['network'] = _lookup(fp['network'], F._NET_BY_NAME, 'network')
    fp.update(sim=sim, env=env, initial=initial, net=net, veh=veh, tls=tls)
    return fp


def dump_flow_params(flow_params, path=None):
    """The text `examples/train.py` stores in `config['env_config']['flow_params']`."""
    txt = json.dumps(flow_params, cls=FlowParamsEncoder, sort_keys=True, indent=4)
    if path:
        with open(path, 'w') as f:
            f.write(txt)
    return txt


class B200VectorEnv:
    """`num_envs` Flow environments behind RLlib's VectorEnv interface, stepped by
    ONE kernel launch per `vector_step`:

        vector_reset() -> [obs_0 .. obs_{n-1}]
        reset_at(i)    -> obs_i
        vector_step(actions) -> ([obs], [reward], [done], [info])
        get_unwrapped() -> []        (no per-env Python objects exist)

    Observations / rewards cross to the host as ONE copy per call and are handed
    out as rows of that array. A sub-env whose episode ended is reset on the next
    `vector_step` the way RLlib does (`reset_at`), through a masked engine reset
    (fresh ring length, placement and warm-up for that env only)."""

    def __init__(self, flow_params, num_envs, dtype='f32', device=0, seed=None, env_id_base=0):
        create_env, self.env_name = F.make_create_env(
            flow_params, n_envs=num_envs, dtype=dtype, device=device, seed=seed,
            env_id_base=env_id_base)
        self.env = create_env()
        if self.env.engine.act_dim < 1:
            raise ValueError('a vector env for RL needs at least one RL vehicle')
        self.num_envs = num_envs
        self.observation_space = self.env.observation_space
        self.action_space = self.env.action_space
        self._obs = None

    def _host_obs(self):
        self._obs = self.env.engine.obs.cpu().numpy()
        return self._obs

    def vector_reset(self):
        self.env.reset()
        return list(self._host_obs())

    def reset_at(self, index):
        mask = np.zeros(self.num_envs, np.uint8)
        mask[index] = 1
        self.env.engine.reset(mask=mask)
        return self._host_obs()[index]

    def reset_done(self):
        """Masked reset of every finished env; returns their indices."""
        done = self.env.engine.done.cpu().numpy()
        idx = np.nonzero(done)[0]
        if idx.size:
            self.env.engine.reset(mask=done)
            self._host_obs()
        return idx

    def vector_step(self, actions):
        a = np.asarray(actions, dtype=np.float64).reshape(self.num_envs, self.env.engine.act_dim)
        obs, rew, done, _ = Env_step(self.env, a)
        o = self._host_obs()
        r = rew.cpu().numpy()
        d = done.cpu().numpy().astype(bool)
        return list(o), list(r), list(d), [{} for _ in range(self.num_envs)]

    def get_unwrapped(self):
        return []

    def close(self):
        self.env.terminate()


def Env_step(env, a):
    """Base-class step (array actions) whatever the env front is."""
    return F.Env.step(env, a)
