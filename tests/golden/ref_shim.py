"""Make the Python reference (/root/reference) importable in the build
container so golden vectors can be generated from its own code.

The reference needs third-party packages that are absent here (gym, traci,
sumolib, lxml, pyglet, matplotlib, imutils, pyprind, ...). None of them does
arithmetic on the step path, so inert stand-ins are placed in sys.modules.
Used by make_golden.py, by tests that skip when no reference package is importable, and by
bench.py's CPU arm (kind "reference"). On the GPU box /root/reference does not exist; the
reference's package is then imported from baseline/_ref, the unmodified offline install of
/root/reference (`pip install --no-index --no-deps --target baseline/_ref`).
"""
import os
import sys
import types

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def _find():
    """The read-only checkout in the build container, else the offline install of the same
    package under baseline/_ref (pip install --target, see __graft_entry__.build): git-ignored,
    NOT gpurun-ignored, so it travels to the GPU box."""
    for p in (os.environ.get('FLOW_REFERENCE'), '/root/reference',
              os.path.join(_ROOT, 'baseline', '_ref')):
        if p and os.path.isdir(os.path.join(p, 'flow', 'controllers')):
            return p
    return '/root/reference'


REF = _find()


def available():
    return os.path.isdir(os.path.join(REF, 'flow'))


class _Anything(types.ModuleType):
    """Module whose every attribute is another inert module/callable."""

    def __getattr__(self, name):
        if name.startswith('__'):
            raise AttributeError(name)
        m = _Anything(self.__name__ + '.' + name)
        setattr(self, name, m)
        return m

    def __call__(self, *a, **k):
        return self


class Box:
    def __init__(self, low, high, shape=None, dtype=None):
        import numpy as np
        if shape is None:
            shape = np.shape(low)
        self.low = np.full(shape, low, dtype=float) if np.isscalar(low) \
            else np.asarray(low, dtype=float)
        self.high = np.full(shape, high, dtype=float) if np.isscalar(high) \
            else np.asarray(high, dtype=float)
        self.shape = tuple(shape)
        self.dtype = dtype


class Discrete:
    def __init__(self, n):
        self.n = n


class Tuple(tuple):
    def __new__(cls, spaces):
        return super().__new__(cls, spaces)


class _GymEnv:
    pass


class _MultiAgentEnv:
    """ray.rllib.env.MultiAgentEnv stand-in (distinct from gym.Env so that
    `class MultiEnv(MultiAgentEnv, Env)` has a consistent MRO)."""


def install():
    if not available():
        raise RuntimeError('reference checkout not found at %s' % REF)
    os.environ.setdefault('SUMO_HOME', '/nonexistent-sumo')
    if REF not in sys.path:
        sys.path.insert(0, REF)
    names = ['traci', 'traci.constants', 'traci.exceptions',
             'traci._trafficlight', 'sumolib', 'lxml', 'lxml.etree', 'pyglet',
             'matplotlib', 'matplotlib.cm', 'matplotlib.pyplot',
             'matplotlib.colors', 'imutils', 'pyprind', 'cached_property',
             'ray', 'ray.rllib', 'ray.rllib.env', 'ray.cloudpickle', 'cv2',
             'ray.rllib.env.multi_agent_env', 'redis']
    for n in names:
        if n not in sys.modules:
            sys.modules[n] = _Anything(n)
    sys.modules['traci.exceptions'].FatalTraCIError = type(
        'FatalTraCIError', (Exception,), {})
    sys.modules['traci.exceptions'].TraCIException = type(
        'TraCIException', (Exception,), {})
    sys.modules['ray.rllib.env'].MultiAgentEnv = _MultiAgentEnv
    sys.modules['ray.rllib.env.multi_agent_env'].MultiAgentEnv = _MultiAgentEnv
    sys.modules['cached_property'].cached_property = property
    if 'gym' not in sys.modules:
        gym = _Anything('gym')
        gym.Env = _GymEnv
        spaces = _Anything('gym.spaces')
        spaces.Box, spaces.Discrete, spaces.Tuple = Box, Discrete, Tuple
        box = _Anything('gym.spaces.box')
        box.Box = Box
        tup = _Anything('gym.spaces.tuple')
        tup.Tuple = Tuple
        tups = _Anything('gym.spaces.tuple_space')
        tups.Tuple = Tuple
        disc = _Anything('gym.spaces.discrete')
        disc.Discrete = Discrete
        dct = _Anything('gym.spaces.dict_space')
        gym.spaces = spaces
        sys.modules.update({
            'gym': gym, 'gym.spaces': spaces, 'gym.spaces.box': box,
            'gym.spaces.tuple': tup, 'gym.spaces.tuple_space': tups,
            'gym.spaces.discrete': disc, 'gym.spaces.dict_space': dct,
            'gym.envs': _Anything('gym.envs'),
            'gym.envs.registration': _Anything('gym.envs.registration')})
