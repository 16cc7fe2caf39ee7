"""B200VecEnv — the batched front of the engine: N independent Flow
environments stepped together on one GPU.

Mirrors the gym contract of flow/envs/base.py (reset() -> obs,
step(rl_actions) -> (obs, reward, done, info)) with a leading env axis:
obs [E, obs_dim], reward [E], done [E] as torch CUDA tensors that alias the
engine's buffers (zero copy, DLPack). `rollout` runs K fused steps from an
action sequence; `*_host` variants take/return pinned host arrays.
"""
import numpy as np

from . import engine as EN


class B200VecEnv:
    def __init__(self, types, type_of_vehicle, n_envs, dtype='f32', device=0, **cfg):
        self.engine = EN.Engine(types, type_of_vehicle, n_envs, dtype=dtype,
                                device=device, **cfg)
        self.n_envs = n_envs
        self.obs_dim, self.act_dim = self.engine.obs_dim, self.engine.act_dim
        self.dtype = self.engine.dtype
        self._pinned = {}

    # -- gym-like, device tensors -------------------------------------------
    def reset(self, mask=None, lengths=None):
        return self.engine.reset(mask, lengths)

    def step(self, rl_actions=None):
        """One Env.step for every env. rl_actions: [E] / [act_dim, E] device
        tensor (any DLPack producer) or None (RL vehicles follow the simulator
        model, as in the reference's warm-up)."""
        obs, rew, done = self.engine.step(rl_actions, k=1)
        return obs, rew, done, {}

    def rollout(self, actions, obs_seq=None, reward_seq=None, done_seq=None):
        """K = actions.shape[0] fused steps in one launch. actions [K, act_dim, E]."""
        k = int(actions.shape[0])
        self.engine.step(actions, k=k, obs_seq=obs_seq, reward_seq=reward_seq,
                         done_seq=done_seq)
        return obs_seq, reward_seq, done_seq

    # -- host buffers (the e2e path: H2D actions, D2H obs/reward/done) ---------
    def _pinned_buf(self, name, shape, dtype):
        import torch
        key = (name, tuple(shape))
        if key not in self._pinned:
            self._pinned[key] = torch.empty(shape, dtype=dtype).pin_memory()
        return self._pinned[key]

    def pinned_actions(self, k):
        import torch
        td = torch.float64 if self.dtype == 'f64' else torch.float32
        return self._pinned_buf('act', (k, self.act_dim, self.n_envs), td)

    def rollout_host(self, actions_host, k=None):
        """actions_host: numpy / pinned torch CPU tensor [K, act_dim, E].
        Returns numpy views (pinned) of obs [K,E,obs_dim], reward [K,E], done [K,E]."""
        import torch
        td = torch.float64 if self.dtype == 'f64' else torch.float32
        if isinstance(actions_host, torch.Tensor):
            a = actions_host.numpy()
        else:
            a = np.ascontiguousarray(actions_host, dtype=self.engine.np_dtype)
        k = int(a.shape[0]) if k is None else k
        E = self.n_envs
        obs = self._pinned_buf('obs', (k, E, self.obs_dim), td).numpy()
        rew = self._pinned_buf('rew', (k, E), td).numpy()
        done = self._pinned_buf('done', (k, E), torch.uint8).numpy()
        e = self.engine
        e._check(e.lib.fs_step_host(e.h, a.ctypes.data, k, obs.ctypes.data,
                                    rew.ctypes.data, done.ctypes.data))
        return obs, rew, done

    def rollout_host_async(self, actions_host, obs_out, rew_out, done_out):
        """Pipelined open-loop rollout: enqueue K = actions_host.shape[0] fused steps
        with PINNED host tensors (actions [K, act_dim, E] in; obs [K, E, obs_dim],
        reward [K, E], done [K, E] out) and return at once; the upload of the next
        call overlaps this call's kernel and the previous call's download
        (fs_step_host_async). Call `wait()` before reading the outputs."""
        e = self.engine
        k = int(actions_host.shape[0])
        e._check(e.lib.fs_step_host_async(e.h, actions_host.data_ptr(), k, obs_out.data_ptr(),
                                          rew_out.data_ptr(), done_out.data_ptr()))

    def wait(self):
        self.engine.sync()

    def step_host(self, actions_host=None):
        """One Env.step of every env with HOST data, the gym cadence obs = env.step(policy(obs)):
        `actions_host` ([E] / [act_dim, E], numpy or CPU tensor; None = simulator model) is
        copied into the engine's pinned staging block, the step runs as one replayed CUDA graph
        (upload -> kernel -> one download) and the call returns when obs [E, obs_dim],
        reward [E], done [E] are on the host. The returned arrays are views of the staging
        block: valid until the next call."""
        e = self.engine
        st = getattr(e, '_staging', None) or e.host_staging(1)
        if actions_host is not None:
            # [E] broadcasts onto [act_dim = 1, E]; [act_dim, E] copies row for row
            np.copyto(st['actions'][0], actions_host, casting='same_kind')
        e.step_staged(1, with_actions=actions_host is not None)
        return st['obs'][0], st['reward'][0], st['done'][0], {}

    def stats(self, reset=False):
        return self.engine.stats(reset)

    def close(self):
        self.engine.close()
