"""Environment objects over the batched env-step kernel (K1).

``VecEnv`` steps N classic-control environments at once on the GPU.  ``GymEnv`` is the N = 1 view
with the slice of the gym API the reference touches -- ``reset``, ``step`` (4-tuple), ``seed``,
``render``, ``spec.max_episode_steps``, ``observation_space.shape``, ``action_space`` (``Discrete.n``
or ``Box.high/low/shape``) -- so the reference-shaped ``rollout`` (``A3C/process.py:73-98``) runs
unchanged on it.  ``make(env_id)`` replaces ``gym.make`` (``A3C/process.py:104``).  Rendering is
unavailable on this box (no graphics stack): ``render`` is accepted and does nothing.
"""
import ctypes

import numpy as np
import torch

from . import _lib as L

ENV_IDS = {"Pendulum-v0": L.ENV_PENDULUM, "CartPole-v0": L.ENV_CARTPOLE, "Acrobot-v1": L.ENV_ACROBOT}


class Discrete(object):
    def __init__(self, n):
        self.n = n
        self.shape = ()


class Box(object):
    def __init__(self, low, high):
        self.low, self.high = np.asarray(low, dtype=np.float32), np.asarray(high, dtype=np.float32)
        self.shape = self.low.shape


class _Spec(object):
    def __init__(self, env_id, max_episode_steps):
        self.id, self.max_episode_steps = env_id, max_episode_steps


class VecEnv(object):
    """N environments in structure-of-arrays device buffers; auto-reset on done (include/acx.h, K1)."""

    def __init__(self, env_id, n, seed=12321, index0=0, device=None):
        if env_id not in ENV_IDS:
            raise L.AcxError("unsupported env %r (the CUDA path covers %s)" % (env_id, sorted(ENV_IDS)))
        if not torch.cuda.is_available():
            raise L.AcxError("VecEnv needs a CUDA device: the hot path has no CPU fallback")
        self.env_id, self.env, self.n, self._seed, self.index0 = env_id, ENV_IDS[env_id], int(n), int(seed), int(index0)
        s, o, a, m = (ctypes.c_int() for _ in range(4))
        L.call("acx_env_dims", self.env, ctypes.byref(s), ctypes.byref(o), ctypes.byref(a), ctypes.byref(m))
        self.state_dim, self.obs_dim, self.n_actions, self.max_episode_steps = s.value, o.value, a.value, m.value
        self.dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.state = torch.zeros((self.state_dim, self.n), device=self.dev)
        self.elapsed = torch.zeros(self.n, device=self.dev, dtype=torch.int32)
        self.obs = torch.zeros((self.obs_dim, self.n), device=self.dev)
        self.rew = torch.zeros(self.n, device=self.dev)
        self.done = torch.zeros(self.n, device=self.dev, dtype=torch.uint8)
        self.t = 0

    def seed(self, seed):
        self._seed = int(seed)

    def reset(self, reset_u=None):
        """-> obs [obs_dim, n] device tensor.  reset_u: optional [state_dim, n] uniforms in [0,1)."""
        ru = None if reset_u is None else torch.as_tensor(reset_u, device=self.dev, dtype=torch.float32).contiguous()
        L.call("acx_env_reset", self.env, self.n, L.ptr(self.state), L.ptr(self.elapsed), L.ptr(self.obs), L.ptr(ru),
               self._seed, self.index0, 0xFFFFFFFF - self.t, L.cur_stream())
        self.t += 1
        return self.obs

    def step(self, action, reset_u=None):
        """action: device tensor [n] (float32 for Pendulum, int32 otherwise) -> (obs, rew, done) device tensors."""
        ru = None if reset_u is None else torch.as_tensor(reset_u, device=self.dev, dtype=torch.float32).contiguous()
        L.call("acx_env_step", self.env, self.n, L.ptr(self.state), L.ptr(self.elapsed), L.ptr(action), L.ptr(self.obs),
               L.ptr(self.rew), L.ptr(self.done), L.ptr(ru), self._seed, self.index0, self.t, None, 0, L.cur_stream())
        self.t += 1
        return self.obs, self.rew, self.done


class GymEnv(object):
    """One environment with the gym-0.9 calls the reference makes."""

    def __init__(self, env_id, seed=12321):
        self.vec = VecEnv(env_id, 1, seed)
        self.spec = _Spec(env_id, self.vec.max_episode_steps)
        self.observation_space = Box(-np.ones(self.vec.obs_dim), np.ones(self.vec.obs_dim))
        if self.vec.n_actions:
            self.action_space = Discrete(self.vec.n_actions)
        else:
            self.action_space = Box([-2.0], [2.0])          # Pendulum max_torque
        self._needs_reset = True

    def seed(self, seed):
        self.vec.seed(seed)
        return [seed]

    def reset(self):
        self._needs_reset = False
        return self.vec.reset()[:, 0].cpu().numpy().astype(np.float64)

    def step(self, ac):
        if self.vec.n_actions:
            a = torch.tensor([int(np.asarray(ac).reshape(-1)[0])], device=self.vec.dev, dtype=torch.int32)
        else:
            a = torch.tensor([float(np.asarray(ac).reshape(-1)[0])], device=self.vec.dev, dtype=torch.float32)
        obs, rew, done = self.vec.step(a)
        # after a done the kernel has already started the next episode; the terminal observation is not
        # kept (the reference never uses it: no bootstrap after a terminal, A3C/process.py:63)
        return obs[:, 0].cpu().numpy().astype(np.float64), float(rew.item()), bool(done.item()), {}

    def render(self, *a, **k):
        return None      # unavailable on this box (no graphics stack)

    def close(self):
        return None


def make(env_id, seed=12321):
    """Stand-in for ``gym.make`` (A3C/process.py:104)."""
    return GymEnv(env_id, seed)
