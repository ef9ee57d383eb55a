"""Stand-in for `gym` (not installed here): just enough for the reference's drivers and
models to import and run on synthetic environments.  TEST INFRASTRUCTURE ONLY.

  gym.make("Chain-v0")      4 float32 observations, 3 discrete actions   (DuelingDQN / discrete AQL)
  gym.make("ChainCont-v0")  3 float32 observations, 1 continuous action  (continuous AQL)
Episodes are deterministic functions of the seed and the actions taken, 15-30 steps long."""
import numpy as np


class _Box(object):
    def __init__(self, low, high, shape=None, dtype=np.float32):
        self.low = np.full(shape, low, dtype=dtype) if shape is not None and np.isscalar(low) else np.asarray(low, dtype=dtype)
        self.high = np.full(shape, high, dtype=dtype) if shape is not None and np.isscalar(high) else np.asarray(high, dtype=dtype)
        self.shape = tuple(self.low.shape)
        self.dtype = dtype


class _Discrete(object):
    def __init__(self, n):
        self.n = n
        self.shape = ()


class spaces(object):
    Box = _Box
    Discrete = _Discrete


class _Spec(object):
    def __init__(self, env_id):
        self.id = env_id


class ChainEnv(object):
    def __init__(self, env_id, continuous):
        self.spec = _Spec(env_id)
        self.unwrapped = self
        self.continuous = continuous
        self.obs_dim = 3 if continuous else 4
        self.observation_space = _Box(-10.0, 10.0, shape=(self.obs_dim,))
        self.action_space = _Box(-2.0, 2.0, shape=(1,)) if continuous else _Discrete(3)
        self._seed, self._t, self._ep = 0, 0, 0

    def seed(self, s=None):
        self._seed = int(s or 0)

    def _obs(self):
        t, s = self._t, self._seed + 7 * self._ep
        base = [np.sin(0.3 * t + s), np.cos(0.2 * t), 0.05 * t, ((s * 31 + t * 17) % 13) / 13.0]
        return np.asarray(base[: self.obs_dim], dtype=np.float32)

    def reset(self):
        self._t = 0
        self._ep += 1
        return self._obs()

    def step(self, action):
        a = float(np.asarray(action, dtype=np.float64).ravel()[0])
        self._t += 1
        reward = float(np.cos(a + 0.1 * self._t))
        done = self._t >= 15 + (self._seed + self._ep) % 16
        return self._obs(), reward, done, {}

    def render(self):
        raise RuntimeError("no graphics stack on this box")

    def close(self):
        pass


def make(env_id):
    if env_id == "Chain-v0":
        return ChainEnv(env_id, False)
    if env_id == "ChainCont-v0":
        return ChainEnv(env_id, True)
    raise ValueError("stub gym knows Chain-v0 and ChainCont-v0, not %r" % (env_id,))
