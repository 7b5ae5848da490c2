"""Tiny gymnasium stand-in (see ../README.md). API subset used by /root/reference."""
from . import spaces  # noqa: F401
from .spaces import Space  # noqa: F401
import numpy as _np

__version__ = "0.28.1"


class Env:
    metadata = {"render_modes": []}
    render_mode = None
    spec = None
    observation_space = None
    action_space = None
    _np_random = None

    @property
    def unwrapped(self):
        return self

    @property
    def np_random(self):
        if self._np_random is None:
            self._np_random = _np.random.default_rng()
        return self._np_random

    @np_random.setter
    def np_random(self, v):
        self._np_random = v

    def reset(self, *, seed=None, options=None):
        if seed is not None:
            self._np_random = _np.random.default_rng(seed)
        return None

    def step(self, action):
        raise NotImplementedError

    def render(self):
        return None

    def close(self):
        return None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
        return False


class Wrapper(Env):
    def __init__(self, env):
        self.env = env

    def __getattr__(self, name):
        if name.startswith("_"):
            raise AttributeError(name)
        return getattr(self.env, name)

    @property
    def unwrapped(self):
        return self.env.unwrapped

    @property
    def observation_space(self):
        return self.env.observation_space

    @property
    def action_space(self):
        return self.env.action_space

    def reset(self, **kw):
        return self.env.reset(**kw)

    def step(self, action):
        return self.env.step(action)

    def render(self):
        return self.env.render()

    def close(self):
        return self.env.close()


class ObservationWrapper(Wrapper):
    def reset(self, **kw):
        obs, info = self.env.reset(**kw)
        return self.observation(obs), info

    def step(self, action):
        o, r, te, tr, i = self.env.step(action)
        return self.observation(o), r, te, tr, i


class ActionWrapper(Wrapper):
    def step(self, action):
        return self.env.step(self.action(action))


class RewardWrapper(Wrapper):
    def step(self, action):
        o, r, te, tr, i = self.env.step(action)
        return o, self.reward(r), te, tr, i


_registry = {}


def register(id, entry_point=None, **kw):
    _registry[id] = (entry_point, kw)


def make(id, **kw):
    ep, k = _registry[id]
    k = dict(k.get("kwargs", {}), **kw)
    if isinstance(ep, str):
        import importlib
        mod, cls = ep.split(":")
        ep = getattr(importlib.import_module(mod), cls)
    return ep(**k)


class _Registration:
    registry = _registry


envs = type("envs", (), {"registration": _Registration, "registry": _registry})
vector = type("vector", (), {})
