"""Test infrastructure: a restatement of the reference's test env MoveToRightEnv (test/base/env.py:12-171, default scalar state)
and a minimal host vector env with the protocol the Collector uses (venvs.py:195-322: reset(env_id) / step(action, id) /
action_space / is_async).  The expected buffers come from the reference itself (tests/golden/collector.npz)."""
from __future__ import annotations

from typing import Any

import numpy as np


class _Discrete:
    def __init__(self, n: int) -> None:
        self.n = n

    def sample(self) -> int:
        return int(np.random.randint(self.n))


class MoveToRightEnv:
    """Go right `size` steps: obs = index, action 1 moves right, arriving at `size` terminates with reward 1."""

    def __init__(self, size: int) -> None:
        self.size, self.index, self.terminated = size, 0, False
        self.action_space = _Discrete(2)

    def reset(self) -> tuple[np.ndarray, dict]:
        self.terminated, self.index = False, 0
        return np.array([self.index], dtype=np.float32), {}

    def step(self, action: Any):
        if self.terminated:
            raise ValueError("step after done !!!")
        if self.index == self.size:
            self.terminated = True
        elif int(action) == 0:
            self.index = max(self.index - 1, 0)
        else:
            self.index += 1
            self.terminated = self.index == self.size
        return np.array([self.index], dtype=np.float32), float(self.terminated), self.terminated, False, {}


class HostVectorEnv:
    is_async = False

    def __init__(self, envs: list) -> None:
        self.envs = envs
        self.action_space = [e.action_space for e in envs]

    def __len__(self) -> int:
        return len(self.envs)

    def reset(self, env_id: Any = None, **kw: Any):
        ids = range(len(self.envs)) if env_id is None else np.atleast_1d(np.asarray(env_id))
        obs = [self.envs[int(i)].reset()[0] for i in ids]
        return np.stack(obs), np.array([{} for _ in obs], dtype=object)

    def step(self, action: Any, id: Any = None):
        ids = range(len(self.envs)) if id is None else np.atleast_1d(np.asarray(id))
        out = [self.envs[int(i)].step(a) for i, a in zip(ids, np.asarray(action))]
        obs, rew, term, trunc, info = zip(*out)
        return np.stack(obs), np.asarray(rew, dtype=np.float64), np.asarray(term, dtype=bool), np.asarray(trunc, dtype=bool), np.array(info, dtype=object)

    def close(self) -> None:
        return None
