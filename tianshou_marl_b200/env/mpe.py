"""Batched MPE vector environment on the device (K1).

Stands where the reference puts `DummyVectorEnv` / `SubprocVectorEnv` over PettingZoo MPE
workers (tianshou/env/venvs.py:130-473, env/pettingzoo_env.py:20-132,
env/enhanced_pettingzoo_env.py:17-237): same `reset(env_id)` / `step(action, id)` /
`__len__` / `action_space` / `is_async` protocol the Collector uses, but one kernel launch
steps every environment.  Rows are flattened as r = env * n_agents + agent; the
"environments" the Collector sees are ROWS, so `len(env) == n_envs * n_agents` and all rows of
one world truncate together.
"""
from __future__ import annotations

import ctypes as C
from typing import Any

import numpy as np
import torch

from .. import _lib

SPREAD, TAG, ADVERSARY = 0, 1, 2
_SCENARIOS = {"simple_spread": SPREAD, "simple_spread_v3": SPREAD, "simple_tag": TAG, "simple_tag_v3": TAG, "simple_adversary": ADVERSARY,
              "simple_adversary_v3": ADVERSARY, 0: SPREAD, 1: TAG, 2: ADVERSARY}


class Discrete:
    """Minimal stand-in for gymnasium.spaces.Discrete (gymnasium is not a dependency)."""

    def __init__(self, n: int, seed: int | None = None) -> None:
        self.n, self.shape, self.dtype = int(n), (), np.int64
        self._rng = np.random.default_rng(seed)

    def sample(self, mask: Any = None) -> np.int64:
        return np.int64(self._rng.integers(self.n))

    def contains(self, x: Any) -> bool:
        return 0 <= int(x) < self.n

    def seed(self, seed: int | None = None) -> None:
        self._rng = np.random.default_rng(seed)

    def __repr__(self) -> str:
        return f"Discrete({self.n})"


class Box:
    def __init__(self, low: float, high: float, shape: tuple[int, ...], dtype: Any = np.float32, seed: int | None = None) -> None:
        self.low, self.high, self.shape, self.dtype = low, high, tuple(shape), dtype
        self._rng = np.random.default_rng(seed)

    def sample(self, mask: Any = None) -> np.ndarray:
        """Uniform over a bounded box (gymnasium.spaces.Box.sample for finite bounds)."""
        return self._rng.uniform(self.low, self.high, self.shape).astype(self.dtype)

    def seed(self, seed: int | None = None) -> None:
        self._rng = np.random.default_rng(seed)

    def __repr__(self) -> str:
        return f"Box({self.low}, {self.high}, {self.shape}, {self.dtype})"


class _SpaceList:
    """`env.action_space[i]` / `len(env.action_space)` without materialising one object per row
    (collector.py:715-720 indexes it per env)."""

    def __init__(self, space: Any, n: int) -> None:
        self._space, self._n = space, n

    def __len__(self) -> int:
        return self._n

    def __getitem__(self, i: Any) -> Any:
        if isinstance(i, slice | list | np.ndarray):
            return [self._space] * len(np.arange(self._n)[i])
        return self._space

    def __iter__(self):
        return (self._space for _ in range(self._n))


class MPEVectorEnv:
    is_async = False

    def __init__(self, scenario: str | int = "simple_spread_v3", n_envs: int = 1, n_agents: int = 3, max_cycles: int = 25,
                 local_ratio: float = 0.5, pos_first: bool = True, seed: int = 0, num_good: int = 1,
                 num_adversaries: int = 3, num_obstacles: int = 2, device: str | torch.device = "cuda",
                 continuous_actions: bool = False) -> None:
        self.scenario = _SCENARIOS[scenario]
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _lib.TsmError("MPEVectorEnv needs a CUDA device (no CPU fallback)")
        _lib.lib()
        if self.scenario == SPREAD:
            A, L, n_adv = n_agents, n_agents, 0
            O = 6 * A
            self.agents = [f"agent_{i}" for i in range(A)]
        elif self.scenario == ADVERSARY:
            # simple_adversary_v3(N = 2): one adversary, N good agents, N landmarks (one is the goal)
            n_good = n_agents - 1 if n_agents >= 2 and num_good == 1 and num_adversaries == 3 else num_good  # MPEVectorEnv("simple_adversary_v3", n_agents=3)
            A, L, n_adv = 1 + n_good, n_good, 1
            O = 2 + 2 * L + 2 * (A - 1)
            self.agents = ["adversary_0"] + [f"agent_{i}" for i in range(n_good)]
        else:
            A, L, n_adv = num_adversaries + num_good, num_obstacles, num_adversaries
            O = 4 + 2 * L + 2 * (A - 1) + 2 * num_good
            self.agents = [f"adversary_{i}" for i in range(n_adv)] + [f"agent_{i}" for i in range(num_good)]
        self.n_envs, self.n_agents, self.n_landmarks, self.n_adversaries, self.obs_dim = n_envs, A, L, n_adv, O
        self.num_agents, self.possible_agents = A, list(self.agents)
        self.agent_idx = {a: i for i, a in enumerate(self.agents)}
        self.env_num = self.n_rows = n_envs * A
        self.max_cycles, self.local_ratio, self.pos_first, self._seed = max_cycles, local_ratio, pos_first, seed
        dev = self.device
        self.pos = torch.zeros(n_envs * A, 2, dtype=torch.float32, device=dev)
        self.vel = torch.zeros(n_envs * A, 2, dtype=torch.float32, device=dev)
        self.lm = torch.zeros(max(n_envs * L, 1), 2, dtype=torch.float32, device=dev)
        self.steps = torch.zeros(n_envs, dtype=torch.int32, device=dev)
        self.episode = torch.zeros(n_envs, dtype=torch.int64, device=dev)
        d = self._desc = _lib.TsmMpe()
        d.scenario, d.n_agents, d.n_landmarks, d.n_adversaries, d.obs_dim = self.scenario, A, L, n_adv, O
        d.max_cycles, d.pos_first, d.local_ratio, d.n_envs, d.seed = max_cycles, int(pos_first), local_ratio, n_envs, seed
        d.pos, d.vel, d.lm = self.pos.data_ptr(), self.vel.data_ptr(), self.lm.data_ptr()
        d.steps, d.episode = self.steps.data_ptr(), self.episode.data_ptr()
        # continuous_actions=True: Box(0, 1, (5,)) per agent (simple_env: space_dim = dim_p * 2 + 1, every agent here is silent)
        self.continuous_actions = bool(continuous_actions)
        self.single_action_space = Box(0.0, 1.0, (5,), seed=seed) if self.continuous_actions else Discrete(5, seed)
        self.single_observation_space = Box(-np.inf, np.inf, (O,))
        self.action_space = _SpaceList(self.single_action_space, self.n_rows)
        self.observation_space = _SpaceList(self.single_observation_space, self.n_rows)
        self.agent_id_of_row = np.arange(self.n_rows) % A  # obs.agent_id[r] = agents[r % A]

    def __len__(self) -> int:
        return self.n_rows

    def seed(self, seed: int | None = None) -> list[int]:
        if seed is not None:
            self._seed = self._desc.seed = int(seed)
        return [self._seed] * self.n_rows

    def close(self) -> None:
        pass

    def render(self, **kwargs: Any) -> list:
        return []  # no graphics stack on the box; accepted and ignored

    # -- reset / step -----------------------------------------------------------------------
    def _env_ids(self, row_ids: Any) -> torch.Tensor:
        rows = torch.as_tensor(row_ids, device=self.device).to(torch.int64).reshape(-1)
        envs = torch.unique_consecutive(torch.div(rows, self.n_agents, rounding_mode="floor"))
        if envs.numel() * self.n_agents != rows.numel():
            raise ValueError("rows of one world must be reset together (all agents of an env truncate together)")
        return envs

    def reset(self, env_id: Any = None, init_pos: Any = None, **kwargs: Any) -> tuple[torch.Tensor, Any]:
        """venvs.py:195-235.  env_id are ROW ids (None = all).  Returns (obs (n, O), info)."""
        ids = None if env_id is None else self._env_ids([env_id] if np.isscalar(env_id) else env_id)
        n = self.n_envs if ids is None else ids.numel()
        obs = torch.empty(n * self.n_agents, self.obs_dim, dtype=torch.float32, device=self.device)
        ip = None
        if init_pos is not None:
            ip = torch.as_tensor(init_pos, device=self.device).to(torch.float32).reshape(n, self.n_agents + self.n_landmarks, 2).contiguous()
        _lib.check(_lib.lib().tsm_mpe_reset(C.byref(self._desc), _lib.ptr(ids), n, _lib.ptr(ip), _lib.ptr(obs), _lib.stream_ptr()))
        return obs, EmptyInfo(n * self.n_agents)

    def step(self, action: Any, id: Any = None, autoreset: bool = False,
             out: dict[str, torch.Tensor] | None = None) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor, Any]:
        """venvs.py:237-322.  `action` (R,) ints, or (R, 5) floats with continuous_actions=True.  Returns (obs_next, rew, terminated, truncated, info)
        as device tensors.  `id` must be None or all rows (a batched world cannot step a subset)."""
        if id is not None and len(id) != self.n_rows:
            raise ValueError("MPEVectorEnv steps all rows at once")
        R, O, dev = self.n_rows, self.obs_dim, self.device
        if self.continuous_actions:
            act = torch.as_tensor(action, device=dev).to(torch.float32).reshape(R, 5).contiguous()
        else:
            act = torch.as_tensor(action, device=dev).to(torch.int32).reshape(R).contiguous()
        o = out or {}
        obs_next = o.get("obs_next", None)
        if obs_next is None:
            obs_next = torch.empty(R, O, dtype=torch.float32, device=dev)
        rew = o["rew"] if "rew" in o else torch.empty(R, dtype=torch.float32, device=dev)
        term = o["terminated"] if "terminated" in o else torch.empty(R, dtype=torch.uint8, device=dev)
        trunc = o["truncated"] if "truncated" in o else torch.empty(R, dtype=torch.uint8, device=dev)
        carry = o.get("obs_carry", None)
        fn = _lib.lib().tsm_mpe_step_continuous if self.continuous_actions else _lib.lib().tsm_mpe_step
        _lib.check(fn(C.byref(self._desc), _lib.ptr(act), _lib.ptr(obs_next), _lib.ptr(rew),
                      _lib.ptr(term), _lib.ptr(trunc), int(autoreset), _lib.ptr(carry), _lib.stream_ptr()))
        return obs_next, rew, term.view(torch.bool), trunc.view(torch.bool), EmptyInfo(R)


class EmptyInfo:
    """Sentinel for the per-env info dicts (the reference builds an object array of dicts,
    venvs.py:311-322; MPE infos are empty, so nothing is materialised)."""

    def __init__(self, n: int) -> None:
        self._n = n

    def __len__(self) -> int:
        return self._n

    def __getitem__(self, i: Any) -> Any:
        return {} if np.isscalar(i) else EmptyInfo(len(np.arange(self._n)[i]) if not isinstance(i, torch.Tensor) else int(i.numel() if i.dtype != torch.bool else i.sum()))

    def __setitem__(self, i: Any, v: Any) -> None:
        pass
