"""The reference arm of bench.py: the UNMODIFIED reference classes (installed into baseline/_ref by baseline/install_ref.py)
timed on the host cores.  Nothing of the product runs here.

    DummyVectorEnv / SubprocVectorEnv (tianshou/env/venvs.py:365-424) over one CPU world per worker
      -> RowFlatten (a VectorEnvWrapper, env/venv_wrappers.py:11-67): (E, A, ...) -> (E*A, ...) rows, row = env * A + agent
      -> Collector.collect(n_step = T * R)            tianshou/data/collector.py:461-1098
      -> VectorReplayBuffer(T * R, R)                 tianshou/data/buffer/vecbuf.py:15-37
      -> PPO.update(buffer, batch_size, repeat)       tianshou/algorithm/modelfree/ppo.py:146-224 (+ a2c.py:113-151, _gae)
      -> collector.reset_buffer(keep_statistics=True)

The world behind each worker is the oracle's restatement of PettingZoo 1.24.2 simple_spread (oracle/mpe.py; the simulator is
third-party and not installable here, see DESIGN.md "Oracle").  It is test infrastructure used as the env of the CPU baseline
only.  The reference's missing third-party imports are served by oracle/shims/.
"""
from __future__ import annotations

import os
import sys
import time
from typing import Any

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def available() -> bool:
    return os.path.isfile(os.path.join(HERE, "_ref", "tianshou", "__init__.py"))


def _paths() -> None:
    for p in (os.path.join(ROOT, "oracle", "shims"), os.path.join(HERE, "_ref"), ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)


def make_world_fn(n_agents: int, seed: int):
    """Factory of one single-world gymnasium-style env (runs inside a DummyEnvWorker / SubprocEnvWorker)."""

    def fn():
        _paths()
        import gymnasium as gym

        from oracle.mpe import SPREAD, OracleMPE

        class SpreadWorld(gym.Env):
            def __init__(self) -> None:
                self.w = OracleMPE(SPREAD, 1, n_agents=n_agents, seed=seed)
                self.observation_space = gym.spaces.Box(-np.inf, np.inf, (n_agents, self.w.obs_dim), np.float32)
                self.action_space = gym.spaces.MultiDiscrete([5] * n_agents)

            def reset(self, *, seed: int | None = None, options: Any = None):
                return self.w.reset()[0], {}

            def step(self, action):
                obs, rew, term, trunc = self.w.step(np.asarray(action).reshape(1, -1))
                return obs[0], rew[0], bool(term[0, 0]), bool(trunc[0, 0]), {}

            def seed(self, s: int | None = None):
                return [s]

            def render(self):
                return None

            def close(self) -> None:
                return None

        return SpreadWorld()

    return fn


def build(E: int, A: int, T: int, hidden: int, subproc: bool):
    _paths()
    import gymnasium as gym
    import torch
    from tianshou.algorithm import PPO
    from tianshou.algorithm.modelfree.reinforce import DiscreteActorPolicy
    from tianshou.algorithm.optim import AdamOptimizerFactory
    from tianshou.data import Collector, VectorReplayBuffer
    from tianshou.env import DummyVectorEnv, SubprocVectorEnv
    from tianshou.env.venv_wrappers import VectorEnvWrapper
    from tianshou.utils.net.common import ActorCritic, Net
    from tianshou.utils.net.discrete import DiscreteActor, DiscreteCritic

    class RowFlatten(VectorEnvWrapper):
        """(E, A, ...) outputs of per-world workers -> (E*A, ...) rows (row = env * A + agent)."""

        def __init__(self, venv: Any, n_agents: int) -> None:
            super().__init__(venv)
            self.A = n_agents
            self.rows = len(venv) * n_agents
            self._row_space = [gym.spaces.Discrete(5)] * self.rows

        def __len__(self) -> int:
            return self.rows

        def __getattribute__(self, key: str) -> Any:
            if key == "action_space":
                return object.__getattribute__(self, "_row_space")
            return super().__getattribute__(key)

        def _worlds(self, ids: Any) -> tuple[np.ndarray, np.ndarray]:
            rows = np.arange(self.rows) if ids is None else np.atleast_1d(np.asarray(ids, dtype=np.int64))
            worlds = np.unique(rows // self.A)
            if len(worlds) * self.A != len(rows):
                raise ValueError("rows of one world step / reset together")
            return rows, worlds

        def reset(self, env_id: Any = None, **kw: Any):
            rows, worlds = self._worlds(env_id)
            obs, _ = self.venv.reset(worlds, **kw)  # (n_worlds, A, O)
            pos = np.searchsorted(worlds, rows // self.A)
            return obs[pos, rows % self.A], np.array([{"env_id": int(r)} for r in rows], dtype=object)

        def step(self, action: Any, id: Any = None):
            rows, worlds = self._worlds(id)
            act = np.empty((len(worlds), self.A), np.int64)
            pos = np.searchsorted(worlds, rows // self.A)
            act[pos, rows % self.A] = np.asarray(action).reshape(-1)
            obs, rew, term, trunc, _ = self.venv.step(act, worlds)
            a = rows % self.A
            return (obs[pos, a], rew[pos, a], term[pos], trunc[pos], np.array([{"env_id": int(r)} for r in rows], dtype=object))

    O, nA = 6 * A, 5
    torch.manual_seed(0)
    net = Net(state_shape=(O,), hidden_sizes=[hidden, hidden])
    actor = DiscreteActor(preprocess_net=net, action_shape=nA, softmax_output=False)
    critic = DiscreteCritic(preprocess_net=net)
    for m in ActorCritic(actor, critic).modules():  # test/discrete/test_ppo_discrete.py:101-105
        if isinstance(m, torch.nn.Linear):
            torch.nn.init.orthogonal_(m.weight)
            torch.nn.init.zeros_(m.bias)
    policy = DiscreteActorPolicy(actor=actor, action_space=gym.spaces.Discrete(nA))
    alg = PPO(policy=policy, critic=critic, optim=AdamOptimizerFactory(lr=3e-4))
    fns = [make_world_fn(A, seed=i) for i in range(E)]
    venv = (SubprocVectorEnv if subproc else DummyVectorEnv)(fns)
    env = RowFlatten(venv, A)
    R = E * A
    buf = VectorReplayBuffer(T * R, R)
    col = Collector(alg, env, buf)
    col.reset()
    return alg, col, env


def run(E: int, A: int, T: int, batch_size: int, repeat: int, hidden: int, steps: int, warmup: int, subproc: bool = False,
        update: bool = True) -> dict:
    """Times `steps` steps (collect T vector steps + PPO.update + reset_buffer) after `warmup`; returns agent-steps/s."""
    _paths()
    import torch
    from tianshou.utils.torch_utils import policy_within_training_step, torch_train_mode

    torch.set_num_threads(os.cpu_count() or 1)
    alg, col, env = build(E, A, T, hidden, subproc)
    R = E * A
    t_col = t_upd = 0.0

    def one(timed: bool) -> None:
        nonlocal t_col, t_upd
        t0 = time.perf_counter()
        col.collect(n_step=T * R)
        t1 = time.perf_counter()
        if update:
            with policy_within_training_step(alg.policy), torch_train_mode(alg):  # trainer.py:1079-1095
                alg.update(buffer=col.buffer, batch_size=batch_size, repeat=repeat)
        col.reset_buffer(keep_statistics=True)
        t2 = time.perf_counter()
        if timed:
            t_col += t1 - t0
            t_upd += t2 - t1

    try:
        for _ in range(warmup):
            one(False)
        t0 = time.perf_counter()
        for _ in range(steps):
            one(True)
        dt = (time.perf_counter() - t0) / steps
    finally:
        env.close()
    kind = "SubprocVectorEnv" if subproc else "DummyVectorEnv"
    return {"value": R * T / dt, "ms_per_step": dt * 1e3, "collect_ms": t_col / steps * 1e3, "update_ms": t_upd / steps * 1e3,
            "cores": os.cpu_count() or 1, "workers": min(E, os.cpu_count() or 1) if subproc else 1,
            "sample": f"reference classes: {kind} x{E} simple_spread worlds (A={A}) -> RowFlatten -> Collector.collect(n_step={T * R}) -> "
                      f"VectorReplayBuffer({T * R}, {R}) -> PPO.update(batch_size={batch_size}, repeat={repeat if update else 0}), T={T}"}


def gae_numba(n: int = 1 << 24, reps: int = 3) -> dict:
    """The reference's numba `_gae` (algorithm_base.py:1079-1134; serial, one core) at B = n: GB/s on SURVEY 8(d)'s 22 B/element."""
    _paths()
    from tianshou.algorithm.algorithm_base import _gae

    rng = np.random.default_rng(0)
    v_s, v_n, rew = rng.normal(size=n), rng.normal(size=n), rng.normal(size=n)
    end = rng.random(n) < 0.04
    _gae(v_s[:1024], v_n[:1024], rew[:1024], end[:1024], 0.99, 0.95)  # JIT
    best = 1e30
    for _ in range(reps):
        t0 = time.perf_counter()
        _gae(v_s, v_n, rew, end, 0.99, 0.95)
        best = min(best, time.perf_counter() - t0)
    return {"n": n, "ms": best * 1e3, "gbs": 22 * n / best / 1e9, "cores": 1}
