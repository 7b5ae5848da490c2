"""The synthetic workload bench.py times: one 'step' = collect T vector steps of simple_spread over
E worlds (policy -> MPE step -> buffer add) + GAE (+ PPO update).  Product code only: everything
here goes through libtsm_b200 (no oracle, no CPU fallback)."""
from __future__ import annotations

import ctypes as C
from typing import Any

import torch

from . import _lib
from .algorithm import gae_time_major
from .data import Batch, VectorReplayBuffer
from .env import MPEVectorEnv


class StageTimer:
    """CUDA-event pairs around stages of the timed region, on the launching (current) stream."""

    def __init__(self) -> None:
        self.pairs: list[tuple[str, torch.cuda.Event, torch.cuda.Event]] = []
        self.on = False

    def start(self, name: str) -> Any:
        if not self.on:
            return None
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        self.pairs.append((name, a, b))
        return b

    @staticmethod
    def stop(tok: Any) -> None:
        if tok is not None:
            tok.record()

    def totals(self) -> dict[str, tuple[float, int]]:
        out: dict[str, tuple[float, int]] = {}
        for name, a, b in self.pairs:
            ms, n = out.get(name, (0.0, 0))
            out[name] = (ms + a.elapsed_time(b), n + 1)
        return out


class Workload:
    def __init__(self, args: Any, rank: int, world: int) -> None:
        self.args, self.rank, self.world = args, rank, world
        self.E, self.A, self.T = args.envs, args.agents, args.horizon
        self.R = self.E * self.A
        self.env = MPEVectorEnv("simple_spread_v3", n_envs=self.E, n_agents=self.A, seed=1000 + rank)
        self.O = self.env.obs_dim
        self.buf = VectorReplayBuffer(self.T * self.R, self.R, index_type="torch")
        dev = "cuda"
        self.act = torch.zeros(self.R, dtype=torch.int32, device=dev)
        self.obs_a = torch.zeros(self.R, self.O, device=dev)
        self.obs_b = torch.zeros(self.R, self.O, device=dev)
        self.obs_next = torch.zeros(self.R, self.O, device=dev)
        self.rew = torch.zeros(self.R, device=dev)
        self.term = torch.zeros(self.R, dtype=torch.uint8, device=dev)
        self.trunc = torch.zeros(self.R, dtype=torch.uint8, device=dev)
        self.adv = torch.zeros(self.T, self.R, device=dev)
        self.ret = torch.zeros(self.T, self.R, device=dev)
        self.vs = torch.zeros(self.T, self.R, device=dev)
        self.vn = torch.zeros(self.T, self.R, device=dev)
        obs, _ = self.env.reset()
        self.obs_a.copy_(obs)
        # allocate buffer storage with one add, then reset
        self.buf.add(Batch(obs=self.obs_a, act=self.act, rew=self.rew, terminated=self.term.view(torch.bool),
                           truncated=self.trunc.view(torch.bool), obs_next=self.obs_next))
        self.buf.reset()
        self.timer = StageTimer()
        self.step_counter = 0
        self.agent_steps_per_step = self.R * self.T
        self.dtype = "f32"
        self.launches_per_step = self.T * 4 + 2  # actions 1 + mpe_step 1 + buf_add 2 per vector step; buf_reset + gae
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self.pinned_stats = torch.zeros(4, dtype=torch.float32).pin_memory()

    # -- device-resident step ---------------------------------------------------------------
    def step(self, timed: bool = False) -> None:
        L, P, S = _lib.lib(), _lib.ptr, _lib.stream_ptr()
        tm = self.timer
        tm.on = timed
        env_d, buf_d = C.byref(self.env._desc), C.byref(self.buf._desc)
        _lib.check(L.tsm_buf_reset(buf_d, 1, S))
        cur, nxt = self.obs_a, self.obs_b
        for t in range(self.T):
            k = tm.start("policy")
            _lib.check(L.tsm_random_actions(P(self.act), self.R, 5, 1234, self.step_counter, S))
            tm.stop(k)
            self.step_counter += 1
            k = tm.start("mpe_step")
            _lib.check(L.tsm_mpe_step(env_d, P(self.act), P(self.obs_next), P(self.rew), P(self.term), P(self.trunc), 1, P(nxt), S))
            tm.stop(k)
            k = tm.start("buf_add")
            _lib.check(L.tsm_buf_add(buf_d, P(cur), P(self.act), P(self.rew), P(self.term), P(self.trunc), P(self.obs_next),
                                     None, None, None, self.R, None, None, None, None, S))
            tm.stop(k)
            cur, nxt = nxt, cur
        self.obs_a, self.obs_b = cur, nxt
        st = self.buf._store
        k = tm.start("gae")
        _lib.check(L.tsm_gae_tm(P(self.vs), P(self.vn), P(st["rew"]), P(st["terminated"]), P(st["truncated"]), self.T, self.R, 0,
                                self.T, 0.99, 0.95, P(self.adv), P(self.ret), S))
        tm.stop(k)
        tm.on = False

    # -- the same step through the public classes, host round trips included -------------------
    def step_e2e(self) -> None:
        env, buf = self.env, self.buf
        buf.reset(keep_statistics=True)
        obs = self.obs_a
        L, P, S = _lib.lib(), _lib.ptr, _lib.stream_ptr()
        for t in range(self.T):
            _lib.check(L.tsm_random_actions(P(self.act), self.R, 5, 1234, self.step_counter, S))
            self.step_counter += 1
            obs_next, rew, term, trunc, _ = env.step(self.act, autoreset=True, out={"obs_carry": self.obs_b})
            buf.add(Batch(obs=obs, act=self.act, rew=rew, terminated=term, truncated=trunc, obs_next=obs_next))
            self.obs_a, self.obs_b = self.obs_b, self.obs_a
            obs = self.obs_a
        st = buf._store
        ret, adv = gae_time_major(self.vs, self.vn, st["rew"], st["terminated"], st["truncated"], 0, self.T, 0.99, 0.95, self.adv, self.ret)
        stats = torch.stack([st["rew"].mean(), adv.mean(), ret.mean(), adv.std()])
        self.pinned_stats.copy_(stats, non_blocking=False)
        self.d2h_bytes = self.pinned_stats.numel() * 4

    # -- reporting ------------------------------------------------------------------------------
    def reset_timers(self) -> None:
        self.timer = StageTimer()

    def stage_ms(self, steps: int) -> dict[str, dict[str, float]]:
        return {k: {"ms_per_step": ms / steps, "launch_ms": ms / n, "calls_per_step": n / steps}
                for k, (ms, n) in self.timer.totals().items()}

    def roofline(self, stages: dict, hbm_peak: float, peak_src: str) -> dict:
        """Dominant kernel by time in the step.  achieved = algorithmic bytes per launch / CUDA-event
        duration per launch (DESIGN.md lists the per-unit byte figures)."""
        per_unit = {"gae": 22.0, "buf_add": 8.0 * self.O + 11.0, "mpe_step": (16 * self.A * 2 + 8 * self.A + 4 * self.A + 4 * self.A * self.O + 6 * self.A) / self.A,
                    "policy": 4.0}
        units = {"gae": self.R * self.T, "buf_add": self.R, "mpe_step": self.R, "policy": self.R}
        dom = max(stages, key=lambda k: stages[k]["ms_per_step"]) if stages else "gae"
        ach = per_unit[dom] * units[dom] / (stages[dom]["launch_ms"] * 1e-3) / 1e9 if stages else 0.0
        out = {"kernel": dom, "bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
               "traffic": None, "peak_source": peak_src, "bytes_per_unit": per_unit[dom], "units_per_launch": units[dom]}
        if "gae" in stages:
            g = per_unit["gae"] * units["gae"] / (stages["gae"]["launch_ms"] * 1e-3) / 1e9
            out["gae"] = {"achieved": g, "frac": g / hbm_peak, "unit": "GB/s"}
        return out

    def config(self) -> dict:
        a = self.args
        return {"workload": f"C2 simple_spread_v3 A={self.A} E={self.E}/GPU T={self.T}: collect(random-uniform policy)+buffer add+GAE "
                            f"(PPO update not in this build yet)", "envs_per_gpu": self.E, "agents": self.A, "horizon": self.T,
                "rows_per_gpu": self.R, "l2": "per-step working set %.0f MB > 126 MB L2" % (self.R * self.T * (8 * self.O + 30) / 1e6),
                "batch_size": a.batch_size, "repeat": a.repeat}
