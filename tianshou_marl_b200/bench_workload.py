"""The synthetic workload bench.py times.  One 'step' = Collector.collect(n_step = T * R) over E
simple_spread worlds (policy forward -> MPE world step -> replay-buffer write), then
PPO.update(buffer, batch_size, repeat) (critic forward, GAE, logp_old, minibatch loop with fused
loss + Adam), then reset_buffer(keep_statistics=True) -- the reference's OnPolicyTrainer step
(trainer.py:878-906, 1079-1109).  Product code only: everything goes through libtsm_b200."""
from __future__ import annotations

from typing import Any

from . import _lib
from .algorithm import PPO, AdamOptimizerFactory, DiscreteActorPolicy
from .data import Collector, VectorReplayBuffer
from .env import MPEVectorEnv
from .utils import ActorCriticNet

# kernels launched per public call (for gpu_launches): see the C entry points
K_POLICY, K_ROLLOUT_STEP = 1, 1  # tsm_net_forward; tsm_rollout_step (world step + ring counters + episode statistics)


class Workload:
    def __init__(self, args: Any, rank: int, world: int) -> None:
        self.args, self.rank, self.world = args, rank, world
        self.E, self.A, self.T = args.envs, args.agents, args.horizon
        self.R = self.E * self.A
        self.env = MPEVectorEnv("simple_spread_v3", n_envs=self.E, n_agents=self.A, seed=1000 + rank)
        self.O = self.env.obs_dim
        self.buf = VectorReplayBuffer(self.T * self.R, self.R, index_type="torch")
        self.net = ActorCriticNet(self.O, args.hidden, 5, seed=0, engine=args.engine)  # same init on every rank
        self.policy = DiscreteActorPolicy(actor=self.net, seed=77 + rank)
        self.alg = PPO(policy=self.policy, critic=self.net, optim=AdamOptimizerFactory(lr=3e-4), max_grad_norm=0.5,
                       minibatch_order="device", seed=rank, reuse_rollout_policy_outputs=bool(getattr(args, "reuse_rollout", False)))
        self.collector = Collector(self.alg, self.env, self.buf, episode_list=False)
        self.collector.reset()
        self.policy.is_within_training_step = True
        self.update = not args.no_update
        self.agent_steps_per_step = self.R * self.T
        self.dtype = "f16 operands / f32 accumulate (tcgen05)" if args.engine == "tc" else "f32"
        B = self.R * self.T
        self.n_mb = len(PPO.split_bounds(B, args.batch_size)) - 1
        parts = 2 if self.E >= 8192 else 1  # tsm_collect_lockstep: environment ranges on their own streams (TSM_COLLECT_STREAMS)
        self.launches_per_step = self.T * parts * (K_POLICY + K_ROLLOUT_STEP) + 1  # + buffer reset
        if self.update:
            # critic x2, logp_old, gae; per epoch: adv stats + either ONE persistent epoch kernel (tc engine, uniform
            # minibatches) or a (minibatch, optimizer) launch pair per minibatch (fp32 engine)
            per_epoch = 2 if args.engine == "tc" else 1 + 2 * self.n_mb
            # reuse_rollout: values_next + critic over ended rows; else: value + logp_old pass, values_next, critic over ended rows; + gae
            self.launches_per_step += (2 if getattr(args, "reuse_rollout", False) else 3) + 1 + args.repeat * per_epoch
            self.launches_per_step += 1 if args.engine == "tc" else 0  # pack_records
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self.last_stats: Any = None

    def step(self, timed: bool = False) -> None:
        """Device-resident step: the same public calls, stats left on the device except the one small
        read each call makes (9 doubles per collect, 4 doubles per minibatch per update)."""
        self.collector.collect(n_step=self.R * self.T)
        if self.update:
            self.last_stats = self.alg.update(self.buf, self.args.batch_size, self.args.repeat)
        self.collector.reset_buffer(keep_statistics=True)

    def step_e2e(self) -> None:
        """The call a user makes, host results included: CollectStats (episode returns / lengths copied
        to the host) and the training stats dataclass."""
        self.collector.episode_list = True
        cs = self.collector.collect(n_step=self.R * self.T)
        self.d2h_bytes = 9 * 8 + cs.returns.nbytes + cs.lens.nbytes
        if self.update:
            ts = self.alg.update(self.buf, self.args.batch_size, self.args.repeat)
            self.d2h_bytes += ts.gradient_steps * 4 * 8
            self.last_stats = ts
        self.collector.reset_buffer(keep_statistics=True)
        self.collector.episode_list = False
        self.h2d_bytes = (self.n_mb + 1) * 8 * (self.args.repeat if self.update else 0)  # minibatch bounds

    # -- reporting ------------------------------------------------------------------------------
    def nets(self) -> list:
        return [self.net] + ([self.critic] if getattr(self, "critic", None) is not None and self.critic is not self.net else [])

    def grad_sync(self) -> str:
        return "peer (inside the optimizer / epoch kernel over NVLink)" if getattr(self.alg, "_peers", None) is not None else "nccl all_reduce per minibatch"

    def reset_timers(self) -> None:
        _lib.timer = _lib.StageTimer()

    def stage_ms(self, steps: int) -> dict[str, dict[str, float]]:
        t = _lib.timer
        _lib.timer = None
        if t is None:
            return {}
        return {k: {"ms_per_step": ms / steps, "launch_ms": ms / n, "calls_per_step": n / steps} for k, (ms, n) in t.totals().items()}

    def roofline(self, stages: dict, hbm_peak: float, peak_src: str, tf_peak: float) -> dict:
        """Dominant kernel by time in the step (DESIGN.md lists the per-unit figures)."""
        O, A, H = self.O, self.A, self.args.hidden
        fwd_flop = 2.0 * (O * H + H * H + H * 6)
        per_unit_bytes = {"gae": 22.0, "collect": (16 * A * 2 + 8 * A + 4 * A + 8 * A * O + 6 * A) / A + 4.0 * O + 12.0 + 8.0, "mpe_step": (16 * A * 2 + 8 * A + 4 * A + 8 * A * O + 6 * A) / A, "buf_counters": 8.0,
                          "policy_forward": 4.0 * O + 12.0, "critic_forward": 4.0 * O + 4.0, "logp_old_forward": 4.0 * O + 8.0,
                          "adv_stats": 4.0, "pack_records": 4.0 * O + 20.0 + 96.0}
        B = self.R * self.T
        units = {"gae": B, "collect": B, "pack_records": B, "mpe_step": self.R, "buf_counters": self.R, "policy_forward": self.R, "critic_forward": B,
                 "logp_old_forward": B, "adv_stats": B}
        if not stages:
            return {}
        dom = max(stages, key=lambda k: stages[k]["ms_per_step"])
        out: dict[str, Any] = {"kernel": dom, "peak_source": peak_src}
        if dom == "ppo_epoch":
            # algorithmic FLOPs fwd + 2x bwd = 33.8 kFLOP/row, algorithmic bytes = one 96-byte record per row (tc engine) or
            # 4*O + 20 gathered bytes (fp32 engine).  tc engine, uniform minibatches: ONE persistent launch (tc_ppo_epoch_kernel)
            # per epoch = per stage call, optimizer steps and grid barriers inside (mlp_tc_epoch4.cuh); fp32 engine: a launch pair per minibatch.
            persistent = self.args.engine == "tc" and B % self.args.batch_size == 0
            rows = B if persistent else min(self.args.batch_size, B)
            t = stages[dom]["launch_ms"] * 1e-3 / (1 if persistent else self.n_mb)
            flop = 3.0 * fwd_flop * rows
            ach = flop / t / 1e12
            rec_bytes = 96.0 if self.args.engine == "tc" else 4.0 * O + 20.0
            out.update({"kernel": "tc_ppo_epoch4_kernel" if persistent else "ppo_minibatch_kernel",
                        "bound": "tensor", "achieved": ach, "peak": tf_peak, "unit": "TFLOP/s", "frac": ach / tf_peak,
                        # dram__bytes_read.sum + dram__bytes_write.sum of one tc_ppo_epoch_kernel launch over 25 165 824 rows,
                        # profiles/r02_ncu_full_epoch.csv (ncu --set full; records at a 128-byte pitch: 1.38x the 96 algorithmic
                        # bytes, 2.0x at the 96-byte pitch of round 1); null for other shapes / engines
                        "traffic": 3.326e9 if (persistent and rows == 25165824 and O == 18) else None,
                        "flop_per_unit": 3.0 * fwd_flop, "bytes_per_unit": rec_bytes, "units_per_launch": rows,
                        "launch_us": t * 1e6, "minibatch_us": stages[dom]["launch_ms"] * 1e3 / self.n_mb,
                        "hbm_gbs": rec_bytes * rows / t / 1e9,
                        "note": "engine=%s; latency bound: per minibatch one round of three or four 128-row tiles per SM (four tile groups, "
                                "5 dependent tcgen05 GEMM stages + loss each), gradient flush, one grid barrier, optimizer step in every CTA; "
                                "%d minibatches per launch" % (self.args.engine, self.n_mb if persistent else 1)})
        elif dom in ("policy_forward", "critic_forward", "logp_old_forward"):
            flop = fwd_flop * units[dom]
            t = stages[dom]["launch_ms"] * 1e-3
            ach = flop / t / 1e12
            out.update({"bound": "tensor", "achieved": ach, "peak": tf_peak, "unit": "TFLOP/s", "frac": ach / tf_peak, "traffic": None,
                        "flop_per_unit": fwd_flop, "units_per_launch": units[dom],
                        "hbm_gbs": per_unit_bytes[dom] * units[dom] / t / 1e9})
        else:
            ach = per_unit_bytes[dom] * units[dom] / (stages[dom]["launch_ms"] * 1e-3) / 1e9
            out.update({"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak, "traffic": None,
                        "bytes_per_unit": per_unit_bytes[dom], "units_per_launch": units[dom]})
        if "gae" in stages:
            g = 22.0 * B / (stages["gae"]["launch_ms"] * 1e-3) / 1e9
            out["gae"] = {"achieved": g, "frac": g / hbm_peak, "unit": "GB/s", "peak": hbm_peak}
        return out

    def config(self) -> dict:
        a = self.args
        what = "collect+GAE+PPO update" if self.update else "collect+buffer write (no update)"
        return {"workload": f"C2 simple_spread_v3 A={self.A} E={self.E}/GPU T={self.T}: {what}; shared actor-critic MLP "
                            f"{self.O}->{a.hidden}->{a.hidden}->(5 logits, 1 value)", "envs_per_gpu": self.E, "agents": self.A,
                "horizon": self.T, "rows_per_gpu": self.R, "batch_size": a.batch_size, "repeat": a.repeat, "hidden": a.hidden, "engine": a.engine, "reuse_rollout_policy_outputs": bool(getattr(a, "reuse_rollout", False)),
                "minibatches_per_epoch": self.n_mb,
                "l2": "inputs larger than L2: per-step working set %.0f MB > 126 MB" % (self.R * self.T * (8 * self.O + 30) / 1e6)}


class WorkloadC3(Workload):
    """BASELINE.json configs[2]: MAPPO / CTDE on simple_spread with N = 10 agents (obs 60), 262 144 worlds over 8 GPUs = 32 768 per
    GPU: DecentralizedActor 60 -> 128 -> 128 -> 5 on every agent row, CentralizedCritic 600 -> 128 -> 128 -> 10 on every
    (step, world) pair, engine "tcw" (csrc/mlp_tcw.cu).  One step = collect T vector steps + MAPPO.update + reset_buffer."""

    def __init__(self, args: Any, rank: int, world: int) -> None:
        from .algorithm.ctde import CentralizedCritic, DecentralizedActor
        from .algorithm.mappo import MAPPO, TILE

        self.args, self.rank, self.world = args, rank, world
        self.E, self.A, self.T = args.envs, args.agents, args.horizon
        self.R = self.E * self.A
        self.env = MPEVectorEnv("simple_spread_v3", n_envs=self.E, n_agents=self.A, seed=1000 + rank)
        self.O = self.env.obs_dim
        self.buf = VectorReplayBuffer(self.T * self.R, self.R, index_type="torch")
        self.net = DecentralizedActor(self.O, 5, 128, engine="tcw", orthogonal=True, seed=0)  # same init on every rank
        self.critic = CentralizedCritic(self.A * self.O, self.A, 128, engine="tcw", orthogonal=True, seed=1)
        self.policy = DiscreteActorPolicy(actor=self.net, seed=77 + rank)
        self.alg = MAPPO(policy=self.policy, critic=self.critic, n_agents=self.A, optim=AdamOptimizerFactory(lr=3e-4), max_grad_norm=0.5)
        self.collector = Collector(self.alg, self.env, self.buf, episode_list=False)
        self.collector.reset()
        self.policy.is_within_training_step = True
        self.update = not args.no_update
        self.agent_steps_per_step = self.R * self.T
        self.dtype = "f16 operands / f32 accumulate (tcgen05)"
        n_super = (self.T * self.E + TILE - 1) // TILE
        self.mb_super = max(1, int(round(args.batch_size / (TILE * self.A))))
        self.n_mb = (n_super + self.mb_super - 1) // self.mb_super
        parts = 2 if self.E >= 8192 else 1
        self.launches_per_step = self.T * parts * (2 * K_POLICY + K_ROLLOUT_STEP) + 1  # pack + forward + rollout step per range
        if self.update:
            # pack x2, critic forward, values_next + pack + critic on ended worlds, logp_old, gae; per epoch adv stats + 3 launches per minibatch
            self.launches_per_step += 8 + args.repeat * (1 + 3 * self.n_mb)
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self.last_stats = None

    def step_e2e(self) -> None:
        super().step_e2e()
        n_super = self.n_mb * self.mb_super
        self.h2d_bytes = (4 * n_super + 8 * self.n_mb) * (self.args.repeat if self.update else 0)  # super-tile order + rows per minibatch
        if self.update:
            self.d2h_bytes += self.last_stats.gradient_steps * 4 * 8  # second loss-sum array (critic)

    def roofline(self, stages: dict, hbm_peak: float, peak_src: str, tf_peak: float) -> dict:
        O, A, H = self.O, self.A, 128
        mac_a, mac_a_dx = O * H + H * H + H * 5, H * H + H * 5          # actor: forward MACs; data-gradient MACs (no dX of layer 1)
        mac_c, mac_c_dx = A * O * H + H * H + H * A, H * H + H * A      # critic, per (step, world) pair
        flop_row = 2.0 * (2 * mac_a + mac_a_dx) + 2.0 * (2 * mac_c + mac_c_dx) / A  # forward + weight gradients + data gradients, per agent row
        B = self.R * self.T
        if not stages:
            return {}
        dom = max(stages, key=lambda k: stages[k]["ms_per_step"])
        out: dict[str, Any] = {"kernel": dom, "peak_source": peak_src}
        if dom == "ppo_epoch":
            t = stages[dom]["launch_ms"] * 1e-3  # one epoch: n_mb x (critic launch, actor launch, optimizer launch)
            ach = flop_row * B / t / 1e12
            bytes_row = 2.0 * (O + 4) + 12.0 + (2.0 * (A * O + 8) + 8.0 * A) / A  # packed fp16 inputs + row scalars, per agent row
            out.update({"kernel": "tcw_kernel<resident|streaming, train> (actor + centralized critic) + tcw_adam_kernel", "bound": "tensor",
                        "achieved": ach, "peak": tf_peak, "unit": "TFLOP/s", "frac": ach / tf_peak, "traffic": None,
                        "flop_per_unit": flop_row, "bytes_per_unit": bytes_row, "units_per_launch": B, "launch_us": t * 1e6,
                        "minibatch_us": t * 1e6 / self.n_mb, "hbm_gbs": bytes_row * B / t / 1e9,
                        "note": "per minibatch: critic tiles (K = 608 streamed, W1 gradient flushed per tile), actor tiles (weights resident), "
                                "joint-clip Adam over both nets; %d minibatches of %d super-tiles per epoch" % (self.n_mb, self.mb_super)})
        else:
            per_unit_bytes = {"collect": (16 * A * 2 + 8 * A + 4 * A + 8 * A * O + 6 * A) / A + 4.0 * O + 12.0 + 8.0, "gae": 22.0,
                              "pack_inputs": 4.0 * O + 2.0 * (O + 4) + (4.0 * O * A + 2.0 * (A * O + 8)) / A,
                              "critic_forward": 2.0 * (A * O + 8) / A + 4.0, "logp_old_forward": 2.0 * (O + 4) + 8.0, "adv_stats": 4.0}
            ach = per_unit_bytes.get(dom, 0.0) * B / (stages[dom]["launch_ms"] * 1e-3) / 1e9
            out.update({"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak, "traffic": None,
                        "bytes_per_unit": per_unit_bytes.get(dom), "units_per_launch": B})
        if "gae" in stages:
            g = 22.0 * B / (stages["gae"]["launch_ms"] * 1e-3) / 1e9
            out["gae"] = {"achieved": g, "frac": g / hbm_peak, "unit": "GB/s", "peak": hbm_peak}
        return out

    def config(self) -> dict:
        a = self.args
        what = "collect+GAE+MAPPO update" if self.update else "collect+buffer write (no update)"
        return {"workload": f"C3 simple_spread N={self.A} agents E={self.E}/GPU T={self.T}: {what}; DecentralizedActor {self.O}->128->128->5, "
                            f"CentralizedCritic {self.A * self.O}->128->128->{self.A} on the concatenated global state",
                "envs_per_gpu": self.E, "agents": self.A, "horizon": self.T, "rows_per_gpu": self.R, "batch_size": a.batch_size,
                "rows_per_minibatch": self.mb_super * 128 * self.A, "repeat": a.repeat, "hidden": 128, "engine": "tcw",
                "minibatches_per_epoch": self.n_mb,
                "l2": "inputs larger than L2: per-step working set %.0f MB > 126 MB" % (self.R * self.T * (8 * self.O + 30) / 1e6)}


class WorkloadC4(Workload):
    """BASELINE.json configs[3]: simple_tag_v3 with grouped policies (three adversaries share one PPO policy, the good agent has its
    own), 131 072 worlds IN TOTAL at 1 / 2 / 4 / 8 GPUs (strong scaling: --envs is the total, each rank takes total / world).
    FlexibleMultiAgentPolicyManager inside the fused rollout loop (tsm_collect_lockstep_groups: one forward launch per policy
    group and step), then one PPO.update per group over its strided row set (persistent epoch kernel with explicit row lists)."""

    def __init__(self, args: Any, rank: int, world: int) -> None:
        from .algorithm import FlexibleMultiAgentPolicyManager

        self.args, self.rank, self.world = args, rank, world
        self.E, self.T = args.envs // world, args.horizon
        self.env = MPEVectorEnv("simple_tag_v3", n_envs=self.E, seed=1000 + rank)
        self.A, self.O = self.env.n_agents, self.env.obs_dim
        self.R = self.E * self.A
        self.buf = VectorReplayBuffer(self.T * self.R, self.R, index_type="torch")
        self.algs = {}
        for name, seed in (("adversaries", 1), ("good", 2)):
            net = ActorCriticNet(self.O, args.hidden, 5, seed=seed, engine=args.engine)  # same init on every rank
            pol = DiscreteActorPolicy(actor=net, seed=77 + 10 * seed + rank)
            self.algs[name] = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=3e-4), max_grad_norm=0.5, minibatch_order="device", seed=rank)
        ag = self.env.agents
        groups = {"adversaries": [a for a in ag if a.startswith("adversary")], "good": [a for a in ag if a.startswith("agent")]}
        self.manager = FlexibleMultiAgentPolicyManager(self.algs, self.env, mode="grouped", agent_groups=groups)
        self.net = self.algs["adversaries"].net
        self.critic = None
        self.alg = self.algs["adversaries"]
        self.collector = Collector(self.manager, self.env, self.buf, episode_list=False)
        self.collector.reset()
        for a in self.algs.values():
            a.policy.is_within_training_step = True
        self.manager.is_within_training_step = True
        self.update = not args.no_update
        self.agent_steps_per_step = self.R * self.T
        self.dtype = "f16 operands / f32 accumulate (tcgen05)" if args.engine == "tc" else "f32"
        n_adv = len(groups["adversaries"])
        self.n_mb = sum(len(PPO.split_bounds(self.E * k * self.T, args.batch_size)) - 1 for k in (n_adv, self.A - n_adv))
        self.launches_per_step = self.T * (2 * K_POLICY + K_ROLLOUT_STEP) + 1
        if self.update:
            per_epoch = 2 if args.engine == "tc" else 1 + 2 * self.n_mb
            self.launches_per_step += 2 * (3 + 1 + 1) + args.repeat * 2 * per_epoch
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self.last_stats = None

    def nets(self) -> list:
        return [a.net for a in self.algs.values()]

    def grad_sync(self) -> str:
        return "peer (inside the epoch kernel over NVLink)" if all(getattr(a, "_peers", None) is not None for a in self.algs.values()) else "nccl all_reduce per minibatch"

    def step(self, timed: bool = False) -> None:
        self.collector.collect(n_step=self.R * self.T)
        if self.update:
            self.last_stats = self.manager.update(self.buf, self.args.batch_size, self.args.repeat)
        self.collector.reset_buffer(keep_statistics=True)

    def step_e2e(self) -> None:
        self.collector.episode_list = True
        cs = self.collector.collect(n_step=self.R * self.T)
        self.d2h_bytes = 9 * 8 + cs.returns.nbytes + cs.lens.nbytes
        if self.update:
            ts = self.manager.update(self.buf, self.args.batch_size, self.args.repeat)
            self.d2h_bytes += sum(t.gradient_steps for t in ts.values()) * 4 * 8
            self.last_stats = ts
        self.collector.reset_buffer(keep_statistics=True)
        self.collector.episode_list = False
        self.h2d_bytes = (self.n_mb + 2) * 8 * (self.args.repeat if self.update else 0)

    def config(self) -> dict:
        a = self.args
        return {"workload": f"C4 simple_tag_v3 grouped policies (3 adversaries share one PPO policy + 1 good agent) {a.envs} worlds in total, "
                            f"{self.E}/GPU, T={self.T}: collect+GAE+PPO update per group; MLP {self.O}->{a.hidden}->{a.hidden}->(5 logits, 1 value) per group",
                "envs_total": a.envs, "envs_per_gpu": self.E, "agents": self.A, "horizon": self.T, "rows_per_gpu": self.R, "batch_size": a.batch_size,
                "repeat": a.repeat, "hidden": a.hidden, "engine": a.engine, "minibatches_per_epoch": self.n_mb,
                "l2": "inputs larger than L2: per-step working set %.0f MB > 126 MB" % (self.R * self.T * (8 * self.O + 30) / 1e6)}
