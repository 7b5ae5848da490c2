"""PPO on the device (K4/K5/K6): drop-in for the reference's on-policy update path

    OnPolicyAlgorithm.update            tianshou/algorithm/algorithm_base.py:852-863
    Algorithm._update                   tianshou/algorithm/algorithm_base.py:584-629
    PPO._preprocess_batch               tianshou/algorithm/modelfree/ppo.py:146-162
    A2C-base._add_returns_and_advantages tianshou/algorithm/modelfree/a2c.py:113-151
    PPO._update_with_batch              tianshou/algorithm/modelfree/ppo.py:164-224
    Algorithm.Optimizer.step            tianshou/algorithm/algorithm_base.py:485-498
    DiscreteActorPolicy / ProbabilisticActorPolicy.forward   reinforce.py:167-215

All arithmetic runs in libtsm_b200 (csrc/mlp.cu, gae.cu); this module orders the launches.
The update batch is never gathered: kernels address the time-major device buffer in place.

Minibatch order: `minibatch_order="numpy"` draws np.random.permutation(B) per epoch from the
global numpy RNG exactly like Batch.split (batch.py:1219) -- bit-exact minibatch index lists;
`"device"` (default) uses the keyed Feistel permutation evaluated inside the kernels (no index
arrays, different but equally uniform order).
"""
from __future__ import annotations

import ctypes as C
import time
from dataclasses import dataclass
from typing import Any

import numpy as np
import torch

from .. import _lib
from ..data.batch import Batch
from ..utils.net import ActorCriticNet
from ..utils.stats import A2CTrainingStats, SequenceSummaryStats
from .returns import gae_flat, gae_time_major


@dataclass
class AdamOptimizerFactory:
    """tianshou/algorithm/optim.py:89-110 (defaults lr 1e-3, betas (0.9, 0.999), eps 1e-8, no weight decay)."""

    lr: float = 1e-3
    betas: tuple[float, float] = (0.9, 0.999)
    eps: float = 1e-8
    weight_decay: float = 0.0

    def __post_init__(self) -> None:
        if self.weight_decay != 0.0:
            raise NotImplementedError("weight_decay != 0 is not on this path")


class DiscreteActorPolicy:
    """reinforce.py:195-247 / :167-192: logits -> Categorical -> sample (or mode when
    deterministic_eval and not within a training step).  One fused kernel per call."""

    def __init__(self, *, actor: ActorCriticNet, action_space: Any = None, observation_space: Any = None,
                 deterministic_eval: bool = False, seed: int = 0, **kwargs: Any) -> None:
        self.actor = actor
        self.action_space, self.observation_space = action_space, observation_space
        self.deterministic_eval = deterministic_eval
        self.is_within_training_step = False
        self.training = False
        self.seed = int(seed)
        self._counter = 0

    def train(self, mode: bool = True) -> "DiscreteActorPolicy":
        self.training = mode
        return self

    def eval(self) -> "DiscreteActorPolicy":
        return self.train(False)

    def map_action(self, act: Any) -> Any:
        return act  # identity for Discrete (algorithm_base.py:255-288)

    def forward(self, batch: Any, state: Any = None, **kwargs: Any) -> Batch:
        obs = batch.obs if hasattr(batch, "obs") else batch["obs"]
        if isinstance(obs, Batch):
            obs = obs["obs"]
        if not (isinstance(obs, torch.Tensor) and obs.is_cuda):
            obs = torch.as_tensor(np.ascontiguousarray(obs), dtype=torch.float32, device=self.actor.device)
        obs = obs.to(torch.float32).contiguous()
        mode = "mode" if (self.deterministic_eval and not self.is_within_training_step) else "sample"
        o = self.actor.forward(obs, mode=mode, seed=self.seed, counter=self._counter, want_logits=True)
        self._counter += 1
        out = Batch()
        out.__dict__.update(logits=o["logits"], act=o["act"].to(torch.int64), state=None, logp=o["logp"])
        if "value" in o:
            out.__dict__["value"] = o["value"]
        out.__dict__["dist"] = torch.distributions.Categorical(logits=o["logits"]) if kwargs.get("want_dist") else None
        return out

    __call__ = forward


class ContinuousActorPolicy(DiscreteActorPolicy):
    """ProbabilisticActorPolicy over a Box action space (reinforce.py:118-215) with the reference's standard continuous head:
    ContinuousActorProbabilistic(conditioned_sigma=False), unbounded or tanh-bounded means (utils/net/continuous.py:196-247), and
    dist_fn = Independent(Normal(mu, sigma), 1) (test/continuous/test_ppo.py:104-110).  `actor` is an ActorCriticNet(gaussian=True):
    its n_act outputs are the means, `log_sigma` the state-independent log standard deviations (or, with conditioned_sigma=True,
    n_act more head outputs per row).  forward samples
    mu + sigma * N(0, 1) (the mean when deterministic_eval and not in a training step); map_action clips / scales like
    Policy.map_action (algorithm_base.py:255-288, tsm_map_action).  The network pass is the library's forward kernel; the
    sampling and the log-density are elementwise device ops on its [n, A] output."""

    def __init__(self, *, actor: ActorCriticNet, action_space: Any = None, action_scaling: bool = True,
                 action_bound_method: str | None = "clip", **kwargs: Any) -> None:
        if not getattr(actor, "gaussian", False):
            raise ValueError("ContinuousActorPolicy needs ActorCriticNet(gaussian=True)")
        super().__init__(actor=actor, action_space=action_space, **kwargs)
        self._mapper = None
        if action_space is not None and hasattr(action_space, "low"):
            from .actions import BoxActionMapper

            self._mapper = BoxActionMapper(action_space.low, action_space.high, action_scaling=action_scaling,
                                           action_bound_method=action_bound_method, device=actor.device)
        self._gen = torch.Generator(device=actor.device).manual_seed(int(kwargs.get("seed", 0)))

    def map_action(self, act: Any) -> Any:
        return act if self._mapper is None else self._mapper.map_action(act)

    @staticmethod
    def log_prob(mu: torch.Tensor, log_sigma: torch.Tensor, act: torch.Tensor) -> torch.Tensor:
        """torch.distributions.Normal.log_prob summed over the action dimensions (Independent(..., 1))."""
        var = (2.0 * log_sigma).exp()
        return (-((act - mu) ** 2) / (2.0 * var) - log_sigma - 0.9189385332046727).sum(-1)

    def forward(self, batch: Any, state: Any = None, **kwargs: Any) -> Batch:
        obs = batch.obs if hasattr(batch, "obs") else batch["obs"]
        if isinstance(obs, Batch):
            obs = obs["obs"]
        if not (isinstance(obs, torch.Tensor) and obs.is_cuda):
            obs = torch.as_tensor(np.ascontiguousarray(obs), dtype=torch.float32, device=self.actor.device)
        net = self.actor
        o = net.forward(obs.to(torch.float32).contiguous(), mode="raw")["out"]
        mu = net.gaussian_mean(o)
        ls = net.gaussian_log_sigma(o)
        if self.deterministic_eval and not self.is_within_training_step:
            act = mu.clone()
        else:
            act = mu + ls.exp() * torch.randn(mu.shape, device=mu.device, generator=self._gen)
        out = Batch()
        out.__dict__.update(logits=(mu, ls.exp().expand_as(mu)), act=act, state=None, logp=self.log_prob(mu, ls, act))
        if net.value_index >= 0:
            out.__dict__["value"] = o[:, net.value_index]
        out.__dict__["dist"] = (torch.distributions.Independent(torch.distributions.Normal(mu, ls.exp().expand_as(mu)), 1)
                                if kwargs.get("want_dist") else None)
        return out

    __call__ = forward


class PPO:
    def __init__(self, *, policy: DiscreteActorPolicy, critic: ActorCriticNet | None = None,
                 optim: AdamOptimizerFactory | None = None, eps_clip: float = 0.2, dual_clip: float | None = None,
                 value_clip: bool = False, advantage_normalization: bool = True, recompute_advantage: bool = False,
                 vf_coef: float = 0.5, ent_coef: float = 0.01, max_grad_norm: float | None = None, gae_lambda: float = 0.95,
                 max_batchsize: int = 256, gamma: float = 0.99, return_scaling: bool = False,
                 minibatch_order: str = "device", reuse_rollout_policy_outputs: bool = False, dedup_next_values: bool = True, seed: int = 0,
                 grad_sync: str = "auto") -> None:
        assert dual_clip is None or dual_clip > 1.0, f"Dual-clip PPO parameter should greater than 1.0 but got {dual_clip}"  # ppo.py
        assert 0.0 <= gamma <= 1.0 and 0.0 <= gae_lambda <= 1.0
        if critic is not None and critic is not policy.actor:
            raise NotImplementedError("separate critic network: use the CTDE path")
        self._mc_returns = bool(getattr(self, "_mc_returns_flag", False))  # Reinforce: Monte-Carlo returns, no critic
        if policy.actor.value_index < 0 and not self._mc_returns:
            raise ValueError("the shared actor-critic net needs a value head (with_value=True)")
        if minibatch_order not in ("device", "numpy"):
            raise ValueError("minibatch_order must be 'device' or 'numpy'")
        self.policy, self.net = policy, policy.actor
        self.critic = self.net
        self.optim = optim or AdamOptimizerFactory()
        self.eps_clip, self.dual_clip, self.value_clip = eps_clip, dual_clip, value_clip
        self.advantage_normalization, self.recompute_adv = advantage_normalization, recompute_advantage
        self.vf_coef, self.ent_coef, self.max_grad_norm = vf_coef, ent_coef, max_grad_norm
        self.gae_lambda, self.gamma, self.max_batchsize = gae_lambda, gamma, max_batchsize
        self.minibatch_order = minibatch_order
        if grad_sync not in ("auto", "peer", "nccl"):
            raise ValueError("grad_sync must be 'auto', 'peer' or 'nccl'")
        self.grad_sync = grad_sync  # world > 1: "peer" = exchange fused into the optimizer kernel, "nccl" = all_reduce per minibatch
        self._peers = None
        self.return_scaling = bool(return_scaling)  # a2c.py:110-111, :132-146
        from ..utils.stats import RunningMeanStd

        self.ret_rms = RunningMeanStd()
        self.reuse_rollout = reuse_rollout_policy_outputs
        self.dedup_next_values = dedup_next_values  # critic(obs_next[t]) = critic(obs[t + 1]) where the rollout kernel vouches for obs_next[t] == obs[t + 1]
        self._eps = 1e-8
        self._epoch_counter = 0
        self._seed = seed
        self.training = False
        self._cfg = _lib.TsmPpoCfg(eps_clip=eps_clip, dual_clip=dual_clip or 0.0, vf_coef=vf_coef, ent_coef=ent_coef,
                                   value_clip=int(value_clip), adv_norm=int(advantage_normalization))
        self._work: dict[str, torch.Tensor] = {}
        self.last_profile: dict[str, float] = {}

    # nn.Module-like surface the reference's trainer / collector touch
    def train(self, mode: bool = True) -> "PPO":
        self.training = mode
        return self

    def eval(self) -> "PPO":
        return self.train(False)

    def __call__(self, batch: Any, state: Any = None, **kw: Any) -> Batch:
        return self.policy(batch, state, **kw)

    def state_dict(self) -> dict[str, Any]:
        """algorithm_base.py:521-541: module parameters + optimizer states under `_optimizers`."""
        n = self.net
        return {"policy.actor.flat": n.flat.detach().clone(),
                "_optimizers": [{"exp_avg": n.adam_m.clone(), "exp_avg_sq": n.adam_v.clone(), "step": n.adam_step.clone()}]}

    def load_state_dict(self, sd: dict[str, Any]) -> None:
        n = self.net
        with torch.no_grad():
            n.flat.copy_(sd["policy.actor.flat"])
            if sd.get("_optimizers"):
                o = sd["_optimizers"][0]
                n.adam_m.copy_(o["exp_avg"]); n.adam_v.copy_(o["exp_avg_sq"]); n.adam_step.copy_(o["step"])

    # -- helpers ------------------------------------------------------------------------------
    def _buf(self, name: str, shape: tuple[int, ...], dtype: torch.dtype = torch.float32) -> torch.Tensor:
        t = self._work.get(name)
        if t is None or tuple(t.shape) != tuple(shape) or t.dtype != dtype:
            t = torch.empty(shape, dtype=dtype, device=self.net.device)
            self._work[name] = t
        return t

    @staticmethod
    def split_bounds(length: int, size: int | None) -> list[int]:
        """Batch.split(size, merge_last=True) boundaries (batch.py:1209-1226)."""
        if size is None or size <= 0 or size == -1:
            size = length
        merge = length % size > 0
        b = [0]
        for idx in range(0, length, size):
            if merge and idx + size + size >= length:
                b.append(length)
                break
            b.append(min(idx + size, length))
        return b

    # -- the update ---------------------------------------------------------------------------
    def update(self, buffer: Any, batch_size: int | None, repeat: int, rows: torch.Tensor | None = None) -> A2CTrainingStats:
        """`rows` (device int64, optional) restricts the update to those sub-buffers (agent rows): the policy group of a
        FlexibleMultiAgentPolicyManager.  GAE runs per row, so restricting rows only shrinks the batch."""
        if not self.policy.is_within_training_step:
            raise RuntimeError(
                f"update() was called outside of a training step as signalled by {self.policy.is_within_training_step=} "
                "If you want to update the policy without a Trainer, you will have to manage the above-mentioned flag yourself.")
        if buffer is None:
            return A2CTrainingStats(*(SequenceSummaryStats.from_sequence([]) for _ in range(4)))
        if rows is not None and self.return_scaling:
            # the scaled GAE kernel reduces the return moments over ALL rows of the buffer; with a row subset the other rows hold
            # another group's (or no) values and would contaminate ret_rms
            raise NotImplementedError("return_scaling / return_standardization with a row subset (policy groups) is not supported")
        t0 = time.time()
        net, L, P, S = self.net, _lib.lib(), _lib.ptr, _lib.stream_ptr()
        st = buffer._store
        size, R, O = buffer._size, buffer.buffer_num, net.in_dim
        lock_len = buffer._lockstep_len
        lockstep = lock_len is not None and lock_len > 0
        obs_all = st["obs"].view(size * R, O)
        v_s = self._buf("v_s", (size, R))
        v_n = self._buf("v_n", (size, R))
        logp = self._buf("logp_old", (size, R))
        adv = self._buf("adv", (size, R))
        ret = self._buf("ret", (size, R))
        # ---- _add_returns_and_advantages (a2c.py:113-151): critic(obs), critic(obs_next) ----
        slots = None  # storage slots (t * R + row) this update touches; None = all
        if rows is not None:
            rows = rows.to(device=net.device, dtype=torch.int64).contiguous()
            slots = (torch.arange(size, device=net.device, dtype=torch.int64)[:, None] * R + rows[None, :]).reshape(-1).contiguous()
        # reuse_rollout_policy_outputs: the fused collector stored value = critic(obs) and logp = log pi(act | obs) of EVERY valid
        # slot, computed by the same forward kernel with the parameters this update starts from (buffer._policy_version), so
        # critic(obs) and logp_old need no second pass, and critic(obs_next[t]) = value[t + 1] wherever obs_next[t] is the
        # next stored obs; only rows whose episode ended at t and the last slot get a critic evaluation of obs_next.
        # the rollout kernel wrote obs_next[t] and obs[t + 1] from the same registers wherever the episode went on: critic(obs_next[t])
        # IS critic(obs[t + 1]) there; only rows that ended an episode at t, and the last slot, need an evaluation of obs_next
        dedup_ok = bool(self.dedup_next_values and lockstep and rows is None and buffer._obs_chain is True)

        def next_values(v_cur: torch.Tensor, dedup: bool) -> None:
            """critic(obs_next) -> v_n, given critic(obs) = v_cur under the same parameters (a2c.py:124-127)."""
            if dedup:
                start0 = buffer._lockstep_ins if lock_len == size else 0
                need = self._buf("need_list", (size * R,), torch.int64)
                cnt = self._buf("need_count", (1,), torch.int64)
                _lib.check(L.tsm_values_next_prepare(P(v_cur), P(st["done"]), size, R, start0, lock_len, P(v_n), P(need), P(cnt), S))
                n_need = int(cnt.item())
                if n_need > 0:
                    net.forward(st["obs_next"].view(size * R, O), mode="raw", out={"value": v_n.view(-1)}, rows=need[:n_need], scatter=True)
            else:
                net.forward(st["obs_next"].view(size * R, O), mode="raw", out={"value": v_n.view(-1)}, rows=slots, scatter=slots is not None)

        reuse = (self.reuse_rollout and lockstep and rows is None and "value" in st and "logp" in st and net.value_index >= 0
                 and getattr(self.policy, "actor", None) is net and buffer._policy_version is not None
                 and buffer._policy_version == net.param_version)
        if self._mc_returns:
            # Reinforce (reinforce.py:297-312): no critic.  The reference passes v_s_ = ret_rms.mean everywhere and v_s = roll(v_s_);
            # in the lambda = 1 recurrence the two cancel except where an episode ends (done, or the last collected slot), where
            # the return bootstraps with gamma * ret_rms.mean * (1 - terminated).  So: v_s = 0, v_n = mean at those slots only.
            reuse = False
            v_s.zero_()
            boot = float(self.ret_rms.mean)
            if self.return_scaling:  # the scaled GAE kernel multiplies the values by sqrt(var + eps) on load: pre-divide
                boot /= float(np.sqrt(self.ret_rms.var + self._eps))
            if boot == 0.0:
                v_n.zero_()
            else:
                torch.mul(st["done"].to(torch.float32), boot, out=v_n)
                v_n[((buffer._lockstep_ins if lock_len == size else 0) + lock_len - 1) % size].fill_(boot)
            logp.zero_()
        elif reuse:
            tk = _lib.t_start("critic_forward")
            v_s, logp = st["value"], st["logp"]
            next_values(v_s, True)
            _lib.t_stop(tk)
        elif net.gaussian:
            # the same single pass for a Gaussian head: the kernel yields [mu | value]; the log-density of the stored float
            # actions is elementwise (Independent(Normal(mu, sigma), 1).log_prob, ppo.py:157-161)
            if rows is not None or "act_f" not in st:
                raise NotImplementedError("Gaussian policies: whole-buffer updates over float action vectors (buffer key act_f)")
            tk = _lib.t_start("logp_old_forward")
            raw = net.forward(obs_all, mode="raw")["out"]
            v_s.view(-1).copy_(raw[:, net.value_index])
            A_ = net.n_act
            logp.view(-1).copy_(ContinuousActorPolicy.log_prob(net.gaussian_mean(raw), net.gaussian_log_sigma(raw), st["act_f"].view(-1, A_)))
            _lib.t_stop(tk)
            tk = _lib.t_start("critic_forward")
            next_values(v_s, dedup_ok)
            _lib.t_stop(tk)
        else:
            # critic(obs) (a2c.py:120-123) and logp_old = log pi(act | obs) (ppo.py:157-161) read the same rows through the
            # same trunk: ONE evaluation pass yields both (the reference runs the network twice)
            tk = _lib.t_start("logp_old_forward")
            net.forward(obs_all, mode="eval", act=st["act"].view(-1), out={"logp": logp.view(-1), "value": v_s.view(-1)},
                        rows=slots, scatter=slots is not None)
            _lib.t_stop(tk)
            tk = _lib.t_start("critic_forward")
            next_values(v_s, dedup_ok)
            _lib.t_stop(tk)
        tk = _lib.t_start("gae")
        flat_ctx: dict[str, Any] = {}

        def run_gae(v_cur: torch.Tensor) -> None:
            """compute_episodic_return over the stored rollout -> adv, ret (algorithm_base.py:651-717)."""
            if lockstep and self._mc_returns and self.return_scaling:  # return_standardization: (ret - mean) / sqrt(var + eps)
                sums_rs = self._buf("ret_sums", (2,), torch.float64)
                gae_time_major(v_cur, v_n, st["rew"], st["terminated"], st["truncated"], buffer._lockstep_ins if flat_ctx["full"] else 0,
                               flat_ctx["len"], self.gamma, 1.0, adv, ret, v_scale=float(np.sqrt(self.ret_rms.var + self._eps)),
                               ret_sums=sums_rs, ret_shift=float(self.ret_rms.mean))
                s1, s2 = (float(x) for x in sums_rs.cpu().numpy())
                n_el = flat_ctx["len"] * R
                self.ret_rms.update_from_moments(s1 / n_el, max(s2 / n_el - (s1 / n_el) ** 2, 0.0), n_el)
            elif lockstep and self.return_scaling:
                scale = float(np.sqrt(self.ret_rms.var + self._eps))
                sums_rs = self._buf("ret_sums", (2,), torch.float64)
                gae_time_major(v_cur, v_n, st["rew"], st["terminated"], st["truncated"], buffer._lockstep_ins if flat_ctx["full"] else 0,
                               flat_ctx["len"], self.gamma, self.gae_lambda, adv, ret, v_scale=scale, ret_sums=sums_rs)
                s1, s2 = (float(x) for x in sums_rs.cpu().numpy())
                n_el = flat_ctx["len"] * R
                self.ret_rms.update_from_moments(s1 / n_el, max(s2 / n_el - (s1 / n_el) ** 2, 0.0), n_el)  # statistics.py:96-114
            elif self.return_scaling:
                raise NotImplementedError("return_scaling needs a lock-step rollout buffer (time-major GAE path)")
            elif lockstep:
                gae_time_major(v_cur, v_n, st["rew"], st["terminated"], st["truncated"], buffer._lockstep_ins if flat_ctx["full"] else 0,
                               flat_ctx["len"], self.gamma, self.gae_lambda, adv, ret)
            else:
                g = lambda t: t.view(-1)[flat_ctx["slot"]]  # noqa: E731
                r_f, a_f = gae_flat(g(v_cur), g(v_n), g(st["rew"]), g(st["terminated"]), g(st["truncated"]), flat_ctx["force"], 0,
                                    self.gamma, self.gae_lambda)
                adv.view(-1)[flat_ctx["slot"]] = a_f
                ret.view(-1)[flat_ctx["slot"]] = r_f

        if lockstep:
            start = buffer._lockstep_ins if lock_len == size else 0
            B = lock_len * R
            flat_ctx.update(full=lock_len == size, len=lock_len)
            run_gae(v_s)
            idx_all = None
            if rows is not None:  # env-major buffer indices of the group's rows, in buffer.sample(0) order
                tt = (start + torch.arange(lock_len, device=net.device, dtype=torch.int64)) % size
                idx_all = (rows[:, None] * size + tt[None, :]).reshape(-1).contiguous()
                B = int(idx_all.numel())
        else:
            idx_all = buffer._all_indices_device()
            if rows is not None:
                member = torch.zeros(R, dtype=torch.bool, device=net.device)
                member[rows] = True
                idx_all = idx_all[member[idx_all // size]].contiguous()
            B = int(idx_all.numel())
            if B == 0:
                return A2CTrainingStats(*(SequenceSummaryStats.from_sequence([]) for _ in range(4)))
            slot = (idx_all % size) * R + idx_all // size
            unf = buffer._unfinished_device()
            mask = torch.zeros(buffer.maxsize, dtype=torch.uint8, device=net.device)
            mask[unf] = 1
            flat_ctx.update(slot=slot, force=mask[idx_all])
            run_gae(v_s)
            start, lock_len = 0, size
        _lib.t_stop(tk)
        # ---- logp_old (ppo.py:157-161) ----
        # ---- minibatch loop (ppo.py:173-216) ----
        bounds = self.split_bounds(B, batch_size)
        n_mb = len(bounds) - 1
        cached = self._work.get("bounds_cache")  # (the same B / batch_size every update: no host -> device copy per call)
        if cached is None or cached[0] != (B, batch_size):
            cached = ((B, batch_size), torch.tensor(bounds, dtype=torch.int64, device=net.device), (C.c_int64 * len(bounds))(*bounds))
            self._work["bounds_cache"] = cached
        bounds_dev, h_bounds = cached[1], cached[2]
        stats = self._buf("mb_stats", (repeat * n_mb, 2))
        sums = self._buf("loss_sums", (repeat * n_mb, 4), torch.float64)
        sums.zero_()
        pb = _lib.TsmPpoBatch(obs=P(st["obs"]), act=P(st["act"]), adv=P(ret if self._mc_returns else adv), ret=P(ret), logp_old=P(logp), v_old=P(v_s),
                              indices=None, size=size, n_rows=R, start=start, len=lock_len, perm_key=0, perm_n=0, records=None)
        if net.gaussian:
            pb.act, pb.act_f = None, P(st["act_f"])
        if net.engine == "tc":  # one 96-byte record per stored row: the minibatch kernels gather 3 sectors / 1 page per row
            tk = _lib.t_start("pack_records")
            rec = self._buf("records", (size * R, _lib.PPO_RECORD_BYTES), torch.uint8)

            def pack() -> None:
                if rows is not None:  # a policy group: only its sub-buffers (the rest of the array is never gathered)
                    _lib.check(L.tsm_ppo_pack_records_rows(net.desc, C.byref(pb), P(rec), P(rows), int(rows.numel()), S))
                else:
                    _lib.check(L.tsm_ppo_pack_records(net.desc, C.byref(pb), P(rec), S))

            pack()
            pb.records = P(rec)
            _lib.t_stop(tk)
            if size * R < 2**31:  # slot table of the epoch's permutation: filled by adv_stats as a by-product, read by the tiles
                pb.slots = P(self._buf("slots", (B,), torch.int32))
        world = torch.distributed.get_world_size() if torch.distributed.is_available() and torch.distributed.is_initialized() else 1
        o = self.optim
        mgn = float(self.max_grad_norm) if self.max_grad_norm else 0.0
        keep = []  # keep explicit index tensors alive until the stream has consumed them
        # Data-parallel runs: the in-kernel exchange gives up on a peer after ~2 s and finishes the epoch with whatever it has
        # (a hung device would be worse); the update is then void.  Keep what it takes to undo it (4 x 23 KB for the C2 net).
        snap = None
        if world > 1 and self.grad_sync != "nccl":
            snap = net.snapshot()
        for ep in range(repeat):
            if self.recompute_adv and ep > 0:
                # ppo.py:174-178 -> a2c.py:113-151 with the CURRENT parameters: v_s, returns and advantages (logp_old stays)
                tk = _lib.t_start("recompute_advantage")
                v_s = self._buf("v_s", (size, R))
                net.forward(obs_all, mode="raw", out={"value": v_s.view(-1)}, rows=slots, scatter=slots is not None)
                next_values(v_s, dedup_ok)
                run_gae(v_s)
                pb.v_old = P(v_s)
                if net.engine == "tc":
                    pack()
                _lib.t_stop(tk)
            if self.minibatch_order == "numpy":
                order = torch.as_tensor(np.random.permutation(B), device=net.device)  # batch.py:1219
                explicit = order if idx_all is None and start == 0 and lock_len == size else (
                    idx_all[order] if idx_all is not None else self._lockstep_indices(order, start, lock_len, size))
            elif idx_all is not None:
                gen = torch.Generator(device=net.device).manual_seed(self._seed * 1000003 + self._epoch_counter)
                explicit = idx_all[torch.randperm(B, device=net.device, generator=gen)]
            else:
                explicit = None
            if explicit is not None:
                keep.append(explicit)
                pb.indices, pb.perm_n = P(explicit), 0
            else:
                pb.indices, pb.perm_n = None, B
                pb.perm_key = (self._seed * 0x9E3779B97F4A7C15 + self._epoch_counter * 0xD1B54A32D192ED03 + 0x632BE59BD9B4E019) & (2**64 - 1)
            self._epoch_counter += 1
            st_ep, sums_ep = stats[ep * n_mb:], sums[ep * n_mb:]
            if self.advantage_normalization:
                tk = _lib.t_start("adv_stats")
                _lib.check(L.tsm_ppo_adv_stats(C.byref(pb), P(bounds_dev), n_mb, self._eps, P(st_ep), S))
                _lib.t_stop(tk)
            elif pb.slots:
                _lib.check(L.tsm_ppo_slot_table(C.byref(pb), 0, B, S))
            tk = _lib.t_start("ppo_epoch")
            if world > 1 and self.grad_sync != "nccl" and self._peers is None:
                try:
                    from .dist import PeerGradients
                    self._peers = PeerGradients(net.n_params, net.device)
                except Exception:  # no symmetric memory on this system: per-minibatch NCCL all-reduce
                    if self.grad_sync == "peer":
                        raise
                    self.grad_sync = "nccl"
            if world > 1 and self._peers is not None:
                pg = self._peers
                _lib.check(L.tsm_ppo_epoch_dist(net.desc, C.byref(pb), C.byref(self._cfg), h_bounds, n_mb,
                                                P(st_ep) if self.advantage_normalization else None, o.lr, o.betas[0], o.betas[1],
                                                o.eps, mgn, P(sums_ep), pg.byref(), pg.seq + 1, S))
                pg.seq += n_mb
            elif world == 1:
                _lib.check(L.tsm_ppo_epoch(net.desc, C.byref(pb), C.byref(self._cfg), h_bounds, n_mb,
                                           P(st_ep) if self.advantage_normalization else None, o.lr, o.betas[0], o.betas[1], o.eps,
                                           mgn, P(sums_ep), S))
            else:
                for m in range(n_mb):
                    _lib.check(L.tsm_ppo_minibatch(net.desc, C.byref(pb), C.byref(self._cfg), bounds[m], bounds[m + 1] - bounds[m],
                                                   P(st_ep[m:]) if self.advantage_normalization else None, P(sums_ep[m:]), S))
                    torch.distributed.all_reduce(net.grads)  # NCCL sum over ranks
                    _lib.check(L.tsm_adam_step(net.desc, o.lr, o.betas[0], o.betas[1], o.eps, mgn, 1.0 / world, None, S))
            _lib.t_stop(tk)
        net.kernel_updates += repeat * n_mb
        # ---- stats: one device -> host copy for the whole update (the reference does 4 .item() per minibatch) ----
        h = sums.cpu().numpy()
        if self._peers is not None:
            try:
                self._peers.check()
            except _lib.TsmError:
                if snap is not None:  # parameters and optimizer state as they were before this update; the error still reaches the caller
                    net.restore(snap)
                raise
        n_per = np.tile(np.diff(np.asarray(bounds)), repeat).astype(np.float64)
        clip_l, vf_l, ent_l = h[:, 0] / n_per, h[:, 1] / n_per, h[:, 2] / n_per
        loss = clip_l + self.vf_coef * vf_l - self.ent_coef * ent_l
        del keep
        out = A2CTrainingStats(loss=SequenceSummaryStats.from_sequence(loss), actor_loss=SequenceSummaryStats.from_sequence(clip_l),
                               vf_loss=SequenceSummaryStats.from_sequence(vf_l), ent_loss=SequenceSummaryStats.from_sequence(ent_l),
                               gradient_steps=repeat * n_mb)
        out.train_time = time.time() - t0
        self.last_sequences = {"loss": loss, "clip": clip_l, "vf": vf_l, "ent": ent_l}
        return out

    @staticmethod
    def _lockstep_indices(order: torch.Tensor, start: int, length: int, size: int) -> torch.Tensor:
        """batch position -> env-major buffer index for a lock-step ring view."""
        row, j = order // length, order % length
        return row * size + (start + j) % size


class A2C(PPO):
    """Advantage actor-critic (reference tianshou/algorithm/modelfree/a2c.py:154-288): the PPO machinery with the plain
    policy-gradient epilogue of the fused minibatch kernels -- actor_loss = -mean(log pi(a | s) * adv), vf_loss =
    mean((returns - value)^2), loss = actor_loss + vf_coef * vf_loss - ent_coef * entropy -- no ratio, no clipping, no
    advantage normalisation, no logp_old.  Same preprocessing (critic values, GAE, optional return_scaling)."""

    def __init__(self, *, policy: DiscreteActorPolicy, critic: Any = None, optim: AdamOptimizerFactory | None = None, vf_coef: float = 0.5,
                 ent_coef: float = 0.01, max_grad_norm: float | None = None, gae_lambda: float = 0.95, max_batchsize: int = 256,
                 gamma: float = 0.99, return_scaling: bool = False, **kwargs: Any) -> None:
        super().__init__(policy=policy, critic=critic, optim=optim, vf_coef=vf_coef, ent_coef=ent_coef, max_grad_norm=max_grad_norm,
                         gae_lambda=gae_lambda, max_batchsize=max_batchsize, gamma=gamma, return_scaling=return_scaling,
                         advantage_normalization=False, value_clip=False, dual_clip=None, recompute_advantage=False, **kwargs)
        self._cfg.plain_pg = 1


class Reinforce(PPO):
    """REINFORCE (reference tianshou/algorithm/modelfree/reinforce.py:249-382): Monte-Carlo discounted returns (lambda = 1, no
    critic; truncated / unfinished episodes bootstrap with ret_rms.mean, :297-304), optional return standardisation with the
    running mean / std (:305-309), loss = -mean(log pi(a | s) * return) through the plain policy-gradient epilogue of the fused
    minibatch kernels.  The net is actor-only (ActorCriticNet(..., with_value=False)).  Lock-step rollout buffers only."""

    _mc_returns_flag = True

    def __init__(self, *, policy: DiscreteActorPolicy, gamma: float = 0.99, return_standardization: bool = False,
                 optim: AdamOptimizerFactory | None = None, **kwargs: Any) -> None:
        if getattr(policy.actor, "gaussian", False):
            raise NotImplementedError("Reinforce over a Gaussian actor: not recorded against the reference (PPO and A2C are: ppo_gauss.npz)")
        super().__init__(policy=policy, critic=None, optim=optim, vf_coef=0.0, ent_coef=0.0, max_grad_norm=None, gae_lambda=1.0, gamma=gamma,
                         return_scaling=return_standardization, advantage_normalization=False, value_clip=False, dual_clip=None,
                         recompute_advantage=False, **kwargs)
        self._cfg.plain_pg = 1
