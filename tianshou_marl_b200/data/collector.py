"""Device-resident Collector: drop-in for tianshou/data/collector.py:548-1098 (`Collector`) on the
on-policy path.

    reset / reset_env / reset_buffer / reset_stat   collector.py:404-442
    collect(n_step | n_episode, random, ...)        collector.py:461-535
    _collect loop                                   collector.py:770-1098
    _compute_action_policy_hidden                   collector.py:704-767

Two code paths, both through libtsm_b200 only:
 * fused (MPEVectorEnv + DiscreteActorPolicy + lock-step VectorReplayBuffer): per vector step the
   policy kernel writes act / logp / value straight into the buffer's time-major slot, the MPE
   step kernel writes obs_next / rew / flags into the slot and the next observation into the
   NEXT slot's obs (so obs is written once, by the kernel that produces it), and one counter
   kernel updates the ring state.  No host round trip inside the loop; episode statistics are
   reduced on the device and read once per collect() call.
 * generic: any vector env with the reference protocol (reset(env_id) / step(action, id)), one
   policy call + env.step + buffer.add(Batch) per step, like the reference loop.
"""
from __future__ import annotations

import ctypes as C
import os
import time
import warnings
from typing import Any

import numpy as np
import torch

from .. import _lib
from ..env.mpe import MPEVectorEnv
from ..utils.stats import CollectStats, SequenceSummaryStats
from .batch import Batch
from .buffer import VectorReplayBuffer


class Collector:
    def __init__(self, policy: Any, env: Any, buffer: VectorReplayBuffer | None = None, exploration_noise: bool = False,
                 on_episode_done_hook: Any = None, on_step_hook: Any = None, raise_on_nan_in_buffer: bool = False,
                 collect_stats_class: type = CollectStats, episode_list: bool = True, random_seed: int = 0) -> None:
        self.algorithm = policy if hasattr(policy, "policy") else None  # accepts Policy or Algorithm (collector.py:358)
        self.policy = policy.policy if hasattr(policy, "policy") else policy
        self.env = env
        self.env_num = len(env)
        if buffer is None:
            buffer = VectorReplayBuffer(self.env_num, self.env_num, index_type="torch")
        if buffer.buffer_num < self.env_num:  # collector.py:373-376
            raise ValueError(
                f"Buffer has only {buffer.buffer_num} buffers, but at least {self.env_num=} are needed.")
        self.buffer = buffer
        self.exploration_noise = exploration_noise
        # hooks (collector.py:583-689): accepted; a collector with hooks runs the reference's per-step host loop (_reference_loop)
        # instead of the fused device loop, because the hooks are Python callables that see every step / episode batch
        self.on_episode_done_hook, self.on_step_hook = on_episode_done_hook, on_step_hook
        self.raise_on_nan_in_buffer = raise_on_nan_in_buffer
        self.collect_stats_class = collect_stats_class
        self.episode_list = episode_list
        self.collect_step, self.collect_episode, self.collect_time = 0, 0, 0.0
        self._pre_collect_obs: Any = None
        self._random_seed, self._random_counter = int(random_seed), 0
        self._is_closed = False
        dev = getattr(env, "device", getattr(buffer, "device", torch.device("cuda")))
        self.device = dev
        R = self.env_num
        self._ep_ret = torch.zeros(R, dtype=torch.float64, device=dev)
        self._ep_len = torch.zeros(R, dtype=torch.int64, device=dev)
        self._acc = torch.zeros(9, dtype=torch.float64, device=dev)
        self._list_cap = 0
        self._list_ret = self._list_len = None
        self._list_count = torch.zeros(1, dtype=torch.int64, device=dev)
        self._carry: torch.Tensor | None = None
        self._pinned: dict[str, torch.Tensor] = {}

    # -- reference surface ------------------------------------------------------------------
    def close(self) -> None:
        self.env.close()
        self._pre_collect_obs = None
        self._is_closed = True

    def reset(self, reset_buffer: bool = True, reset_stats: bool = True, gym_reset_kwargs: dict | None = None) -> tuple[Any, Any]:
        obs, info = self.reset_env(gym_reset_kwargs)
        if reset_buffer:
            self.reset_buffer()
        if reset_stats:
            self.reset_stat()
        self._is_closed = False
        return obs, info

    def reset_stat(self) -> None:
        self.collect_step, self.collect_episode, self.collect_time = 0, 0, 0.0

    def reset_buffer(self, keep_statistics: bool = False) -> None:
        self.buffer.reset(keep_statistics=keep_statistics)

    def reset_env(self, gym_reset_kwargs: dict | None = None) -> tuple[Any, Any]:
        obs, info = self.env.reset(**(gym_reset_kwargs or {}))
        self._pre_collect_obs = obs
        self._carry = None
        b = self.buffer
        if b is not None and getattr(b, "_lockstep_len", None):
            # the buffer holds a rollout whose last obs_next is NOT the next stored obs any more: PPO.update must evaluate
            # critic(obs_next) itself at that seam (dedup_next_values / reuse_rollout_policy_outputs are off for this buffer)
            b._obs_chain = False
            b._policy_version = None
        return obs, info

    # -- collect ------------------------------------------------------------------------------
    def collect(self, n_step: int | None = None, n_episode: int | None = None, random: bool = False, render: float | None = None,
                reset_before_collect: bool = False, gym_reset_kwargs: dict | None = None) -> CollectStats:
        # collector.py:526-535
        if n_step is not None and n_episode is not None or (n_step is None and n_episode is None):
            raise ValueError(f"Only one of n_step and n_episode should be set, but got {n_step=} and {n_episode=}.")
        if n_step is not None and n_step < 1 or n_episode is not None and n_episode < 1:
            raise ValueError("n_step / n_episode must be positive")
        if reset_before_collect:
            self.reset(reset_buffer=False, gym_reset_kwargs=gym_reset_kwargs)
        if self._pre_collect_obs is None:  # collector.py:827-831
            raise ValueError("Initial obs and info should not be None. Either reset the collector (using reset or reset_env) "
                             "or pass reset_before_collect=True to collect.")
        t0 = time.time()
        was_training = getattr(self.policy, "training", False)
        if hasattr(self.policy, "train"):
            self.policy.train(False)  # torch_train_mode(policy, False), collector.py:506
        try:
            stats = self._collect(n_step, n_episode, random)
        finally:
            if hasattr(self.policy, "train"):
                self.policy.train(was_training)
        stats.set_collect_time(max(time.time() - t0, 1e-9))
        # returns_stat / lens_stat come from the device accumulators of tsm_episode_stats (count, sum, sum of squares, min,
        # max in float64) rather than from a numpy pass over up to millions of episodes
        summ = self.last_episode_summary
        if summ["count"] > 0:
            n = float(summ["count"])
            for key, pre in (("returns_stat", "ret"), ("lens_stat", "len")):
                mean = summ[pre + "_sum"] / n
                var = max(summ[pre + "_sqsum"] / n - mean * mean, 0.0)
                setattr(stats, key, SequenceSummaryStats(mean=float(mean), std=float(var ** 0.5), max=float(summ[pre + "_max"]),
                                                         min=float(summ[pre + "_min"])))
        self.collect_step += stats.n_collected_steps
        self.collect_episode += stats.n_collected_episodes
        self.collect_time += stats.collect_time
        return stats

    def _fused_ok(self, random: bool) -> bool:
        from ..algorithm.ppo import DiscreteActorPolicy

        b = self.buffer
        if getattr(self.env, "continuous_actions", False) or getattr(getattr(self.policy, "actor", None), "gaussian", False):
            return False  # Box actions: the generic step (float action vectors, Policy.map_action); the fused kernels sample Categorical
        if not random and self._manager_groups() is not None:
            return isinstance(self.env, MPEVectorEnv) and b._lockstep_len is not None and b.buffer_num == self.env_num and b._save_obs_next
        return (isinstance(self.env, MPEVectorEnv) and (random or isinstance(self.policy, DiscreteActorPolicy))
                and b._lockstep_len is not None and b.buffer_num == self.env_num and b._save_obs_next)

    def _manager_groups(self) -> list | None:
        """[(DiscreteActorPolicy, rows)] when the policy is a FlexibleMultiAgentPolicyManager over fused-capable policies."""
        from ..algorithm.multiagent import FlexibleMultiAgentPolicyManager
        from ..algorithm.ppo import DiscreteActorPolicy

        man = self.policy
        if not isinstance(man, FlexibleMultiAgentPolicyManager):
            return None
        out = []
        for p, rows in man.groups(self.env_num, self.device):
            pol = man._actor_policy(p)
            if not isinstance(pol, DiscreteActorPolicy) or getattr(pol.actor, "engine", None) == "tcw":
                return None
            if rows is None:
                rows = torch.arange(self.env_num, device=self.device, dtype=torch.int64)
            out.append((pol, rows))
        return out

    def _collect(self, n_step: int | None, n_episode: int | None, random: bool) -> CollectStats:
        R = self.env_num
        dev = self.device
        if getattr(self, "_acc_init", None) is None:  # (a device copy instead of a host tensor + pageable H2D on every call)
            self._acc_init = torch.tensor([0, 0, 0, np.inf, -np.inf, 0, 0, np.inf, -np.inf], dtype=torch.float64, device=dev)
        self._acc.copy_(self._acc_init)
        self._list_count.zero_()
        n_vec = None if n_step is None else -(-n_step // R)  # ceil(n_step / R) vector steps (collector.py:917,1066)
        if self.episode_list:
            horizon = n_vec if n_vec is not None else 4 * getattr(self.env, "max_cycles", 25)
            cap = R * max(1, horizon // max(1, getattr(self.env, "max_cycles", 25)) + 1)
            if cap > self._list_cap:
                self._list_ret = torch.zeros(cap, dtype=torch.float64, device=dev)
                self._list_len = torch.zeros(cap, dtype=torch.int64, device=dev)
                self._list_cap = cap
        fused = self._fused_ok(random)
        hooks = self.on_episode_done_hook is not None or self.on_step_hook is not None
        if hooks or not isinstance(self.env, MPEVectorEnv):
            # any other vector env, or Python hooks: the reference's loop itself (ready-env subset, surplus-env rule, episode
            # extraction, hooks, end-of-collect reset), one host iteration per vector step
            return self._reference_loop(n_step, n_episode, random)
        steps = 0
        if fused and n_vec is not None and not os.environ.get("TSM_COLLECT_PY"):
            self._fused_run(n_vec, random)  # the whole loop in one C call (csrc/collect.cu)
            steps = n_vec
        while steps == 0 or n_vec is None or steps < n_vec:
            if fused and (random or hasattr(self.policy, "actor")):
                self._fused_step(random, last=(n_vec is not None and steps == n_vec - 1) or n_vec is None)
            else:
                self._generic_step(random)
            steps += 1
            if n_vec is None and int(self._acc[0].item()) >= n_episode:  # one small D2H per step only in n_episode mode
                break
        acc = self._acc.cpu().numpy()
        n_ep = int(acc[0])
        st = self.collect_stats_class(n_collected_episodes=n_ep, n_collected_steps=steps * R)
        if self.episode_list and n_ep > 0:
            k = min(n_ep, self._list_cap)
            st.returns = self._to_host("ret", self._list_ret[:k])
            st.lens = self._to_host("len", self._list_len[:k])
            torch.cuda.current_stream().synchronize()
        self.last_episode_summary = {"count": n_ep, "ret_sum": acc[1], "ret_sqsum": acc[2], "ret_min": acc[3], "ret_max": acc[4],
                                     "len_sum": acc[5], "len_sqsum": acc[6], "len_min": acc[7], "len_max": acc[8]}
        return st

    def _to_host(self, name: str, t: torch.Tensor) -> np.ndarray:
        """Asynchronous device -> pinned host copy (caller synchronises); the array aliases a pinned buffer that the next
        collect() call with episode_list=True overwrites."""
        n = t.numel()
        h = self._pinned.get(name)
        if h is None or h.numel() < n or h.dtype != t.dtype:
            # sized for the whole episode list: the number of episodes differs from call to call, and a pinned allocation
            # costs milliseconds
            h = torch.empty(max(n, self._list_cap, 1), dtype=t.dtype, pin_memory=True)
            self._pinned[name] = h
        h[:n].copy_(t, non_blocking=True)
        return h[:n].numpy()

    # -- n vector steps, fused, one C call ------------------------------------------------------
    def _fused_run(self, n_vec: int, random: bool) -> None:
        env, buf, L, P, S = self.env, self.buffer, _lib.lib(), _lib.ptr, _lib.stream_ptr()
        R, O = self.env_num, env.obs_dim
        if "obs" not in buf._store:
            buf._alloc((O,), want_policy=True)
        st = buf._store
        size = buf._size
        slot = buf._lockstep_ins
        obs_slot = st["obs"][slot]
        src = self._pre_collect_obs
        if not (isinstance(src, torch.Tensor) and src.data_ptr() == obs_slot.data_ptr()):
            obs_slot.copy_(torch.as_tensor(src, device=self.device).reshape(R, O))
        if self._carry is None:
            self._carry = torch.empty(R, O, dtype=torch.float32, device=self.device)
        ro = _lib.TsmRollout(obs=P(st["obs"]), act=P(st["act"]), rew=P(st["rew"]), terminated=P(st["terminated"]),
                             truncated=P(st["truncated"]), obs_next=P(st["obs_next"]), logp=P(st.get("logp")), value=P(st.get("value")),
                             carry=P(self._carry), size=size, n_rows=R, obs_dim=O, slot=slot, ep_ret=P(self._ep_ret), ep_len=P(self._ep_len),
                             acc=P(self._acc), list_ret=P(self._list_ret) if self.episode_list else None,
                             list_len=P(self._list_len) if self.episode_list else None,
                             list_count=P(self._list_count) if self.episode_list else None, list_cap=self._list_cap)
        # which parameters produced the stored logp / value (PPO's reuse_rollout_policy_outputs checks this)
        single = hasattr(self.policy, "actor")
        ver = None if random or not single or st.get("value") is None or self.policy.actor.value_index < 0 else self.policy.actor.param_version
        buf._policy_version = ver if (buf._lockstep_len == 0 or buf._policy_version == ver) else None
        buf._obs_chain = bool(buf._lockstep_len == 0 or buf._obs_chain)  # the kernel writes obs_next[t] and obs[t + 1] from the same registers
        slot_o, len_o, carry_o = C.c_int64(0), C.c_int64(0), C.c_int(0)
        tk = _lib.t_start("collect")
        if random:
            _lib.check(L.tsm_collect_lockstep(None, C.byref(env._desc), C.byref(buf._desc), C.byref(ro), n_vec, 0, self._random_seed,
                                              self._random_counter, buf._lockstep_len, int(env.single_action_space.n),
                                              C.byref(slot_o), C.byref(len_o), C.byref(carry_o), S))
            self._random_counter += n_vec
        elif self._manager_groups() is not None:
            # grouped policies: one fused forward launch per policy group and step, no Python per step (csrc/collect.cu)
            man, groups = self.policy, self._manager_groups()
            G = len(groups)
            det = all(p.deterministic_eval and not (p.is_within_training_step or man.is_within_training_step) for p, _ in groups)
            nets = (C.c_void_p * G)(*[C.addressof(p.actor._desc) for p, _ in groups])
            for p, _ in groups:
                p.actor.desc  # refresh (packs the fp16 image when the parameters changed)
            rows_a = (C.c_void_p * G)(*[r.data_ptr() for _, r in groups])
            n_rows = (C.c_int64 * G)(*[int(r.numel()) for _, r in groups])
            seeds = (C.c_uint64 * G)(*[int(p.seed) for p, _ in groups])
            ctrs = (C.c_uint64 * G)(*[int(p._counter) for p, _ in groups])
            buf._policy_version = None
            _lib.check(L.tsm_collect_lockstep_groups(nets, rows_a, n_rows, seeds, ctrs, G, C.byref(env._desc), C.byref(buf._desc), C.byref(ro), n_vec,
                                                     2 if det else 1, buf._lockstep_len, C.byref(slot_o), C.byref(len_o), C.byref(carry_o), S))
            for p, _ in groups:
                p._counter += n_vec
        else:
            pol = self.policy
            mode = 2 if (pol.deterministic_eval and not pol.is_within_training_step) else 1
            if pol.actor.engine == "tcw":
                pol.actor.ensure_scratch(R)
            _lib.check(L.tsm_collect_lockstep(pol.actor.desc, C.byref(env._desc), C.byref(buf._desc), C.byref(ro), n_vec, mode, pol.seed,
                                              pol._counter, buf._lockstep_len, 0, C.byref(slot_o), C.byref(len_o), C.byref(carry_o), S))
            pol._counter += n_vec
        _lib.t_stop(tk)
        buf._lockstep_ins, buf._lockstep_len = int(slot_o.value), int(len_o.value)
        self._pre_collect_obs = self._carry if carry_o.value else st["obs"][buf._lockstep_ins]

    # -- one vector step, fused ---------------------------------------------------------------
    def _fused_step(self, random: bool, last: bool) -> None:
        env, buf, L, P, S = self.env, self.buffer, _lib.lib(), _lib.ptr, _lib.stream_ptr()
        R, O = self.env_num, env.obs_dim
        if "obs" not in buf._store:
            buf._alloc((O,), want_policy=True)
        st = buf._store
        size = buf._size
        slot = buf._lockstep_ins
        buf._policy_version = None  # (only the fused multi-step loop vouches for the stored policy outputs)
        buf._obs_chain = bool(buf._lockstep_len == 0 or buf._obs_chain)
        nxt = (slot + 1) % size
        obs_slot = st["obs"][slot]
        # the observation to act on: already in the slot if the previous fused step put it there
        src = self._pre_collect_obs
        if not (isinstance(src, torch.Tensor) and src.data_ptr() == obs_slot.data_ptr()):
            obs_slot.copy_(torch.as_tensor(src, device=self.device).reshape(R, O))
        act_slot = st["act"][slot]
        if random:
            _lib.check(L.tsm_random_actions(P(act_slot), R, env.single_action_space.n, self._random_seed, self._random_counter, S))
            self._random_counter += 1
            if "logp" in st:
                st["logp"][slot].zero_()
                st["value"][slot].zero_()
        else:
            pol = self.policy
            mode = "mode" if (pol.deterministic_eval and not pol.is_within_training_step) else "sample"
            out = {"act": act_slot, "logp": st["logp"][slot]} if "logp" in st else {"act": act_slot}
            if "value" in st and pol.actor.value_index >= 0:
                out["value"] = st["value"][slot]
            tk = _lib.t_start("policy_forward")
            pol.actor.forward(obs_slot, mode=mode, seed=pol.seed, counter=pol._counter, out=out)
            _lib.t_stop(tk)
            pol._counter += 1
        # world step: obs_next / rew / flags into this slot; the observation to act on next goes to the
        # next slot's obs unless this is the last step of the call or the ring would wrap onto live data
        if last or buf._lockstep_len + 1 >= size:
            if self._carry is None:
                self._carry = torch.empty(R, O, dtype=torch.float32, device=self.device)
            carry = self._carry
        else:
            carry = st["obs"][nxt]
        tk = _lib.t_start("mpe_step")
        _lib.check(L.tsm_mpe_step(C.byref(env._desc), P(act_slot), P(st["obs_next"][slot]), P(st["rew"][slot]),
                                  P(st["terminated"][slot]), P(st["truncated"][slot]), 1, P(carry), S))
        _lib.t_stop(tk)
        tk = _lib.t_start("buf_counters")
        # ring counters + done flag + episode return/length (payload pointers alias the slot: self-copy)
        _lib.check(L.tsm_buf_add(C.byref(buf._desc), None, P(act_slot), P(st["rew"][slot]), P(st["terminated"][slot]),
                                 P(st["truncated"][slot]), None, None, None, None, R, None, P(self._ep_ret), P(self._ep_len), None, S))
        buf._lockstep_len = min(buf._lockstep_len + 1, size)
        buf._lockstep_ins = nxt
        _lib.check(L.tsm_episode_stats(P(self._ep_ret), P(self._ep_len), R, P(self._acc),
                                       P(self._list_ret) if self.episode_list else None, P(self._list_len) if self.episode_list else None,
                                       P(self._list_count) if self.episode_list else None, self._list_cap, S))
        _lib.t_stop(tk)
        self._pre_collect_obs = carry

    # -- the reference's loop (collector.py:770-1098), step numbers as in its docstring -------------------------------------
    def run_on_episode_done(self, episode_batch: Any) -> dict | None:
        """collector.py:663-675."""
        return self.on_episode_done_hook(episode_batch) if self.on_episode_done_hook is not None else None

    def run_on_step_hook(self, action_batch: Any, rollout_batch: Any) -> None:
        """collector.py:677-689."""
        if self.on_step_hook is not None:
            self.on_step_hook(action_batch, rollout_batch)

    def _reference_loop(self, n_step: int | None, n_episode: int | None, random: bool) -> CollectStats:
        env, buf, N = self.env, self.buffer, self.env_num
        if getattr(env, "is_async", False):  # collector.py:804-809
            raise ValueError(f"Please use AsyncCollector for asynchronous environments. Env class: {env.__class__.__name__}.")
        whole = isinstance(env, MPEVectorEnv)  # a batched world steps all rows at once
        if n_step is not None:
            ready = np.arange(N)
        else:
            if N > n_episode:  # collector.py:816-822 (a warning, not an error)
                warnings.warn(f"Number of episodes ({n_episode}) is smaller than the number of environments ({N}). This means that "
                              f"{N - n_episode} environments (or, equivalently, parallel workers) will not be used!")
            ready = np.arange(min(N, n_episode))
        if whole and len(ready) != N:
            raise NotImplementedError("MPEVectorEnv steps whole worlds: n_episode must be >= the number of rows")
        to_host = lambda x: x.detach().cpu().numpy() if isinstance(x, torch.Tensor) else np.asarray(x)  # noqa: E731
        take = lambda x, ids: x[torch.as_tensor(ids, device=x.device)] if isinstance(x, torch.Tensor) else np.asarray(x)[ids]  # noqa: E731
        step_count, n_ep = 0, 0
        returns: list[float] = []
        lens: list[int] = []
        last_obs = take(self._pre_collect_obs, ready)
        while True:
            # Step 2: actions for the R ready envs (collector.py:704-767)
            if random:
                act = np.array([env.action_space[int(i)].sample() for i in ready])
                pol_out = Batch(act=act)
            else:
                pol_out = self.policy(Batch(obs=last_obs, info=Batch()), None)
                act = pol_out.act
            act_env = self.policy.map_action(act) if hasattr(self.policy, "map_action") else act
            # Step 3
            if whole:
                obs_next, rew, term, trunc, info = env.step(act_env)
            else:
                obs_next, rew, term, trunc, info = env.step(to_host(act_env), ready)
            done = to_host(term).astype(bool) | to_host(trunc).astype(bool)
            step_b = Batch(obs=last_obs, act=act, rew=rew, terminated=term, truncated=trunc, obs_next=obs_next)
            step_b.__dict__["done"] = done
            # Step 4
            self.run_on_step_hook(pol_out, step_b)
            # Step 5
            n_done = int(done.sum())
            n_ep += n_done
            step_count += len(ready)
            # Step 6
            add_b = Batch(obs=last_obs, act=act, rew=rew, terminated=term, truncated=trunc, obs_next=obs_next)
            if "logp" in pol_out and "value" in pol_out:
                add_b.__dict__["policy"] = Batch(logp=pol_out.logp, value=pol_out.value)
            idx, ep_ret, ep_len, ep_start = buf.add(add_b, buffer_ids=None if (len(ready) == buf.buffer_num) else ready)
            last_obs = obs_next.clone() if isinstance(obs_next, torch.Tensor) else np.array(obs_next, copy=True)
            if n_done > 0:
                # Steps 7-8: the D envs that ended an episode, reset them
                d_local = np.where(done)[0]
                idx_h, ret_h, len_h, start_h = to_host(idx), to_host(ep_ret), to_host(ep_len), to_host(ep_start)
                lens.extend(int(x) for x in len_h[d_local])
                returns.extend(float(x) for x in ret_h[d_local])
                d_global = ready[d_local]
                obs_reset, _ = env.reset(torch.as_tensor(d_global, device=self.device) if whole else d_global)
                # Steps 9-11: the episode batch of every finished episode, hook, additions written back (n_episode only)
                if self.on_episode_done_hook is not None:
                    for li in d_local:
                        ep_idx = buf.get_buffer_indices(int(start_h[li]), int(idx_h[li]) + 1)
                        ep_batch = buf[np.asarray(ep_idx)]
                        add = self.run_on_episode_done(ep_batch)
                        if add is not None:
                            if n_episode is None:
                                raise ValueError(
                                    "An on_episode_done_hook with non-empty returns is not supported for n_step collection."
                                    "Such hooks should only be used when collecting full episodes. Got a on_episode_done_hook "
                                    f"that would add the following fields to the buffer: {list(add)}.")
                            for key, arr in add.items():
                                buf.set_array_at_key(arr, key, index=np.asarray(ep_idx))
                # Step 12
                if isinstance(last_obs, torch.Tensor):
                    last_obs[torch.as_tensor(d_local, device=last_obs.device)] = torch.as_tensor(obs_reset, device=last_obs.device)
                else:
                    last_obs[d_local] = to_host(obs_reset)
                # Step 13: surplus envs leave the ready set (only envs that just finished an episode can be removed)
                if n_episode:
                    surplus = len(ready) - (n_episode - n_ep)
                    if surplus > 0:
                        if whole:
                            raise NotImplementedError("MPEVectorEnv steps whole worlds: rows cannot leave the ready set")
                        keep = np.ones(len(ready), dtype=bool)
                        keep[d_local[:surplus]] = False
                        ready = ready[keep]
                        last_obs = take(last_obs, np.where(keep)[0])
            if (n_step and step_count >= n_step) or (n_episode and n_ep >= n_episode):
                break
        st = self.collect_stats_class(n_collected_episodes=n_ep, n_collected_steps=step_count)
        if n_ep > 0:
            st.returns, st.lens = np.asarray(returns, dtype=float), np.asarray(lens, dtype=int)
        r, l = np.asarray(returns, dtype=float), np.asarray(lens, dtype=float)
        self.last_episode_summary = {"count": n_ep, "ret_sum": r.sum(), "ret_sqsum": (r * r).sum(), "ret_min": r.min() if n_ep else np.inf,
                                     "ret_max": r.max() if n_ep else -np.inf, "len_sum": l.sum(), "len_sqsum": (l * l).sum(),
                                     "len_min": l.min() if n_ep else np.inf, "len_max": l.max() if n_ep else -np.inf}
        # Step 15
        if n_step:
            self._pre_collect_obs = last_obs
        else:
            self.reset_env()
        return st

    # -- one vector step, generic (reference loop) ---------------------------------------------
    def _generic_step(self, random: bool) -> None:
        env, buf = self.env, self.buffer
        obs = self._pre_collect_obs
        R = self.env_num
        if random:
            act = np.array([env.action_space[i].sample() for i in range(R)])
            pol_out = None
        else:
            pol_out = self.policy(Batch(obs=obs, info=Batch()), None)
            act = pol_out.act
        act_env = self.policy.map_action(act) if hasattr(self.policy, "map_action") else act
        if isinstance(env, MPEVectorEnv):
            obs_next, rew, term, trunc, info = env.step(act_env)
        else:
            a = act_env.cpu().numpy() if isinstance(act_env, torch.Tensor) else act_env
            obs_next, rew, term, trunc, info = env.step(a, np.arange(R))
        b = Batch(obs=obs, act=act, rew=rew, terminated=term, truncated=trunc, obs_next=obs_next)
        if pol_out is not None and "logp" in pol_out and "value" in pol_out:
            b.__dict__["policy"] = Batch(logp=pol_out.logp, value=pol_out.value)
        _, ep_ret, ep_len, _ = buf.add(b, buffer_ids=None if buf.buffer_num == R else np.arange(R))
        er = torch.as_tensor(ep_ret, device=self.device).to(torch.float64).contiguous()
        el = torch.as_tensor(ep_len, device=self.device).to(torch.int64).contiguous()
        P = _lib.ptr
        _lib.check(_lib.lib().tsm_episode_stats(P(er), P(el), R, P(self._acc), P(self._list_ret) if self.episode_list else None,
                                                P(self._list_len) if self.episode_list else None,
                                                P(self._list_count) if self.episode_list else None, self._list_cap, _lib.stream_ptr()))
        done = torch.as_tensor(np.asarray(term) | np.asarray(trunc)) if not isinstance(term, torch.Tensor) else (term | trunc)
        if bool(done.any()):
            ids = torch.nonzero(done).reshape(-1)
            ids_h = ids.cpu().numpy()
            obs_reset, _ = env.reset(ids if isinstance(env, MPEVectorEnv) else ids_h)
            if isinstance(obs_next, torch.Tensor):
                obs_next = obs_next.clone()
                obs_next[ids.to(obs_next.device)] = torch.as_tensor(obs_reset, device=obs_next.device)
            else:
                obs_next = np.array(obs_next, copy=True)
                obs_next[ids_h] = obs_reset
        self._pre_collect_obs = obs_next


class EpisodeRolloutHookMCReturn:
    """collector.py:1480-1509: adds the discounted Monte-Carlo return to go and the full-episode return to every transition of
    a finished episode (an on_episode_done_hook; n_episode collection).  The return to go is K3 with lambda = 1 and zero values
    (algorithm.returns.episode_mc_return_to_go -> csrc/gae.cu)."""

    MC_RETURN_TO_GO_KEY = "mc_return_to_go"
    FULL_EPISODE_MC_RETURN_KEY = "full_episode_mc_return"

    def __init__(self, gamma: float = 0.99) -> None:
        if not 0 <= gamma <= 1:
            raise ValueError(f"Expected 0 <= gamma <= 1, but got {gamma=}.")
        self.gamma = gamma

    def __call__(self, episode_batch: Any) -> dict[str, np.ndarray]:
        from ..algorithm.returns import episode_mc_return_to_go

        rew = episode_batch.rew
        rew = rew.detach().cpu().numpy() if isinstance(rew, torch.Tensor) else np.asarray(rew)
        to_go = np.asarray(episode_mc_return_to_go(rew.astype(np.float64), self.gamma), dtype=np.float64)
        return {self.MC_RETURN_TO_GO_KEY: to_go, self.FULL_EPISODE_MC_RETURN_KEY: np.full_like(to_go, to_go[0])}
