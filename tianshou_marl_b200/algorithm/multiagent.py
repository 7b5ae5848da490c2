"""Multi-agent policy routing on the device.

    FlexibleMultiAgentPolicyManager     tianshou/algorithm/multiagent/flexible_policy.py:12-231
    MultiAgentPolicy.forward            tianshou/algorithm/multiagent/marl.py:108-185

The reference routes rows to policies by comparing `obs.agent_id` strings per agent in a Python loop.
With the row flattening of this package (row r = env * A + agent) the agent of a row is
`r % A`, so every policy owns a fixed, strided set of rows: routing is a precomputed index list per
unique policy and one fused forward launch per policy *group* (tsm_net_forward with a row list and
scatter), each writing its rows of the shared act / logp / value outputs.  Modes, validation rules
and error messages follow the reference class; `update` runs each group's PPO over its own rows."""
from __future__ import annotations

from typing import Any, Callable

import numpy as np
import torch

from ..data.batch import Batch

MODES = ("independent", "shared", "grouped", "custom")


class FlexibleMultiAgentPolicyManager:
    def __init__(self, policies: Any, env: Any, mode: str = "independent", agent_groups: dict[str, list[str]] | None = None,
                 policy_mapping_fn: Callable[[str], str] | None = None, **kwargs: Any) -> None:
        self.mode = mode
        self.agent_groups = agent_groups or {}
        self.policy_mapping_fn = policy_mapping_fn
        self.env = env
        self.agents = list(env.agents)
        self._validate_config(policies)
        self.policy_map = self._build_policy_map(policies, self.agents)
        self._original_policies = policies
        self.policy_mapping = self.policy_map
        if self.mode == "shared":
            self.policies = {"shared": next(iter(self.policy_map.values()))}
        elif self.mode == "grouped":
            self.policies = {g: self.policy_map[self.agent_groups[g][0]] for g in self.agent_groups}
        else:
            unique: dict[str, Any] = {}
            for agent, pol in self.policy_map.items():
                if not any(pol is q for q in unique.values()):
                    unique[agent] = pol
            self.policies = unique
        self.agent_idx = {a: i for i, a in enumerate(self.agents)}
        self.is_within_training_step = False
        self._rows_cache: dict[tuple[int, int], list[tuple[Any, torch.Tensor | None]]] = {}

    # -- configuration (flexible_policy.py:108-186) ---------------------------------------------
    @staticmethod
    def _is_single(p: Any) -> bool:
        return not isinstance(p, (list, tuple, dict))

    def _validate_config(self, policies: Any) -> None:
        if self.mode not in MODES:
            raise ValueError(f"Unknown mode: {self.mode}")
        if self.mode == "independent" and self._is_single(policies):
            raise ValueError("Independent mode requires list or dict of policies, got single policy")
        if self.mode == "grouped":
            if not self.agent_groups:
                raise ValueError("Grouped mode requires agent_groups to be specified")
            if self._is_single(policies):
                raise ValueError("Grouped mode requires dict of policies mapped to group names")
        if self.mode == "custom" and not self.policy_mapping_fn:
            raise ValueError("Custom mode requires policy_mapping_fn to be specified")

    def _build_policy_map(self, policies: Any, agents: list[str]) -> dict[str, Any]:
        if self.mode == "shared":
            if self._is_single(policies):
                shared = policies
            elif isinstance(policies, (list, tuple)):
                shared = policies[0]
            else:
                shared = next(iter(policies.values()))
            return {a: shared for a in agents}
        if self.mode == "grouped":
            if not isinstance(policies, dict):
                raise ValueError("Grouped mode requires dict of policies")
            pm: dict[str, Any] = {}
            for g, members in self.agent_groups.items():
                if g not in policies:
                    raise ValueError(f"No policy found for group {g}")
                for a in members:
                    if a in agents:
                        pm[a] = policies[g]
            unassigned = set(agents) - set(pm)
            if unassigned:
                raise ValueError(f"Agents {unassigned} are not assigned to any group")
            return pm
        if self.mode == "custom":
            if not isinstance(policies, dict):
                raise ValueError("Custom mode requires dict of policies")
            pm = {}
            for a in agents:
                pid = self.policy_mapping_fn(a)
                if pid not in policies:
                    raise ValueError(f"Policy {pid} not found for agent {a}")
                pm[a] = policies[pid]
            return pm
        if isinstance(policies, (list, tuple)):
            if len(policies) != len(agents):
                raise ValueError(f"Number of policies ({len(policies)}) must match number of agents ({len(agents)})")
            return dict(zip(agents, policies))
        if isinstance(policies, dict):
            missing = set(agents) - set(policies)
            if missing:
                raise ValueError(f"Missing policies for agents: {missing}")
            return policies
        raise ValueError("Independent mode requires list or dict of policies")

    def get_shared_parameters(self) -> bool:
        return len({id(p) for p in self.policy_map.values()}) < len(self.policy_map)

    def get_policy_groups(self) -> dict[str, list[str]]:
        by_pol: dict[int, list[str]] = {}
        for a, p in self.policy_map.items():
            by_pol.setdefault(id(p), []).append(a)
        return {(f"group_{i}" if len(by_pol) > 1 else "shared"): ag for i, ag in enumerate(by_pol.values())}

    # -- static routing ---------------------------------------------------------------------------
    @staticmethod
    def _actor_policy(p: Any) -> Any:
        """The object holding `.actor` (a PPO algorithm wraps its policy)."""
        return p.policy if hasattr(p, "policy") and hasattr(p.policy, "actor") else p

    def groups(self, n_rows: int, device: torch.device) -> list[tuple[Any, torch.Tensor | None]]:
        """[(policy, rows)] per unique policy; rows is None when one policy owns every row."""
        A = len(self.agents)
        key = (n_rows, A)
        if key not in self._rows_cache:
            uniq: list[Any] = []
            members: list[list[int]] = []
            for i, a in enumerate(self.agents):
                p = self.policy_map[a]
                for k, q in enumerate(uniq):
                    if q is p:
                        members[k].append(i)
                        break
                else:
                    uniq.append(p)
                    members.append([i])
            out: list[tuple[Any, torch.Tensor | None]] = []
            for p, m in zip(uniq, members):
                if len(m) == A:
                    out.append((p, None))
                else:
                    r = torch.arange(n_rows, device=device, dtype=torch.int64)
                    sel = torch.zeros(A, dtype=torch.bool, device=device)
                    sel[torch.tensor(m, device=device)] = True
                    out.append((p, r[sel[r % A]].contiguous()))
            self._rows_cache[key] = out
        return self._rows_cache[key]

    # -- forward (flexible_policy.py:189-231, marl.py:108-185) -----------------------------------
    def forward(self, batch: Any, state: Any = None, **kwargs: Any) -> Batch:
        obs = batch.obs if hasattr(batch, "obs") else batch["obs"]
        if isinstance(obs, Batch):
            obs = obs["obs"]
        first = self._actor_policy(next(iter(self.policy_map.values())))
        dev = first.actor.device
        if not (isinstance(obs, torch.Tensor) and obs.is_cuda):
            obs = torch.as_tensor(np.ascontiguousarray(obs), dtype=torch.float32, device=dev)
        obs = obs.to(torch.float32).contiguous()
        n = obs.shape[0]
        if n % len(self.agents):
            raise ValueError("the batch must hold whole environments: rows are env * n_agents + agent")
        act = torch.empty(n, dtype=torch.int32, device=dev)
        logp = torch.empty(n, dtype=torch.float32, device=dev)
        value = torch.zeros(n, dtype=torch.float32, device=dev)
        per_policy: dict[int, Batch] = {}
        for p, rows in self.groups(n, dev):
            pol = self._actor_policy(p)
            mode = "mode" if (pol.deterministic_eval and not (pol.is_within_training_step or self.is_within_training_step)) else "sample"
            out = {"act": act, "logp": logp}
            if pol.actor.value_index >= 0:
                out["value"] = value
            pol.actor.forward(obs, mode=mode, seed=pol.seed, counter=pol._counter, rows=rows, scatter=rows is not None, out=out)
            pol._counter += 1
            b = Batch()
            b.__dict__.update(act=act if rows is None else act[rows], logp=logp if rows is None else logp[rows], state=None)
            per_policy[id(p)] = b
        holder = Batch()
        holder.__dict__["act"] = act.to(torch.int64)
        holder.__dict__["logp"] = logp
        holder.__dict__["value"] = value
        holder.__dict__["out"] = {a: per_policy[id(self.policy_map[a])] for a in self.agents}
        holder.__dict__["state"] = {a: Batch() for a in self.agents}
        return holder

    __call__ = forward

    def map_action(self, act: Any) -> Any:
        return act

    def train(self, mode: bool = True) -> "FlexibleMultiAgentPolicyManager":
        for p in self.policies.values():
            q = self._actor_policy(p)
            if hasattr(q, "train"):
                q.train(mode)
        return self

    # -- update: every unique policy learns from its own rows -------------------------------------
    def update(self, buffer: Any, batch_size: int | None, repeat: int) -> dict[str, Any]:
        """`policies` must be PPO algorithms here (they own the optimizer state); returns {name: training stats}."""
        dev = None
        stats: dict[str, Any] = {}
        names = {id(p): k for k, p in self.policies.items()}
        for p in {id(q): q for q in self.policy_map.values()}.values():
            if not hasattr(p, "update"):
                raise TypeError("update() needs PPO algorithms as policies")
            dev = p.net.device
            rows = next(r for q, r in self.groups(buffer.buffer_num, dev) if q is p)
            stats[names.get(id(p), str(len(stats)))] = p.update(buffer, batch_size, repeat, rows=rows)
        return stats
