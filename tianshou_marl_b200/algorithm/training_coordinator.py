"""Multi-agent training coordination: which policy learns from which part of a multi-agent batch.

Mirrors the reference's `SimultaneousTrainer` (tianshou/algorithm/multiagent/training_coordinator.py:266-339; base
`MATrainer.__init__` :48-71): per call the step counter advances, every agent whose training frequency divides the
counter and that has an entry in the batch gets `policy.learn(agent_batch)` -- with `global_obs` / `global_obs_next`
of the joint batch attached for centralized critics -- and the losses come back keyed by agent.  In "shared" mode all
agents train the one shared policy.  Pure host routing: the arithmetic happens inside `policy.learn` (CTDEPolicy.learn
runs the fused device kernels)."""
from __future__ import annotations

from typing import Any

VALID_MODES = ["simultaneous", "sequential", "self_play", "league"]


class SimultaneousTrainer:
    def __init__(self, policy_manager: Any, agent_train_freq: dict[str, int] | None = None, replay_buffers: dict[str, Any] | None = None,
                 **kwargs: Any) -> None:
        self.policy_manager = policy_manager
        self.training_mode = "simultaneous"
        self.step_count = 0
        self.kwargs = kwargs
        self.agent_train_freq = agent_train_freq or {}
        self.replay_buffers = replay_buffers or {}

    @staticmethod
    def _has(batch: Any, key: str) -> bool:
        try:
            return key in batch
        except TypeError:
            return hasattr(batch, key)

    def train_step(self, batch: Any) -> dict[str, Any]:
        """training_coordinator.py:292-339."""
        self.step_count += 1
        pm = self.policy_manager
        losses: dict[str, Any] = {}
        agents_to_train = pm.agents if pm.mode == "shared" else list(pm.policies.keys())
        for agent_id in agents_to_train:
            if self.step_count % self.agent_train_freq.get(agent_id, 1) != 0:
                continue
            if not self._has(batch, agent_id):
                continue
            agent_batch = batch[agent_id]
            for k in ("global_obs", "global_obs_next"):  # centralized critics read the joint observation
                if hasattr(batch, k) or self._has(batch, k):
                    setattr(agent_batch, k, batch[k] if self._has(batch, k) else getattr(batch, k))
            policy = pm.policies["shared"] if pm.mode == "shared" else pm.policies.get(agent_id)
            if policy is not None:
                if hasattr(policy, "train"):
                    policy.train()
                losses[agent_id] = policy.learn(agent_batch)
        return losses
