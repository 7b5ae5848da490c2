"""CPU restatement (torch fp32) of the CTDE row.  TEST INFRASTRUCTURE ONLY: imported by tests/ (and nothing
in the product path); pinned to fixtures recorded from the unmodified reference (tests/golden/ctde.npz <- make_golden_ctde.py).

    GlobalStateConstructor.build   tianshou/algorithm/multiagent/ctde.py:282-300 (concatenate, mean)
    DecentralizedActor.forward     ctde.py:366-379      CentralizedCritic.forward   ctde.py:402-414
    CTDEPolicy.learn               ctde.py:121-199      (torch.optim.Adam, defaults betas (0.9, 0.999), eps 1e-8)"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F


def global_concat(obs_rows: np.ndarray, n_agents: int) -> np.ndarray:
    """obs_rows [E*A, O], row = env*A + agent -> [E, A*O] = cat(obs of agent 0..A-1, dim=-1) (ctde.py:291-294)."""
    n, O = obs_rows.shape
    return np.ascontiguousarray(obs_rows.reshape(n // n_agents, n_agents * O))


def global_mean(obs_rows: np.ndarray, n_agents: int) -> np.ndarray:
    """stack over agents, mean(dim=0) (ctde.py:296-300) -> [E, O]."""
    n, O = obs_rows.shape
    return torch.from_numpy(np.ascontiguousarray(obs_rows.reshape(n // n_agents, n_agents, O).transpose(1, 0, 2))).mean(dim=0).numpy()


class Mlp3:
    """fc1 -> relu -> fc2 -> relu -> fc3 with torch.optim.Adam state (m, v, t)."""

    def __init__(self, W1, b1, W2, b2, W3, b3) -> None:
        self.p = [torch.tensor(np.asarray(a), dtype=torch.float32, requires_grad=True) for a in (W1, b1, W2, b2, W3, b3)]
        self.m = [torch.zeros_like(x) for x in self.p]
        self.v = [torch.zeros_like(x) for x in self.p]
        self.t = 0

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        W1, b1, W2, b2, W3, b3 = self.p
        return F.linear(F.relu(F.linear(F.relu(F.linear(x, W1, b1)), W2, b2)), W3, b3)

    def adam(self, lr: float, b1: float = 0.9, b2: float = 0.999, eps: float = 1e-8) -> None:
        self.t += 1
        with torch.no_grad():
            for x, m, v in zip(self.p, self.m, self.v):
                g = x.grad
                m.lerp_(g, 1 - b1)
                v.mul_(b2).addcmul_(g, g, value=1 - b2)
                denom = (v.sqrt() / np.sqrt(1 - b2 ** self.t)).add_(eps)
                x.addcdiv_(m, denom, value=-lr / (1 - b1 ** self.t))
                x.grad = None


def learn(actor: Mlp3, critic: Mlp3, obs, act, rew, obs_next, term, critic_in, critic_in_next, gamma: float, lr_a: float,
          lr_c: float) -> tuple[float, float]:
    """One CTDEPolicy.learn call (ctde.py:121-199)."""
    T = lambda a, dt=torch.float32: torch.from_numpy(np.ascontiguousarray(a)).to(dt)  # noqa: E731
    obs, rew, ci, cn = T(obs), T(rew).unsqueeze(-1), T(critic_in), T(critic_in_next)
    not_term = (~torch.from_numpy(np.asarray(term, bool))).float().unsqueeze(-1)
    values = critic(ci)
    values_next = critic(cn)
    if values.shape[1] > 1:
        values = values.mean(dim=1, keepdim=True)
        values_next = values_next.mean(dim=1, keepdim=True)
    td = rew + gamma * values_next * not_term
    critic_loss = F.mse_loss(values, td.detach())
    logits = actor(obs)
    adv = (td - values).detach()
    logp = -F.cross_entropy(logits, T(act, torch.int64), reduction="none")
    actor_loss = -(logp * adv).mean()   # note: [n] * [n, 1] broadcasts to [n, n] in the reference (ctde.py:182-183)
    critic_loss.backward()
    critic.adam(lr_c)
    actor_loss.backward()
    actor.adam(lr_a)
    return float(actor_loss), float(critic_loss)
