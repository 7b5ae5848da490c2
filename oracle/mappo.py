"""Oracle: PPO with a centralized critic ("MAPPO", BASELINE.json configs[2]) in plain torch float32 on the CPU.  Test
infrastructure only (tests/, __graft_entry__.smoke(), bench.py's CPU arm).

Restates the reference PPO arithmetic over the reference's CTDE networks, rows flattened row = world * A + agent:
  tianshou/algorithm/multiagent/ctde.py:346-379    DecentralizedActor  fc1-relu-fc2-relu-fc3 on the row's local observation
  tianshou/algorithm/multiagent/ctde.py:382-414    CentralizedCritic   fc1-relu-fc2-relu-fc3 on the world's global state, one head per agent
  tianshou/algorithm/multiagent/ctde.py:291-294    GlobalStateConstructor("concatenate"): cat(obs_agent_0 .. obs_agent_{A-1})
  tianshou/algorithm/modelfree/a2c.py:113-151      v_s = critic(obs), v_s_ = critic(obs_next), GAE
  tianshou/algorithm/modelfree/ppo.py:146-224      logp_old; clipped surrogate / value / entropy loss per minibatch
  tianshou/algorithm/algorithm_base.py:485-498     clip_grad_norm_ over actor + critic parameters together, Adam
PINNED: tests/test_oracle_mappo.py checks it against tests/golden/mappo.npz, recorded from the unmodified reference PPO run over
the reference's DecentralizedActor / CentralizedCritic (tests/golden/make_golden_mappo.py).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

from . import ppo as oppo


class Mlp3:
    def __init__(self, W1, b1, W2, b2, W3, b3) -> None:
        self.p = [torch.tensor(np.asarray(a), dtype=torch.float32).clone().requires_grad_(True) for a in (W1, b1, W2, b2, W3, b3)]
        self.m = [torch.zeros_like(x) for x in self.p]
        self.v = [torch.zeros_like(x) for x in self.p]

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        W1, b1, W2, b2, W3, b3 = self.p
        return F.linear(F.relu(F.linear(F.relu(F.linear(x, W1, b1)), W2, b2)), W3, b3)

    def flat(self) -> np.ndarray:
        return np.concatenate([x.detach().numpy().reshape(-1) for x in self.p])


def global_state(obs_rows: np.ndarray, A: int) -> np.ndarray:
    """[E * A, O] rows (row = world * A + agent) -> [E, A * O]: concatenation in agent order is a reshape (ctde.py:291-294)."""
    n, O = obs_rows.shape
    return np.ascontiguousarray(obs_rows.reshape(n // A, A * O))


def values(critic: Mlp3, obs_rows: np.ndarray, A: int) -> np.ndarray:
    """Row values: head `agent` of the critic on the row's world state -> [E * A]."""
    with torch.no_grad():
        return critic.forward(torch.from_numpy(global_state(np.ascontiguousarray(obs_rows, np.float32), A))).numpy().reshape(-1)


def preprocess(actor: Mlp3, critic: Mlp3, A: int, obs, obs_next, act_, rew, term, trunc, indices, unfinished, gamma, lam):
    """Inputs flat in buffer.sample(0) order (env-major: all steps of row 0, then row 1, ...) EXCEPT obs / obs_next, which are
    [T][R][O] time-major so that whole worlds are adjacent; returns v_s, returns, adv, logp_old in buffer.sample(0) order."""
    T, R, O = obs.shape
    em = lambda a: np.swapaxes(a, 0, 1).reshape(R * T, *a.shape[2:])  # noqa: E731  [T][R] -> env-major
    v_s = em(values(critic, obs.reshape(T * R, O), A).reshape(T, R))
    v_n = em(values(critic, obs_next.reshape(T * R, O), A).reshape(T, R))
    ret, adv = oppo._gae(rew, term, trunc, indices, unfinished, v_n, v_s, gamma, lam)
    with torch.no_grad():
        logits = actor.forward(torch.from_numpy(np.ascontiguousarray(em(obs), np.float32)))
        logp_old = torch.distributions.Categorical(logits=logits).log_prob(torch.from_numpy(np.asarray(act_)).float()).numpy()
    return v_s, ret.astype(np.float32), adv.astype(np.float32), logp_old


def adam_step_joint(nets: list[Mlp3], t: int, lr: float, max_grad_norm: float, b1=0.9, b2=0.999, eps=1e-8) -> None:
    params = [x for n in nets for x in n.p]
    if max_grad_norm and max_grad_norm > 0:
        torch.nn.utils.clip_grad_norm_(params, max_grad_norm)
    with torch.no_grad():
        for n in nets:
            for x, m, v in zip(n.p, n.m, n.v):
                g = x.grad
                m.lerp_(g, 1 - b1)
                v.mul_(b2).addcmul_(g, g, value=1 - b2)
                bc1, bc2 = 1 - b1 ** t, 1 - b2 ** t
                x.addcdiv_(m, (v.sqrt() / np.sqrt(bc2)).add_(eps), value=-lr / bc1)
                x.grad = None


def update_whole_batch(actor: Mlp3, critic: Mlp3, A: int, obs_rows, gstate_rows, agent, act_, adv, ret, logp_old, v_old, cfg: dict,
                       repeat: int, lr: float, max_grad_norm: float) -> dict[str, list[float]]:
    """ppo.py:164-224 with batch_size = the whole batch: `repeat` optimizer steps over all rows.  obs_rows [n, O] local
    observations, gstate_rows [n, A * O] the row's world state, agent [n] head index."""
    T = lambda a, dt=torch.float32: torch.from_numpy(np.ascontiguousarray(a)).to(dt)  # noqa: E731
    obs, gs, ag = T(obs_rows), T(gstate_rows), T(agent, torch.int64)
    act_t, adv_t, ret_t, lp_t, vo_t = T(act_), T(adv), T(ret), T(logp_old), T(v_old)
    log = {"loss": [], "clip": [], "vf": [], "ent": []}
    eps_c = cfg.get("eps_clip", 0.2)
    for t in range(1, repeat + 1):
        dist = torch.distributions.Categorical(logits=actor.forward(obs))
        a = adv_t
        if cfg.get("adv_norm", True):
            a = (a - a.mean()) / (a.std() + 1e-8)
        ratio = (dist.log_prob(act_t) - lp_t).exp().float()
        s1, s2 = ratio * a, ratio.clamp(1 - eps_c, 1 + eps_c) * a
        if cfg.get("dual_clip", 0.0):
            c1 = torch.min(s1, s2)
            clip_loss = -torch.where(a < 0, torch.max(c1, cfg["dual_clip"] * a), c1).mean()
        else:
            clip_loss = -torch.min(s1, s2).mean()
        value = critic.forward(gs).gather(1, ag[:, None]).flatten()
        if cfg.get("value_clip", False):
            vc = vo_t + (value - vo_t).clamp(-eps_c, eps_c)
            vf_loss = torch.max((ret_t - value).pow(2), (ret_t - vc).pow(2)).mean()
        else:
            vf_loss = (ret_t - value).pow(2).mean()
        ent = dist.entropy().mean()
        loss = clip_loss + cfg.get("vf_coef", 0.5) * vf_loss - cfg.get("ent_coef", 0.01) * ent
        loss.backward()
        adam_step_joint([actor, critic], t, lr, max_grad_norm)
        for k, x in zip(("loss", "clip", "vf", "ent"), (loss, clip_loss, vf_loss, ent)):
            log[k].append(float(x.detach()))
    return log
