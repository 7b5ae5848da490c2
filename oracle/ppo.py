"""Oracle: PPO preprocess + minibatch update in plain torch float32 on the CPU.  Test
infrastructure only (tests/, __graft_entry__.smoke(), bench.py's CPU arm).

Restates, for the shared-trunk actor-critic  in -> H -> H -> (n_act logits, 1 value):
  tianshou/utils/net/common.py:173-181,343-369   MLP / Net forward (Linear + ReLU)
  tianshou/utils/net/discrete.py:67-89,116-124   DiscreteActor.last / DiscreteCritic.last
  tianshou/algorithm/modelfree/a2c.py:113-151    v_s, v_s_ and GAE -> returns, adv
  tianshou/algorithm/modelfree/ppo.py:146-162    logp_old
  tianshou/algorithm/modelfree/ppo.py:164-224    clipped-surrogate minibatch loop
  tianshou/data/batch.py:1209-1226               split(size, shuffle=True, merge_last=True)
  tianshou/algorithm/algorithm_base.py:485-498   zero_grad / backward / clip_grad_norm_ / Adam
PINNED: tests/test_oracle_ppo.py checks it against tests/golden/ppo.npz, recorded from the
unmodified reference PPO (tests/golden/make_golden_more.py).
Continuous actions (Net(log_sigma=...)): the head's n_act outputs are the means and log_sigma the state-independent log
standard deviations of Independent(Normal, 1) -- ContinuousActorProbabilistic(unbounded=True, conditioned_sigma=False),
tianshou/utils/net/continuous.py:196-247, test/continuous/test_ppo.py:104-110; pinned to tests/golden/ppo_gauss.npz.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

from . import gae as ogae


class Net:
    """Parameters in torch.nn.Linear convention; W3/b3 stack the actor rows then the critic row."""

    def __init__(self, W1, b1, W2, b2, W3, b3, n_act: int, log_sigma=None, mu_bound: float = 0.0, cond_sigma: bool = False) -> None:
        arrs = (W1, b1, W2, b2, W3, b3) + ((np.asarray(log_sigma).reshape(-1),) if log_sigma is not None else ())
        self.p = [torch.tensor(np.asarray(a), dtype=torch.float32).clone().requires_grad_(True) for a in arrs]
        self.n_act = n_act
        # cond_sigma (conditioned_sigma=True, continuous.py:228-233,242): head rows [mu | log sigma | value], sigma = exp(clamp(., -20, 2))
        self.cond_sigma = bool(cond_sigma)
        self._ls = None  # per-row log sigma of the last forward (cond_sigma)
        self.gaussian = log_sigma is not None or self.cond_sigma
        self.mu_bound = float(mu_bound)  # > 0: unbounded=False, mu = max_action * tanh(head output) (continuous.py:239-241)
        self.m = [torch.zeros_like(x) for x in self.p]
        self.v = [torch.zeros_like(x) for x in self.p]
        self.t = 0

    def forward(self, x: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
        W1, b1, W2, b2, W3, b3 = self.p[:6]
        h = F.relu(F.linear(x, W1, b1))
        h = F.relu(F.linear(h, W2, b2))
        o = F.linear(h, W3, b3)
        head = o[:, : self.n_act]
        if self.gaussian and self.mu_bound > 0:
            head = self.mu_bound * torch.tanh(head)
        vi = self.n_act * (2 if self.cond_sigma else 1)
        if self.cond_sigma:
            self._ls = torch.clamp(o[:, self.n_act: vi], min=-20.0, max=2.0)
        return head, (o[:, vi] if o.shape[1] > vi else torch.zeros(o.shape[0]))  # (actor-only net: no value row)

    def flat(self) -> np.ndarray:
        return np.concatenate([x.detach().numpy().reshape(-1) for x in self.p])


def make_net(in_dim: int, hidden: int, n_act: int, seed: int = 0) -> Net:
    g = torch.Generator().manual_seed(seed)
    mk = lambda *s: (torch.randn(*s, generator=g) / np.sqrt(s[-1])).numpy()  # noqa: E731
    return Net(mk(hidden, in_dim), np.zeros(hidden), mk(hidden, hidden), np.zeros(hidden), mk(n_act + 1, hidden), np.zeros(n_act + 1), n_act)


def act(net: Net, obs: np.ndarray) -> np.ndarray:
    with torch.no_grad():
        logits, _ = net.forward(torch.from_numpy(np.ascontiguousarray(obs, np.float32)))
        return torch.distributions.Categorical(logits=logits).sample().numpy()


def values(net: Net, obs: np.ndarray, chunk: int = 1 << 16) -> np.ndarray:
    out = []
    with torch.no_grad():
        for i in range(0, len(obs), chunk):
            out.append(net.forward(torch.from_numpy(np.ascontiguousarray(obs[i:i + chunk], np.float32)))[1])
    return torch.cat(out).numpy()


def policy_dist(net: Net, logits: torch.Tensor, probs_param: bool):
    """dist_fn of the policy: Categorical for discrete heads, Independent(Normal(mu, sigma), 1) for a Gaussian net
    (sigma = exp(sigma_param) broadcast over the rows, continuous.py:244-246)."""
    if net.gaussian:
        sigma = net._ls.exp() if net.cond_sigma else (net.p[6].view(1, -1) + torch.zeros_like(logits)).exp()
        return torch.distributions.Independent(torch.distributions.Normal(logits, sigma), 1)
    return log_softmax_dist(logits, probs_param)


def log_softmax_dist(logits: torch.Tensor, probs_param: bool) -> torch.distributions.Categorical:
    if probs_param:  # DiscreteActor(softmax_output=True) + Categorical(probs) (test_ppo_discrete.py:97-113)
        return torch.distributions.Categorical(probs=F.softmax(logits, dim=-1))
    return torch.distributions.Categorical(logits=logits)


def split_bounds(length: int, size: int, merge_last: bool = True) -> list[tuple[int, int]]:
    """batch.py:1209-1226 minibatch [start, stop) positions."""
    if size == -1 or size is None:
        size = length
    merge = merge_last and length % size > 0
    out = []
    for idx in range(0, length, size):
        if merge and idx + size + size >= length:
            out.append((idx, length))
            break
        out.append((idx, min(idx + size, length)))
    return out


def rms_update(rms: dict, data: np.ndarray) -> None:
    """RunningMeanStd.update (tianshou/utils/statistics.py:96-114) on rms = dict(mean, var, count)."""
    b_mean, b_var, b_n = float(np.mean(data)), float(np.var(data)), len(data)
    delta, tot = b_mean - rms["mean"], rms["count"] + b_n
    m_2 = rms["var"] * rms["count"] + b_var * b_n + delta**2 * rms["count"] * b_n / tot
    rms["mean"], rms["var"], rms["count"] = rms["mean"] + delta * b_n / tot, m_2 / tot, tot


def preprocess(net: Net, obs, obs_next, act_, rew, term, trunc, indices, unfinished, gamma, lam, probs_param=False, ret_rms: dict | None = None):
    """a2c.py:113-151 + ppo.py:146-162 over a flat batch in buffer.sample(0) order.  ret_rms = dict(mean, var, count):
    return_scaling=True (a2c.py:132-146); the dict is updated like self.ret_rms."""
    v_s, v_n = values(net, obs), values(net, obs_next)
    if ret_rms is not None:
        sc = np.sqrt(ret_rms["var"] + 1e-8)
        ret_u, adv = _gae(rew, term, trunc, indices, unfinished, v_n * sc, v_s * sc, gamma, lam)
        ret = ret_u / sc
        rms_update(ret_rms, ret_u)
    else:
        ret, adv = _gae(rew, term, trunc, indices, unfinished, v_n, v_s, gamma, lam)
    with torch.no_grad():
        logits, _ = net.forward(torch.from_numpy(np.ascontiguousarray(obs, np.float32)))
        logp_old = policy_dist(net, logits, probs_param).log_prob(torch.from_numpy(np.asarray(act_)).float()).numpy()
    return v_s, ret.astype(np.float32), adv.astype(np.float32), logp_old


def _gae(rew, term, trunc, indices, unfinished, v_n, v_s, gamma, lam):
    v_n = v_n.astype(np.float64) * (1.0 - np.asarray(term, np.float64))  # value_mask (algorithm_base.py:709)
    end = np.logical_or(term, trunc)
    end[np.isin(indices, unfinished)] = True  # :712-713
    adv = ogae.gae(v_s.astype(np.float64), v_n, np.asarray(rew, np.float64), end, gamma, lam)
    return adv + v_s, adv


def adam_step(net: Net, lr: float, max_grad_norm: float, b1=0.9, b2=0.999, eps=1e-8) -> float:
    grads = [x.grad for x in net.p]
    norm = float(torch.sqrt(sum((g.double() ** 2).sum() for g in grads)).float())
    if max_grad_norm and max_grad_norm > 0:
        torch.nn.utils.clip_grad_norm_(net.p, max_grad_norm)
    net.t += 1
    with torch.no_grad():
        for x, m, v in zip(net.p, net.m, net.v):
            g = x.grad
            m.lerp_(g, 1 - b1)
            v.mul_(b2).addcmul_(g, g, value=1 - b2)
            bc1, bc2 = 1 - b1 ** net.t, 1 - b2 ** net.t
            denom = (v.sqrt() / np.sqrt(bc2)).add_(eps)
            x.addcdiv_(m, denom, value=-lr / bc1)
            x.grad = None
    return norm


def minibatch_loss(net: Net, obs, act_, adv, ret, logp_old, v_old, cfg: dict, probs_param=False):
    """ppo.py:181-211 for one minibatch (tensors)."""
    logits, value = net.forward(obs)
    dist = policy_dist(net, logits, probs_param)
    if cfg.get("a2c", False):  # A2C._update_with_batch (a2c.py:262-272)
        actor_loss = -(dist.log_prob(act_) * adv).mean()
        vf_loss = (ret - value).pow(2).mean()
        ent = dist.entropy().mean()
        return actor_loss + cfg.get("vf_coef", 0.5) * vf_loss - cfg.get("ent_coef", 0.01) * ent, actor_loss, vf_loss, ent
    if cfg.get("adv_norm", True):
        adv = (adv - adv.mean()) / (adv.std() + 1e-8)
    ratio = (dist.log_prob(act_) - logp_old).exp().float()
    eps = cfg.get("eps_clip", 0.2)
    s1, s2 = ratio * adv, ratio.clamp(1 - eps, 1 + eps) * adv
    if cfg.get("dual_clip", 0.0):
        c1 = torch.min(s1, s2)
        c2 = torch.max(c1, cfg["dual_clip"] * adv)
        clip_loss = -torch.where(adv < 0, c2, c1).mean()
    else:
        clip_loss = -torch.min(s1, s2).mean()
    if cfg.get("value_clip", False):
        vc = v_old + (value - v_old).clamp(-eps, eps)
        vf_loss = torch.max((ret - value).pow(2), (ret - vc).pow(2)).mean()
    else:
        vf_loss = (ret - value).pow(2).mean()
    ent = dist.entropy().mean()
    loss = clip_loss + cfg.get("vf_coef", 0.5) * vf_loss - cfg.get("ent_coef", 0.01) * ent
    return loss, clip_loss, vf_loss, ent


def update_with_orders(net: Net, obs, act_, adv, ret, logp_old, v_old, cfg: dict, orders: list[np.ndarray], batch_size: int,
                       lr: float, max_grad_norm: float, probs_param=False, recompute: dict | None = None) -> dict[str, list[float]]:
    """ppo.py:164-224 with the per-epoch permutations given (np.random.permutation(B) in the reference).
    recompute (recompute_advantage=True, ppo.py:174-178): dict(obs_next, rew, term, trunc, indices, unfinished, gamma, lam);
    before every epoch after the first, v_s / returns / advantages are re-evaluated with the current parameters."""
    obs_np = np.ascontiguousarray(obs, np.float32)
    T = lambda a, dt=torch.float32: torch.from_numpy(np.ascontiguousarray(a)).to(dt)  # noqa: E731
    obs, adv, ret, logp_old, v_old, act_t = T(obs), T(adv), T(ret), T(logp_old), T(v_old), T(act_)
    log = {"loss": [], "clip": [], "vf": [], "ent": [], "gnorm": []}
    B = len(adv)
    for ep, order in enumerate(orders):
        if recompute is not None and ep > 0:
            r = recompute
            v_new, v_nn = values(net, obs_np), values(net, r["obs_next"])
            ret_n, adv_n = _gae(r["rew"], r["term"], r["trunc"], r["indices"], r["unfinished"], v_nn, v_new, r["gamma"], r["lam"])
            adv, ret, v_old = T(adv_n.astype(np.float32)), T(ret_n.astype(np.float32)), T(v_new)
        for s, e in split_bounds(B, batch_size, True):
            ix = torch.from_numpy(np.asarray(order[s:e], np.int64))
            loss, c, v, en = minibatch_loss(net, obs[ix], act_t[ix], adv[ix], ret[ix], logp_old[ix], v_old[ix], cfg, probs_param)
            loss.backward()
            log["gnorm"].append(adam_step(net, lr, max_grad_norm))
            for k, x in zip(("loss", "clip", "vf", "ent"), (loss, c, v, en)):
                log[k].append(float(x))
    return log


def update(net: Net, store: dict, adv: np.ndarray, ret: np.ndarray, batch_size: int, repeat: int, lr: float = 3e-4) -> None:
    """bench.py CPU arm: logp_old + `repeat` shuffled epochs with the reference's default hyper-parameters."""
    obs = store["obs"]
    with torch.no_grad():
        logits, v_old = net.forward(torch.from_numpy(obs))
        logp_old = torch.distributions.Categorical(logits=logits).log_prob(torch.from_numpy(store["act"]).float()).numpy()
    orders = [np.random.permutation(len(adv)) for _ in range(repeat)]
    update_with_orders(net, obs, store["act"], adv.astype(np.float32), ret.astype(np.float32), logp_old, v_old.numpy(),
                       dict(adv_norm=True), orders, batch_size, lr, 0.0)


def reinforce_returns(rew, term, trunc, indices, unfinished, gamma, ret_rms: dict | None = None, standardize: bool = False):
    """DiscountedReturnComputation.add_discounted_returns (reinforce.py:297-312): lambda = 1 returns whose truncated /
    unfinished ends bootstrap with ret_rms.mean; v_s = roll(masked v_s_, 1) as in algorithm_base.py:710."""
    mean = ret_rms["mean"] if ret_rms is not None else 0.0
    v_n = np.full(len(rew), mean, np.float64)
    v_s = np.roll(v_n * (1.0 - np.asarray(term, np.float64)), 1)
    ret_u, _ = _gae(rew, term, trunc, indices, unfinished, v_n, v_s, gamma, 1.0)
    if standardize:
        ret = (ret_u - ret_rms["mean"]) / np.sqrt(ret_rms["var"] + 1e-8)
        rms_update(ret_rms, ret_u)
        return ret
    return ret_u


def reinforce_update(net: Net, obs, act_, ret, orders: list[np.ndarray], batch_size: int, lr: float) -> list[float]:
    """Reinforce._update_with_batch (reinforce.py:362-382): loss = -(log pi(a | s) * return).mean(), Adam, no clipping."""
    T = lambda a, dt=torch.float32: torch.from_numpy(np.ascontiguousarray(a)).to(dt)  # noqa: E731
    obs, ret, act_t = T(obs), T(ret), T(act_)
    losses = []
    for order in orders:
        for s_, e_ in split_bounds(len(ret), batch_size, True):
            ix = torch.from_numpy(np.asarray(order[s_:e_], np.int64))
            logits, _ = net.forward(obs[ix])
            loss = -(log_softmax_dist(logits, False).log_prob(act_t[ix]) * ret[ix]).mean()
            loss.backward()
            adam_step(net, lr, 0.0)
            losses.append(float(loss))
    return losses
