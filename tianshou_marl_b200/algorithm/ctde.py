"""Centralized training, decentralized execution on the device.

    CTDEPolicy.__init__/forward/learn   tianshou/algorithm/multiagent/ctde.py:47-199
    GlobalStateConstructor.build        ctde.py:282-343   (modes "concatenate", "mean")
    DecentralizedActor                  ctde.py:346-379   CentralizedCritic   ctde.py:382-414

With rows flattened as row = env * A + agent the observations of one world are contiguous, so the concatenated global state
of a world is a *view* (E, A*O) of the obs block; `GlobalStateConstructor.build_rows` materialises the per-row copy the
critic kernels read (tsm_global_state).  `learn` is the reference's one-step actor-critic: critic forward on the (global)
state and next state, TD target and advantage (tsm_td_target), then one fused forward + loss + backward + Adam launch pair
per network (tsm_ppo_minibatch with vf-only / plain policy-gradient epilogues, tsm_adam_step).  No torch arithmetic."""
from __future__ import annotations

import ctypes as C
from typing import Any

import numpy as np
import torch

from .. import _lib
from ..data.batch import Batch
from ..utils.net import ActorCriticNet
from .ppo import AdamOptimizerFactory


def DecentralizedActor(obs_dim: int, action_dim: int, hidden_dim: int = 128, **kw: Any) -> ActorCriticNet:
    """fc1-relu-fc2-relu-fc3 -> action logits on LOCAL observations (ctde.py:346-379)."""
    return ActorCriticNet(obs_dim, hidden_dim, action_dim, with_value=False, orthogonal=kw.pop("orthogonal", False), **kw)


def CentralizedCritic(global_obs_dim: int, n_agents: int, hidden_dim: int = 128, **kw: Any) -> ActorCriticNet:
    """fc1-relu-fc2-relu-fc3 -> one value per agent on the GLOBAL state (ctde.py:382-414)."""
    return ActorCriticNet(global_obs_dim, hidden_dim, 0, n_value_only=n_agents, orthogonal=kw.pop("orthogonal", False), **kw)


def load_linear3(net: ActorCriticNet, sd: dict[str, Any], prefix: str = "") -> ActorCriticNet:
    """Copy a reference DecentralizedActor / CentralizedCritic state_dict (fc1, fc2, fc3) into the flat parameter buffer."""
    g = lambda k: np.asarray(sd[prefix + k].detach().cpu() if hasattr(sd[prefix + k], "detach") else sd[prefix + k])  # noqa: E731
    return net.load_arrays(g("fc1.weight"), g("fc1.bias"), g("fc2.weight"), g("fc2.bias"), g("fc3.weight"), g("fc3.bias"))


class GlobalStateConstructor:
    def __init__(self, mode: str = "concatenate", obs_dim: int | None = None, n_agents: int | None = None, hidden_dim: int = 64,
                 adjacency_matrix: Any = None, custom_fn: Any = None) -> None:
        if mode in ("attention", "graph"):
            raise NotImplementedError(f"GlobalStateConstructor mode {mode!r} (learned aggregation) is not on the hot path")
        self.mode, self.obs_dim, self.n_agents, self.custom_fn = mode, obs_dim, n_agents, custom_fn

    def build(self, observations: dict[str, torch.Tensor]) -> torch.Tensor:
        """Reference signature: {agent_id: [batch, O]} -> [batch, A*O] ("concatenate", dict order) or [batch, O] ("mean")."""
        if self.mode == "custom" and self.custom_fn:
            return self.custom_fn(observations)
        obs = [torch.as_tensor(o) for o in observations.values()]
        if obs[0].is_cuda:
            A, (n, O) = len(obs), obs[0].shape
            rows = torch.stack([o.to(torch.float32) for o in obs], dim=1).reshape(n * A, O).contiguous()  # row = batch * A + agent
            g = self.build_rows(rows, A)
            return g.view(n, A, -1)[:, 0].contiguous()
        return torch.stack(obs, 0).mean(0) if self.mode == "mean" else torch.cat(obs, dim=-1)

    def build_rows(self, obs_rows: torch.Tensor, n_agents: int | None = None) -> torch.Tensor:
        """Device layout: obs_rows [E*A, O] (row = env * A + agent) -> [E*A, global_dim]; every agent row of a world receives
        that world's global state (what CTDEPolicy.learn reads as batch.global_obs)."""
        A = int(n_agents or self.n_agents)
        n, O = obs_rows.shape
        if n % A:
            raise ValueError("rows must hold whole worlds (row = env * n_agents + agent)")
        x = obs_rows.to(torch.float32).contiguous()
        mode = 1 if self.mode == "mean" else 0
        out = torch.empty(n, O if mode else A * O, dtype=torch.float32, device=x.device)
        _lib.check(_lib.lib().tsm_global_state(_lib.ptr(x), n // A, A, O, mode, _lib.ptr(out), _lib.stream_ptr()))
        return out


class CTDEPolicy:
    def __init__(self, actor: ActorCriticNet, critic: ActorCriticNet, optim_actor: AdamOptimizerFactory,
                 optim_critic: AdamOptimizerFactory, observation_space: Any = None, action_space: Any = None,
                 enable_global_info: bool = True, discount_factor: float = 0.99, broadcast_advantage: bool = True,
                 seed: int = 0, **kwargs: Any) -> None:
        """`broadcast_advantage=True` reproduces the reference's learn() literally: its `log_probs [n] * advantage [n, 1]`
        broadcasts to [n, n], so every row's log-probability is weighted by the batch-mean advantage (ctde.py:181-183);
        False uses the per-row advantage."""
        self.tau = kwargs.pop("tau", 0.005)
        if actor.n_act < 1 or critic.n_act != 0:
            raise ValueError("actor must emit action logits and critic must be a value-only net (CentralizedCritic)")
        self.actor, self.critic = actor, critic
        self.optim_actor, self.optim_critic = optim_actor, optim_critic
        self.observation_space, self.action_space = observation_space, action_space
        self.enable_global_info = enable_global_info
        self.discount_factor = float(discount_factor)
        self.broadcast_advantage = broadcast_advantage
        self.device = actor.device
        self.seed, self._counter = int(seed), 0
        self._work: dict[str, torch.Tensor] = {}

    def _buf(self, name: str, shape: tuple[int, ...], dtype: torch.dtype = torch.float32) -> torch.Tensor:
        t = self._work.get(name)
        if t is None or tuple(t.shape) != tuple(shape) or t.dtype != dtype:
            t = torch.empty(shape, dtype=dtype, device=self.device)
            self._work[name] = t
        return t

    def _dev(self, a: Any, dtype: torch.dtype = torch.float32) -> torch.Tensor:
        if not (isinstance(a, torch.Tensor) and a.is_cuda):
            a = torch.as_tensor(np.ascontiguousarray(a), device=self.device)
        return a.to(dtype).contiguous()

    # -- decentralized execution (ctde.py:84-119): the reference returns the actor's raw output as `act` ----------------
    def forward(self, batch: Any, state: Any = None, sample: bool = False, **kwargs: Any) -> Batch:
        obs = batch.obs if hasattr(batch, "obs") else batch["obs"]
        obs = self._dev(obs["obs"] if isinstance(obs, Batch) else obs)
        out = Batch()
        if sample:  # rollout use: categorical sample from the logits in the same launch
            o = self.actor.forward(obs, mode="sample", seed=self.seed, counter=self._counter, want_logits=True)
            self._counter += 1
            out.__dict__.update(act=o["act"].to(torch.int64), logits=o["logits"], logp=o["logp"], state=state)
        else:
            out.__dict__.update(act=self.actor.forward(obs, mode="raw")["out"], state=state)
        return out

    __call__ = forward

    # -- centralized training (ctde.py:121-199) ------------------------------------------------------------------------
    def learn(self, batch: Any, **kwargs: Any) -> dict[str, float]:
        L, P, S = _lib.lib(), _lib.ptr, _lib.stream_ptr()
        get = lambda k: getattr(batch, k) if hasattr(batch, k) else batch[k]  # noqa: E731
        has = lambda k: hasattr(batch, k) or (hasattr(batch, "__contains__") and k in batch)  # noqa: E731
        obs, obs_next = self._dev(get("obs")), self._dev(get("obs_next"))
        act, rew = self._dev(get("act"), torch.int32).view(-1), self._dev(get("rew")).view(-1)
        term = self._dev(get("terminated"), torch.uint8).view(-1)
        n = obs.shape[0]
        if self.enable_global_info and has("global_obs"):
            cin, cin_next = self._dev(get("global_obs")), self._dev(get("global_obs_next"))
        else:
            cin, cin_next = obs, obs_next
        heads = self.critic.out_dim
        v_cur, v_next = self._buf("v_cur", (n, heads)), self._buf("v_next", (n, heads))
        self.critic.forward(cin, mode="raw", out={"out": v_cur})
        self.critic.forward(cin_next, mode="raw", out={"out": v_next})
        td, adv = self._buf("td", (n,)), self._buf("adv", (n,))
        _lib.check(L.tsm_td_target(P(v_cur), P(v_next), heads, P(rew), P(term), self.discount_factor, n, P(td), P(adv), None, S))
        sums = self._buf("sums", (2, 4), torch.float64)
        sums.zero_()
        zeros = self._buf("zeros", (n,))
        zeros.zero_()
        # critic: mse(mean over heads, td) -> backward -> Adam
        pb = _lib.TsmPpoBatch(obs=P(cin), act=None, adv=None, ret=P(td), logp_old=None, v_old=None, indices=None, size=1, n_rows=n,
                              start=0, len=1, perm_key=0, perm_n=0, records=None)
        cfg_c = _lib.TsmPpoCfg(eps_clip=0.2, dual_clip=0.0, vf_coef=1.0, ent_coef=0.0, value_clip=0, adv_norm=0, plain_pg=0,
                               value_heads=heads)
        self._minibatch(self.critic, pb, cfg_c, n, None, sums[0], self.optim_critic)
        # actor: -(logp(act) * adv).mean() -> backward -> Adam
        stats = None
        if self.broadcast_advantage:
            stats = self._buf("stats", (2,))
            bounds = self._buf("bounds", (2,), torch.int64)
            bounds.copy_(torch.tensor([0, n], dtype=torch.int64))
            pa0 = _lib.TsmPpoBatch(obs=None, act=None, adv=P(adv), ret=None, logp_old=None, v_old=None, indices=None, size=1, n_rows=n,
                                   start=0, len=1, perm_key=0, perm_n=0, records=None)
            _lib.check(L.tsm_ppo_adv_stats(C.byref(pa0), P(bounds), 1, 1e-8, P(stats), S))
        pa = _lib.TsmPpoBatch(obs=P(obs), act=P(act), adv=P(adv), ret=None, logp_old=P(zeros), v_old=None, indices=None, size=1,
                              n_rows=n, start=0, len=1, perm_key=0, perm_n=0, records=None)
        cfg_a = _lib.TsmPpoCfg(eps_clip=0.2, dual_clip=0.0, vf_coef=0.0, ent_coef=0.0, value_clip=0, adv_norm=0,
                               plain_pg=2 if self.broadcast_advantage else 1, value_heads=0)
        self._minibatch(self.actor, pa, cfg_a, n, stats, sums[1], self.optim_actor)
        h = sums.cpu().numpy()
        return {"actor_loss": float(h[1, 0] / n), "critic_loss": float(h[0, 1] / n)}

    def _minibatch(self, net: ActorCriticNet, pb: Any, cfg: Any, n: int, stats: torch.Tensor | None, sums: torch.Tensor,
                   opt: AdamOptimizerFactory) -> None:
        L, P, S = _lib.lib(), _lib.ptr, _lib.stream_ptr()
        if net.engine == "tcw":  # the wide tcgen05 engine (csrc/mlp_tcw.cu): packed input image, one launch pair
            x = pb.obs  # device address of the [n, in_dim] float32 rows (kept alive by the caller)
            ximg = self._buf("ximg_%d" % id(net), (net.x_bytes(n),), torch.uint8)
            _lib.check(L.tsm_tcw_pack_x(x, net.in_dim, net.in_dim, None, n, P(ximg), S))
            tiles = _lib.TsmTcwTiles(ximg=P(ximg), n_rows=n, super_ids=None, sub=1, tile0=0, n_tiles=(n + 127) // 128)
            if cfg.value_heads:  # critic: mse(mean over heads, td) (ctde.py:153-166)
                loss = _lib.TsmTcwLoss(kind=2, act=None, adv=None, logp_old=None, target=pb.ret, v_old=None, weight=None, heads=cfg.value_heads)
            else:
                loss = _lib.TsmTcwLoss(kind=0, act=pb.act, adv=pb.adv, logp_old=None, target=None, v_old=None, weight=None, heads=0)
            gs = float(min(2 ** int(np.ceil(np.log2(max(n, 2)))), 1048576))
            _lib.check(L.tsm_tcw_train(net.desc, C.byref(tiles), C.byref(loss), C.byref(cfg), 1.0 / n, gs, P(stats), P(sums), S))
            step = int(getattr(net, "_tcw_steps", 0)) + 1
            net._tcw_steps = step
            _lib.check(L.tsm_tcw_adam(net.desc, 1, step, opt.lr, opt.betas[0], opt.betas[1], opt.eps, 0.0, 1.0, None, 0, None, S))
            net.kernel_updates += 1
            return
        if net.engine == "tc":
            rec = self._buf("rec_%d" % id(net), (n, _lib.PPO_RECORD_BYTES), torch.uint8)
            _lib.check(L.tsm_ppo_pack_records(net.desc, C.byref(pb), P(rec), S))
            pb.records = P(rec)
        _lib.check(L.tsm_ppo_minibatch(net.desc, C.byref(pb), C.byref(cfg), 0, n, P(stats), P(sums), S))
        _lib.check(L.tsm_adam_step(net.desc, opt.lr, opt.betas[0], opt.betas[1], opt.eps, 0.0, 1.0, None, S))
        net.kernel_updates += 1

    def soft_update_targets(self) -> None:
        """ctde.py:200-208: no target networks are attached on this path (the reference only updates them when present)."""
        return None
