"""PPO with a centralized critic on the device -- the build's "MAPPO" (the reference has no MAPPO class; it is the
reference's PPO arithmetic over its CTDE networks):

    PPO._preprocess_batch / _update_with_batch     tianshou/algorithm/modelfree/ppo.py:146-224
    A2C-base._add_returns_and_advantages            tianshou/algorithm/modelfree/a2c.py:113-151
    Algorithm.compute_episodic_return (GAE)         tianshou/algorithm/algorithm_base.py:651-717
    DecentralizedActor / CentralizedCritic          tianshou/algorithm/multiagent/ctde.py:346-414
    GlobalStateConstructor("concatenate")           ctde.py:291-294
    Optimizer.step over ActorCritic parameters      algorithm_base.py:485-498, a2c.py:102-105

Rows are flattened row = world * A + agent, so the concatenated global state of (step t, world e) is the A * O consecutive
floats obs[t][e * A : (e + 1) * A] of the time-major rollout buffer: the critic's input is a VIEW (T * E, A * O) of the stored
observations, and its output (T * E, A) -- one value head per agent -- IS the [T][R] value array GAE reads.  The critic therefore
runs once per WORLD and step, not once per agent row.

Minibatches are sets of SUPER-TILES: 128 consecutive (step, world) pairs = one critic tile = A actor tiles.  An epoch draws a
permutation of the super-tiles (np.random.permutation of the global numpy RNG, like Batch.split draws its row permutation,
data/batch.py:1209-1226) and cuts it into minibatches of round(batch_size / (128 * A)) super-tiles.  Rows inside a super-tile
are different worlds at the same step, i.e. independent samples; the loss of a minibatch is the reference's: clipped surrogate
and entropy over its agent rows, value loss over the same agent rows (head a of the world's critic output against the row's
return), one Adam step over both nets with one gradient norm.  With batch_size >= the batch (one minibatch per epoch) the
update equals the reference PPO's on the same rows (tests/golden/mappo.npz).

All arithmetic runs in libtsm_b200 (csrc/mlp_tcw.cu engine 2, gae.cu); torch is used for memory and index plumbing only."""
from __future__ import annotations

import ctypes as C
import time
from typing import Any

import numpy as np
import torch

from .. import _lib
from ..utils.net import ActorCriticNet
from ..utils.stats import A2CTrainingStats, SequenceSummaryStats
from .ppo import AdamOptimizerFactory, DiscreteActorPolicy
from .returns import gae_time_major

TILE = 128


class MAPPO:
    def __init__(self, *, policy: DiscreteActorPolicy, critic: ActorCriticNet, n_agents: int, optim: AdamOptimizerFactory | None = None,
                 eps_clip: float = 0.2, dual_clip: float | None = None, value_clip: bool = False, advantage_normalization: bool = True,
                 vf_coef: float = 0.5, ent_coef: float = 0.01, max_grad_norm: float | None = None, gae_lambda: float = 0.95,
                 gamma: float = 0.99, minibatch_order: str = "numpy", dedup_next_values: bool = True, grad_sync: str = "auto") -> None:
        assert dual_clip is None or dual_clip > 1.0, f"Dual-clip PPO parameter should greater than 1.0 but got {dual_clip}"  # ppo.py
        actor = policy.actor
        if actor.engine != "tcw" or critic.engine != "tcw":
            raise ValueError("MAPPO runs on engine 'tcw' nets (DecentralizedActor / CentralizedCritic with engine='tcw')")
        if actor.n_act < 1 or critic.n_act != 0 or critic.out_dim != n_agents:
            raise ValueError("actor must emit logits, critic must be value-only with one head per agent")
        if critic.in_dim != n_agents * actor.in_dim:
            raise ValueError("the centralized critic reads the concatenated global state: in_dim == n_agents * obs_dim")
        if minibatch_order not in ("numpy", "sequential"):
            raise ValueError("minibatch_order must be 'numpy' or 'sequential'")
        if grad_sync not in ("auto", "peer", "nccl"):
            raise ValueError("grad_sync must be 'auto', 'peer' or 'nccl'")
        self.grad_sync = grad_sync  # world > 1: "peer" = exchange inside the optimizer kernel over NVLink peer loads; "nccl" = all_reduce per minibatch
        self._peers = None
        self.policy, self.actor, self.critic, self.n_agents = policy, actor, critic, int(n_agents)
        self.optim = optim or AdamOptimizerFactory()
        self.eps_clip, self.dual_clip, self.value_clip = eps_clip, dual_clip, value_clip
        self.advantage_normalization = advantage_normalization
        self.vf_coef, self.ent_coef, self.max_grad_norm = vf_coef, ent_coef, max_grad_norm
        self.gae_lambda, self.gamma = gae_lambda, gamma
        self.minibatch_order, self.dedup_next_values = minibatch_order, dedup_next_values
        self._eps = 1e-8
        self.training = False
        self._cfg = _lib.TsmPpoCfg(eps_clip=eps_clip, dual_clip=dual_clip or 0.0, vf_coef=vf_coef, ent_coef=ent_coef, value_clip=int(value_clip),
                                   adv_norm=int(advantage_normalization), plain_pg=0, value_heads=0)
        self._work: dict[str, torch.Tensor] = {}
        dev = actor.device
        # one gradient arena for both nets: a single all-reduce across ranks, 16-byte aligned slices
        na = (actor.n_params + 3) // 4 * 4
        self._grads = torch.zeros(na + critic.n_params, dtype=torch.float32, device=dev)
        actor.grads, critic.grads = self._grads[: actor.n_params], self._grads[na:]
        self._adam_ws = torch.zeros(2048, dtype=torch.uint8, device=dev)  # TSM_TCW_ADAM_WS_BYTES
        self._adam_seq = 0
        self._steps = 0
        self.last_sequences: dict[str, np.ndarray] = {}

    # nn.Module-like surface the trainer / collector touch
    def train(self, mode: bool = True) -> "MAPPO":
        self.training = mode
        return self

    def eval(self) -> "MAPPO":
        return self.train(False)

    def __call__(self, batch: Any, state: Any = None, **kw: Any) -> Any:
        return self.policy(batch, state, **kw)

    def state_dict(self) -> dict[str, Any]:
        out: dict[str, Any] = {"_optimizers": []}
        for name, n in (("actor", self.actor), ("critic", self.critic)):
            out[f"{name}.flat"] = n.flat.detach().clone()
            out["_optimizers"].append({"exp_avg": n.adam_m.clone(), "exp_avg_sq": n.adam_v.clone(), "step": n.adam_step.clone()})
        return out

    def load_state_dict(self, sd: dict[str, Any]) -> None:
        with torch.no_grad():
            for i, (name, n) in enumerate((("actor", self.actor), ("critic", self.critic))):
                n.flat.copy_(sd[f"{name}.flat"])
                if sd.get("_optimizers"):
                    o = sd["_optimizers"][i]
                    n.adam_m.copy_(o["exp_avg"]); n.adam_v.copy_(o["exp_avg_sq"]); n.adam_step.copy_(o["step"])
        self._steps = int(self.actor.adam_step.item())

    def _buf(self, name: str, shape: tuple[int, ...], dtype: torch.dtype = torch.float32, zero: bool = False) -> torch.Tensor:
        t = self._work.get(name)
        if t is None or tuple(t.shape) != tuple(shape) or t.dtype != dtype:
            t = (torch.zeros if zero else torch.empty)(shape, dtype=dtype, device=self.actor.device)
            self._work[name] = t
        return t

    # -- the update ---------------------------------------------------------------------------
    def update(self, buffer: Any, batch_size: int | None, repeat: int) -> A2CTrainingStats:
        if not self.policy.is_within_training_step:
            raise RuntimeError(
                f"update() was called outside of a training step as signalled by {self.policy.is_within_training_step=} "
                "If you want to update the policy without a Trainer, you will have to manage the above-mentioned flag yourself.")
        t0 = time.time()
        actor, critic, A = self.actor, self.critic, self.n_agents
        L, P, S = _lib.lib(), _lib.ptr, _lib.stream_ptr()
        st = buffer._store
        size, R, O = buffer._size, buffer.buffer_num, actor.in_dim
        T = buffer._lockstep_len
        if not T or R % A:
            raise NotImplementedError("MAPPO.update needs a lock-step (time-major) rollout buffer whose rows are whole worlds")
        start = buffer._lockstep_ins if T == size else 0
        if start != 0:
            raise NotImplementedError("MAPPO.update expects the rollout to start at slot 0 (reset_buffer before collect)")
        E = R // A
        n_rows, n_es = T * R, T * E
        dev = actor.device
        obs = st["obs"][:T].view(n_rows, O)
        # ---- packed inputs: local observations per agent row, global states per (step, world) ----
        tk = _lib.t_start("pack_inputs")
        n_super = (n_es + TILE - 1) // TILE
        xa = self._buf("ximg_actor", (actor.x_bytes(n_super * TILE * A),), torch.uint8, zero=True)  # whole super-tiles (a ragged tail stays zero)
        xc = self._buf("ximg_critic", (critic.x_bytes(n_es),), torch.uint8)
        actor.pack_x(obs, out=xa)
        critic.pack_x(obs.view(n_es, A * O), out=xc)
        _lib.t_stop(tk)
        # ---- _add_returns_and_advantages (a2c.py:113-151): critic(obs), critic(obs_next), GAE; logp_old (ppo.py:157-161) ----
        v_s, v_n = self._buf("v_s", (T, R)), self._buf("v_n", (T, R))
        logp, adv, ret = self._buf("logp_old", (T, R)), self._buf("adv", (T, R)), self._buf("ret", (T, R))
        tk = _lib.t_start("critic_forward")
        _lib.check(L.tsm_tcw_forward(critic.desc, P(xc), n_es, 0, 0, 0, 0, P(v_s), None, None, None, S))
        done = st["done"][:T]
        if self.dedup_next_values and buffer._obs_chain is True:
            # critic(obs_next[t]) = critic(obs[t + 1]) wherever the rollout kernel wrote both from the same registers; only
            # (step, world) pairs whose episode ended, and the last step, need the critic on obs_next
            need = self._buf("need_list", (n_rows,), torch.int64)
            cnt = self._buf("need_count", (1,), torch.int64)
            _lib.check(L.tsm_values_next_prepare(P(v_s), P(done), T, R, 0, T, P(v_n), P(need), P(cnt), S))
            k = int(cnt.item())
            if k > 0:
                rows = need[:k]
                es = (rows[rows % A == 0] // A).contiguous()  # all agents of a world end together (MPE truncation)
                xn = critic.pack_x(st["obs_next"][:T].view(n_es, A * O), idx=es)
                vn = self._buf("v_need", (int(es.numel()), A))
                _lib.check(L.tsm_tcw_forward(critic.desc, P(xn), int(es.numel()), 0, 0, 0, 0, P(vn), None, None, None, S))
                v_n.view(n_es, A)[es] = vn
        else:
            xn = critic.pack_x(st["obs_next"][:T].view(n_es, A * O))
            _lib.check(L.tsm_tcw_forward(critic.desc, P(xn), n_es, 0, 0, 0, 0, P(v_n), None, None, None, S))
        _lib.t_stop(tk)
        tk = _lib.t_start("logp_old_forward")
        act = st["act"][:T].view(-1)
        _lib.check(L.tsm_tcw_forward(actor.desc, P(xa), n_rows, 3, 0, 0, 0, None, P(act), P(logp), None, S))
        _lib.t_stop(tk)
        tk = _lib.t_start("gae")
        gae_time_major(v_s, v_n, st["rew"], st["terminated"], st["truncated"], 0, T, self.gamma, self.gae_lambda, adv, ret)
        _lib.t_stop(tk)
        # ---- minibatches of super-tiles ----
        rows_per_super = TILE * A
        if batch_size is None or batch_size <= 0 or batch_size >= n_rows:
            mb_super = n_super
        else:
            mb_super = max(1, int(round(batch_size / rows_per_super)))
        n_mb = (n_super + mb_super - 1) // mb_super
        stats = self._buf("mb_stats", (repeat * n_mb, 2))
        sums_a = self._buf("sums_actor", (repeat * n_mb, 4), torch.float64)
        sums_c = self._buf("sums_critic", (repeat * n_mb, 4), torch.float64)
        sums_a.zero_(); sums_c.zero_()
        world = torch.distributed.get_world_size() if torch.distributed.is_available() and torch.distributed.is_initialized() else 1
        o = self.optim
        mgn = float(self.max_grad_norm) if self.max_grad_norm else 0.0
        rows_seq: list[np.ndarray] = []
        keep = []
        # what it takes to undo the update if the in-kernel gradient exchange loses a peer (checked after the epochs)
        snaps = (actor.snapshot(), critic.snapshot(), (self._steps, self._adam_seq)) if world > 1 and self.grad_sync != "nccl" else None
        for ep in range(repeat):
            if self.minibatch_order == "numpy" and n_mb > 1:
                ids_h = np.random.permutation(n_super).astype(np.int32)  # the epoch's order of super-tiles (global numpy RNG, batch.py:1219)
                ids = torch.as_tensor(ids_h, device=dev)
                keep.append(ids)
            else:
                ids_h, ids = np.arange(n_super, dtype=np.int32), None
            # agent rows of every minibatch (host): full super-tiles except possibly the last physical one
            per_super = (np.minimum(TILE, n_es - ids_h.astype(np.int64) * TILE) * A).astype(np.int64)
            mb_rows = np.add.reduceat(per_super, np.arange(0, n_super, mb_super)).astype(np.int64)
            rows_seq.append(mb_rows)
            st_ep, sa, sc = stats[ep * n_mb:], sums_a[ep * n_mb:], sums_c[ep * n_mb:]
            if self.advantage_normalization:
                tk = _lib.t_start("adv_stats")
                _lib.check(L.tsm_tcw_adv_stats(P(adv), n_rows, P(ids), rows_per_super, n_super, mb_super, self._eps, P(st_ep), S))
                _lib.t_stop(tk)
            tk = _lib.t_start("ppo_epoch")
            h_rows = (C.c_int64 * n_mb)(*[int(x) for x in mb_rows])
            if world > 1 and self.grad_sync != "nccl" and self._peers is None:
                try:
                    self._peers = _TcwPeers(self)
                except Exception:  # no symmetric memory on this system: per-minibatch NCCL all-reduce
                    if self.grad_sync == "peer":
                        raise
                    self.grad_sync = "nccl"
            if world == 1 or self._peers is not None:
                actor.desc, critic.desc  # (refresh the descriptors: parameter image current, pointers valid)
                e = _lib.TsmTcwEpoch(actor=C.addressof(actor._desc), critic=C.addressof(critic._desc), ximg_actor=P(xa), ximg_critic=P(xc),
                                     n_rows=n_rows, n_agents=A, super_ids=P(ids), n_super=n_super, mb_super=mb_super, act=P(act), adv=P(adv),
                                     logp_old=P(logp), ret=P(ret), v_old=P(v_s), h_mb_rows=C.cast(h_rows, C.c_void_p),
                                     mb_stats=P(st_ep) if self.advantage_normalization else None, sums_actor=P(sa), sums_critic=P(sc),
                                     lr=o.lr, beta1=o.betas[0], beta2=o.betas[1], eps=o.eps, max_grad_norm=mgn, step0=self._steps,
                                     adam_ws=P(self._adam_ws), adam_seq0=self._adam_seq,
                                     peers=C.addressof(self._peers.desc) if self._peers is not None else None,
                                     xseq0=self._peers.xseq if self._peers is not None else 0)
                _lib.check(L.tsm_tcw_ppo_epoch(C.byref(e), C.byref(self._cfg), S))
                if self._peers is not None:
                    self._peers.xseq += n_mb
            else:
                self._epoch_across_ranks(xa, xc, n_rows, ids, n_super, mb_super, mb_rows, act, adv, logp, ret, v_s,
                                         st_ep if self.advantage_normalization else None, sa, sc, world)
            self._steps += n_mb
            if mgn > 0:
                self._adam_seq += n_mb
            _lib.t_stop(tk)
        actor.kernel_updates += repeat * n_mb
        critic.kernel_updates += repeat * n_mb
        ha, hc = sums_a.cpu().numpy(), sums_c.cpu().numpy()
        if self._peers is not None:
            try:
                self._peers.check()
            except _lib.TsmError:  # the exchange lost a peer: the update is void (see PPO.update)
                if snaps is None:
                    raise
                actor.restore(snaps[0])
                critic.restore(snaps[1])
                self._steps, self._adam_seq = snaps[2]
                raise
        del keep
        n_per = np.concatenate(rows_seq).astype(np.float64)
        clip_l, ent_l, vf_l = ha[:, 0] / n_per, ha[:, 2] / n_per, hc[:, 1] / n_per
        loss = clip_l + self.vf_coef * vf_l - self.ent_coef * ent_l
        out = A2CTrainingStats(loss=SequenceSummaryStats.from_sequence(loss), actor_loss=SequenceSummaryStats.from_sequence(clip_l),
                               vf_loss=SequenceSummaryStats.from_sequence(vf_l), ent_loss=SequenceSummaryStats.from_sequence(ent_l),
                               gradient_steps=repeat * n_mb)
        out.train_time = time.time() - t0
        self.last_sequences = {"loss": loss, "clip": clip_l, "vf": vf_l, "ent": ent_l}
        return out

    def _epoch_across_ranks(self, xa, xc, n_rows, ids, n_super, mb_super, mb_rows, act, adv, logp, ret, v_s, st_ep, sa, sc, world) -> None:
        """Data-parallel ranks: per minibatch the two backward launches, ONE all-reduce of the gradient arena
        (actor + critic, NCCL sum over NVLink), then the optimizer kernel with grad_scale = 1 / world."""
        L, P, S, A = _lib.lib(), _lib.ptr, _lib.stream_ptr(), self.n_agents
        o = self.optim
        mgn = float(self.max_grad_norm) if self.max_grad_norm else 0.0
        nets = (_lib.TsmNet * 2)()
        for i, n in enumerate((self.actor, self.critic)):
            n.desc
            C.memmove(C.byref(nets, i * C.sizeof(_lib.TsmNet)), C.byref(n._desc), C.sizeof(_lib.TsmNet))
        lc = _lib.TsmTcwLoss(kind=1, act=None, adv=None, logp_old=None, target=P(ret), v_old=P(v_s), weight=None, heads=A)
        la = _lib.TsmTcwLoss(kind=0, act=P(act), adv=P(adv), logp_old=P(logp), target=None, v_old=None, weight=None, heads=0)
        for m in range(len(mb_rows)):
            s0, cnt, rows = m * mb_super, min(mb_super, n_super - m * mb_super), int(mb_rows[m])
            gs = float(min(2 ** int(np.ceil(np.log2(max(rows, 1)))), 1048576))
            tc = _lib.TsmTcwTiles(ximg=P(xc), n_rows=n_rows // A, super_ids=P(ids), sub=1, tile0=s0, n_tiles=cnt)
            ta = _lib.TsmTcwTiles(ximg=P(xa), n_rows=n_rows, super_ids=P(ids), sub=A, tile0=s0 * A, n_tiles=cnt * A)
            _lib.check(L.tsm_tcw_train(self.critic.desc, C.byref(tc), C.byref(lc), C.byref(self._cfg), 1.0 / rows, gs, None, P(sc[m:]), S))
            _lib.check(L.tsm_tcw_train(self.actor.desc, C.byref(ta), C.byref(la), C.byref(self._cfg), 1.0 / rows, gs,
                                       P(st_ep[m:]) if st_ep is not None else None, P(sa[m:]), S))
            torch.distributed.all_reduce(self._grads)
            _lib.check(L.tsm_tcw_adam(nets, 2, self._steps + m + 1, o.lr, o.betas[0], o.betas[1], o.eps, mgn, 1.0 / world, P(self._adam_ws),
                                      self._adam_seq + (m if mgn > 0 else 0), None, S))


class _TcwPeers:
    """Symmetric-memory gradient arenas ([2 parities][stride] floats) + flag pads of all ranks as a tsm_tcw_peers struct
    (csrc/mlp_tcw.cu tcw_adam_kernel reads every rank's gradient over NVLink; torch.distributed only does the rendezvous)."""

    FLAG_WORDS = 64

    def __init__(self, alg: MAPPO) -> None:
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem

        actor, critic = alg.actor, alg.critic
        dev = actor.device
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        if self.world > 8:
            raise _lib.TsmError("the fused gradient exchange covers one node (<= 8 ranks)")
        na = (actor.n_params + 3) // 4 * 4
        self.stride = (na + critic.n_params + 3) // 4 * 4
        self.block = symm_mem.empty(2 * self.stride + self.FLAG_WORDS, dtype=torch.float32, device=dev)
        self.block.zero_()
        self.handle = symm_mem.rendezvous(self.block, group=dist.group.WORLD.group_name)
        self.gsum = torch.zeros(self.stride, dtype=torch.float32, device=dev)
        self.status = torch.zeros(1, dtype=torch.int32, device=dev)
        self.xseq = 0  # exchanges done
        p = _lib.TsmTcwPeers()
        p.world, p.rank, p.stride = self.world, self.rank, self.stride
        p.net_off[0], p.net_off[1] = 0, na
        for q in range(self.world):
            base = int(self.handle.buffer_ptrs[q])
            p.arena[q] = base
            p.flags[q] = base + 4 * 2 * self.stride
        p.gsum, p.status = self.gsum.data_ptr(), self.status.data_ptr()
        self.desc = p
        # the nets' gradient views point into parity 0 of the local arena (the epoch call selects the parity per minibatch)
        actor.grads, critic.grads = self.block[: actor.n_params], self.block[na: na + critic.n_params]
        torch.cuda.synchronize(dev)
        dist.barrier()  # every rank has zeroed its block before anyone's kernels touch it

    def check(self) -> None:
        if int(self.status.item()) != 0:
            raise _lib.TsmError("cross-rank gradient barrier timed out (a peer rank did not reach the optimizer step)")
