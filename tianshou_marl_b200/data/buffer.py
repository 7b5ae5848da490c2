"""Device-resident VectorReplayBuffer (drop-in for tianshou/data/buffer/vecbuf.py:15-37,
manager.py:13-229 and buffer_base.py:20-656 on the on-policy path).

All bookkeeping runs in libtsm_b200 kernels (csrc/buffer.cu); this class only owns the torch
tensors and mirrors the reference's method surface.  Index semantics are bit-exact with the
reference: indices are numbered env-major (row * size + slot) although the payload lives
time-major ([slot][row]) on the device.

Host <-> device convention: methods return numpy arrays when they were handed numpy / Python
data (the reference's convention) and device tensors when handed device tensors; methods
without array arguments follow `index_type` ("numpy" by default, "torch" for the fused
collector).  rew is stored float32 (the reference stores float64, manager.py:183).
"""
from __future__ import annotations

import ctypes as C
from typing import Any

import numpy as np
import torch

from .. import _lib
from .batch import Batch

_RESERVED = ("obs", "act", "rew", "terminated", "truncated", "done", "obs_next", "info", "policy")


def _is_host(x: Any) -> bool:
    return not (isinstance(x, torch.Tensor) and x.is_cuda)


class VectorReplayBuffer:
    def __init__(self, total_size: int, buffer_num: int, stack_num: int = 1, ignore_obs_next: bool = False,
                 save_only_last_obs: bool = False, sample_avail: bool = False, random_seed: int = 42,
                 device: str | torch.device | None = None, index_type: str = "numpy", **kwargs: Any) -> None:
        assert buffer_num > 0
        if stack_num != 1 or save_only_last_obs or sample_avail:
            # Atari-only options of the reference (buffer_base.py:462-469); not on this path
            raise NotImplementedError("stack_num > 1, save_only_last_obs and sample_avail are not supported")
        self.options = {"stack_num": stack_num, "ignore_obs_next": ignore_obs_next,
                        "save_only_last_obs": save_only_last_obs, "sample_avail": sample_avail}
        self.buffer_num = int(buffer_num)
        self._size = int(np.ceil(total_size / buffer_num))  # vecbuf.py:35
        self.maxsize = self._size * self.buffer_num
        self.stack_num = 1
        self._save_obs_next = not ignore_obs_next
        self._seed = random_seed
        self._random_state = np.random.RandomState(random_seed)  # buffer_base.py:101
        self._child_rs: dict[int, np.random.RandomState] = {}
        self.index_type = index_type
        self.device = torch.device(device if device is not None else "cuda")
        if self.device.type != "cuda":
            raise _lib.TsmError("VectorReplayBuffer needs a CUDA device (no CPU fallback)")
        _lib.lib()  # fail loudly now if the extension is missing
        R = self.buffer_num
        i64 = dict(dtype=torch.int64, device=self.device)
        self._ins = torch.zeros(R, **i64)
        self._sz = torch.zeros(R, **i64)
        self._last = torch.arange(R, **i64) * self._size
        self._ep_len = torch.zeros(R, **i64)
        self._ep_start = torch.zeros(R, **i64)
        self._ep_ret = torch.zeros(R, dtype=torch.float64, device=self.device)
        self._offset = np.arange(R, dtype=np.int64) * self._size
        self._extend_offset = np.concatenate([self._offset, [self.maxsize]])
        self._store: dict[str, torch.Tensor] = {}
        self._obs_shape: tuple[int, ...] | None = None
        self._act_dtype = np.int64
        self._ws = torch.empty(_lib.lib().tsm_buf_indices_workspace_bytes(R) // 8 + 1, **i64)
        self._count = torch.zeros(1, **i64)
        self._desc = _lib.TsmBuf()
        self._lockstep_len: int | None = 0  # host-known common length while every add was a full lock-step add
        self._lockstep_ins = 0  # host-known common insertion index in lock-step use
        self._obs_chain = None  # True while EVERY valid slot was written by the fused rollout: obs[t + 1] is obs_next[t] bit for bit unless done[t]
        self._policy_version = None  # param_version of the net whose logp / value the fused collector stored for EVERY valid slot, else None
        self._refresh_desc()

    # -- plumbing -----------------------------------------------------------------------------
    def _refresh_desc(self) -> None:
        d, s = self._desc, self._store
        d.n_bufs, d.size, d.obs_dim = self.buffer_num, self._size, int(np.prod(self._obs_shape)) if self._obs_shape is not None else 0
        d.ins, d.sz, d.last_index = self._ins.data_ptr(), self._sz.data_ptr(), self._last.data_ptr()
        d.ep_len, d.ep_start, d.ep_ret = self._ep_len.data_ptr(), self._ep_start.data_ptr(), self._ep_ret.data_ptr()
        for k in ("obs", "obs_next", "act", "rew", "terminated", "truncated", "done", "logp", "value"):
            setattr(d, k, s[k].data_ptr() if k in s else None)

    def _alloc(self, obs_shape: tuple[int, ...], want_policy: bool) -> None:
        T, R, dev = self._size, self.buffer_num, self.device
        O = int(np.prod(obs_shape)) if len(obs_shape) else 1
        self._obs_shape = tuple(obs_shape)
        s = self._store
        s["obs"] = torch.zeros(T, R, O, dtype=torch.float32, device=dev)
        if self._save_obs_next:
            s["obs_next"] = torch.zeros(T, R, O, dtype=torch.float32, device=dev)
        s["act"] = torch.zeros(T, R, dtype=torch.int32, device=dev)
        s["rew"] = torch.zeros(T, R, dtype=torch.float32, device=dev)
        for k in ("terminated", "truncated", "done"):
            s[k] = torch.zeros(T, R, dtype=torch.uint8, device=dev)
        if want_policy:
            s["logp"] = torch.zeros(T, R, dtype=torch.float32, device=dev)
            s["value"] = torch.zeros(T, R, dtype=torch.float32, device=dev)
        self._refresh_desc()

    def _dev(self, x: Any, dtype: torch.dtype) -> torch.Tensor:
        if isinstance(x, torch.Tensor):
            return x.to(device=self.device, dtype=dtype).contiguous()
        return torch.as_tensor(np.ascontiguousarray(x), device=self.device).to(dtype).contiguous()

    def _out(self, t: torch.Tensor, host: bool) -> Any:
        return t.cpu().numpy() if host else t

    @property
    def subbuffer_edges(self) -> np.ndarray:
        return self._extend_offset

    @property
    def last_index(self) -> Any:
        return self._out(self._last, self.index_type == "numpy")

    @property
    def _lengths(self) -> Any:
        return self._out(self._sz, self.index_type == "numpy")

    def __len__(self) -> int:
        """manager.py:67-68.  Host-known in lock-step use; otherwise one small D2H read."""
        if self._lockstep_len is not None:
            return self._lockstep_len * self.buffer_num
        return int(self._sz.sum().item())

    def reset(self, keep_statistics: bool = False) -> None:
        _lib.check(_lib.lib().tsm_buf_reset(C.byref(self._desc), int(keep_statistics), _lib.stream_ptr()))
        self._lockstep_len = 0
        self._lockstep_ins = 0
        self._policy_version = None
        self._obs_chain = None

    def update(self, buffer: Any) -> np.ndarray:
        """manager.py:127-129."""
        raise NotImplementedError

    # -- add ----------------------------------------------------------------------------------
    def add(self, batch: Any, buffer_ids: Any = None) -> tuple[Any, Any, Any, Any]:
        """manager.py:131-193.  Returns (insertion_idx, ep_return, ep_len, ep_start_idx)."""
        get = (lambda k: batch[k] if k in batch else None) if isinstance(batch, Batch | dict) else None
        obs, act, rew = get("obs"), get("act"), get("rew")
        term, trunc, obs_next = get("terminated"), get("truncated"), get("obs_next")
        if obs is None or act is None or rew is None or term is None or trunc is None:
            raise ValueError("Input batch must have the keys obs, act, rew, terminated, truncated")
        if isinstance(obs, Batch):
            obs = obs["obs"]  # AEC wrapper dict obs {agent_id, obs, mask}: store the array part
            obs_next = obs_next["obs"] if isinstance(obs_next, Batch) else obs_next
        host = _is_host(rew)
        n = int(rew.shape[0]) if hasattr(rew, "shape") else len(rew)
        obs_t = self._dev(obs, torch.float32)
        if "obs" not in self._store:
            pol = get("policy")
            self._alloc(tuple(obs_t.shape[1:]), want_policy=isinstance(pol, Batch) and "logp" in pol)
            if not isinstance(act, torch.Tensor):
                self._act_dtype = np.asarray(act).dtype
        obs_t = obs_t.reshape(n, -1)
        on_t = self._dev(obs_next, torch.float32).reshape(n, -1) if (obs_next is not None and self._save_obs_next) else None
        # continuous (Box) actions: float vectors per row.  They live in their own time-major array `act_f` [size][rows][A]
        # (what the Gaussian PPO kernels read, tsm_ppo_batch.act_f); the integer action array then holds zeros.
        act_f_t = None
        a_probe = act if isinstance(act, torch.Tensor) else np.asarray(act)
        if a_probe.ndim >= 2 and (a_probe.dtype.is_floating_point if isinstance(a_probe, torch.Tensor) else np.issubdtype(a_probe.dtype, np.floating)):
            act_f_t = self._dev(act, torch.float32).reshape(n, -1)
            act_t = torch.zeros(n, dtype=torch.int32, device=self.device)
        else:
            act_t = self._dev(act, torch.int32).reshape(n)
        rew_t = self._dev(rew, torch.float32).reshape(n)
        te_t = self._dev(term, torch.bool).view(torch.uint8).reshape(n) if not (isinstance(term, torch.Tensor) and term.dtype == torch.uint8) else term.reshape(n)
        tr_t = self._dev(trunc, torch.bool).view(torch.uint8).reshape(n) if not (isinstance(trunc, torch.Tensor) and trunc.dtype == torch.uint8) else trunc.reshape(n)
        pol = get("policy")
        logp_t = value_t = None
        if isinstance(pol, Batch) and "logp" in self._store:
            logp_t = self._dev(pol["logp"], torch.float32).reshape(n) if "logp" in pol else None
            value_t = self._dev(pol["value"], torch.float32).reshape(n) if "value" in pol else None
        ids_t = None
        if buffer_ids is not None:
            ids_t = self._dev(buffer_ids, torch.int64).reshape(-1)
            if ids_t.numel() != n:
                raise ValueError("buffer_ids and batch must have the same length")
        elif n != self.buffer_num:
            raise ValueError("batch length must equal buffer_num when buffer_ids is None")
        i64 = dict(dtype=torch.int64, device=self.device)
        idx, ep_len, ep_start = torch.empty(n, **i64), torch.empty(n, **i64), torch.empty(n, **i64)
        ep_ret = torch.empty(n, dtype=torch.float64, device=self.device)
        P = _lib.ptr
        _lib.check(_lib.lib().tsm_buf_add(
            C.byref(self._desc), P(obs_t), P(act_t), P(rew_t), P(te_t), P(tr_t), P(on_t), P(logp_t), P(value_t),
            P(ids_t), n, P(idx), P(ep_ret), P(ep_len), P(ep_start), _lib.stream_ptr()))
        if act_f_t is not None:
            A = int(act_f_t.shape[1])
            if "act_f" not in self._store or self._store["act_f"].shape[2] != A:
                self._store["act_f"] = torch.zeros(self._size, self.buffer_num, A, dtype=torch.float32, device=self.device)
            self._store["act_f"].view(-1, A)[(idx % self._size) * self.buffer_num + idx // self._size] = act_f_t
        self._policy_version = None  # generic adds: stored logp / value (if any) are not known to cover every slot
        self._obs_chain = False  # nor is obs[t + 1] known to repeat obs_next[t]
        if buffer_ids is None and self._lockstep_len is not None:
            self._lockstep_len = min(self._lockstep_len + 1, self._size)
            self._lockstep_ins = (self._lockstep_ins + 1) % self._size
        else:
            self._lockstep_len = None
        return self._out(idx, host), self._out(ep_ret, host), self._out(ep_len, host), self._out(ep_start, host)

    # -- indices ------------------------------------------------------------------------------
    def _all_indices_device(self) -> torch.Tensor:
        if "done" not in self._store:  # nothing was ever added
            return torch.empty(0, dtype=torch.int64, device=self.device)
        out = torch.empty(self.maxsize, dtype=torch.int64, device=self.device)
        _lib.check(_lib.lib().tsm_buf_indices_all(C.byref(self._desc), _lib.ptr(out), _lib.ptr(self._count),
                                                  _lib.ptr(self._ws), self._ws.numel() * 8, _lib.stream_ptr()))
        n = len(self) if self._lockstep_len is not None else int(self._count.item())
        return out[:n]

    def _unfinished_device(self) -> torch.Tensor:
        out = torch.empty(self.buffer_num, dtype=torch.int64, device=self.device)
        if "done" not in self._store:
            return out[:0]
        _lib.check(_lib.lib().tsm_buf_unfinished(C.byref(self._desc), _lib.ptr(out), _lib.ptr(self._count),
                                                 _lib.ptr(self._ws), self._ws.numel() * 8, _lib.stream_ptr()))
        return out[: int(self._count.item())]

    def unfinished_index(self) -> Any:
        """manager.py:85-91."""
        return self._out(self._unfinished_device(), self.index_type == "numpy")

    def _child_rng(self, b: int) -> np.random.RandomState:
        if b not in self._child_rs:
            self._child_rs[b] = np.random.RandomState(self._seed)  # every child buffer: RandomState(42)
        return self._child_rs[b]

    def sample_indices(self, batch_size: int | None) -> Any:
        """manager.py:195-229.  batch_size 0/None -> all valid indices (device kernel, bit-exact
        order).  batch_size > 0 -> the reference's numpy RandomState draws, replayed on the host
        from the device lengths (exact by construction; MT19937 is not re-implemented on the GPU)."""
        host = self.index_type == "numpy"
        if batch_size is not None and batch_size < 0:
            return np.array([], int) if host else torch.empty(0, dtype=torch.int64, device=self.device)
        if batch_size == 0 or batch_size is None:
            return self._out(self._all_indices_device(), host)
        lengths = self._sz.cpu().numpy()
        buffer_idx = self._random_state.choice(self.buffer_num, batch_size, p=lengths / lengths.sum())
        sample_num = np.bincount(buffer_idx, minlength=self.buffer_num)
        parts = [self._child_rng(int(b)).choice(lengths[b], int(k)) + self._offset[b] for b, k in enumerate(sample_num) if k > 0]
        out = np.concatenate(parts).astype(np.int64) if parts else np.array([], np.int64)
        return out if host else torch.as_tensor(out, device=self.device)

    def _neighbour(self, index: Any, nxt: bool) -> Any:
        scalar = not isinstance(index, list | np.ndarray | torch.Tensor)
        host = _is_host(index)
        idx = self._dev(np.array([index]) if scalar else index, torch.int64).reshape(-1)
        out = torch.empty_like(idx)
        fn = _lib.lib().tsm_buf_next if nxt else _lib.lib().tsm_buf_prev
        if "done" not in self._store:
            raise ValueError("buffer is empty")
        _lib.check(fn(C.byref(self._desc), _lib.ptr(idx), idx.numel(), _lib.ptr(out), _lib.stream_ptr()))
        res = self._out(out, host)
        return res[0] if scalar else res

    def prev(self, index: Any) -> Any:
        """manager.py:93-108 / :306-331."""
        return self._neighbour(index, False)

    def next(self, index: Any) -> Any:
        """manager.py:110-125 / :334-358."""
        return self._neighbour(index, True)

    def get_buffer_indices(self, start: int, stop: int) -> np.ndarray:
        """buffer_base.py:173-228 (host integer arithmetic on two scalars; no device work)."""
        edges = self._extend_offset
        s_edge = int(np.searchsorted(edges, start, side="right")) - 1
        e_edge = int(np.searchsorted(edges, stop - 1, side="right")) - 1
        if s_edge != e_edge:
            raise ValueError(
                f"Start and stop indices must be within the same subbuffer. Got {start=} in subbuffer edge "
                f"{s_edge} and {stop=} in subbuffer edge {e_edge}.")
        if stop >= start:
            return np.arange(start, stop, dtype=int)
        return np.concatenate((np.arange(start, edges[s_edge + 1], dtype=int), np.arange(edges[s_edge], stop, dtype=int)))

    # -- data access ----------------------------------------------------------------------------
    def _gather(self, idx: torch.Tensor, keys: tuple[str, ...], extras: dict[str, torch.Tensor] | None = None) -> dict[str, torch.Tensor]:
        n, dev, s = idx.numel(), self.device, self._store
        O = self._desc.obs_dim
        out: dict[str, torch.Tensor] = {}
        mk = {"obs": (lambda: torch.empty(n, O, dtype=torch.float32, device=dev)),
              "obs_next": (lambda: torch.empty(n, O, dtype=torch.float32, device=dev)),
              "act": (lambda: torch.empty(n, dtype=torch.int32, device=dev)),
              "rew": (lambda: torch.empty(n, dtype=torch.float32, device=dev)),
              "logp": (lambda: torch.empty(n, dtype=torch.float32, device=dev)),
              "value": (lambda: torch.empty(n, dtype=torch.float32, device=dev))}
        for k in keys:
            if k in s:
                out[k] = mk[k]() if k in mk else torch.empty(n, dtype=torch.uint8, device=dev)
        extras = extras or {}
        ex_out = {k: torch.empty(n, dtype=torch.float32, device=dev) for k in extras}
        ne = len(extras)
        arr_in = (C.c_void_p * max(ne, 1))(*[v.data_ptr() for v in extras.values()])
        arr_out = (C.c_void_p * max(ne, 1))(*[v.data_ptr() for v in ex_out.values()])
        P = lambda k: _lib.ptr(out.get(k))  # noqa: E731
        _lib.check(_lib.lib().tsm_buf_gather(
            C.byref(self._desc), _lib.ptr(idx), n, P("obs"), P("obs_next"), P("act"), P("rew"), P("terminated"),
            P("truncated"), P("done"), P("logp"), P("value"), C.cast(arr_in, C.c_void_p), C.cast(arr_out, C.c_void_p),
            ne, _lib.stream_ptr()))
        out.update(ex_out)
        return out

    def _finish(self, k: str, t: torch.Tensor, host: bool) -> Any:
        if k in ("obs", "obs_next") and self._obs_shape is not None:
            t = t.reshape(t.shape[0], *self._obs_shape)
        elif k in ("terminated", "truncated", "done"):
            t = t.view(torch.bool)
        elif k == "act":
            t = t.to(torch.int64)
        if host:
            a = t.cpu().numpy()
            return a.astype(self._act_dtype) if k == "act" else a
        return t

    def __getitem__(self, index: Any) -> Batch:
        """buffer_base.py:591-635: gather every stored key at env-major `index`."""
        if isinstance(index, str):
            return self.__getattr__(index)
        host = _is_host(index)
        if isinstance(index, slice):
            index = np.arange(self.maxsize)[index]
        scalar = np.isscalar(index)
        idx = self._dev(np.atleast_1d(index) if host else index, torch.int64).reshape(-1)
        keys = tuple(k for k in ("obs", "obs_next", "act", "rew", "terminated", "truncated", "done", "logp", "value") if k in self._store)
        g = self._gather(idx, keys)
        out = Batch()
        for k in keys:
            if k in ("logp", "value"):
                continue
            v = self._finish(k, g[k], host)
            out.__dict__[k] = v[0] if scalar else v
        if "act_f" in self._store:  # continuous actions (see add)
            af = self._store["act_f"]
            v = af.view(-1, af.shape[2])[(idx % self._size) * self.buffer_num + idx // self._size]
            v = v.cpu().numpy() if host else v
            out.__dict__["act"] = v[0] if scalar else v
        if not self._save_obs_next and "obs" in g:
            nxt = self._dev(self.next(idx), torch.int64)
            v = self._finish("obs", self._gather(nxt, ("obs",))["obs"], host)
            out.__dict__["obs_next"] = v[0] if scalar else v
        if "logp" in g:
            pol = Batch()
            for k in ("logp", "value"):
                v = self._finish(k, g[k], host)
                pol.__dict__[k] = v[0] if scalar else v
            out.__dict__["policy"] = pol
        for k, t in self.__dict__.get("_extra", {}).items():
            v = t[idx]
            v = v.cpu().numpy() if host else v
            out.__dict__[k] = v[0] if scalar else v
        out.__dict__.setdefault("info", Batch())
        return out

    def set_array_at_key(self, seq: Any, key: str, index: Any = None, default_value: float | None = None) -> None:
        """buffer_base.py:637-644 / batch.py set_array_at_key: write `seq` into the extra key `key` at env-major `index` (None: the
        whole key); a key that does not exist yet is created for all slots, filled with default_value (NaN when None) -- what
        the collector's episode hooks use to attach per-transition results such as MC returns (collector.py:1000-1011)."""
        if key in _RESERVED:
            raise ValueError(f"{key!r} is a reserved buffer key")
        extra = self.__dict__.setdefault("_extra", {})
        arr = torch.as_tensor(np.asarray(seq) if not isinstance(seq, torch.Tensor) else seq).to(device=self.device, dtype=torch.float64)
        if index is None:
            if arr.shape[0] != self.maxsize:
                raise ValueError("a whole-key assignment must cover every slot")
            extra[key] = arr.clone()
            return
        if key not in extra:
            fill = float("nan") if default_value is None else float(default_value)
            extra[key] = torch.full((self.maxsize, *arr.shape[1:]), fill, dtype=torch.float64, device=self.device)
        idx = self._dev(np.atleast_1d(index) if _is_host(index) else index, torch.int64).reshape(-1)
        extra[key][idx] = arr

    def sample(self, batch_size: int | None) -> tuple[Batch, Any]:
        """buffer_base.py:533-541."""
        indices = self.sample_indices(batch_size)
        return self[indices], indices

    def __getattr__(self, key: str) -> Any:
        """buf.obs / buf.done / ... (buffer_base.py:237-241): the whole stored array, numbered
        env-major like the reference.  This is a gathered COPY of the time-major device storage."""
        store = self.__dict__.get("_store")
        extra = self.__dict__.get("_extra")
        if extra is not None and key in extra:
            return extra[key].cpu().numpy() if self.index_type == "numpy" else extra[key]
        if store is None or key not in _RESERVED or key not in store:
            raise AttributeError(key)
        t = store[key]
        t = t.permute(1, 0, 2).reshape(self.maxsize, -1) if t.ndim == 3 else t.t().reshape(self.maxsize)
        return self._finish(key, t.contiguous(), self.index_type == "numpy")


class ReplayBuffer(VectorReplayBuffer):
    """Single sub-buffer (tianshou/data/buffer/buffer_base.py:20-656): accepts one unbatched
    transition per `add` like the reference (`buf.add(Batch(obs=..., act=..., rew=..., ...))`)."""

    def __init__(self, size: int, **kwargs: Any) -> None:
        super().__init__(size, 1, **kwargs)

    def add(self, batch: Any, buffer_ids: Any = None) -> tuple[Any, Any, Any, Any]:
        if buffer_ids is None:
            b = Batch()
            for k, v in batch.items():
                if isinstance(v, Batch):
                    if len(v.__dict__):
                        b.__dict__[k] = Batch({kk: (vv[None] if hasattr(vv, "shape") else vv) for kk, vv in v.items()})
                elif isinstance(v, torch.Tensor):
                    b.__dict__[k] = v[None]
                else:
                    b.__dict__[k] = np.asarray(v)[None]
            batch = b
        return super().add(batch, buffer_ids=None if buffer_ids is None else buffer_ids)
