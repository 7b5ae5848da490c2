"""GAE / discounted returns on the device (K3).

Drop-in for `Algorithm.compute_episodic_return` (tianshou/algorithm/algorithm_base.py:651-717),
`Algorithm.value_mask` (:631-649), numba `_gae` (:1079-1134) and `episode_mc_return_to_go`
(:1137-1151).  The arithmetic runs in csrc/gae.cu; this module only moves arguments.
"""
from __future__ import annotations

from typing import Any

import numpy as np
import torch

from .. import _lib


def _device_of(*xs: Any) -> torch.device:
    for x in xs:
        if isinstance(x, torch.Tensor) and x.is_cuda:
            return x.device
    return torch.device("cuda")


_WS: dict[tuple[str, int], torch.Tensor] = {}


def _workspace(dev: torch.device, nbytes: int) -> torch.Tensor:
    key = (str(dev), 0)
    ws = _WS.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 1 << 16), dtype=torch.uint8, device=dev)
        _WS[key] = ws
    return ws


def gae_flat(v_s: torch.Tensor | None, v_next: torch.Tensor | None, rew: torch.Tensor, terminated: torch.Tensor,
             truncated: torch.Tensor, force_end: torch.Tensor | None = None, seg_len: int = 0, gamma: float = 0.99,
             gae_lambda: float = 0.95, want_returns: bool = True) -> tuple[torch.Tensor | None, torch.Tensor]:
    """One-pass segmented reverse scan over a flat device vector.  All value tensors share one
    dtype (float32 or float64); flags are uint8/bool.  Returns (returns, advantage)."""
    n = rew.numel()
    f64 = rew.dtype == torch.float64
    dt = torch.float64 if f64 else torch.float32
    dev = rew.device
    c = lambda t: None if t is None else t.to(dt).contiguous().view(-1)  # noqa: E731
    u8 = lambda t: None if t is None else (t.view(torch.uint8) if t.dtype == torch.bool else t.to(torch.uint8)).contiguous().view(-1)  # noqa: E731
    v_s, v_next, rew = c(v_s), c(v_next), c(rew)
    te, tr, fe = u8(terminated), u8(truncated), u8(force_end)
    adv = torch.empty(n, dtype=dt, device=dev)
    ret = torch.empty(n, dtype=dt, device=dev) if want_returns else None
    need = _lib.lib().tsm_gae_flat_workspace_bytes(n)
    ws = _workspace(dev, need)
    P = _lib.ptr
    _lib.check(_lib.lib().tsm_gae_flat(P(v_s), P(v_next), P(rew), int(f64), P(te), P(tr), P(fe), n, int(seg_len),
                                       float(gamma), float(gae_lambda), P(adv), P(ret), P(ws), ws.numel(),
                                       _lib.stream_ptr()))
    return ret, adv


def gae_time_major(v_s: torch.Tensor | None, v_next: torch.Tensor | None, rew: torch.Tensor, terminated: torch.Tensor,
                   truncated: torch.Tensor, start: int, length: int, gamma: float, gae_lambda: float,
                   adv_out: torch.Tensor | None = None, ret_out: torch.Tensor | None = None, v_scale: float | None = None,
                   ret_sums: torch.Tensor | None = None, ret_shift: float = 0.0) -> tuple[torch.Tensor, torch.Tensor]:
    """GAE over the device buffer's native [size][rows] float32 layout (tsm_gae_tm).  With v_scale (return_scaling,
    a2c.py:132-146) values are un-normalised on load, the returned returns are normalised again and ret_sums (2 device
    doubles) receives sum / sum of squares of the un-normalised returns (tsm_gae_tm_scaled)."""
    size, rows = rew.shape
    adv = adv_out if adv_out is not None else torch.empty_like(rew)
    ret = ret_out if ret_out is not None else torch.empty_like(rew)
    P = _lib.ptr
    u8 = lambda t: t.view(torch.uint8) if t.dtype == torch.bool else t  # noqa: E731
    if v_scale is not None:
        _lib.check(_lib.lib().tsm_gae_tm_scaled(P(v_s), P(v_next), P(rew), P(u8(terminated)), P(u8(truncated)), size, rows,
                                                int(start), int(length), float(gamma), float(gae_lambda), float(v_scale), float(ret_shift), P(adv),
                                                P(ret), P(ret_sums), _lib.stream_ptr()))
        return ret, adv
    _lib.check(_lib.lib().tsm_gae_tm(P(v_s), P(v_next), P(rew), P(u8(terminated)), P(u8(truncated)), size, rows,
                                     int(start), int(length), float(gamma), float(gae_lambda), P(adv), P(ret),
                                     _lib.stream_ptr()))
    return ret, adv


def compute_episodic_return(batch: Any, buffer: Any, indices: Any, v_s_: Any = None, v_s: Any = None,
                            gamma: float = 0.99, gae_lambda: float = 0.95) -> tuple[Any, Any]:
    """`Algorithm.compute_episodic_return` (algorithm_base.py:651-717).

    numpy / CPU inputs  -> float64 arithmetic on the device, numpy float64 results (like numba);
    device-tensor inputs -> float32 I/O with a float64 carry, device-tensor results.
    """
    rew = batch.rew
    host = not (isinstance(rew, torch.Tensor) and rew.is_cuda)
    dev = _device_of(rew, v_s_, v_s, getattr(buffer, "device", None) and torch.empty(0, device=buffer.device))
    dt = torch.float64 if host else torch.float32
    T = lambda x, d=dt: torch.as_tensor(x if isinstance(x, torch.Tensor) else np.ascontiguousarray(x), device=dev).to(d).reshape(-1)  # noqa: E731
    rew_t = T(rew)
    n = rew_t.numel()
    if v_s_ is None:
        assert np.isclose(gae_lambda, 1.0)  # algorithm_base.py:705
        vn_t = None
    else:
        vn_t = T(v_s_.flatten() if hasattr(v_s_, "flatten") else v_s_)
    term_t = T(batch.terminated, torch.bool)
    trunc_t = T(batch.truncated, torch.bool)
    if v_s is None:
        # :710  v_s = np.roll(v_s_, 1) with the value mask already applied (:709)
        if vn_t is None:
            vs_t = None
        else:
            vs_t = torch.roll(vn_t * (~term_t).to(dt), 1)
    else:
        vs_t = T(v_s.flatten() if hasattr(v_s, "flatten") else v_s)
    # :712-713  end_flag[np.isin(indices, buffer.unfinished_index())] = True
    unf = buffer.unfinished_index()
    unf_t = torch.as_tensor(unf, device=dev).to(torch.int64)
    force = None
    if unf_t.numel() > 0 and n > 0:
        idx_t = torch.as_tensor(indices, device=dev).to(torch.int64).reshape(-1)
        mask = torch.zeros(int(buffer.maxsize), dtype=torch.uint8, device=dev)
        mask[unf_t] = 1
        force = mask[idx_t]
    ret, adv = gae_flat(vs_t, vn_t, rew_t, term_t, trunc_t, force, 0, gamma, gae_lambda)
    if host:
        return ret.cpu().numpy(), adv.cpu().numpy()
    return ret, adv


def episode_mc_return_to_go(rewards: Any, gamma: float = 0.99) -> Any:
    """algorithm_base.py:1137-1151: K3 with lambda = 1 and zero values over one episode."""
    host = not (isinstance(rewards, torch.Tensor) and rewards.is_cuda)
    r = torch.as_tensor(rewards, device="cuda").to(torch.float64 if host else torch.float32).reshape(-1)
    z = torch.zeros(r.numel(), dtype=torch.uint8, device=r.device)
    _, adv = gae_flat(None, None, r, z, z, None, 0, gamma, 1.0, want_returns=False)
    return adv.cpu().numpy() if host else adv
