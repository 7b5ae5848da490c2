"""Oracle: GAE / discounted returns.  Test infrastructure only (see oracle/__init__.py).

Follows tianshou/algorithm/algorithm_base.py:
  _gae                          :1079-1134
  Algorithm.value_mask          :631-649
  Algorithm.compute_episodic_return :651-717
  episode_mc_return_to_go       :1137-1151
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def _clib():
    """liboracle.so built by oracle/Makefile (optional accelerator for large sizes)."""
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "_build", "liboracle.so")
        if not os.path.exists(path):
            _LIB = False
        else:
            lib = ctypes.CDLL(path)
            lib.orc_gae.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_int64, ctypes.c_double, ctypes.c_double, ctypes.c_void_p]
            lib.orc_gae.restype = None
            _LIB = lib
    return _LIB or None


def gae(v_s, v_s_, rew, end_flag, gamma: float, gae_lambda: float, use_c: bool = True) -> np.ndarray:
    """`_gae` (algorithm_base.py:1126-1134): serial reverse recurrence, float64 result.

    delta = rew + v_s_*gamma - v_s ; discount = (1-end)*gamma*lambda ;
    gae_i = delta_i + discount_i * gae_{i+1}.
    Input dtypes are kept as the reference has them (v_s, v_s_ float32 or float64, rew
    float64); numpy promotion then reproduces numba's float64 arithmetic exactly.
    """
    v_s = np.ascontiguousarray(v_s)
    v_s_ = np.ascontiguousarray(v_s_)
    rew = np.ascontiguousarray(rew)
    end_flag = np.ascontiguousarray(end_flag)
    n = rew.shape[0]
    lib = _clib() if use_c else None
    if lib is not None and n > 64:
        v64 = v_s.astype(np.float64)
        vn64 = v_s_.astype(np.float64)
        r64 = rew.astype(np.float64)
        e8 = end_flag.astype(np.uint8)
        out = np.zeros(n, np.float64)
        lib.orc_gae(v64.ctypes.data, vn64.ctypes.data, r64.ctypes.data, e8.ctypes.data, n, float(gamma), float(gae_lambda), out.ctypes.data)
        return out
    returns = np.zeros(rew.shape)
    delta = rew + v_s_ * np.float64(gamma) - v_s
    discount = (1.0 - end_flag) * (gamma * gae_lambda)
    g = 0.0
    for i in range(n - 1, -1, -1):
        g = delta[i] + discount[i] * g
        returns[i] = g
    return returns


def compute_episodic_return(rew, terminated, truncated, indices, unfinished_index, buffer_terminated,
                            v_s_=None, v_s=None, gamma: float = 0.99, gae_lambda: float = 0.95):
    """`Algorithm.compute_episodic_return` (algorithm_base.py:700-717) on plain arrays.

    `buffer_terminated` is the buffer's whole `terminated` array (value_mask indexes it by
    `indices`, :649); `unfinished_index` is `buffer.unfinished_index()`.
    Returns (returns, advantage), both float64.
    """
    rew = np.asarray(rew)
    if v_s_ is None:
        assert np.isclose(gae_lambda, 1.0)
        v_s_ = np.zeros_like(rew)
    else:
        v_s_ = np.asarray(v_s_).flatten()
        v_s_ = v_s_ * ~np.asarray(buffer_terminated).astype(bool)[indices]
    v_s = np.roll(v_s_, 1) if v_s is None else np.asarray(v_s).flatten()
    end_flag = np.logical_or(terminated, truncated)
    end_flag[np.isin(indices, unfinished_index)] = True
    advantage = gae(v_s, v_s_, rew, end_flag, gamma, gae_lambda)
    returns = advantage + v_s
    return returns, advantage


def episode_mc_return_to_go(rewards, gamma: float = 0.99) -> np.ndarray:
    """algorithm_base.py:1137-1151."""
    rewards = np.asarray(rewards, dtype=np.float64)
    n = len(rewards)
    out = np.zeros(n)
    out[-1] = rewards[-1]
    for j in range(n - 2, -1, -1):
        out[j] = rewards[j] + gamma * out[j + 1]
    return out
