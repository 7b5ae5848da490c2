"""Oracle: Philox4x32-10 counter-based RNG, restated in numpy.  Test infrastructure only.

This is NOT reference code: the reference draws reset positions from numpy's PCG64 and
actions from torch's global generator, neither of which a GPU kernel can reproduce.  The
CUDA kernels (csrc/philox.cuh) use Philox4x32-10 keyed by (seed, stream) with counter
(index_lo, index_hi, sub, block); this file restates that bit-for-bit so tests can compare
device draws with host draws.
"""
from __future__ import annotations

import numpy as np

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = np.uint32(0x9E3779B9)
_W1 = np.uint32(0xBB67AE85)
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10.  All args broadcastable uint32 arrays.  Returns 4 uint32 arrays."""
    c0, c1, c2, c3, k0, k1 = (np.asarray(x).astype(np.uint32) for x in (c0, c1, c2, c3, k0, k1))
    c0, c1, c2, c3, k0, k1 = np.broadcast_arrays(c0, c1, c2, c3, k0, k1)
    c0, c1, c2, c3, k0, k1 = (x.copy() for x in (c0, c1, c2, c3, k0, k1))
    with np.errstate(over="ignore"):
        for r in range(10):
            p0 = _M0 * c0.astype(np.uint64)
            p1 = _M1 * c2.astype(np.uint64)
            hi0 = (p0 >> np.uint64(32)).astype(np.uint32)
            lo0 = (p0 & _MASK).astype(np.uint32)
            hi1 = (p1 >> np.uint64(32)).astype(np.uint32)
            lo1 = (p1 & _MASK).astype(np.uint32)
            c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
            if r < 9:
                k0 = k0 + _W0
                k1 = k1 + _W1
    return c0, c1, c2, c3


def u01(x) -> np.ndarray:
    """uint32 -> float32 uniform in [0,1) with 24 bits: (x >> 8) * 2^-24 (exact in f32)."""
    return (np.asarray(x, dtype=np.uint32) >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)


STREAM_RESET = 0x52455345  # 'RESE'  key1 for env reset draws
STREAM_ACT = 0x41435421  # 'ACT!'  key1 for action sampling draws
STREAM_PERM = 0x5045524D  # 'PERM'  key1 for minibatch permutation keys


# ----------------------------------------------------------------------------------------------
# keyed Feistel permutation of [0, n) (csrc/feistel.cuh), restated for tests
def _mix32(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.uint32)
    with np.errstate(over="ignore"):
        x = x ^ (x >> np.uint32(16))
        x = x * np.uint32(0x7FEB352D)
        x = x ^ (x >> np.uint32(15))
        x = x * np.uint32(0x846CA68B)
        x = x ^ (x >> np.uint32(16))
    return x


def feistel_bits(n: int) -> int:
    bits = 2
    while bits < 62 and (1 << bits) < n:
        bits += 1
    return bits


def feistel_permute(x, n: int, key: int, rounds: int = 6) -> np.ndarray:
    """Alternating unbalanced Feistel network over ceil(log2 n) bits with cycle walking."""
    x = np.asarray(x, dtype=np.uint64).copy()
    bits = feistel_bits(n)
    a = bits >> 1
    b = bits - a
    mask_a, mask_b = np.uint32((1 << a) - 1), np.uint32((1 << b) - 1)
    k0, k1 = np.uint32(key & 0xFFFFFFFF), np.uint32((key >> 32) & 0xFFFFFFFF)
    todo = np.ones(x.shape, bool)
    while todo.any():
        v = x[todo]
        lo = (v & np.uint64(int(mask_a))).astype(np.uint32)
        hi = ((v >> np.uint64(a)) & np.uint64(int(mask_b))).astype(np.uint32)
        for r in range(rounds):
            with np.errstate(over="ignore"):
                kr = _mix32(np.asarray(k0 + np.uint32(0x9E3779B9) * np.uint32(r + 1), dtype=np.uint32)) ^ k1
            if r & 1:
                lo = lo ^ (_mix32(hi ^ kr) & mask_a)
            else:
                hi = hi ^ (_mix32(lo ^ kr) & mask_b)
        v = (hi.astype(np.uint64) << np.uint64(a)) | lo.astype(np.uint64)
        x[todo] = v
        todo[todo] = v >= np.uint64(n)
    return x.astype(np.int64)
