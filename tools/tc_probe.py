"""Exploration of tcgen05 no-swizzle descriptor semantics (results pinned by tests/test_gpu_tc.py)."""
import ctypes as C
import itertools
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from tianshou_marl_b200 import _lib  # noqa: E402


def act_img(a):  # [rows][F] -> [F/4][rows][4]
    r, f = a.shape
    return np.ascontiguousarray(a.reshape(r, f // 4, 4).transpose(1, 0, 2)).reshape(-1)


def w_img(w):  # [out][in] -> [in/4][out][4]
    return act_img(w)


def run(a_img, b_img, desc, a_mn, b_mn, n, ksteps):
    d = torch.full((128, n), float("nan"), device="cuda")
    A = torch.as_tensor(a_img, device="cuda")
    B = torch.as_tensor(b_img, device="cuda")
    h = (C.c_uint32 * 6)(*desc)
    _lib.check(_lib.lib().tsm_tc_selftest(A.data_ptr(), A.numel(), B.data_ptr(), B.numel(), C.cast(h, C.c_void_p), a_mn, b_mn, n, ksteps,
                                          d.data_ptr(), _lib.stream_ptr()))
    torch.cuda.synchronize()
    return d.cpu().numpy()


def tf32(x):
    return (x.view(np.uint32) & np.uint32(0xFFFFE000)).view(np.float32)


rng = np.random.default_rng(0)
which = sys.argv[1:] or ["1", "2", "3"]
K, N = 32, 64
if "1" in which:
    A = rng.normal(size=(128, K)).astype(np.float32)
    W = rng.normal(size=(N, K)).astype(np.float32)
    want = tf32(A).astype(np.float64) @ tf32(W).astype(np.float64).T
    d = run(act_img(A), w_img(W), [2048, 128, 4096, N * 16, 128, 2 * N * 16], 0, 0, N, K // 8)
    print(f"GEMM1 K-major A x K-major B: max err {np.abs(d - want).max():.4g}")
if "2" in which:
    O = 16
    A2 = rng.normal(size=(128, O)).astype(np.float32)
    W2 = rng.normal(size=(O, N)).astype(np.float32)  # [out=o][in=j]
    want2 = tf32(A2).astype(np.float64) @ tf32(W2).astype(np.float64)
    d = run(act_img(A2), w_img(W2), [2048, 128, 4096, 128, O * 16, 128], 0, 1, N, O // 8)
    print(f"GEMM2 K-major A x MN-major B (lbo=128, sbo={O*16}): max err {np.abs(d - want2).max():.4g}")
if "3" in which:
    P = rng.normal(size=(128, 128)).astype(np.float32)
    Q = rng.normal(size=(128, N)).astype(np.float32)
    want3 = tf32(P).astype(np.float64).T @ tf32(Q).astype(np.float64)
    d = run(act_img(P), act_img(Q), [128, 2048, 128, 128, 2048, 128], 1, 1, N, 128 // 8)
    print(f"GEMM3 MN-major A x MN-major B (lbo=128, sbo=2048): max err {np.abs(d - want3).max():.4g}")
if "2s" in which:  # swapped roles, to see which reading is right if "2" fails
    O = 16
    A2 = rng.normal(size=(128, O)).astype(np.float32)
    W2 = rng.normal(size=(O, N)).astype(np.float32)
    want2 = tf32(A2).astype(np.float64) @ tf32(W2).astype(np.float64)
    d = run(act_img(A2), w_img(W2), [2048, 128, 4096, O * 16, 128, 128], 0, 1, N, O // 8)
    print(f"GEMM2s (lbo={O*16}, sbo=128): max err {np.abs(d - want2).max():.4g}")
if "3s" in which:
    P = rng.normal(size=(128, 128)).astype(np.float32)
    Q = rng.normal(size=(128, N)).astype(np.float32)
    want3 = tf32(P).astype(np.float64).T @ tf32(Q).astype(np.float64)
    d = run(act_img(P), act_img(Q), [2048, 128, 128, 2048, 128, 128], 1, 1, N, 128 // 8)
    print(f"GEMM3s (lbo=2048, sbo=128): max err {np.abs(d - want3).max():.4g}")
