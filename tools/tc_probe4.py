"""16-bit operands, no-swizzle: K-major and MN-major views of the same [f/8][r][8] image."""
import ctypes as C
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from tianshou_marl_b200 import _lib  # noqa: E402


def img16(a, dt):  # [rows][F] -> [F/8][rows][8] 16-bit, returned as float32 words
    r, f = a.shape
    t = torch.as_tensor(a).to(dt).reshape(r, f // 8, 8).permute(1, 0, 2).contiguous()
    return t.view(torch.float32).reshape(-1).numpy(), t.permute(1, 0, 2).reshape(r, f).float().numpy()


def run(a_img, b_img, desc, a_mn, b_mn, n, ksteps, a_bf16, b_bf16):
    d = torch.full((128, n), float("nan"), device="cuda")
    A, B = torch.as_tensor(a_img, device="cuda"), torch.as_tensor(b_img, device="cuda")
    h = (C.c_uint32 * 6)(*desc)
    _lib.check(_lib.lib().tsm_tc_selftest(A.data_ptr(), A.numel(), B.data_ptr(), B.numel(), C.cast(h, C.c_void_p), a_mn, b_mn, n, ksteps,
                                          1, a_bf16, b_bf16, d.data_ptr(), _lib.stream_ptr()))
    torch.cuda.synchronize()
    return d.cpu().numpy()


rng = np.random.default_rng(0)
which = sys.argv[1:]
GD = torch.bfloat16 if "bf" in which else torch.float16
GB = 1 if "bf" in which else 0
N = 64
if "1" in which:  # forward: A [128 rows][K=32] fp16 K-major, B weights [N out][K in] fp16 K-major
    K = 32
    ai, a = img16(rng.normal(size=(128, K)).astype(np.float32), torch.float16)
    bi, w = img16(rng.normal(size=(N, K)).astype(np.float32), torch.float16)
    d = run(ai, bi, [2048, 128, 4096, N * 16, 128, 2 * N * 16], 0, 0, N, K // 16, 0, 0)
    print("f16 K-major x K-major: max err", np.abs(d - a.astype(np.float64) @ w.astype(np.float64).T).max())
if "2" in which:  # data-grad: A = dOut [128][16] bf16 K-major, B = weights image [in=64 /8][out=16][8] fp16 read MN-major (N = in, K = out)
    O = 16
    ai, a = img16(rng.normal(size=(128, O)).astype(np.float32), GD)
    bi, w = img16(rng.normal(size=(O, N)).astype(np.float32), torch.float16)  # W[out][in]
    d = run(ai, bi, [2048, 128, 4096, 128, O * 16, 256], 0, 1, N, O // 16, GB, 0)
    print("bf16 K-major x f16 MN-major: max err", np.abs(d - a.astype(np.float64) @ w.astype(np.float64)).max())
if "3" in which:  # weight-grad: A = P [128 rows][128 feats] bf16 MN-major, B = Q [128 rows][64 feats] fp16 MN-major, K = rows
    pi, p = img16(rng.normal(size=(128, 128)).astype(np.float32), GD)
    qi, q = img16(rng.normal(size=(128, N)).astype(np.float32), torch.float16)
    d = run(pi, qi, [128, 2048, 256, 128, 2048, 256], 1, 1, N, 128 // 16, GB, 0)
    print("bf16 MN-major x f16 MN-major (K = rows): max err", np.abs(d - p.astype(np.float64).T @ q.astype(np.float64)).max())
