import sys; sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import ctypes as C, numpy as np, torch
from tianshou_marl_b200 import _lib
from test_gpu_tc import _h
rng = np.random.default_rng(0)
bad = 0
for trial in range(200):
    N = 64
    pi, p = _h(rng.normal(size=(128, 128)).astype(np.float32))
    qi, q = _h(rng.normal(size=(128, N)).astype(np.float32))
    d = torch.full((128, N), float("nan"), device="cuda")
    A, B = torch.as_tensor(pi, device="cuda"), torch.as_tensor(qi, device="cuda")
    h = (C.c_uint32 * 6)(128, 2048, 256, 128, 2048, 256)
    _lib.check(_lib.lib().tsm_tc_selftest(A.data_ptr(), A.numel(), B.data_ptr(), B.numel(), C.cast(h, C.c_void_p), 1, 1, N, 8, 3, 0, 0, d.data_ptr(), _lib.stream_ptr()))
    torch.cuda.synchronize()
    want = 64.0 * (p.T @ q)
    err = np.abs(d.cpu().numpy() - want).max() / np.abs(want).max()
    if err > 1e-4: bad += 1; print("trial", trial, "rel err", err)
print("bad trials:", bad, "of 200")
