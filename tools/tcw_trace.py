"""clock64 timeline of one tcw_kernel launch (engine 2), CTA 0: `python tools/tcw_trace.py build` here (cross-compiles a
-DTSM_TCW_TRACE variant into tools/_variants/), `python tools/tcw_trace.py run [actor|critic]` on the GPU box."""
import ctypes as C
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
VLIB = os.path.join(ROOT, "tools", "_variants", "libtsm_tcw_trace.so")

if sys.argv[1] == "build":
    os.makedirs(os.path.dirname(VLIB), exist_ok=True)
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "tianshou_marl_b200", "csrc"), "-j8", "BUILD=build_tcwtrace", f"OUT={VLIB}",
                           "EXTRA=-DTSM_TCW_TRACE" + (" -DTSM_TCW_EXPERIMENT_STORE" if len(sys.argv) > 2 and sys.argv[2] == "store" else "")])
    sys.exit(0)

os.environ["TSM_B200_LIB"] = VLIB
import numpy as np  # noqa: E402
import torch  # noqa: E402

from tianshou_marl_b200 import _lib  # noqa: E402
from tianshou_marl_b200.utils import ActorCriticNet  # noqa: E402

which = sys.argv[2] if len(sys.argv) > 2 else "actor"
L = _lib.lib()
L.tsm_tcw_trace_read.restype, L.tsm_tcw_trace_read.argtypes = C.c_int, [C.c_void_p]
P = _lib.ptr
if which == "actor":
    n = 128 * 510
    net = ActorCriticNet(60, 128, 5, with_value=False, engine="tcw", seed=0)
    x = torch.randn(n, 60, device="cuda")
    loss = _lib.TsmTcwLoss(kind=0, act=P(torch.randint(0, 5, (n,), device="cuda", dtype=torch.int32)), adv=P(torch.randn(n, device="cuda")),
                           logp_old=P(torch.full((n,), -1.6, device="cuda")), target=None, v_old=None, weight=None, heads=0)
else:
    n = 128 * 51
    net = ActorCriticNet(600, 128, 0, n_value_only=10, engine="tcw", seed=0)
    x = torch.randn(n, 600, device="cuda")
    loss = _lib.TsmTcwLoss(kind=1, act=None, adv=None, logp_old=None, target=P(torch.randn(n, 10, device="cuda")), v_old=None, weight=None, heads=10)
ximg = net.pack_x(x)
cfg = _lib.TsmPpoCfg(eps_clip=0.2, dual_clip=0.0, vf_coef=0.5, ent_coef=0.01, value_clip=0, adv_norm=0, plain_pg=0, value_heads=0)
tiles = _lib.TsmTcwTiles(ximg=P(ximg), n_rows=n, super_ids=None, sub=1, tile0=0, n_tiles=n // 128)
sums = torch.zeros(4, dtype=torch.float64, device="cuda")
for _ in range(3):
    _lib.check(L.tsm_tcw_train(net.desc, C.byref(tiles), C.byref(loss), C.byref(cfg), 1.0 / n, 65536.0, None, P(sums), _lib.stream_ptr()))
h = (C.c_longlong * 192)()
_lib.check(L.tsm_tcw_trace_read(h))
t = np.array(h[:], dtype=np.int64)
t0 = t[0]
print(f"{which}: setup {t[1] - t0} cycles; kernel end (all threads past the last sync) {t[191] - t0}; flush wait {t[63] - t[62]}, flush+exit {t[191] - t[63]}")
names = ["tile start", "L1 done seen", "H1 written", "L2 done seen", "H2 written", "L3 done seen", "G written", "DH2 done seen", "D2 written",
         "DH1 done seen", "D1 written"]
for i in range(5):
    b = 2 + 12 * i
    if t[b] == 0 or t[b + 10] == 0:
        break
    print(f"epilogue tile {i}: start +{t[b] - t0}: " + " ".join(f"{names[k + 1]} {t[b + k + 1] - t[b + k]}" for k in range(10)) + f" | total {t[b + 10] - t[b]}")
inames = ["tile start", "L1 issued", "H1 ready", "L2 issued", "H2 ready", "L3 issued", "bwd issued (from L3 issued)", "D1 ready"]
for i in range(5):
    b = 64 + 12 * i
    if t[b] == 0:
        break
    print(f"issuer tile {i}: start +{t[b] - t0}: " + " ".join(f"{inames[k + 1]} {t[b + k + 1] - t[b + k]}" for k in range(7)))
print("issuer done +", t[127] - t0)
if which == "critic":
    for k in range(10):
        e, q = 128 + 3 * k, 160 + 3 * k
        print(f"W1-gradient block {k}: issuer full seen +{t[q] - t0} free seen +{t[q + 1] - t0} issued +{t[q + 2] - t0} | epilogue done seen +{t[e] - t0} staged +{t[e + 1] - t0} reduced +{t[e + 2] - t0}")
