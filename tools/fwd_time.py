"""Engine 1 policy forward (tc_forward_kernel) at rollout-step sizes: GPU time per launch against the number of 128-row tiles.
python tools/fwd_time.py [rows ...]   (default: one wave of tiles at 3 / 5 CTAs per SM, the C2 rollout range, two waves, a full step)"""
import ctypes as C
import sys

import torch

sys.path.insert(0, ".")
from tianshou_marl_b200 import _lib  # noqa: E402
from tianshou_marl_b200.utils import ActorCriticNet  # noqa: E402

dev = torch.device("cuda")
O, NA = 18, 5
sizes = [int(a) for a in sys.argv[1:]] or [148 * 128, 444 * 128, 740 * 128, 98304, 888 * 128, 196608]
net = ActorCriticNet(O, 64, NA, seed=0, engine="tc")
L, P = _lib.lib(), _lib.ptr
st = _lib.stream_ptr()
for n in sizes:
    x = torch.randn(n, O, device=dev)
    act = torch.empty(n, dtype=torch.int32, device=dev)
    logp, val = torch.empty(n, device=dev), torch.empty(n, device=dev)
    reps = 200
    for it in range(2):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(reps):
            _lib.check(L.tsm_net_forward(net.desc, P(x), None, n, 1, 0, 7, k, None, P(act), P(logp), P(val), None, st))
        e1.record()
        torch.cuda.synchronize()
    print(f"rows {n:7d} tiles {(n + 127) // 128:5d}: {1e3 * e0.elapsed_time(e1) / reps:.2f} us per launch (back to back)")
