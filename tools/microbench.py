"""Kernel micro-timings (CUDA events) used while developing; bench.py is the contract bench."""
import json
import sys

import torch

sys.path.insert(0, ".")
from tianshou_marl_b200.algorithm import gae_flat, gae_time_major  # noqa: E402


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def main():
    T, R = 128, 1 << 20
    n = T * R
    g = torch.Generator(device="cuda").manual_seed(0)
    rew, v_s, v_n = (torch.randn(T, R, device="cuda", generator=g) for _ in range(3))
    trunc = (torch.rand(T, R, device="cuda", generator=g) < 0.04).to(torch.uint8)
    term = (torch.rand(T, R, device="cuda", generator=g) < 0.01).to(torch.uint8)
    adv, ret = torch.empty_like(rew), torch.empty_like(rew)
    out = {}
    med, best = timeit(lambda: gae_time_major(v_s, v_n, rew, term, trunc, 0, T, 0.99, 0.95, adv, ret))
    out["gae_tm"] = {"ms": med, "best_ms": best, "GBps": 22 * n / med / 1e6}
    f = lambda x: x.view(-1)  # noqa: E731
    med, best = timeit(lambda: gae_flat(f(v_s), f(v_n), f(rew), f(term), f(trunc), None, T, 0.99, 0.95))
    out["gae_flat"] = {"ms": med, "best_ms": best, "GBps": 22 * n / med / 1e6}
    fe = torch.zeros(n, dtype=torch.uint8, device="cuda")
    fe[T - 1::T] = 1  # what compute_episodic_return passes: no seg_len, the unfinished slots as force_end
    med, best = timeit(lambda: gae_flat(f(v_s), f(v_n), f(rew), f(term), f(trunc), fe, 0, 0.99, 0.95))
    out["gae_flat_force_end"] = {"ms": med, "best_ms": best, "GBps": 23 * n / med / 1e6}
    a, b = torch.empty(1 << 29, device="cuda", dtype=torch.float32), torch.empty(1 << 29, device="cuda", dtype=torch.float32)
    med, best = timeit(lambda: b.copy_(a))
    out["copy_4GiB"] = {"ms": med, "GBps": 2 * a.numel() * 4 / med / 1e6}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
