"""GPU idle time inside one device-resident C2 step: kernel timeline from CUPTI (torch.profiler), gaps between consecutive
kernels / copies on the device listed with the kernels either side.  python tools/step_gaps.py [min_gap_us]"""
import argparse
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, ".")
from tianshou_marl_b200.bench_workload import Workload  # noqa: E402

min_gap = float(sys.argv[1]) if len(sys.argv) > 1 else 15.0
a = argparse.Namespace(envs=65536, agents=3, horizon=128, hidden=64, engine="tc", no_update=False, batch_size=65536, repeat=4)
wl = Workload(a, 0, 1)
for _ in range(3):
    wl.step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(2):
        wl.step()
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
t0, t1 = ev[0].time_range.start, max(e.time_range.end for e in ev)
busy_end = ev[0].time_range.end
idle = 0.0
rows = []
for prev, e in zip(ev, ev[1:]):
    gap = e.time_range.start - busy_end
    if gap > 0:
        idle += gap
        if gap >= min_gap:
            rows.append((gap, prev.name[:60], e.name[:60], e.time_range.start - t0))
    busy_end = max(busy_end, e.time_range.end)
print(f"2 steps: span {(t1 - t0) / 1e3:.2f} ms, device idle {idle / 1e3:.2f} ms in total")
for gap, a_, b_, at in rows:
    print(f"  at {at / 1e3:8.2f} ms: {gap:7.1f} us idle between [{a_}] and [{b_}]")
