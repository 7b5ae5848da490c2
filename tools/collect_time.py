"""Lock-step rollout of the C2 workload: wall time of Collector.collect (host sync included) against the host time spent inside
tsm_collect_lockstep (the enqueue of 2 launches per world range and step).  TSM_COLLECT_NO_PDL=1 / TSM_COLLECT_STREAMS=n apply."""
import argparse
import sys
import time

import torch

sys.path.insert(0, ".")
from tianshou_marl_b200 import _lib  # noqa: E402
from tianshou_marl_b200.bench_workload import Workload  # noqa: E402

a = argparse.Namespace(envs=65536, agents=3, horizon=128, hidden=64, engine="tc", no_update=True, batch_size=65536, repeat=4)
wl = Workload(a, 0, 1)
L = _lib.lib()
orig = L.tsm_collect_lockstep
spent = [0.0]


def timed(*args):
    t0 = time.perf_counter()
    rc = orig(*args)
    spent[0] += time.perf_counter() - t0
    return rc


L.tsm_collect_lockstep = timed
for _ in range(3):
    wl.step()
torch.cuda.synchronize()
for rep in range(3):
    spent[0] = 0.0
    t0 = time.perf_counter()
    n = 5
    for _ in range(n):
        wl.step()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / n
    print(f"collect-only step: wall {1e3 * wall:.2f} ms, host inside tsm_collect_lockstep {1e3 * spent[0] / n:.2f} ms")
