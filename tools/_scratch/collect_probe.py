import sys, time, torch
sys.argv = ["bench.py"]
sys.path.insert(0, "/root/repo")
import bench
from tianshou_marl_b200.bench_workload import Workload
args = bench.parse()
wl = Workload(args, 0, 1)
c = wl.collector
def coll(): c.collect(n_step=wl.R * wl.T); c.reset_buffer(keep_statistics=True)
def T(f, n=5):
    for _ in range(2): f()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
print("collect ms", T(coll))
