import sys, time, torch
sys.argv = ["bench.py"]
sys.path.insert(0, "/root/repo")
import bench
from tianshou_marl_b200.bench_workload import Workload
args = bench.parse()
wl = Workload(args, 0, 1)
def T(f, n=3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
for _ in range(3): wl.step()
print("step        ", T(wl.step))
for _ in range(2): wl.step_e2e()
print("step_e2e    ", T(wl.step_e2e))
c = wl.collector
def coll(): c.collect(n_step=wl.R * wl.T); c.reset_buffer(keep_statistics=True)
print("collect no list", T(coll))
c.episode_list = True
print("collect list   ", T(coll))
def coll2():
    cs = c.collect(n_step=wl.R * wl.T)
    t0 = time.perf_counter(); wl.alg.update(wl.buf, args.batch_size, args.repeat); torch.cuda.synchronize(); coll2.u += time.perf_counter() - t0
    c.reset_buffer(keep_statistics=True)
coll2.u = 0.0
print("collect list + update", T(coll2), "update part", coll2.u / 3 * 1e3)
c.episode_list = False
coll2.u = 0.0
print("collect nolist + update", T(coll2), "update part", coll2.u / 3 * 1e3)
