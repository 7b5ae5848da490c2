"""cProfile of Workload.step_e2e (host-side overhead of the public API path)."""
import argparse, cProfile, pstats, torch
from tianshou_marl_b200.bench_workload import Workload
a = argparse.Namespace(envs=65536, agents=3, horizon=128, hidden=64, engine="tc", no_update=False, batch_size=65536, repeat=4)
wl = Workload(a, 0, 1)
for _ in range(2):
    wl.step_e2e()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(3):
    wl.step_e2e()
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(30)
