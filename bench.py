#!/usr/bin/env python
"""bench.py -- the contract bench of the hot path (see DESIGN.md "Measurement").

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # CPU arm (oracle port, host cores)

A "step" is one pass of the hot path over one batch of synthetic input: collect T vector steps
of `simple_spread_v3` over E environments per GPU (policy forward -> MPE world step -> replay
buffer add), GAE over the collected [T, E*A] block, and the PPO minibatch update (`repeat`
epochs).  Metric (BASELINE.json): agent-steps/s, whole job.  One process per GPU; environments
shard across ranks with no data-path collective during collect/GAE (scaling: weak); the PPO
update exchanges the flat gradient between ranks inside its persistent kernel over NVLink peer
memory (torch.distributed / NCCL only for the rendezvous and the timing all-reduce).

Nothing here reads /root/reference.  oracle/ is executed only by the `cpu_baseline` leg and
by `--impl reference`.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "agent-steps/sec collect+PPO update, simple_spread"
UNIT = "agent-steps/s"
GAE_BYTES_PER_ELT = 22  # read v_s 4 + v_next 4 + rew 4 + term 1 + trunc 1, write adv 4 + ret 4


def parse() -> argparse.Namespace:
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--envs", type=int, default=65536, help="environments per GPU (C2: 65536)")
    p.add_argument("--agents", type=int, default=3)
    p.add_argument("--horizon", type=int, default=128, help="T: vector steps per collect")
    p.add_argument("--batch-size", type=int, default=65536)
    p.add_argument("--repeat", type=int, default=4)
    p.add_argument("--hidden", type=int, default=64)
    p.add_argument("--engine", default="tc", choices=["tc", "fp32"], help="MLP kernels: tcgen05 (fp16 operands, fp32 accumulate) or exact fp32")
    p.add_argument("--no-update", action="store_true", help="collect + GAE only")
    p.add_argument("--reuse-rollout", action="store_true",
                   help="PPO(reuse_rollout_policy_outputs=True): critic(obs) / logp_old come from the collect-time forward (off in the headline run)")
    p.add_argument("--workload", default="c2", choices=["c2", "c3", "c4", "c5"],
                   help="c2: the headline collect+GAE+PPO step; c3: MAPPO / CTDE, N = 10 agents, 32 768 worlds per GPU (BASELINE.json "
                        "configs[2]); c4: simple_tag grouped policies, 131 072 worlds in total over the ranks (configs[3], strong scaling); c5: GAE + buffer add/sample microbench (T=128, 1M rows), one GPU")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-c5", action="store_true", help="skip the configs[4] microbench that rides on the default N = 1 line")
    p.add_argument("--cpu-envs", type=int, default=4096, help="reference arm: environments in the bounded sample")
    p.add_argument("--cpu-horizon", type=int, default=64)
    p.add_argument("--ref-quick", action="store_true", help="--impl reference: only the C1 point of the reference classes (the cpu_baseline leg of the own arm)")
    p.add_argument("--ref-port", action="store_true", help="--impl reference: time the oracle port even when baseline/_ref is installed")
    return p.parse_args()


# ------------------------------------------------------------------------------------------
# clocks sampler (B200_PROFILING.md "clocks DURING the timed region")
class Clocks:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int) -> None:
        self.index, self.rows, self._stop, self._th = index, [], threading.Event(), None

    def _run(self) -> None:
        try:  # NVML in-process: ~0.1 ms per sample (nvidia-smi takes >100 ms per call)
            import pynvml as nv

            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            bits = [("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4)]
            while not self._stop.is_set():
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.rows.append([str(sm), str(mx)] + ["Active" if r & b else "Not Active" for _, b in bits])
                self._stop.wait(0.005)
            return
        except Exception:
            pass
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.1)

    def __enter__(self) -> "Clocks":
        self._th = threading.Thread(target=self._run, daemon=True)
        self._th.start()
        return self

    def __exit__(self, *a) -> None:
        self._stop.set()
        if self._th:
            self._th.join(timeout=6)

    def summary(self) -> dict:
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


# ------------------------------------------------------------------------------------------
# CPU arm: the oracle port (numpy / C restatement of the reference path) on the host cores
def cpu_arm(args: argparse.Namespace, steps: int, warmup: int) -> dict:
    """One 'step' = collect T_s vector steps over E_s simple_spread worlds (random-uniform
    policy stand-in is NOT used: the policy MLP runs in torch fp32 on the CPU like the reference's
    Net), buffer add state machine, GAE (serial float64 loop like numba `_gae`), PPO update."""
    import torch

    from oracle import gae as ogae
    from oracle.gae import _clib

    torch.set_num_threads(os.cpu_count() or 1)
    lib = _clib()
    E, A, T = args.cpu_envs, args.agents, args.cpu_horizon
    R, O = E * A, 6 * A
    rng = np.random.default_rng(0)
    pos = rng.uniform(-1, 1, (E, 2 * A, 2))
    vel = np.zeros((E, A, 2))
    obs = np.zeros((R, O), np.float32)
    rew = np.zeros(R)
    ins, sz, ep_len, ep_start, last, lens = (np.zeros(R, np.int64) for _ in range(6))
    ep_ret = np.zeros(R)
    outs = [np.zeros(R, np.int64), np.zeros(R), np.zeros(R, np.int64), np.zeros(R, np.int64)]
    store = {"obs": np.zeros((R * T, O), np.float32), "obs_next": np.zeros((R * T, O), np.float32),
             "act": np.zeros(R * T, np.int64), "rew": np.zeros(R * T), "done": np.zeros(R * T, bool)}
    try:
        from oracle import ppo as oppo
    except ImportError:
        oppo = None
    net = oppo.make_net(O, args.hidden, 5, seed=0) if (oppo is not None and not args.no_update) else None
    P = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731

    def one_step() -> None:
        ins[:] = 0
        sz[:] = 0
        cur = obs.copy()
        for t in range(T):
            if net is not None:
                act = oppo.act(net, cur)
            else:
                act = rng.integers(0, 5, R)
            a32 = act.astype(np.int32)
            lib.orc_mpe_spread_step(P(pos), P(vel), P(a32), C.c_int64(E), C.c_int(A), C.c_double(0.5), C.c_int(1), P(obs), P(rew))
            done = np.full(R, (t + 1) % 25 == 0, np.uint8)
            lib.orc_buf_add(P(ins), P(sz), P(ep_ret), P(ep_len), P(ep_start), P(last), P(lens), C.c_int64(T), None,
                            P(rew), P(done), C.c_int64(R), *(P(o) for o in outs))
            slot = outs[0]
            store["obs"][slot] = cur
            store["obs_next"][slot] = obs
            store["act"][slot] = act
            store["rew"][slot] = rew
            store["done"][slot] = done.astype(bool)
            cur = obs.copy()
            if done[0]:
                pos[:] = rng.uniform(-1, 1, pos.shape)
                vel[:] = 0
        B = R * T
        if net is not None:
            v_s, v_n = oppo.values(net, store["obs"]), oppo.values(net, store["obs_next"])
        else:
            v_s, v_n = np.zeros(B), np.zeros(B)
        end = store["done"].copy()
        end[T - 1::T] = True
        adv = ogae.gae(v_s.astype(np.float64), v_n.astype(np.float64), store["rew"], end, 0.99, 0.95)
        if net is not None:
            oppo.update(net, store, adv, adv + v_s, batch_size=max(256, args.batch_size * R // (args.envs * A)), repeat=args.repeat)

    for _ in range(warmup):
        one_step()
    t0 = time.perf_counter()
    for _ in range(steps):
        one_step()
    dt = (time.perf_counter() - t0) / steps
    return {"value": R * T / dt, "ms_per_step": dt * 1e3, "cores": os.cpu_count() or 1,
            "sample": f"simple_spread E={E} A={A} T={T} per step, {'collect+GAE+PPO update' if net is not None else 'collect(random actions)+GAE'}, "
                      f"oracle port (C/OpenMP MPE step + buffer add, serial f64 GAE, torch-fp32 CPU MLP)"}


def reference_classes_arm(args: argparse.Namespace, steps: int, warmup: int, quick: bool) -> dict | None:
    """The reference's own classes (baseline/_ref, installed by baseline/install_ref.py at build time) on the host cores:
    BASELINE.json configs[0] exactly (DummyVectorEnv x8, A agents, T = --horizon) as the timed value, plus -- unless `quick` --
    SubprocVectorEnv x8, a DummyVectorEnv x1024 scaling point (one step, T = 16) and the numba `_gae` at B = 2^24 on one core.
    Returns None when baseline/_ref is absent (the caller then times the oracle port)."""
    from baseline import ref_arm

    if not ref_arm.available():
        return None
    bs = 256  # the reference's own PPO scripts use batch_size 64..256 at this scale (test/discrete/test_ppo_discrete.py:45)
    r = ref_arm.run(8, args.agents, args.horizon, bs, args.repeat, args.hidden, steps, warmup, update=not args.no_update)
    if not quick:
        extra = {}
        try:
            s = ref_arm.run(8, args.agents, args.horizon, bs, args.repeat, args.hidden, max(1, steps // 2), 1, subproc=True, update=not args.no_update)
            extra["subproc_x8"] = {k: s[k] for k in ("value", "ms_per_step", "collect_ms", "update_ms", "workers")}
        except Exception as e:  # noqa: BLE001  (fork restrictions of the box must not lose the main number)
            extra["subproc_x8"] = {"unavailable": repr(e)[:200]}
        s = ref_arm.run(1024, args.agents, 16, bs, args.repeat, args.hidden, 1, 0, update=not args.no_update)
        extra["dummy_x1024_T16"] = {k: s[k] for k in ("value", "ms_per_step", "collect_ms", "update_ms")}
        extra["gae_numba_2p24"] = ref_arm.gae_numba(1 << 24)
        r["extra"] = extra
    return r


def cpu_baseline_leg(args: argparse.Namespace) -> dict:
    """cpu_baseline of the own arm (rank 0, N = 1): the reference classes in a fresh process (no CUDA context, own torch thread
    pool), about 5 s; the oracle port beside it as cpu_baseline_port."""
    out: dict = {}
    try:
        cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--ref-quick", "--steps", "3", "--warmup", "1",
               "--agents", str(args.agents), "--horizon", str(args.horizon), "--repeat", str(args.repeat), "--hidden", str(args.hidden)]
        env = dict(os.environ)
        for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
            env.pop(k, None)
        p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
        ref = json.loads(p.stdout.strip().splitlines()[-1])["cpu_baseline"]
    except Exception as e:  # noqa: BLE001
        ref = None
        print("cpu_baseline: reference process failed:", repr(e)[:300], file=sys.stderr)
    r = cpu_arm(args, 3, 1)
    port = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]}
    if ref is not None and ref.get("kind") == "reference":
        out["cpu_baseline"], out["cpu_baseline_port"] = ref, port
    else:
        out["cpu_baseline"] = port
    return out


def reference_main(args: argparse.Namespace) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    r = None if args.ref_port else reference_classes_arm(args, steps, warmup, args.ref_quick)
    kind = "reference"
    port = None
    if r is None:
        r, kind = cpu_arm(args, steps, warmup), "port"
    elif not args.ref_quick:
        port = cpu_arm(args, 2, 1)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"C2 simple_spread_v3 A={args.agents}, bounded CPU sample: {r['sample']}"},
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": kind, "sample": r["sample"]},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    for k in ("collect_ms", "update_ms", "extra"):
        if k in r:
            line["cpu_baseline"][k] = r[k]
    if port is not None:
        line["cpu_baseline_port"] = {"value": port["value"], "unit": UNIT, "cores": port["cores"], "kind": "port", "sample": port["sample"]}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------
def c5_main(args: argparse.Namespace) -> None:
    line = c5_measure(args.steps, max(args.warmup, 3))
    print(json.dumps(line))


def c5_measure(steps: int, warmup: int) -> dict:
    """BASELINE.json configs[4]: GAE + buffer add / sample at T = 128, envs x agents = 1 048 576, gamma 0.99, lambda 0.95.
    One step = T vector adds of R rows (obs dim 18) through VectorReplayBuffer.add, GAE over the [T, R] block in place,
    sample_indices(0) over the full buffer.  Inputs per SURVEY 8(d): done every 25th step with a per-row phase, 1 % terminated."""
    import torch

    from tianshou_marl_b200 import _lib
    from tianshou_marl_b200.algorithm import gae_flat, gae_time_major
    from tianshou_marl_b200.data import Batch, VectorReplayBuffer

    torch.cuda.set_device(0)
    _lib.lib()
    try:
        hbm_peak, src = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured"
    except Exception:
        hbm_peak, src = 6650.0, "fallback"
    T, R, O = 128, 1 << 20, 18
    g = torch.Generator(device="cuda").manual_seed(0)
    obs = torch.randn(R, O, device="cuda", generator=g)
    obs2 = torch.randn(R, O, device="cuda", generator=g)
    act = torch.randint(0, 5, (R,), device="cuda", generator=g, dtype=torch.int32)
    rew = torch.randn(R, device="cuda", generator=g)
    phase = torch.randint(0, 25, (R,), device="cuda", generator=g)
    trunc = [(((t + phase) % 25) == 24).to(torch.uint8) for t in range(25)]
    term = (torch.rand(R, device="cuda", generator=g) < 0.01).to(torch.uint8)
    buf = VectorReplayBuffer(T * R, R, index_type="torch")
    v_s, v_n = torch.randn(T, R, device="cuda", generator=g), torch.randn(T, R, device="cuda", generator=g)
    adv, ret = torch.empty_like(v_s), torch.empty_like(v_s)
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    tot = {"add": 0.0, "gae": 0.0, "sample_indices": 0.0, "gae_flat": 0.0}
    # what compute_episodic_return passes for a flat index list: no segment length, the unfinished slots as force_end
    fe = torch.zeros(T * R, dtype=torch.uint8, device="cuda")
    fe[T - 1::T] = 1

    def step(timed: bool) -> None:
        buf.reset()
        e = [ev() for _ in range(5)]
        e[0].record()
        for t in range(T):
            buf.add(Batch(obs=obs, act=act, rew=rew, terminated=term, truncated=trunc[t % 25], obs_next=obs2))
        e[1].record()
        st = buf._store
        gae_time_major(v_s, v_n, st["rew"], st["terminated"], st["truncated"], 0, T, 0.99, 0.95, adv, ret)
        e[2].record()
        idx = buf.sample_indices(0)
        e[3].record()
        # the same 134 M elements as one flat array of 128-element episodes (the layout compute_episodic_return receives)
        gae_flat(v_s.view(-1), v_n.view(-1), st["rew"].view(-1), st["terminated"].view(-1), st["truncated"].view(-1), fe, 0, 0.99, 0.95)
        e[4].record()
        torch.cuda.synchronize()
        assert idx.numel() == T * R
        if timed:
            for k, (a, b) in zip(tot, ((0, 1), (1, 2), (2, 3), (3, 4))):
                tot[k] += e[a].elapsed_time(e[b])

    for _ in range(warmup):
        step(False)
    with Clocks(0) as clk:
        for _ in range(steps):
            step(True)
    n = T * R
    ms = {k: v / steps for k, v in tot.items()}
    gbs = {"gae": GAE_BYTES_PER_ELT * n / ms["gae"] / 1e6, "add": (2 * 4 * O + 4 + 4 + 3) * n / ms["add"] / 1e6,
           "sample_indices": 8 * n / ms["sample_indices"] / 1e6, "gae_flat": (GAE_BYTES_PER_ELT + 1) * n / ms["gae_flat"] / 1e6,
           "add_moved": (2 * (2 * 4 * O + 11) + 80) * n / ms["add"] / 1e6}
    line = {"metric": "GAE GB/s vs HBM peak", "value": gbs["gae"], "unit": "GB/s", "n_gpus": 1, "steps": steps, "warmup": warmup,
            "ms_per_step": sum(ms.values()), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32 I/O, f64 carry",
            "data": "synthetic", "config": {"workload": "C5 GAE + buffer add/sample microbench T=128 rows=1048576 gamma=0.99 lambda=0.95 obs_dim=18",
                                            "l2": "inputs larger than L2: 2.95 GB per GAE call, 20.8 GB per 128 adds"},
            "clocks": clk.summary(), "gpu_launches": steps * (1 + 2 * T + 1 + 1 + 1),
            "stages_ms": ms,
            "roofline": {"kernel": "gae_tm_kernel", "bound": "hbm", "achieved": gbs["gae"], "peak": hbm_peak, "peak_source": src, "unit": "GB/s",
                         "frac": gbs["gae"] / hbm_peak, "traffic": None, "bytes_per_unit": GAE_BYTES_PER_ELT, "units_per_launch": n,
                         "buffer_add": {"achieved": gbs["add"], "frac": gbs["add"] / hbm_peak, "bytes_per_unit": 2 * 4 * O + 11,
                                        "moved_bytes_per_unit": 2 * (2 * 4 * O + 11) + 80, "moved_frac": gbs["add_moved"] / hbm_peak,
                                        "note": "128 VectorReplayBuffer.add calls of 1M rows (payload copy + counters kernel each); bytes_per_unit is "
                                                "SURVEY 8(d)'s write-only figure, moved_bytes_per_unit adds the reads of the caller's arrays and the "
                                                "five 8-byte per-row counters read and written"},
                         "gae_flat": {"kernel": "gae_flat_kernel", "achieved": gbs["gae_flat"], "frac": gbs["gae_flat"] / hbm_peak,
                                      "bytes_per_unit": GAE_BYTES_PER_ELT + 1, "note": "flat layout with force_end flags, output allocation included"},
                         "sample_indices": {"achieved": gbs["sample_indices"], "frac": gbs["sample_indices"] / hbm_peak, "bytes_per_unit": 8}}}
    del buf, obs, obs2, v_s, v_n, adv, ret
    torch.cuda.empty_cache()
    return line


def main() -> None:
    args = parse()
    # stdout carries exactly ONE JSON line: anything a library prints there (NCCL's version banner under torchrun) goes to stderr
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    global print

    def print(*a, **k):  # noqa: A001  (the result line only)
        os.write(json_fd, (" ".join(str(x) for x in a) + "\n").encode())

    if args.impl == "reference":
        reference_main(args)
        return
    if args.workload == "c5":
        if int(os.environ.get("RANK", "0")) == 0:
            c5_main(args)
        return
    import torch
    import torch.distributed as dist

    from tianshou_marl_b200 import _lib
    from tianshou_marl_b200.bench_workload import Workload, WorkloadC3, WorkloadC4

    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    _lib.lib()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak, peak_src = (peaks["hbm_gbs"], "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
    tf_peak = peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1590.0))

    if args.workload == "c3":
        if args.envs == 65536 and args.agents == 3:  # the c2 defaults: switch to configs[2]
            args.envs, args.agents = 32768, 10
        wl = WorkloadC3(args, rank, world)
    elif args.workload == "c4":
        if args.envs == 65536 and args.agents == 3:  # the c2 defaults: switch to configs[3]
            args.envs = 131072
        wl = WorkloadC4(args, rank, world)
    else:
        wl = Workload(args, rank, world)

    def barrier() -> None:
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing ("value") ----
    for _ in range(max(args.warmup, 3)):
        wl.step()
    barrier()
    wl.reset_timers()
    with Clocks(local) as clk:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            wl.step(timed=True)
        e1.record()
        barrier()
    ms = e0.elapsed_time(e1) / args.steps
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    stages = wl.stage_ms(args.steps)
    launches = wl.launches_per_step * args.steps

    # ---- end to end through the public API with host round trips ("e2e") ----
    for _ in range(2):
        wl.step_e2e()
    barrier()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        wl.step_e2e()
    e1.record()
    barrier()
    ms_e2e = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3) / args.steps
    if world > 1:
        t = torch.tensor([ms_e2e], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_e2e = float(t.item())

    ident = None
    if world > 1:  # the correctness claim of the data-parallel update travels with the line (costs one small all_gather)
        flat = torch.cat([n.flat.detach().view(-1) for n in wl.nets()]).contiguous()
        got = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(got, flat)
        ident = {"params_identical_across_ranks": bool(all(torch.equal(got[0], g) for g in got[1:])),
                 "params_max_abs_diff_across_ranks": float(max((got[0] - g).abs().max().item() for g in got[1:])),
                 "grad_sync": wl.grad_sync()}
    if rank == 0:
        units = wl.agent_steps_per_step * world
        roof = wl.roofline(stages, hbm_peak, peak_src, tf_peak)
        line = {"metric": METRIC, "value": units / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": wl.dtype, "data": "synthetic", "config": wl.config(),
                "clocks": clk.summary(), "gpu_launches": launches,
                "e2e": {"value": units / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
                        "h2d_bytes_per_step": wl.h2d_bytes, "d2h_bytes_per_step": wl.d2h_bytes},
                "stages_ms": stages, "roofline": roof}
        if ident is not None:
            line.update(ident)
        if args.workload == "c4":
            line["metric"] = "agent-steps/sec collect+PPO update, simple_tag grouped policies"
            line["scaling"] = "strong"
        if args.workload == "c3":
            line["metric"] = "agent-steps/sec collect+MAPPO update, simple_spread N=10 (centralized critic)"
        if world == 1 and args.workload == "c2" and not args.no_c5:
            # BASELINE.json configs[4] (GAE + buffer add / sample microbench, T = 128, 1 048 576 rows) rides on the default line so
            # that the driver's record carries its HBM fractions too (python bench.py --workload c5 prints it alone)
            del wl
            torch.cuda.empty_cache()
            c5 = c5_measure(2, 3)
            r5 = c5["roofline"]
            line["roofline"]["c5"] = {"gae": {"achieved": r5["achieved"], "frac": r5["frac"], "unit": "GB/s", "bytes_per_unit": r5["bytes_per_unit"]},
                                      "gae_flat": r5["gae_flat"], "buffer_add": r5["buffer_add"], "sample_indices": r5["sample_indices"], "peak": r5["peak"],
                                      "workload": c5["config"]["workload"], "stages_ms": c5["stages_ms"]}
        if not args.no_cpu_baseline and world == 1 and args.workload == "c2":
            line.update(cpu_baseline_leg(args))
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
