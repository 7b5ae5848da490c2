"""CPU cost of the epoch launch loop: wall time of the tsm_ppo_epoch call (asynchronous) vs GPU time."""
import ctypes as C, time, torch, numpy as np
from tianshou_marl_b200 import _lib
from tianshou_marl_b200.utils import ActorCriticNet
import os
DIST = "RANK" in os.environ
if DIST:
    import torch.distributed as dist
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
dev = torch.device("cuda")
R, T, O = 196608, 128, 18
B = R * T
obs = torch.randn(T, R, O, device=dev); act = torch.randint(0, 5, (T, R), device=dev, dtype=torch.int32)
adv = torch.randn(T, R, device=dev); ret = torch.randn(T, R, device=dev); logp = torch.full((T, R), -1.6, device=dev)
net = ActorCriticNet(O, 64, 5, seed=0, engine="tc")
L = _lib.lib(); P = _lib.ptr
b = _lib.TsmPpoBatch()
b.obs, b.act, b.adv, b.ret, b.logp_old, b.v_old, b.indices = P(obs), P(act), P(adv), P(ret), P(logp), None, None
b.size, b.n_rows, b.start, b.len, b.perm_key, b.perm_n = T, R, 0, T, 12345, B
cfg = _lib.TsmPpoCfg(); cfg.eps_clip, cfg.dual_clip, cfg.vf_coef, cfg.ent_coef, cfg.value_clip, cfg.adv_norm = 0.2, 0.0, 0.5, 0.01, 0, 0
rec = torch.empty(T * R, _lib.PPO_RECORD_BYTES, dtype=torch.uint8, device=dev)
st = _lib.stream_ptr()
_lib.check(L.tsm_ppo_pack_records(net.desc, C.byref(b), P(rec), st)); b.records = P(rec)
n_mb = 384
bounds = np.arange(0, 65536 * (n_mb + 1), 65536, dtype=np.int64)
hb = bounds.ctypes.data_as(C.c_void_p)
for rep in range(4):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    _lib.check(L.tsm_ppo_epoch(net.desc, C.byref(b), C.byref(cfg), hb, n_mb, None, 3e-4, 0.9, 0.999, 1e-8, 0.5, None, st))
    t1 = time.perf_counter(); e1.record()
    torch.cuda.synchronize()
    print(f"epoch of {n_mb}: CPU enqueue {1e3*(t1-t0):.2f} ms ({1e6*(t1-t0)/(2*n_mb):.1f} us per launch), GPU {e0.elapsed_time(e1):.2f} ms ({1e3*e0.elapsed_time(e1)/n_mb:.1f} us per minibatch)")

if DIST:
    from tianshou_marl_b200.algorithm.dist import PeerGradients
    pg = PeerGradients(net.n_params, dev)
    for mode in ("plain epoch again", "dist epoch"):
        for rep in range(3):
            torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter(); e0.record()
            if mode == "dist epoch":
                _lib.check(L.tsm_ppo_epoch_dist(net.desc, C.byref(b), C.byref(cfg), hb, n_mb, None, 3e-4, 0.9, 0.999, 1e-8, 0.5, None, pg.byref(), pg.seq + 1, st))
                pg.seq += n_mb
            else:
                _lib.check(L.tsm_ppo_epoch(net.desc, C.byref(b), C.byref(cfg), hb, n_mb, None, 3e-4, 0.9, 0.999, 1e-8, 0.5, None, st))
            t1 = time.perf_counter(); e1.record()
            torch.cuda.synchronize()
            if dist.get_rank() == 0:
                print(f"{mode}: CPU enqueue {1e3*(t1-t0):.2f} ms, GPU {e0.elapsed_time(e1):.2f} ms ({1e3*e0.elapsed_time(e1)/n_mb:.1f} us per minibatch)", flush=True)
    pg.check()
    dist.destroy_process_group()
