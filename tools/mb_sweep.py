"""Microbenchmark: tc / fp32 PPO minibatch kernel time vs minibatch length (fixed vs per-tile cost)."""
import ctypes as C
import sys
import torch
from tianshou_marl_b200 import _lib
from tianshou_marl_b200.utils import ActorCriticNet

engine = sys.argv[1] if len(sys.argv) > 1 else "tc"
dev = torch.device("cuda")
R, T, O = 196608, 128, 18
B = R * T
torch.manual_seed(0)
obs = torch.randn(T, R, O, device=dev)
act = torch.randint(0, 5, (T, R), device=dev, dtype=torch.int32)
adv = torch.randn(T, R, device=dev)
ret = torch.randn(T, R, device=dev)
logp = torch.full((T, R), -1.6, device=dev)
net = ActorCriticNet(O, 64, 5, seed=0, engine=engine)
L = _lib.lib()
b = _lib.TsmPpoBatch()
P = _lib.ptr
b.obs, b.act, b.adv, b.ret, b.logp_old, b.v_old, b.indices = P(obs), P(act), P(adv), P(ret), P(logp), None, None
b.size, b.n_rows, b.start, b.len, b.perm_key, b.perm_n = T, R, 0, T, 12345, B
cfg = _lib.TsmPpoCfg()
cfg.eps_clip, cfg.dual_clip, cfg.vf_coef, cfg.ent_coef, cfg.value_clip, cfg.adv_norm = 0.2, 0.0, 0.5, 0.01, 0, 0
st = _lib.stream_ptr()
desc = net.desc
rec = torch.empty(T * R, _lib.PPO_RECORD_BYTES, dtype=torch.uint8, device=dev)
_lib.check(L.tsm_ppo_pack_records(desc, C.byref(b), P(rec), st))
b.records = P(rec)


def timeit(fn, n=200):
    """GPU time per call: the n calls are captured into one CUDA graph so that the host launch rate does not matter."""
    global st
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        st = _lib.stream_ptr()
        with torch.cuda.graph(g, stream=side):
            st = _lib.stream_ptr()
            for _ in range(n):
                fn()
    st = _lib.stream_ptr()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


for tiles_per_cta in (1, 2, 4, 8, 16):
    mb = 128 * 296 * tiles_per_cta
    for perm in (B, 0):
        b.perm_n = perm
        t = timeit(lambda: _lib.check(L.tsm_ppo_minibatch(desc, C.byref(b), C.byref(cfg), 0, mb, None, None, st)))
        print(f"minibatch only  mb={mb:8d} tiles/cta={tiles_per_cta:2d} perm={'feistel' if perm else 'identity'}: {t:7.1f} us")
b.perm_n = B
mb = 65536
t = timeit(lambda: _lib.check(L.tsm_ppo_minibatch(desc, C.byref(b), C.byref(cfg), 0, mb, None, None, st)))
print(f"minibatch only mb=65536: {t:.1f} us")
t = timeit(lambda: _lib.check(L.tsm_adam_step(desc, 3e-4, 0.9, 0.999, 1e-8, 0.5, 1.0, None, st)))
print(f"adam only: {t:.1f} us")
import numpy as np
bounds = np.arange(0, 65536 * 33, 65536, dtype=np.int64) + 65536 * 100
t = timeit(lambda: _lib.check(L.tsm_ppo_epoch(desc, C.byref(b), C.byref(cfg), bounds.ctypes.data_as(C.c_void_p), 32, None, 3e-4, 0.9, 0.999, 1e-8, 0.5, None, st)), n=20)
print(f"epoch of 32 minibatches: {t/32:.1f} us per minibatch")
x = obs.view(-1, O)
for n in (128, 128 * 148, 196608, 196608 * 8, B):
    out = torch.empty(n, device=dev)
    xn = x[:n]
    t = timeit(lambda: _lib.check(L.tsm_net_forward(desc, P(xn), None, n, 0, 0, 0, 0, None, None, None, P(out), None, st)), n=20 if n == B else 200)
    print(f"forward raw n={n}: {t:.1f} us  ({n / t / 1e3:.2f} G rows/s)")
