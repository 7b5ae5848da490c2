"""Is a symmetric-memory gradient block slower for the flush atomics than ordinary device memory? (1 rank)"""
import ctypes as C, os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
from tianshou_marl_b200 import _lib
from tianshou_marl_b200.utils import ActorCriticNet
os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29544")
torch.cuda.set_device(0)
dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))
dev = torch.device("cuda")
R, T, O = 196608, 16, 18
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
sym = symm_mem.empty(2 * net.n_params + 64, dtype=torch.float32, device=dev); sym.zero_()
hdl = symm_mem.rendezvous(sym, group=dist.group.WORLD.group_name)
plain = torch.zeros(2 * net.n_params + 64, device=dev)

def run(grads, label):
    net.grads = grads[: net.n_params]
    net._fill_desc()
    d = net.desc
    for _ in range(20):
        _lib.check(L.tsm_ppo_minibatch(d, C.byref(b), C.byref(cfg), 0, 65536, None, None, st))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(200):
        _lib.check(L.tsm_ppo_minibatch(d, C.byref(b), C.byref(cfg), 0, 65536, None, None, st))
    e1.record(); torch.cuda.synchronize()
    print(label, "minibatch", e0.elapsed_time(e1) / 200 * 1e3, "us")
    e0.record()
    for _ in range(200):
        _lib.check(L.tsm_adam_step(d, 3e-4, 0.9, 0.999, 1e-8, 0.5, 1.0, None, st))
    e1.record(); torch.cuda.synchronize()
    print(label, "adam", e0.elapsed_time(e1) / 200 * 1e3, "us")

run(plain, "plain")
run(sym, "symmetric")
run(plain, "plain")
dist.destroy_process_group()
