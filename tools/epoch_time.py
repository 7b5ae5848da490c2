"""Wall time of one tsm_ppo_epoch launch (engine 1) at the C2 shape: `python tools/epoch_time.py [n_mb]` on the GPU box;
TSM_TC_EPOCH_GROUPS=2|4 picks the kernel."""
import ctypes as C, os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tianshou_marl_b200 import _lib
from tianshou_marl_b200.utils import ActorCriticNet
dev = torch.device("cuda")
R, T, O = 196608, 128, 18
B = R * T
n_mb = int(sys.argv[1]) if len(sys.argv) > 1 else 384
adv_norm = int(os.environ.get("ADV_NORM", "1"))
obs = torch.randn(T, R, O, device=dev); act = torch.randint(0, 5, (T, R), device=dev, dtype=torch.int32)
adv = torch.randn(T, R, device=dev); ret = torch.randn(T, R, device=dev); logp = torch.full((T, R), -1.6, device=dev)
net = ActorCriticNet(O, 64, 5, seed=0, engine="tc")
L = _lib.lib(); P = _lib.ptr
b = _lib.TsmPpoBatch()
b.obs, b.act, b.adv, b.ret, b.logp_old, b.v_old, b.indices = P(obs), P(act), P(adv), P(ret), P(logp), None, None
b.size, b.n_rows, b.start, b.len, b.perm_key, b.perm_n = T, R, 0, T, 12345, B
cfg = _lib.TsmPpoCfg(); cfg.eps_clip, cfg.dual_clip, cfg.vf_coef, cfg.ent_coef, cfg.value_clip, cfg.adv_norm = 0.2, 0.0, 0.5, 0.01, 0, adv_norm
rec = torch.empty(T * R, _lib.PPO_RECORD_BYTES, dtype=torch.uint8, device=dev)
st = _lib.stream_ptr()
_lib.check(L.tsm_ppo_pack_records(net.desc, C.byref(b), P(rec), st)); b.records = P(rec)
slots = torch.empty(B, dtype=torch.int32, device=dev); b.slots = P(slots)
bounds = np.arange(0, 65536 * (n_mb + 1), 65536, dtype=np.int64); hb = bounds.ctypes.data_as(C.c_void_p)
bd = torch.as_tensor(bounds, device=dev)
stats = torch.zeros(n_mb, 2, device=dev)
_lib.check(L.tsm_ppo_adv_stats(C.byref(b), P(bd), n_mb, 1e-8, P(stats), st))
sums = torch.zeros(n_mb, 4, dtype=torch.float64, device=dev)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(4):
    if i == 1: e0.record()
    _lib.check(L.tsm_ppo_epoch(net.desc, C.byref(b), C.byref(cfg), hb, n_mb, P(stats) if adv_norm else None, 3e-4, 0.9, 0.999, 1e-8, 0.5, P(sums), st))
e1.record(); torch.cuda.synchronize()
print(f"groups {os.environ.get('TSM_TC_EPOCH_GROUPS', '4')} n_mb {n_mb} adv_norm {adv_norm}: {e0.elapsed_time(e1) / 3 * 1e3 / n_mb:.2f} us per minibatch")
