import ctypes as C, os, numpy as np, torch
from tianshou_marl_b200 import _lib
from tianshou_marl_b200.utils import ActorCriticNet
dev = torch.device("cuda")
R, T, O = 196608, int(os.environ.get("TRACE_T", "32")), 18
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
if not os.environ.get("NO_SLOTS"):
    slots = torch.empty(B, dtype=torch.int32, device=dev); b.slots = P(slots)
    _lib.check(L.tsm_ppo_slot_table(C.byref(b), 0, B, st))
n_mb = 16
MB = int(os.environ.get("TRACE_MB", "65536"))
bounds = np.arange(0, MB * (n_mb + 1), MB, dtype=np.int64); hb = bounds.ctypes.data_as(C.c_void_p)
for i in range(3):
    if i == 2: os.environ["TSM_TC_TRACE_DUMP"] = "1"
    _lib.check(L.tsm_ppo_epoch(net.desc, C.byref(b), C.byref(cfg), hb, n_mb, None, 3e-4, 0.9, 0.999, 1e-8, 0.5, None, st))
torch.cuda.synchronize()
