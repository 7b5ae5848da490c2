"""Run one traced minibatch launch (library built with -DTSM_TC_TRACE) and print per-phase clock deltas."""
import ctypes as C, os, sys
import torch
from tianshou_marl_b200 import _lib
from tianshou_marl_b200.utils import ActorCriticNet
dev = torch.device("cuda")
R, T, O = 196608, 128, 18
B = R * T
obs = torch.randn(T, R, O, device=dev)
act = torch.randint(0, 5, (T, R), device=dev, dtype=torch.int32)
adv = torch.randn(T, R, device=dev); ret = torch.randn(T, R, device=dev); logp = torch.full((T, R), -1.6, device=dev)
net = ActorCriticNet(O, 64, 5, seed=0, engine="tc")
L = _lib.lib(); P = _lib.ptr
b = _lib.TsmPpoBatch()
b.obs, b.act, b.adv, b.ret, b.logp_old, b.v_old, b.indices = P(obs), P(act), P(adv), P(ret), P(logp), None, None
b.size, b.n_rows, b.start, b.len, b.perm_key, b.perm_n = T, R, 0, T, 12345, B
cfg = _lib.TsmPpoCfg()
cfg.eps_clip, cfg.dual_clip, cfg.vf_coef, cfg.ent_coef, cfg.value_clip, cfg.adv_norm = 0.2, 0.0, 0.5, 0.01, 0, 0
st = _lib.stream_ptr(); desc = net.desc
rec = torch.empty(T * R, _lib.PPO_RECORD_BYTES, dtype=torch.uint8, device=dev)
_lib.check(L.tsm_ppo_pack_records(desc, C.byref(b), P(rec), st))
b.records = P(rec)
mb = int(sys.argv[1]) if len(sys.argv) > 1 else 75776
for i in range(5):
    if i == 4: os.environ["TSM_TC_TRACE_DUMP"] = "1"
    _lib.check(L.tsm_ppo_minibatch(desc, C.byref(b), C.byref(cfg), 0, mb, None, None, st))
torch.cuda.synchronize()
