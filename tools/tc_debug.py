import ctypes as C
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from tianshou_marl_b200 import _lib
from tianshou_marl_b200.utils import ActorCriticNet

O, n = 18, int(sys.argv[1]) if len(sys.argv) > 1 else 128
rng = np.random.default_rng(n)
net = ActorCriticNet(O, 64, 5, seed=1)
with torch.no_grad():
    net.view("b1").normal_(0, 0.1); net.view("b2").normal_(0, 0.1); net.view("b3").normal_(0, 0.1)
d = lambda a, dt=torch.float32: torch.as_tensor(a, device="cuda").to(dt).contiguous()
t = dict(obs=d(rng.normal(size=(n, O))), act=d(rng.integers(0, 5, n), torch.int32), adv=d(rng.normal(size=n)), ret=d(rng.normal(size=n)),
         lp=d(np.log(0.2) + 0.3 * rng.normal(size=n)), vo=d(rng.normal(size=n)))
P = _lib.ptr
pb = _lib.TsmPpoBatch(obs=P(t["obs"]), act=P(t["act"]), adv=P(t["adv"]), ret=P(t["ret"]), logp_old=P(t["lp"]), v_old=P(t["vo"]),
                      indices=None, size=1, n_rows=n, start=0, len=1, perm_key=0, perm_n=0)
pc = _lib.TsmPpoCfg(eps_clip=0.2, dual_clip=0.0, vf_coef=0.5, ent_coef=0.01, value_clip=0, adv_norm=1)
stats = d(np.array([float(t["adv"].mean()), 1.0 / (float(t["adv"].std()) + 1e-8)]))
res = {}
for eng in ("fp32", "tc"):
    net.engine = eng
    net.grads.zero_()
    sums = torch.zeros(4, dtype=torch.float64, device="cuda")
    _lib.check(_lib.lib().tsm_ppo_minibatch(net.desc, C.byref(pb), C.byref(pc), 0, n, P(stats), P(sums), _lib.stream_ptr()))
    res[eng] = net.grads.clone().cpu().numpy()
np.set_printoptions(linewidth=220, precision=4, suppress=True)
for name, (o_, shp) in net._views.items():
    k = int(np.prod(shp))
    a, b = res["fp32"][o_:o_ + k].reshape(shp), res["tc"][o_:o_ + k].reshape(shp)
    print(name, "scale", np.abs(a).max(), "max err", np.abs(a - b).max(), "rel", np.abs(a - b).max() / np.abs(a).max())
    if name == "W2":
        e = np.abs(a - b)
        print(" err by row block of 8:", e.reshape(8, 8, 64).max(axis=(1, 2)))
        print(" err by col block of 8:", e.reshape(64, 8, 8).max(axis=(0, 2)))
        i, j = np.unravel_index(e.argmax(), e.shape)
        print(" worst", i, j, a[i, j], b[i, j])
        print(" ratio b/a sample:", (b / a)[:4, :8])
