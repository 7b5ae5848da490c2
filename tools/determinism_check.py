"""Run the tcgen05 minibatch kernel repeatedly on the same input: single-CTA cases must be bit-identical run to run."""
import ctypes as C, sys, numpy as np, torch
from tianshou_marl_b200 import _lib
from tianshou_marl_b200.utils import ActorCriticNet

P, L, S = _lib.ptr, _lib.lib(), _lib.stream_ptr()
for O, n in ((18, 128), (18, 3000), (9, 200), (16, 65536)):
    rng = np.random.default_rng(n)
    net = ActorCriticNet(O, 64, 5, seed=1, engine="tc")
    d = lambda a, dt=torch.float32: torch.as_tensor(a, device="cuda").to(dt).contiguous()
    t = dict(obs=d(rng.normal(size=(n, O))), act=d(rng.integers(0, 5, n), torch.int32), adv=d(rng.normal(size=n)), ret=d(rng.normal(size=n)),
             lp=d(np.log(0.2) + 0.3 * rng.normal(size=n)), vo=d(rng.normal(size=n)))
    pb = _lib.TsmPpoBatch(obs=P(t["obs"]), act=P(t["act"]), adv=P(t["adv"]), ret=P(t["ret"]), logp_old=P(t["lp"]), v_old=P(t["vo"]),
                          indices=None, size=1, n_rows=n, start=0, len=1, perm_key=0xABCDEF1234567, perm_n=n, records=None)
    rec = torch.empty(n, _lib.PPO_RECORD_BYTES, dtype=torch.uint8, device="cuda")
    _lib.check(L.tsm_ppo_pack_records(net.desc, C.byref(pb), P(rec), S))
    pb.records = P(rec)
    pc = _lib.TsmPpoCfg(eps_clip=0.2, dual_clip=0.0, vf_coef=0.5, ent_coef=0.01, value_clip=0, adv_norm=1)
    stats = d(np.array([float(t["adv"].mean()), 1.0 / (float(t["adv"].std()) + 1e-8)]))
    ref, worst, bad = None, 0.0, 0
    for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 200):
        net.grads.zero_()
        sums = torch.zeros(4, dtype=torch.float64, device="cuda")
        _lib.check(L.tsm_ppo_minibatch(net.desc, C.byref(pb), C.byref(pc), 0, n, P(stats), P(sums), S))
        g = net.grads.clone()
        if ref is None:
            ref = g
            print(f"  first-run checksum O={O} n={n}: {g.double().abs().sum().item():.10e}")
        else:
            e = ((g - ref).abs().max() / ref.abs().max()).item()
            worst = max(worst, e)
            bad += e > 1e-4
    print(f"O={O} n={n}: worst relative deviation from run 0 = {worst:.3e}, runs off by > 1e-4: {bad}")
