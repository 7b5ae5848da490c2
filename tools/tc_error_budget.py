"""Error budget of the tcgen05 engine against the REFERENCE fixtures (tests/golden/ppo.npz, ppo_large.npz), one error source at
a time.  Variant libraries are built here (cross-compiled) by `python tools/tc_error_budget.py build` into tools/_variants/ and
travel to the GPU box; `python tools/tc_error_budget.py run` (GPU) re-runs itself once per variant with TSM_B200_LIB pointing at
it and writes gpurun_out/tc_error_budget.json:

    tc            the shipped library: fp16 operands + ex2/lg2/rcp loss math + sqrt.approx/rcp.approx Adam
    exact_loss    -DTSM_TC_EXACT_LOSS: IEEE loss math             -> what the fast loss math costs
    exact_adam    -DTSM_TC_EXACT_ADAM: IEEE optimizer arithmetic  -> what adam_elem_fast costs
    exact_both    both: what remains is the operand rounding (fp16 X / weights / activations / G)
    fp32          the CUDA-core engine of the shipped library (the same-precision arm)
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
VDIR = os.path.join(ROOT, "tools", "_variants")
VARIANTS = {"exact_loss": "-DTSM_TC_EXACT_LOSS", "exact_adam": "-DTSM_TC_EXACT_ADAM", "exact_both": "-DTSM_TC_EXACT_LOSS -DTSM_TC_EXACT_ADAM"}


def build() -> None:
    os.makedirs(VDIR, exist_ok=True)
    for name, flags in VARIANTS.items():
        r = subprocess.run(["make", "-C", os.path.join(ROOT, "tianshou_marl_b200", "csrc"), "-j8", f"BUILD=build_{name}",
                            f"OUT={VDIR}/libtsm_{name}.so", f"EXTRA={flags}"], capture_output=True, text=True)
        if r.returncode:
            print(r.stdout[-2000:], r.stderr[-2000:])
            raise SystemExit(1)
        print("built", name)


def run() -> None:
    res = {}
    libs = {"tc": None, "fp32": None}
    libs.update({n: os.path.join(VDIR, f"libtsm_{n}.so") for n in VARIANTS})
    for name, path in libs.items():
        env = dict(os.environ)
        if path:
            if not os.path.exists(path):
                print("missing", path)
                continue
            env["TSM_B200_LIB"] = path
        r = subprocess.run([sys.executable, __file__, "one", "fp32" if name == "fp32" else "tc"], env=env, capture_output=True, text=True)
        if r.returncode:
            print(name, "FAILED", r.stderr[-1500:])
            continue
        res[name] = json.loads(r.stdout.strip().splitlines()[-1])
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "tc_error_budget.json"), "w") as f:
        json.dump(res, f, indent=1)
    for name, r in res.items():
        for tag, e in r.items():
            print(f"{name:11s} {tag:7s} " + " ".join(f"{k}={v:.2e}" for k, v in e.items()))


def one(engine: str) -> dict:
    import numpy as np

    from _ppo_golden import load_case, load_large
    from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory, DiscreteActorPolicy
    from tianshou_marl_b200.data import Batch, VectorReplayBuffer
    from tianshou_marl_b200.utils import ActorCriticNet

    out = {}
    g_small = np.load(os.path.join(ROOT, "tests", "golden", "ppo.npz"))
    g_large = np.load(os.path.join(ROOT, "tests", "golden", "ppo_large.npz"))
    for tag, c in (("small0", load_case(g_small, 0)), ("small1", load_case(g_small, 1)), ("large0", load_large(g_large, 0)),
                   ("large1", load_large(g_large, 1))):
        net = ActorCriticNet(c["O"], c["H"], c["nA"], engine=engine).load_arrays(c["W1"], c["b1"], c["W2"], c["b2"], c["W3"], c["b3"])
        pol = DiscreteActorPolicy(actor=net)
        k = c["cfg"]
        alg = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=c["lr"]), eps_clip=k["eps_clip"], dual_clip=k["dual_clip"] or None,
                  value_clip=k["value_clip"], advantage_normalization=k["adv_norm"], vf_coef=k["vf_coef"], ent_coef=k["ent_coef"],
                  max_grad_norm=c["max_norm"] or None, gamma=c["gamma"], gae_lambda=c["lam"], minibatch_order="numpy")
        buf = VectorReplayBuffer(c["R"] * c["T"], c["R"], index_type="torch")
        for t in range(c["T"]):
            buf.add(Batch(obs=c["obs"][t], act=c["act"][t], rew=c["rew"][t], terminated=c["term"][t], truncated=c["trunc"][t], obs_next=c["obs_next"][t]))
        pol.is_within_training_step = True
        np.random.seed(c["seed"])
        alg.update(buf, c["bs"], c["repeat"])
        seq = alg.last_sequences
        r = {}
        for kk, ref in (("loss", "loss_seq"), ("clip", "clip_seq"), ("vf", "vf_seq"), ("ent", "ent_seq")):
            d = np.abs(seq[kk] - c[ref])
            r[kk + "_abs"] = float(d.max())
            r[kk + "_rel"] = float((d / (np.abs(c[ref]) + 1e-12)).max())
        v = net.named_views()
        names = ("W1", "b1", "W2", "b2", "W3", "b3")
        r["param_abs"] = float(max(np.abs(v[kk].cpu().numpy() - c["after_" + kk]).max() for kk in names))
        moved = max(np.abs(c[kk] - c["after_" + kk]).max() for kk in names)
        r["param_vs_travel"] = r["param_abs"] / float(moved)  # error relative to how far the reference moved the parameters
        err = np.concatenate([(v[kk].cpu().numpy() - c["after_" + kk]).reshape(-1) for kk in names])
        trav = np.concatenate([(c[kk] - c["after_" + kk]).reshape(-1) for kk in names])
        r["param_rms"] = float(np.sqrt((err ** 2).mean()))
        r["param_rms_vs_travel_rms"] = float(np.sqrt((err ** 2).mean() / (trav ** 2).mean()))
        r["param_abs_vs_lr_steps"] = r["param_abs"] / (c["lr"] * len(c["loss_seq"]))  # Adam moves an entry by at most ~lr per step
        em = lambda t: t.t().contiguous().view(-1).cpu().numpy()  # noqa: E731
        si = c["sample_idx"] if "sample_idx" in c else slice(None)
        for kk, w in (("v_s", "v_s"), ("returns", "ret"), ("adv", "adv"), ("logp_old", "logp_old")):
            r["pre_" + kk] = float(np.abs(em(alg._work[w])[si] - c[kk]).max())
        out[tag] = r
    return out


if __name__ == "__main__":
    cmd = sys.argv[1] if len(sys.argv) > 1 else "run"
    if cmd == "build":
        build()
    elif cmd == "one":
        print(json.dumps(one(sys.argv[2])))
    else:
        run()
