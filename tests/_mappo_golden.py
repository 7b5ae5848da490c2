"""Shared loader of tests/golden/mappo.npz (recorded from the reference: tests/golden/make_golden_mappo.py)."""
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "mappo.npz")


def cases() -> list[dict]:
    z = np.load(GOLD)
    out = []
    for ci in range(int(z["n_cases"])):
        pre = f"m{ci}_"
        c = {k[len(pre):]: z[k] for k in z.files if k.startswith(pre)}
        E, A, T, O, H, nA, repeat = (int(x) for x in c["cfg"])
        h = c["hyper"]
        c.update(E=E, A=A, T=T, O=O, H=H, nA=nA, repeat=repeat, R=E * A,
                 ppo=dict(eps_clip=float(h[0]), dual_clip=float(h[1]), vf_coef=float(h[2]), ent_coef=float(h[3]), value_clip=bool(h[4]),
                          adv_norm=bool(h[5])), max_grad_norm=float(h[6]), lr=float(h[7]), gamma=float(h[8]), lam=float(h[9]))
        out.append(c)
    return out


def net_arrays(c: dict, which: str, when: str = "init") -> list[np.ndarray]:
    return [c[f"{when}_{which}_{k}"] for k in ("fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias", "fc3.weight", "fc3.bias")]


def em(a: np.ndarray) -> np.ndarray:
    """[T][R]... -> env-major flat (buffer.sample(0) order)."""
    return np.swapaxes(a, 0, 1).reshape(a.shape[0] * a.shape[1], *a.shape[2:])
