"""Pins oracle/ctde.py to fixtures recorded from the unmodified reference CTDEPolicy / GlobalStateConstructor
(tests/golden/ctde.npz <- tests/golden/make_golden_ctde.py)."""
import numpy as np
import pytest

from oracle import ctde as oc


def load(golden, ci):
    g = golden("ctde.npz")
    pre = f"c{ci}_"
    return {k[len(pre):]: g[k] for k in g.files if k.startswith(pre)}


def nets(c):
    mk = lambda n, tag: oc.Mlp3(*(c[f"{tag}_{n}_fc{i}.{w}"] for i in (1, 2, 3) for w in ("weight", "bias")))  # noqa: E731
    return mk("actor", "init"), mk("critic", "init")


@pytest.mark.parametrize("ci", [0, 1, 2])
def test_oracle_matches_reference_ctde(golden, ci):
    c = load(golden, ci)
    A = int(c["A"])
    assert np.array_equal(oc.global_concat(c["obs"], A), c["global_concat"])
    assert np.allclose(oc.global_mean(c["obs"], A), c["global_mean"], rtol=0, atol=1e-7)
    actor, critic = nets(c)
    import torch

    with torch.no_grad():
        assert np.allclose(actor(torch.from_numpy(c["obs"])).numpy(), c["fwd_act"], rtol=1e-6, atol=1e-6)
    use = bool(c["use_global"])
    cin = np.repeat(oc.global_concat(c["obs"], A), A, 0) if use else c["obs"]
    cin_next = np.repeat(oc.global_concat(c["obs_next"], A), A, 0) if use else c["obs_next"]
    for s in range(int(c["steps"])):
        la, lc = oc.learn(actor, critic, c["obs"], c["act"], c["rew"], c["obs_next"], c["term"], cin, cin_next, float(c["gamma"]),
                          float(c["lr_a"]), float(c["lr_c"]))
        assert np.allclose([la, lc], c["losses"][s], rtol=1e-5, atol=1e-7)
    for n, net in (("actor", actor), ("critic", critic)):
        for k, (i, w) in enumerate((i, w) for i in (1, 2, 3) for w in ("weight", "bias")):
            assert np.allclose(net.p[k].detach().numpy(), c[f"after_{n}_fc{i}.{w}"], rtol=0, atol=2e-6), (n, i, w)
