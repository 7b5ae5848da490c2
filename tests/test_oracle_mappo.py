"""oracle/mappo.py against tests/golden/mappo.npz (the reference PPO over the reference's CTDE nets, recorded by
tests/golden/make_golden_mappo.py): preprocess (critic values, GAE, logp_old) and `repeat` whole-batch optimizer steps."""
import numpy as np
import pytest

from _mappo_golden import cases, em, net_arrays
from oracle import mappo as om

CASES = cases()


@pytest.mark.parametrize("ci", range(len(CASES)))
def test_oracle_matches_reference(ci):
    c = CASES[ci]
    A, T, R = c["A"], c["T"], c["R"]
    actor, critic = om.Mlp3(*net_arrays(c, "actor")), om.Mlp3(*net_arrays(c, "critic"))
    idx = c["indices"]
    assert np.array_equal(idx, np.arange(R * T))
    unf = np.arange(R) * T + T - 1
    v_s, ret, adv, lp = om.preprocess(actor, critic, A, c["obs"], c["obs_next"], em(c["act"]), em(c["rew"]), em(c["term"]), em(c["trunc"]), idx,
                                      unf, c["gamma"], c["lam"])
    for got, key in ((v_s, "v_s"), (ret, "returns"), (adv, "adv"), (lp, "logp_old")):
        assert np.allclose(got, c[key], rtol=2e-5, atol=2e-5), key
    obs_rows = em(c["obs"])
    g_rows = em(np.repeat(c["obs"].reshape(T, R // A, A * c["O"]), A, axis=1))
    agent = em(np.tile(np.arange(A), (T, R // A)))
    log = om.update_whole_batch(actor, critic, A, obs_rows, g_rows, agent, em(c["act"]), c["adv"], c["returns"], c["logp_old"], c["v_s"], c["ppo"],
                                c["repeat"], c["lr"], c["max_grad_norm"])
    for k, ref in (("loss", "loss_seq"), ("clip", "clip_seq"), ("vf", "vf_seq"), ("ent", "ent_seq")):
        assert np.allclose(log[k], c[ref], rtol=1e-4, atol=1e-6), (k, log[k], c[ref])
    for which, net in (("actor", actor), ("critic", critic)):
        want = np.concatenate([a.reshape(-1) for a in net_arrays(c, which, "after")])
        assert np.abs(net.flat() - want).max() < 2e-5, which  # Adam normalises tiny gradients: a last-bit difference moves a weight by a fraction of lr
