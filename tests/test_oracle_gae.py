"""Pins the GAE oracle (oracle/gae.py, oracle/liboracle.c) against the reference's own
known-answer vectors (test/base/test_returns.py:23-158) and against outputs recorded from the
unmodified reference (tests/golden/gae.npz, made by tests/golden/make_golden.py)."""
import numpy as np
import pytest

from oracle import gae as ogae


def _known(g, ci):
    term, trunc, rew = g[f"k{ci}_term"], g[f"k{ci}_trunc"], g[f"k{ci}_rew"]
    v = g[f"k{ci}_v"]
    gamma, lam = g[f"k{ci}_gl"]
    n = len(rew)
    indices = np.arange(n)
    done = np.logical_or(term, trunc)
    unfinished = np.array([] if done[-1] else [n - 1], int)
    return ogae.compute_episodic_return(
        rew, term.copy(), trunc.copy(), indices, unfinished, term.astype(bool),
        None if len(v) == 0 else v, None, gamma, lam)


@pytest.mark.parametrize("ci", range(5))
def test_known_answers(golden, ci):
    g = golden("gae.npz")
    ret, adv = _known(g, ci)
    assert np.allclose(ret, g[f"k{ci}_ans"])  # hand-computed answers in test_returns.py
    assert np.allclose(ret, g[f"k{ci}_ret"], rtol=1e-13, atol=1e-13)
    assert np.allclose(adv, g[f"k{ci}_adv"], rtol=1e-13, atol=1e-13)


@pytest.mark.parametrize("ri", range(5))
@pytest.mark.parametrize("use_c", [True, False])
def test_random_vector_buffers(golden, ri, use_c):
    g = golden("gae.npz")
    p = f"r{ri}_"
    idx = g[p + "indices"]
    end = np.logical_or(g[p + "term"], g[p + "trunc"])
    end[np.isin(idx, g[p + "unfinished"])] = True
    v_s_ = g[p + "v_s_"] * ~g[p + "buf_term"][idx]
    adv = ogae.gae(g[p + "v_s"], v_s_, g[p + "rew"], end, 0.99, 0.95, use_c=use_c)
    assert np.allclose(adv, g[p + "adv"], rtol=1e-12, atol=1e-12)
    ret, adv2 = ogae.compute_episodic_return(
        g[p + "rew"], g[p + "term"].copy(), g[p + "trunc"].copy(), idx, g[p + "unfinished"], g[p + "buf_term"],
        g[p + "v_s_"], g[p + "v_s"], 0.99, 0.95)
    assert np.allclose(ret, g[p + "ret"], rtol=1e-12, atol=1e-12)
    ret_mc, _ = ogae.compute_episodic_return(
        g[p + "rew"], g[p + "term"].copy(), g[p + "trunc"].copy(), idx, g[p + "unfinished"], g[p + "buf_term"],
        None, None, 0.9, 1.0)
    assert np.allclose(ret_mc, g[p + "ret_mc"], rtol=1e-12, atol=1e-12)


def test_mc_return_to_go(golden):
    g = golden("gae.npz")
    assert np.allclose(ogae.episode_mc_return_to_go(g["mc_rew"], 0.97), g["mc_out"], rtol=1e-13)
