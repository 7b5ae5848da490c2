"""oracle/mpe.py on the CPU: properties the restatement must have whatever PettingZoo version it is checked against later
(MPE parity is unpinned: the package is absent, see the oracle's header)."""
import numpy as np

from oracle.mpe import OracleMPE


def test_one_hot_continuous_action_equals_discrete():
    # simple_env._set_action: u[0] += a[2] - a[1]; u[1] += a[4] - a[3] keeps the directions of discrete actions 1..4
    for scen, kw in ((0, dict(n_agents=3)), (1, dict()), (2, dict(num_good=2, num_adversaries=1))):
        d, c = OracleMPE(scen, 64, seed=5, **kw), OracleMPE(scen, 64, seed=5, **kw)
        assert np.array_equal(d.reset(), c.reset())
        rng = np.random.default_rng(0)
        for _ in range(10):
            a = rng.integers(0, 5, (64, d.A))
            od, rd, _, _ = d.step(a)
            oc, rc, _, _ = c.step(np.eye(5, dtype=np.float32)[a])
            assert np.array_equal(od, oc) and np.array_equal(rd, rc)


def test_continuous_actions_are_clipped_to_the_box():
    a, b = OracleMPE(0, 16, n_agents=3, seed=1), OracleMPE(0, 16, n_agents=3, seed=1)
    a.reset()
    b.reset()
    rng = np.random.default_rng(3)
    act = rng.uniform(-2.0, 3.0, (16, 3, 5)).astype(np.float32)
    oa, ra, _, _ = a.step(act)
    ob, rb, _, _ = b.step(np.clip(act, 0.0, 1.0))
    assert np.array_equal(oa, ob) and np.array_equal(ra, rb)
