"""The reference arm of bench.py (baseline/ref_arm.py): the installed reference classes drive the CPU worlds through the
row-flattening VectorEnvWrapper.  Skipped where baseline/_ref is not installed (it is built by __graft_entry__.build())."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from baseline import ref_arm  # noqa: E402

pytestmark = pytest.mark.skipif(not ref_arm.available(), reason="baseline/_ref not installed")


def test_row_flatten_matches_oracle_worlds():
    """Rows presented to the reference Collector are row = env * A + agent of the oracle worlds, reset per world."""
    from oracle.mpe import SPREAD, OracleMPE

    E, A, T = 3, 3, 30
    alg, col, env = ref_arm.build(E, A, T, 64, subproc=False)
    try:
        worlds = [OracleMPE(SPREAD, 1, n_agents=A, seed=i) for i in range(E)]
        for w in worlds:
            w.reset()  # Collector.reset() inside build() already reset every world once
        want = np.concatenate([w.reset()[0] for w in worlds], 0)
        obs, info = env.reset()
        assert obs.shape == (E * A, 6 * A) and np.array_equal(obs, want)
        rng = np.random.default_rng(0)
        for t in range(T):
            act = rng.integers(0, 5, E * A)
            o, r, term, trunc, _ = env.step(act)
            outs = [w.step(act[i * A:(i + 1) * A][None]) for i, w in enumerate(worlds)]
            assert np.array_equal(o, np.concatenate([x[0][0] for x in outs], 0))
            assert np.array_equal(r, np.concatenate([x[1][0] for x in outs], 0))
            assert trunc.shape == (E * A,) and bool(trunc[0]) == ((t + 1) % 25 == 0)
            if trunc[0]:
                o2, _ = env.reset(np.arange(A, 2 * A))  # one world's rows
                assert np.array_equal(o2, worlds[1].reset()[0])
                env.reset(np.r_[0:A, 2 * A:3 * A])
                worlds[0].reset(), worlds[2].reset()
    finally:
        env.close()


def test_reference_step_runs():
    r = ref_arm.run(2, 3, 26, 64, 2, 64, 1, 0)
    assert r["value"] > 0 and "DummyVectorEnv x2" in r["sample"]
