"""tsm_map_action (Policy.map_action for Box spaces, algorithm_base.py:255-288) against the reference's own outputs
(tests/golden/actions.npz): clip is exact, tanh within 2 ulp of float32."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_map_action_matches_reference():
    from tianshou_marl_b200.algorithm import BoxActionMapper

    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "actions.npz"))
    raw = z["raw"]
    for key in [k for k in z.files if k not in ("low", "high", "raw")]:
        bound, scaling = key.rsplit("_", 1)
        m = BoxActionMapper(z["low"], z["high"], action_scaling=bool(int(scaling)), action_bound_method=None if bound == "None" else bound)
        got = m.map_action(raw)
        assert got.shape == raw.shape
        assert np.allclose(got, z[key], rtol=3e-7, atol=3e-7), (key, np.abs(got - z[key]).max())
        dev = m.map_action(torch.as_tensor(raw, device="cuda"))
        assert dev.is_cuda and np.array_equal(dev.cpu().numpy(), got)
    with pytest.raises(AssertionError):
        BoxActionMapper(z["low"], z["high"], action_scaling=True, action_bound_method=None).map_action(raw)
