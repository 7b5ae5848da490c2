"""Golden fixture `actions.npz`: the UNMODIFIED reference Policy.map_action (algorithm_base.py:255-288) on a Box action space, for
every (action_bound_method, action_scaling) combination.  Run:  cd tests/golden && python make_golden_actions.py"""
from __future__ import annotations

import numpy as np

from make_golden import save  # noqa: E402


def gen() -> dict:
    import gymnasium as gym
    from tianshou.algorithm.algorithm_base import Policy

    class P(Policy):
        def forward(self, batch, state=None, **kw):
            raise NotImplementedError

    low, high = np.array([-2.0, 0.0, -1.0], np.float32), np.array([2.0, 5.0, 1.0], np.float32)
    rng = np.random.default_rng(3)
    raw = (rng.normal(size=(64, 3)) * 1.5).astype(np.float32)
    out = {"low": low, "high": high, "raw": raw}
    for bound in (None, "clip", "tanh"):
        for scaling in (False, True):
            if bound is None and scaling:
                continue  # the reference asserts on raw actions outside [-1, 1]
            p = P(action_space=gym.spaces.Box(low, high), action_scaling=scaling, action_bound_method=bound)
            out[f"{bound}_{int(scaling)}"] = np.asarray(p.map_action(raw), dtype=np.float32)
    return out


if __name__ == "__main__":
    save("actions.npz", **gen())
