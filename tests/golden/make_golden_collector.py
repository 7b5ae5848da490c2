"""Golden fixture `collector.npz`: the UNMODIFIED reference Collector (tianshou/data/collector.py:548-1098) over DummyVectorEnv of the
reference's own MoveToRightEnv (test/base/env.py:12-171), driven through the scenarios its tests pin by hand
(test/base/test_collector.py:155-204 slot layout and n_episode surplus rule, :946-967 episode-length order) plus
EpisodeRolloutHookMCReturn (collector.py:1480-1509).  After every collect call the buffer's flat arrays and the CollectStats are
recorded.  Run in the build container:  cd tests/golden && python make_golden_collector.py"""
from __future__ import annotations

import numpy as np

from make_golden import save  # noqa: E402  (also sets sys.path for the reference + shims + test/base)


def _policy():
    import gymnasium as gym
    from tianshou.algorithm.algorithm_base import Policy
    from tianshou.data import Batch

    class MaxActionPolicy(Policy):  # test/base/test_collector.py:30-80 without its state / dict options
        def __init__(self) -> None:
            super().__init__(action_space=gym.spaces.Discrete(2))

        def forward(self, batch, state=None, **kwargs):
            return Batch(act=np.ones(len(batch.obs)), state=state)

    return MaxActionPolicy()


def _snapshot(out: dict, pre: str, col, stats) -> None:
    buf = col.buffer
    out[pre + "len"] = np.array(len(buf))
    out[pre + "obs"] = np.asarray(buf.obs[:, 0], dtype=np.float64).copy()
    out[pre + "obs_next"] = np.asarray(buf.obs_next[:, 0], dtype=np.float64).copy()
    out[pre + "rew"] = np.asarray(buf.rew, dtype=np.float64).copy()
    out[pre + "act"] = np.asarray(buf.act, dtype=np.float64).copy()
    out[pre + "terminated"] = np.asarray(buf.terminated, dtype=bool).copy()
    out[pre + "done"] = np.asarray(buf.done, dtype=bool).copy()
    out[pre + "indices"] = np.asarray(buf.sample_indices(0), dtype=np.int64).copy()
    out[pre + "n_ep"] = np.array(stats.n_collected_episodes)
    out[pre + "n_st"] = np.array(stats.n_collected_steps)
    out[pre + "lens"] = np.asarray(stats.lens, dtype=np.int64).copy()
    out[pre + "returns"] = np.asarray(stats.returns, dtype=np.float64).copy()
    out[pre + "collect_step"] = np.array(col.collect_step)
    out[pre + "collect_episode"] = np.array(col.collect_episode)


SCENARIOS = {
    # name: (env sizes, buffer total size, [(kind, n), ...])
    "a": ([2, 3, 4, 5], 100, [("n_step", 8), ("n_episode", 4)]),          # test_collector.py:155-204
    "b": ([1, 8, 9, 10], 100, [("n_episode", 2), ("n_episode", 10), ("n_step", 20), ("n_step", 20)]),  # :946-967
    "c": ([2, 3, 4, 5], 100, [("n_episode", 2), ("n_step", 5), ("n_episode", 7)]),  # fewer episodes than envs; ragged n_step
}


def gen() -> dict:
    from env import MoveToRightEnv
    from tianshou.data import Collector, VectorReplayBuffer
    from tianshou.data.collector import EpisodeRolloutHookMCReturn
    from tianshou.env import DummyVectorEnv

    out: dict = {}
    for name, (sizes, total, calls) in SCENARIOS.items():
        venv = DummyVectorEnv([lambda x=i: MoveToRightEnv(size=x, sleep=0) for i in sizes])
        col = Collector(_policy(), venv, VectorReplayBuffer(total_size=total, buffer_num=len(sizes)))
        col.reset()
        out[f"{name}_sizes"] = np.array(sizes)
        out[f"{name}_total"] = np.array(total)
        out[f"{name}_calls"] = np.array([[0 if k == "n_step" else 1, n] for k, n in calls])
        for ci, (kind, n) in enumerate(calls):
            st = col.collect(**{kind: n})
            _snapshot(out, f"{name}{ci}_", col, st)
    # episode hook: MC return to go + full-episode MC return written back into the buffer (n_episode collection)
    sizes = [2, 3, 4, 5]
    venv = DummyVectorEnv([lambda x=i: MoveToRightEnv(size=x, sleep=0) for i in sizes])
    col = Collector(_policy(), venv, VectorReplayBuffer(total_size=100, buffer_num=4), on_episode_done_hook=EpisodeRolloutHookMCReturn(gamma=0.9))
    col.reset()
    st = col.collect(n_episode=6)
    _snapshot(out, "hook0_", col, st)
    out["hook0_mc_return_to_go"] = np.asarray(col.buffer.mc_return_to_go, dtype=np.float64).copy()
    out["hook0_full_episode_mc_return"] = np.asarray(col.buffer.full_episode_mc_return, dtype=np.float64).copy()
    out["hook_gamma"] = np.array(0.9)
    return out


if __name__ == "__main__":
    save("collector.npz", **gen())
