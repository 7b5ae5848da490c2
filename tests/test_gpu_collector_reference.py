"""The Collector's generic loop against the UNMODIFIED reference Collector (tests/golden/collector.npz, recorded by
tests/golden/make_golden_collector.py over DummyVectorEnv of the reference's MoveToRightEnv): slot layout of n_step collection
(test/base/test_collector.py:155-178), the n_episode ready-set / surplus-env rule (:183-204, collector.py:1031-1065), episode
length order (:946-967), end-of-collect reset after n_episode, and EpisodeRolloutHookMCReturn (collector.py:1480-1509).
Every buffer array and CollectStats field is compared bit for bit after every collect call."""
import os

import numpy as np
import pytest

from _move_env import HostVectorEnv, MoveToRightEnv

pytestmark = pytest.mark.gpu
Z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "collector.npz"))


class MaxActionPolicy:
    def __call__(self, batch, state=None, **kw):
        from tianshou_marl_b200.data import Batch

        return Batch(act=np.ones(len(batch.obs)), state=state)

    def map_action(self, act):
        return act


def _check(pre, col, st):
    buf = col.buffer
    assert len(buf) == int(Z[pre + "len"])
    assert np.array_equal(np.asarray(buf.obs)[:, 0].astype(np.float64), Z[pre + "obs"])
    assert np.array_equal(np.asarray(buf.obs_next)[:, 0].astype(np.float64), Z[pre + "obs_next"])
    assert np.array_equal(np.asarray(buf.rew).astype(np.float64), Z[pre + "rew"])
    assert np.array_equal(np.asarray(buf.act).astype(np.float64), Z[pre + "act"])
    assert np.array_equal(np.asarray(buf.terminated), Z[pre + "terminated"])
    assert np.array_equal(np.asarray(buf.done), Z[pre + "done"])
    assert np.array_equal(np.asarray(buf.sample_indices(0)), Z[pre + "indices"])
    assert st.n_collected_episodes == int(Z[pre + "n_ep"]) and st.n_collected_steps == int(Z[pre + "n_st"])
    assert np.array_equal(st.lens, Z[pre + "lens"]) and np.array_equal(st.returns, Z[pre + "returns"])
    assert col.collect_step == int(Z[pre + "collect_step"]) and col.collect_episode == int(Z[pre + "collect_episode"])


@pytest.mark.parametrize("name", ["a", "b", "c"])
def test_generic_loop_matches_reference_collector(name):
    from tianshou_marl_b200.data import Collector, VectorReplayBuffer

    sizes, total = [int(x) for x in Z[name + "_sizes"]], int(Z[name + "_total"])
    env = HostVectorEnv([MoveToRightEnv(s) for s in sizes])
    col = Collector(MaxActionPolicy(), env, VectorReplayBuffer(total, len(sizes)))
    col.reset()
    for ci, (kind, n) in enumerate(Z[name + "_calls"]):
        with pytest.warns(UserWarning) if (kind == 1 and n < len(sizes)) else _nullcontext():
            st = col.collect(**({"n_episode": int(n)} if kind else {"n_step": int(n)}))
        _check(f"{name}{ci}_", col, st)


class _nullcontext:
    def __enter__(self):
        return None

    def __exit__(self, *a):
        return False


def test_episode_hook_mc_return_matches_reference():
    from tianshou_marl_b200.data import Collector, EpisodeRolloutHookMCReturn, VectorReplayBuffer

    env = HostVectorEnv([MoveToRightEnv(s) for s in (2, 3, 4, 5)])
    col = Collector(MaxActionPolicy(), env, VectorReplayBuffer(100, 4), on_episode_done_hook=EpisodeRolloutHookMCReturn(gamma=float(Z["hook_gamma"])))
    col.reset()
    st = col.collect(n_episode=6)
    _check("hook0_", col, st)
    for key in ("mc_return_to_go", "full_episode_mc_return"):
        got, want = np.asarray(getattr(col.buffer, key)), Z["hook0_" + key]
        assert np.array_equal(np.isnan(got), np.isnan(want))
        assert np.allclose(got[~np.isnan(want)], want[~np.isnan(want)], rtol=1e-12)
    # the reference refuses hook additions in n_step collection (collector.py:991-998)
    col2 = Collector(MaxActionPolicy(), HostVectorEnv([MoveToRightEnv(2)]), VectorReplayBuffer(20, 1), on_episode_done_hook=EpisodeRolloutHookMCReturn())
    col2.reset()
    with pytest.raises(ValueError, match="not supported for n_step collection"):
        col2.collect(n_step=4)
