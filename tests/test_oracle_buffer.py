"""Pins the buffer oracle (oracle/buffer.py) against traces recorded from the reference
VectorReplayBuffer and against test/base/test_buffer.py:740-964,1525-1549 known answers."""
import numpy as np
import pytest

from _replay import replay_buffer_trace
from oracle.buffer import OracleVectorReplayBuffer


@pytest.mark.parametrize("ci", range(5))
def test_trace(golden, ci):
    g = golden("buffer.npz")
    replay_buffer_trace(g, ci, OracleVectorReplayBuffer, lambda **kw: kw)


class _Attr:
    """attribute access to stored keys like the reference's buf.done"""


def _wrap(total, n):
    b = OracleVectorReplayBuffer(total, n)
    return b


@pytest.fixture(autouse=True)
def _attr_access(monkeypatch):
    monkeypatch.setattr(OracleVectorReplayBuffer, "__getattr__", lambda self, k: self.__dict__["data"][k], raising=False)


def test_reference_known_answers():
    """test/base/test_buffer.py:740-790: VectorReplayBuffer(20, 4)."""
    buf = OracleVectorReplayBuffer(20, 4)
    ptr, ep_rew, ep_len, ep_idx = buf.add(
        dict(obs=np.array([1, 2, 3]), act=np.array([1, 2, 3]), rew=np.array([1, 2, 3]),
             terminated=np.array([0, 0, 1]), truncated=np.array([0, 0, 0])), buffer_ids=[0, 1, 2])
    assert np.all(ep_len == [0, 0, 1]) and np.all(ep_rew == [0, 0, 3])
    assert np.all(ptr == [0, 5, 10]) and np.all(ep_idx == [0, 5, 10])
    idx = buf.sample_indices(11000)
    assert np.bincount(idx)[[0, 5, 10]].min() >= 3000
    idx = buf.sample_indices(0)
    assert np.array_equal(idx, [0, 5, 10])
    assert np.array_equal(buf.prev(idx), idx) and np.array_equal(buf.next(idx), idx)
    assert np.array_equal(buf.unfinished_index(), [0, 5])
    buf.add(dict(obs=np.array([4]), act=np.array([4]), rew=np.array([4]), terminated=np.array([1]),
                 truncated=np.array([0])), buffer_ids=[3])
    assert np.array_equal(buf.unfinished_index(), [0, 5])
    assert np.array_equal(buf.sample_indices(0), [0, 5, 10, 15])
    assert np.array_equal(buf.sample_indices(-1), [])


def test_get_buffer_indices(golden):
    g = golden("buffer.npz")
    buf = OracleVectorReplayBuffer(10, 2)
    assert np.array_equal(buf.get_buffer_indices(2, 4), g["gbi_2_4"]) and list(g["gbi_2_4"]) == [2, 3]
    assert np.array_equal(buf.get_buffer_indices(4, 2), g["gbi_4_2"]) and list(g["gbi_4_2"]) == [4, 0, 1]
    assert np.array_equal(buf.get_buffer_indices(8, 7), g["gbi_8_7"]) and list(g["gbi_8_7"]) == [8, 9, 5, 6]
    for bad in [(1, 6), (8, 1)]:
        with pytest.raises(ValueError):
            buf.get_buffer_indices(*bad)
