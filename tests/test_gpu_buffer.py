"""K2 parity (GPU): csrc/buffer.cu through the C ABI, bit-exact against traces recorded from the
reference VectorReplayBuffer and against test/base/test_buffer.py known answers."""
import numpy as np
import pytest
import torch

from _replay import replay_buffer_trace
from oracle.buffer import OracleVectorReplayBuffer

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("ci", range(5))
def test_trace_numpy_api(golden, ci):
    from tianshou_marl_b200.data import Batch, VectorReplayBuffer

    replay_buffer_trace(golden("buffer.npz"), ci, VectorReplayBuffer, lambda **kw: Batch(**kw))


@pytest.mark.parametrize("ci", [0, 2])
def test_trace_device_api(golden, ci):
    from tianshou_marl_b200.data import Batch, VectorReplayBuffer

    def mk(**kw):
        return Batch(**{k: torch.as_tensor(v, device="cuda") for k, v in kw.items()})

    replay_buffer_trace(golden("buffer.npz"), ci, lambda t, n: VectorReplayBuffer(t, n, index_type="torch"), mk,
                        check_random=False)


def test_reference_known_answers():
    """test/base/test_buffer.py:740-790 (VectorReplayBuffer(20, 4))."""
    from tianshou_marl_b200.data import Batch, VectorReplayBuffer

    buf = VectorReplayBuffer(20, 4)
    ptr, ep_rew, ep_len, ep_idx = buf.add(
        Batch(obs=[1, 2, 3], act=[1, 2, 3], rew=[1, 2, 3], terminated=[0, 0, 1], truncated=[0, 0, 0]), buffer_ids=[0, 1, 2])
    assert np.all(ep_len == [0, 0, 1]) and np.all(ep_rew == [0, 0, 3])
    assert np.all(ptr == [0, 5, 10]) and np.all(ep_idx == [0, 5, 10])
    with pytest.raises(NotImplementedError):
        buf.update(buf)
    indices = buf.sample_indices(11000)
    assert np.bincount(indices)[[0, 5, 10]].min() >= 3000
    batch, indices = buf.sample(0)
    assert np.allclose(indices, [0, 5, 10])
    assert np.allclose(buf.prev(indices), indices) and np.allclose(buf.next(indices), indices)
    assert np.allclose(buf.unfinished_index(), [0, 5])
    buf.add(Batch(obs=[4], act=[4], rew=[4], terminated=[1], truncated=[0]), buffer_ids=[3])
    assert np.allclose(buf.unfinished_index(), [0, 5])
    batch, indices = buf.sample(10)
    batch, indices = buf.sample(0)
    assert np.allclose(indices, [0, 5, 10, 15])
    data = np.array([0, 0, 0, 0])
    buf.add(Batch(obs=data, act=data, rew=data, terminated=data, truncated=data), buffer_ids=[0, 1, 2, 3])
    buf.add(Batch(obs=data, act=data, rew=data, terminated=1 - data, truncated=data), buffer_ids=[0, 1, 2, 3])
    assert len(buf) == 12
    buf.add(Batch(obs=data, act=data, rew=data, terminated=data, truncated=data), buffer_ids=[0, 1, 2, 3])
    buf.add(Batch(obs=data, act=data, rew=data, terminated=[0, 1, 0, 1], truncated=data), buffer_ids=[0, 1, 2, 3])
    assert len(buf) == 20
    assert np.bincount(buf.sample_indices(120000)).min() >= 5000
    assert np.allclose(buf.sample_indices(0), np.arange(len(buf)))
    indices = buf.sample_indices(0)
    assert np.allclose(buf.done, [0, 0, 1, 0, 0, 0, 0, 1, 0, 1, 1, 0, 1, 0, 0, 1, 0, 1, 0, 1])
    assert np.allclose(buf.prev(indices), [0, 0, 1, 3, 3, 5, 5, 6, 8, 8, 10, 11, 11, 13, 13, 15, 16, 16, 18, 18])
    assert np.allclose(buf.next(indices), [1, 2, 2, 4, 4, 6, 7, 7, 9, 9, 10, 12, 12, 14, 14, 15, 17, 17, 19, 19])
    assert np.allclose(buf.unfinished_index(), [4, 14])
    ptr, ep_rew, ep_len, ep_idx = buf.add(Batch(obs=[1], act=[1], rew=[1], terminated=[1], truncated=[0]), buffer_ids=[2])
    assert np.all(ep_len == [3]) and np.all(ep_rew == [1]) and np.all(ptr == [10]) and np.all(ep_idx == [13])
    assert np.allclose(buf.unfinished_index(), [4])
    assert np.allclose(np.array(sorted(buf.sample_indices(0))), np.arange(len(buf)))
    # corner case: int and -1 (test_buffer.py:957-958)
    assert buf.prev(-1) == buf.prev(np.array([buf.maxsize - 1]))[0]
    assert buf.next(-1) == buf.next(np.array([buf.maxsize - 1]))[0]
    assert buf.sample_indices(-1).tolist() == []
    assert buf.done.shape == (20,) and buf.obs.shape[0] == 20


def test_get_buffer_indices(golden):
    from tianshou_marl_b200.data import VectorReplayBuffer

    g = golden("buffer.npz")
    buf = VectorReplayBuffer(10, 2)
    assert list(buf.get_buffer_indices(2, 4)) == [2, 3] == list(g["gbi_2_4"])
    assert list(buf.get_buffer_indices(4, 2)) == [4, 0, 1] == list(g["gbi_4_2"])
    assert list(buf.get_buffer_indices(8, 7)) == [8, 9, 5, 6] == list(g["gbi_8_7"])
    for bad in [(1, 6), (8, 1)]:
        with pytest.raises(ValueError):
            buf.get_buffer_indices(*bad)


@pytest.mark.parametrize("nbuf,size,steps", [(3000, 7, 20), (1025, 128, 140), (70000, 4, 9)])
def test_random_large_vs_oracle(nbuf, size, steps):
    """Ragged adds on many sub-buffers vs the pinned numpy oracle: every index table bit-exact."""
    from tianshou_marl_b200.data import Batch, VectorReplayBuffer

    rng = np.random.default_rng(nbuf)
    buf = VectorReplayBuffer(nbuf * size, nbuf)
    orc = OracleVectorReplayBuffer(nbuf * size, nbuf)
    for t in range(steps):
        ids = np.nonzero(rng.random(nbuf) < 0.9)[0] if t % 3 else np.arange(nbuf)
        n = len(ids)
        b = dict(obs=rng.normal(size=(n, 6)).astype(np.float32), act=rng.integers(0, 5, n),
                 rew=rng.normal(size=n).astype(np.float32), terminated=rng.random(n) < 0.05, truncated=rng.random(n) < 0.1)
        b["obs_next"] = b["obs"] + 1
        got = buf.add(Batch(**b), buffer_ids=ids)
        want = orc.add(b, buffer_ids=ids)
        for a, w in zip(got, want):
            assert np.allclose(a, w, rtol=1e-6, atol=1e-6)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[2], want[2]) and np.array_equal(got[3], want[3])
    assert len(buf) == len(orc)
    assert np.array_equal(buf.sample_indices(0), orc.sample_indices(0))
    assert np.array_equal(buf.unfinished_index(), orc.unfinished_index())
    probe = rng.integers(0, buf.maxsize, 5000)
    assert np.array_equal(buf.prev(probe), orc.prev(probe))
    assert np.array_equal(buf.next(probe), orc.next(probe))
    idx = orc.sample_indices(0)
    got = buf[idx]
    assert np.array_equal(got.obs, orc.data["obs"][idx]) and np.array_equal(got.obs_next, orc.data["obs_next"][idx])
    assert np.array_equal(got.act, orc.data["act"][idx]) and np.array_equal(got.done, orc.data["done"][idx])
    assert np.array_equal(got.rew, orc.data["rew"][idx].astype(np.float32))


def test_empty_and_errors():
    from tianshou_marl_b200.data import Batch, VectorReplayBuffer

    buf = VectorReplayBuffer(8, 2)
    assert len(buf) == 0 and len(buf.sample_indices(0)) == 0 and len(buf.unfinished_index()) == 0
    with pytest.raises(ValueError):
        buf.add(Batch(obs=[1], act=[1], rew=[1], terminated=[0]))  # missing key
    with pytest.raises(ValueError):
        buf.add(Batch(obs=[1], act=[1], rew=[1], terminated=[0], truncated=[0]))  # n != buffer_num without ids
    with pytest.raises(NotImplementedError):
        VectorReplayBuffer(8, 2, stack_num=4)
