"""K3 parity (GPU): csrc/gae.cu through the C ABI vs the pinned oracle and the reference's own
known-answer vectors.  Tolerance: float64 path 1e-12 (re-association of a float64 scan), float32
path rtol 1e-5 (north_star) with atol 1e-6*max|x| for near-zero advantages."""
import numpy as np
import pytest
import torch

from oracle import gae as ogae

pytestmark = pytest.mark.gpu


class _FakeBuf:
    """what compute_episodic_return needs from a buffer: unfinished_index(), maxsize"""

    def __init__(self, unfinished, maxsize):
        self._u, self.maxsize = np.asarray(unfinished, np.int64), maxsize

    def unfinished_index(self):
        return self._u


def _batch(**kw):
    from tianshou_marl_b200.data import Batch

    return Batch(**kw)


@pytest.mark.parametrize("ci", range(5))
def test_reference_known_answers(golden, ci):
    """test/base/test_returns.py:23-158 through the drop-in API with the device ReplayBuffer."""
    from tianshou_marl_b200.algorithm import compute_episodic_return
    from tianshou_marl_b200.data import Batch, ReplayBuffer

    g = golden("gae.npz")
    term, trunc, rew, v = g[f"k{ci}_term"], g[f"k{ci}_trunc"], g[f"k{ci}_rew"], g[f"k{ci}_v"]
    gamma, lam = g[f"k{ci}_gl"]
    buf = ReplayBuffer(20)
    batch = Batch(terminated=term, truncated=trunc, rew=rew)
    for b in batch:
        b.obs = b.act = 1
        buf.add(b)
    ret, adv = compute_episodic_return(batch, buf, buf.sample_indices(0), None if len(v) == 0 else v, None, gamma, lam)
    assert isinstance(ret, np.ndarray) and ret.dtype == np.float64
    assert np.allclose(ret, g[f"k{ci}_ans"])
    assert np.allclose(ret, g[f"k{ci}_ret"], rtol=1e-12, atol=1e-12)
    assert np.allclose(adv, g[f"k{ci}_adv"], rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("ri", range(5))
def test_golden_vector_buffers(golden, ri):
    from tianshou_marl_b200.algorithm import compute_episodic_return

    g = golden("gae.npz")
    p = f"r{ri}_"
    nbuf, size = g[p + "shape"]
    buf = _FakeBuf(g[p + "unfinished"], int(nbuf * size))
    batch = _batch(rew=g[p + "rew"], terminated=g[p + "term"], truncated=g[p + "trunc"])
    ret, adv = compute_episodic_return(batch, buf, g[p + "indices"], g[p + "v_s_"], g[p + "v_s"], 0.99, 0.95)
    assert np.allclose(ret, g[p + "ret"], rtol=1e-12, atol=1e-12)
    assert np.allclose(adv, g[p + "adv"], rtol=1e-12, atol=1e-12)
    ret_mc, _ = compute_episodic_return(batch, buf, g[p + "indices"], None, None, 0.9, 1.0)
    assert np.allclose(ret_mc, g[p + "ret_mc"], rtol=1e-12, atol=1e-12)
    # device-tensor path: float32 I/O
    dv = lambda x: torch.as_tensor(x, device="cuda")  # noqa: E731
    batch_d = _batch(rew=dv(g[p + "rew"]).float(), terminated=dv(g[p + "term"]), truncated=dv(g[p + "trunc"]))
    ret_d, adv_d = compute_episodic_return(batch_d, buf, dv(g[p + "indices"]), dv(g[p + "v_s_"]), dv(g[p + "v_s"]), 0.99, 0.95)
    assert ret_d.is_cuda and ret_d.dtype == torch.float32
    rew32 = g[p + "rew"].astype(np.float32).astype(np.float64)
    end = np.logical_or(g[p + "term"], g[p + "trunc"])
    end[np.isin(g[p + "indices"], g[p + "unfinished"])] = True
    want = ogae.gae(g[p + "v_s"], g[p + "v_s_"] * ~g[p + "term"], rew32, end, 0.99, 0.95)
    tol = 1e-6 * np.abs(want).max()
    assert np.allclose(adv_d.cpu().numpy(), want, rtol=1e-5, atol=tol)
    assert np.allclose(ret_d.cpu().numpy(), want + g[p + "v_s"], rtol=1e-5, atol=tol)


def _synthetic(n_rows, T, seed, p_term=0.01):
    """SURVEY 8(d) C5 generator at reduced size, env-major flat order."""
    rng = np.random.default_rng(seed)
    rew = rng.normal(size=(n_rows, T)).astype(np.float32)
    v_s = rng.normal(size=(n_rows, T)).astype(np.float32)
    phase = rng.integers(0, 25, size=(n_rows, 1))
    trunc = ((np.arange(T)[None, :] + phase) % 25) == 24
    term = rng.random((n_rows, T)) < p_term
    done = trunc | term
    v_n = np.roll(v_s, -1, axis=1)
    fresh = rng.normal(size=(n_rows, T)).astype(np.float32)
    v_n = np.where(done | (np.arange(T)[None, :] == T - 1), fresh, v_n).astype(np.float32)
    return rew, v_s, v_n, term, trunc


def _oracle_rows(rew, v_s, v_n, term, trunc, gamma, lam):
    end = (term | trunc).copy()
    end[:, -1] = True
    return ogae.gae(v_s.reshape(-1), (v_n * ~term).reshape(-1), rew.reshape(-1).astype(np.float64), end.reshape(-1), gamma, lam)


@pytest.mark.parametrize("n_rows,T", [(1, 1), (3, 7), (257, 128), (4099, 128), (1000, 25), (5, 5000), (1, 300000),
                                        (1, 1023), (1, 1024), (1, 1025), (2, 1024), (1, 2049), (3, 1024)])  # around the 1024-element tile
@pytest.mark.parametrize("mode", ["seg_len", "force_end"])
def test_flat_vs_oracle(n_rows, T, mode):
    from tianshou_marl_b200.algorithm import gae_flat

    rew, v_s, v_n, term, trunc = _synthetic(n_rows, T, 7 * n_rows + T)
    want = _oracle_rows(rew, v_s, v_n, term, trunc, 0.99, 0.95)
    dv = lambda x: torch.as_tensor(x.reshape(-1), device="cuda")  # noqa: E731
    force = None
    if mode == "force_end":
        f = np.zeros((n_rows, T), np.uint8)
        f[:, -1] = 1
        force = dv(f)
    ret, adv = gae_flat(dv(v_s), dv(v_n), dv(rew), dv(term), dv(trunc), force, T if mode == "seg_len" else 0, 0.99, 0.95)
    tol = 1e-6 * max(np.abs(want).max(), 1.0)
    assert np.allclose(adv.cpu().numpy(), want, rtol=1e-5, atol=tol)
    assert np.allclose(ret.cpu().numpy(), want + v_s.reshape(-1), rtol=1e-5, atol=tol)
    # float64 I/O: only re-association error
    ret64, adv64 = gae_flat(dv(v_s).double(), dv(v_n).double(), dv(rew).double(), dv(term), dv(trunc), force,
                            T if mode == "seg_len" else 0, 0.99, 0.95)
    assert np.allclose(adv64.cpu().numpy(), want, rtol=1e-11, atol=1e-11)


def test_flat_long_segments_lookback():
    """No end flag for 200k elements: the decoupled look-back has to chain across ~100 tiles."""
    from tianshou_marl_b200.algorithm import gae_flat

    rng = np.random.default_rng(3)
    n = 700_001
    rew = rng.normal(size=n)
    v_s = rng.normal(size=n).astype(np.float32).astype(np.float64)
    v_n = rng.normal(size=n).astype(np.float32).astype(np.float64)
    term = np.zeros(n, bool)
    trunc = np.zeros(n, bool)
    trunc[[199_999, 200_000, 555_555]] = True
    want = ogae.gae(v_s, v_n, rew, trunc.copy(), 0.999, 1.0)
    dv = lambda x: torch.as_tensor(x, device="cuda")  # noqa: E731
    _, adv = gae_flat(dv(v_s), dv(v_n), dv(rew), dv(term), dv(trunc), None, 0, 0.999, 1.0)
    assert np.allclose(adv.cpu().numpy(), want, rtol=1e-10, atol=1e-9)


@pytest.mark.parametrize("n_rows,T,start,length", [(1, 1, 0, 1), (37, 128, 0, 128), (1000, 25, 7, 25), (5000, 128, 100, 90), (70000, 16, 3, 16)])
def test_time_major_vs_oracle(n_rows, T, start, length):
    from tianshou_marl_b200.algorithm import gae_time_major

    rew, v_s, v_n, term, trunc = _synthetic(n_rows, length, 11 * n_rows + T)
    want = _oracle_rows(rew, v_s, v_n, term, trunc, 0.99, 0.95).reshape(n_rows, length)

    def ring(x):  # [rows][j] -> [size][rows] with slot (start + j) % size
        out = np.zeros((T, n_rows), x.dtype)
        out[(start + np.arange(length)) % T] = x.T
        return torch.as_tensor(out, device="cuda")

    ret, adv = gae_time_major(ring(v_s), ring(v_n), ring(rew), ring(term.astype(np.uint8)), ring(trunc.astype(np.uint8)),
                              start, length, 0.99, 0.95)
    slots = (start + np.arange(length)) % T
    got = adv.cpu().numpy()[slots].T
    # the serial float64 recurrence is the reference's op order: only the final float32 rounding differs
    assert np.array_equal(got, want.astype(np.float32))
    assert np.array_equal(ret.cpu().numpy()[slots].T, (want + v_s).astype(np.float32))


def test_mc_return_to_go(golden):
    from tianshou_marl_b200.algorithm import episode_mc_return_to_go

    g = golden("gae.npz")
    assert np.allclose(episode_mc_return_to_go(g["mc_rew"], 0.97), g["mc_out"], rtol=1e-12)


def test_full_size_properties():
    """BASELINE config C5 size (T=128, 1M rows): size-independent checks -- linearity in rew and
    agreement of the flat and time-major kernels on the same data."""
    from tianshou_marl_b200.algorithm import gae_flat, gae_time_major

    T, R = 128, 1 << 20
    gen = torch.Generator(device="cuda").manual_seed(0)
    rew = torch.randn(T, R, device="cuda", generator=gen)
    v_s = torch.randn(T, R, device="cuda", generator=gen)
    v_n = torch.randn(T, R, device="cuda", generator=gen)
    trunc = (torch.rand(T, R, device="cuda", generator=gen) < 0.04).to(torch.uint8)
    term = (torch.rand(T, R, device="cuda", generator=gen) < 0.01).to(torch.uint8)
    ret_tm, adv_tm = gae_time_major(v_s, v_n, rew, term, trunc, 0, T, 0.99, 0.95)
    em = lambda x: x.t().contiguous().view(-1)  # noqa: E731  time-major -> env-major flat
    ret_f, adv_f = gae_flat(em(v_s), em(v_n), em(rew), em(term), em(trunc), None, T, 0.99, 0.95)
    assert torch.allclose(adv_f.view(R, T).t(), adv_tm, rtol=1e-5, atol=1e-5)
    assert torch.allclose(ret_f.view(R, T).t(), ret_tm, rtol=1e-5, atol=1e-5)
    # linearity: gae(2*rew, 2*v) == 2*gae(rew, v) exactly in floating point (power of two)
    _, adv2 = gae_time_major(2 * v_s, 2 * v_n, 2 * rew, term, trunc, 0, T, 0.99, 0.95)
    assert torch.equal(adv2, 2 * adv_tm)
