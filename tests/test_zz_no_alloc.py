"""tsm_gae_flat_static (GPU): the allocation-free variant of tsm_gae_flat (scratch in a file-scope device array; no counterpart in
the reference) against tsm_gae_flat through the C ABI, two calls in a row with
different sizes (the shared scratch keeps the previous call's contents and must be re-initialised by the entry point), plus the
size limit.  Same kernel, same arithmetic: outputs must be bit-identical; both are also held to the tolerance of the GAE parity
tests (rtol 1e-5 against the oracle's float64 recurrence).  Collected last on purpose (file name)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(n, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    f = lambda: torch.randn(n, device="cuda", generator=g)  # noqa: E731
    v_s, v_n, rew = f(), f(), f()
    term = (torch.rand(n, device="cuda", generator=g) < 0.01).to(torch.uint8)
    trunc = (torch.rand(n, device="cuda", generator=g) < 0.04).to(torch.uint8)
    return v_s, v_n, rew, term, trunc


def _run_both(n, seed, seg_len):
    from tianshou_marl_b200 import _lib

    L, P, S = _lib.lib(), _lib.ptr, _lib.stream_ptr()
    v_s, v_n, rew, term, trunc = _inputs(n, seed)
    ws_bytes = int(L.tsm_gae_flat_workspace_bytes(n))
    ws = torch.empty(ws_bytes + 32, dtype=torch.uint8, device="cuda")
    adv0, ret0 = torch.empty(n, device="cuda"), torch.empty(n, device="cuda")
    adv1, ret1 = torch.full((n,), float("nan"), device="cuda"), torch.full((n,), float("nan"), device="cuda")
    _lib.check(L.tsm_gae_flat(P(v_s), P(v_n), P(rew), 0, P(term), P(trunc), None, n, seg_len, 0.99, 0.95, P(adv0), P(ret0), P(ws), ws_bytes, S))
    _lib.check(L.tsm_gae_flat_static(P(v_s), P(v_n), P(rew), 0, P(term), P(trunc), None, n, seg_len, 0.99, 0.95, P(adv1), P(ret1), S))
    torch.cuda.synchronize()
    return (v_s, v_n, rew, term, trunc), (adv0, ret0), (adv1, ret1)


def test_static_scratch_matches_workspace_variant_twice():
    from oracle import gae as ogae

    for n, seed, seg in ((1_000_003, 1, 0), (70_001, 2, 25), (1_000_003, 3, 128)):  # large, small (stale tile states beyond), large again
        (v_s, v_n, rew, term, trunc), (adv0, ret0), (adv1, ret1) = _run_both(n, seed, seg)
        assert torch.equal(adv0, adv1) and torch.equal(ret0, ret1), (n, seed)
        end = (term | trunc).cpu().numpy().astype(bool)
        if seg > 0:
            end[seg - 1::seg] = True
        want = ogae.gae(v_s.cpu().numpy().astype(np.float64), v_n.cpu().numpy().astype(np.float64) * (1 - term.cpu().numpy()),
                        rew.cpu().numpy().astype(np.float64), end, 0.99, 0.95)
        assert np.allclose(adv1.cpu().numpy(), want, rtol=1e-5, atol=1e-5)
        assert np.allclose(ret1.cpu().numpy(), want + v_s.cpu().numpy(), rtol=1e-5, atol=1e-5)


def test_static_scratch_limit_and_empty():
    from tianshou_marl_b200 import _lib

    L, P, S = _lib.lib(), _lib.ptr, _lib.stream_ptr()
    lim = int(L.tsm_gae_flat_static_max_n())
    assert lim == 16777216
    x = torch.zeros(8, device="cuda")
    b = torch.zeros(8, dtype=torch.uint8, device="cuda")
    with pytest.raises(ValueError, match="TSM_GAE_FLAT_STATIC_MAX_N"):  # rejected before any pointer is touched
        _lib.check(L.tsm_gae_flat_static(P(x), P(x), P(x), 0, P(b), P(b), None, lim + 1, 0, 0.99, 0.95, P(x), P(x), S))
    _lib.check(L.tsm_gae_flat_static(P(x), P(x), P(x), 0, P(b), P(b), None, 0, 0, 0.99, 0.95, P(x), P(x), S))
