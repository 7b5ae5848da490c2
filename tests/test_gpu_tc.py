"""tcgen05 engine (csrc/mlp_tc.cu) vs the exact fp32 engine (csrc/mlp.cu, itself pinned to the
reference fixtures in test_gpu_ppo.py).  Operands are rounded to fp16 (11-bit significand), sums are
fp32: forward outputs agree to ~2e-3 absolute on O(1) values; gradients to ~1% of their scale."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _h(a, dt=torch.float16):  # [rows][F] -> 16-bit image [F/8][rows][8] as float32 words, and the rounded values
    r, f = a.shape
    t = torch.as_tensor(a).to(dt).reshape(r, f // 8, 8).permute(1, 0, 2).contiguous()
    return t.view(torch.float32).reshape(-1).numpy(), t.permute(1, 0, 2).reshape(r, f).double().numpy()


def _gemm(a_img, b_img, desc, a_mn, b_mn, n, ksteps):
    from tianshou_marl_b200 import _lib

    d = torch.full((128, n), float("nan"), device="cuda")
    A, B = torch.as_tensor(a_img, device="cuda"), torch.as_tensor(b_img, device="cuda")
    h = (C.c_uint32 * 6)(*desc)
    _lib.check(_lib.lib().tsm_tc_selftest(A.data_ptr(), A.numel(), B.data_ptr(), B.numel(), C.cast(h, C.c_void_p), a_mn, b_mn, n, ksteps,
                                          1, 0, 0, d.data_ptr(), _lib.stream_ptr()))
    torch.cuda.synchronize()
    return d.cpu().numpy()


def test_descriptor_semantics():
    """The three operand views mlp_tc_kernels.cuh relies on, one tcgen05.mma group each."""
    rng = np.random.default_rng(0)
    N, K, O = 64, 32, 16
    ai, a = _h(rng.normal(size=(128, K)).astype(np.float32))
    bi, w = _h(rng.normal(size=(N, K)).astype(np.float32))
    d = _gemm(ai, bi, [2048, 128, 4096, N * 16, 128, 2 * N * 16], 0, 0, N, K // 16)  # rows x feats times weights[out][in]
    assert np.abs(d - a @ w.T).max() < 1e-4
    ai, a = _h(rng.normal(size=(128, O)).astype(np.float32))
    bi, w = _h(rng.normal(size=(O, N)).astype(np.float32))  # W[out][in] image read as N = in, K = out
    d = _gemm(ai, bi, [2048, 128, 4096, 128, O * 16, 256], 0, 1, N, O // 16)
    assert np.abs(d - a @ w).max() < 1e-4
    pi, p = _h(rng.normal(size=(128, 128)).astype(np.float32))
    qi, q = _h(rng.normal(size=(128, N)).astype(np.float32))
    d = _gemm(pi, qi, [128, 2048, 256, 128, 2048, 256], 1, 1, N, 128 // 16)  # K = rows for both
    assert np.abs(d - p.T @ q).max() < 1e-3


@pytest.mark.parametrize("O,n", [(18, 128), (18, 5000), (16, 777), (7, 130), (31, 1000)])
def test_forward_tc_vs_fp32(O, n):
    from tianshou_marl_b200.utils import ActorCriticNet

    net = ActorCriticNet(O, 64, 5, seed=O)
    with torch.no_grad():
        net.view("b1").normal_(0, 0.1); net.view("b2").normal_(0, 0.1); net.view("b3").normal_(0, 0.1)
    obs = torch.randn(n, O, device="cuda")
    net.engine = "fp32"
    ref = net.forward(obs, mode="mode", want_logits=True, want_entropy=True)
    net.engine = "tc"
    got = net.forward(obs, mode="mode", want_logits=True, want_entropy=True)
    assert (got["logits"] - ref["logits"]).abs().max().item() < 4e-3
    assert (got["value"] - ref["value"]).abs().max().item() < 4e-3
    assert (got["logp"] - ref["logp"]).abs().max().item() < 4e-3
    assert (got["act"] == ref["act"]).float().mean().item() > 0.99  # arg-max flips only on near ties
    # sampling: same Philox stream, so actions agree except where a CDF edge moved by the rounding
    a1 = net.forward(obs, mode="sample", seed=3, counter=9)["act"]
    net.engine = "fp32"
    a0 = net.forward(obs, mode="sample", seed=3, counter=9)["act"]
    assert (a0 == a1).float().mean().item() > 0.99
    # gather / scatter rows
    rows = torch.randperm(n, device="cuda")[: n // 2].contiguous()
    net.engine = "tc"
    sub = net.forward(obs, mode="raw", rows=rows)
    full = net.forward(obs, mode="raw")
    assert torch.allclose(sub["out"], full["out"][rows], atol=1e-6)


@pytest.mark.parametrize("O,n", [(18, 128), (18, 3000), (16, 65536), (9, 200)])
def test_ppo_minibatch_tc_vs_fp32(O, n):
    from tianshou_marl_b200 import _lib
    from tianshou_marl_b200.utils import ActorCriticNet

    rng = np.random.default_rng(n)
    net = ActorCriticNet(O, 64, 5, seed=1)
    with torch.no_grad():
        net.view("b1").normal_(0, 0.1); net.view("b2").normal_(0, 0.1); net.view("b3").normal_(0, 0.1)
    d = lambda a, dt=torch.float32: torch.as_tensor(a, device="cuda").to(dt).contiguous()  # noqa: E731
    t = dict(obs=d(rng.normal(size=(n, O))), act=d(rng.integers(0, 5, n), torch.int32), adv=d(rng.normal(size=n)), ret=d(rng.normal(size=n)),
             lp=d(np.log(0.2) + 0.3 * rng.normal(size=n)), vo=d(rng.normal(size=n)))
    P = _lib.ptr
    key = 0xABCDEF1234567
    for cfgk in (dict(adv_norm=1), dict(adv_norm=0, value_clip=1, dual_clip=2.0, eps_clip=0.1)):
        pb = _lib.TsmPpoBatch(obs=P(t["obs"]), act=P(t["act"]), adv=P(t["adv"]), ret=P(t["ret"]), logp_old=P(t["lp"]), v_old=P(t["vo"]),
                              indices=None, size=1, n_rows=n, start=0, len=1, perm_key=key, perm_n=n, records=None)
        rec = torch.empty(n, _lib.PPO_RECORD_BYTES, dtype=torch.uint8, device="cuda")
        _lib.check(_lib.lib().tsm_ppo_pack_records(net.desc, C.byref(pb), P(rec), _lib.stream_ptr()))
        pb.records = P(rec)
        pc = _lib.TsmPpoCfg(eps_clip=cfgk.get("eps_clip", 0.2), dual_clip=cfgk.get("dual_clip", 0.0), vf_coef=0.5, ent_coef=0.01,
                            value_clip=cfgk.get("value_clip", 0), adv_norm=cfgk["adv_norm"])
        stats = d(np.array([float(t["adv"].mean()), 1.0 / (float(t["adv"].std()) + 1e-8)]))
        res = {}
        for eng in ("fp32", "tc"):
            net.engine = eng
            net.grads.zero_()
            sums = torch.zeros(4, dtype=torch.float64, device="cuda")
            _lib.check(_lib.lib().tsm_ppo_minibatch(net.desc, C.byref(pb), C.byref(pc), 0, n, P(stats), P(sums), _lib.stream_ptr()))
            res[eng] = (net.grads.clone(), sums.cpu().numpy())
        g0, g1 = res["fp32"][0], res["tc"][0]
        off = 0
        for name, (o_, shp) in net._views.items():
            k = int(np.prod(shp))
            a, b = g0[o_: o_ + k], g1[o_: o_ + k]
            scale = a.abs().max().item() + 1e-12
            err = (a - b).abs().max().item()
            # forward outputs carry ~2e-3 absolute rounding (11-bit operands), which moves the PPO ratio / clip decisions of
            # individual rows; the effect on a gradient entry averages out with the minibatch size
            assert err < (0.05 if n >= 3000 else 0.12) * scale + 1e-7, (name, err, scale, cfgk)
        assert np.allclose(res["tc"][1][:3] / n, res["fp32"][1][:3] / n, rtol=5e-3, atol=2e-3)
    net.engine = "fp32"


def test_pack_records_bit_exact():
    """tsm_ppo_pack_records: 96-byte rows at a 128-byte pitch = fp16(obs) zero padded with x[31] = 1 | adv ret logp_old | act | v_old | pad."""
    from tianshou_marl_b200 import _lib
    from tianshou_marl_b200.utils import ActorCriticNet

    rng = np.random.default_rng(5)
    for O, size, R in ((18, 3, 37), (7, 2, 5), (31, 1, 130)):
        n = size * R
        net = ActorCriticNet(O, 64, 5, seed=1)
        obs = rng.normal(size=(n, O)).astype(np.float32) * 3
        act = rng.integers(0, 5, n).astype(np.int32)
        adv, ret, lp, vo = (rng.normal(size=n).astype(np.float32) for _ in range(4))
        d = lambda a: torch.as_tensor(a, device="cuda").contiguous()  # noqa: E731
        t = [d(x) for x in (obs, act, adv, ret, lp, vo)]
        P = _lib.ptr
        for with_vo in (True, False):
            pb = _lib.TsmPpoBatch(obs=P(t[0]), act=P(t[1]), adv=P(t[2]), ret=P(t[3]), logp_old=P(t[4]), v_old=P(t[5]) if with_vo else None,
                                  indices=None, size=size, n_rows=R, start=0, len=size, perm_key=0, perm_n=0, records=None)
            rec = torch.full((n, _lib.PPO_RECORD_BYTES), 0xAB, dtype=torch.uint8, device="cuda")
            _lib.check(_lib.lib().tsm_ppo_pack_records(net.desc, C.byref(pb), P(rec), _lib.stream_ptr()))
            h = rec.cpu().numpy()
            x = h[:, :64].copy().view(np.float16)
            want = np.zeros((n, 32), np.float16)
            want[:, :O] = obs.astype(np.float16)
            want[:, 31] = 1.0
            assert np.array_equal(x.view(np.uint16), want.view(np.uint16))
            sc = h[:, 64:80].copy().view(np.float32)
            assert np.array_equal(sc[:, 0], adv) and np.array_equal(sc[:, 1], ret) and np.array_equal(sc[:, 2], lp)
            assert np.array_equal(h[:, 76:80].copy().view(np.int32)[:, 0], act)
            assert np.array_equal(h[:, 80:84].copy().view(np.float32)[:, 0], vo if with_vo else np.zeros(n, np.float32))
            assert not h[:, 84:96].any()  # (bytes 96..127 of the 128-byte pitch are never written)
            # tsm_ppo_pack_records_rows: the same records for the listed sub-buffers, everything else untouched
            sel = np.sort(rng.choice(R, size=max(1, R // 3), replace=False)).astype(np.int64)
            rec2 = torch.full((n, _lib.PPO_RECORD_BYTES), 0xAB, dtype=torch.uint8, device="cuda")
            _lib.check(_lib.lib().tsm_ppo_pack_records_rows(net.desc, C.byref(pb), P(rec2), P(d(sel)), len(sel), _lib.stream_ptr()))
            h2 = rec2.cpu().numpy().reshape(size, R, -1)
            mask = np.zeros(R, bool)
            mask[sel] = True
            assert np.array_equal(h2[:, mask], h.reshape(size, R, -1)[:, mask]) and (h2[:, ~mask] == 0xAB).all()


def test_minibatch_tc_needs_records():
    from tianshou_marl_b200 import _lib
    from tianshou_marl_b200.utils import ActorCriticNet

    net = ActorCriticNet(18, 64, 5, seed=1, engine="tc")
    x = torch.zeros(128, 18, device="cuda")
    z = torch.zeros(128, device="cuda")
    a = torch.zeros(128, dtype=torch.int32, device="cuda")
    P = _lib.ptr
    pb = _lib.TsmPpoBatch(obs=P(x), act=P(a), adv=P(z), ret=P(z), logp_old=P(z), v_old=None, indices=None, size=1, n_rows=128, start=0,
                          len=1, perm_key=0, perm_n=0, records=None)
    pc = _lib.TsmPpoCfg(eps_clip=0.2, dual_clip=0.0, vf_coef=0.5, ent_coef=0.01, value_clip=0, adv_norm=0)
    with pytest.raises(ValueError):
        _lib.check(_lib.lib().tsm_ppo_minibatch(net.desc, C.byref(pb), C.byref(pc), 0, 128, None, None, _lib.stream_ptr()))


def test_update_tc_tracks_fp32():
    """Whole PPO.update on both engines from the same start: losses track each other and fall."""
    from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory, DiscreteActorPolicy
    from tianshou_marl_b200.data import Batch, VectorReplayBuffer
    from tianshou_marl_b200.utils import ActorCriticNet

    R, T, O = 1024, 32, 18
    out = {}
    for eng in ("fp32", "tc"):
        net = ActorCriticNet(O, 64, 5, seed=4, engine=eng)
        pol = DiscreteActorPolicy(actor=net, seed=5)
        alg = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=1e-3), minibatch_order="device", max_grad_norm=0.5, seed=1)
        buf = VectorReplayBuffer(R * T, R, index_type="torch")
        g = torch.Generator(device="cuda").manual_seed(0)
        obs = torch.randn(R, O, device="cuda", generator=g)
        for t in range(T):
            act = torch.randint(0, 5, (R,), device="cuda", generator=g)
            nxt = obs + 0.1 * torch.randn(R, O, device="cuda", generator=g)
            buf.add(Batch(obs=obs, act=act, rew=(obs[:, 0] * (act == 2).float()).contiguous(), terminated=torch.zeros(R, dtype=torch.bool, device="cuda"),
                          truncated=torch.full((R,), (t + 1) % 8 == 0, device="cuda"), obs_next=nxt))
            obs = nxt
        pol.is_within_training_step = True
        alg.update(buf, 4096, 4)
        out[eng] = alg.last_sequences
    for k in ("loss", "vf", "ent", "clip"):
        assert np.allclose(out["tc"][k], out["fp32"][k], rtol=3e-2, atol=3e-3), (k, np.abs(out["tc"][k] - out["fp32"][k]).max())
    assert out["tc"]["vf"][-1] < out["tc"]["vf"][0]


def test_slot_table_matches_oracle_permutation():
    """tsm_ppo_slot_table / the adv_stats by-product: slots[p] = storage slot of batch position p (lock-step ring view,
    Feistel permutation restated in oracle/philox.py; explicit indices)."""
    from oracle.philox import feistel_permute
    from tianshou_marl_b200 import _lib

    P, L, S = _lib.ptr, _lib.lib(), _lib.stream_ptr()
    size, R, start, ln = 7, 53, 5, 4
    B = ln * R
    key = 0x1234ABCD5678EF
    adv = torch.randn(size * R, device="cuda")
    slots = torch.full((B,), -1, dtype=torch.int32, device="cuda")
    pb = _lib.TsmPpoBatch(obs=None, act=None, adv=P(adv), ret=None, logp_old=None, v_old=None, indices=None, size=size, n_rows=R, start=start,
                          len=ln, perm_key=key, perm_n=B, records=None, slots=P(slots))
    q = feistel_permute(np.arange(B, dtype=np.uint64), B, key).astype(np.int64)
    row, j = q // ln, q % ln
    want = ((start + j) % size) * R + row
    _lib.check(L.tsm_ppo_slot_table(C.byref(pb), 0, B, S))
    assert np.array_equal(slots.cpu().numpy(), want)
    # by-product of adv_stats: only the visited positions are written
    slots.fill_(-1)
    bounds = torch.tensor([10, 100, B - 3], dtype=torch.int64, device="cuda")
    stats = torch.zeros(2, 2, device="cuda")
    _lib.check(L.tsm_ppo_adv_stats(C.byref(pb), P(bounds), 2, 1e-8, P(stats), S))
    got = slots.cpu().numpy()
    assert np.array_equal(got[10:B - 3], want[10:B - 3]) and (got[:10] == -1).all() and (got[B - 3:] == -1).all()
    a = adv.cpu().numpy()[want[10:100]]
    assert np.allclose(stats.cpu().numpy()[0], [a.mean(), 1.0 / (a.std(ddof=1) + 1e-8)], rtol=1e-5)
    # explicit env-major indices (i = row * size + t)
    idx = torch.randperm(size * R, device="cuda")[:B].contiguous()
    pb.indices, pb.perm_n = P(idx), 0
    _lib.check(L.tsm_ppo_slot_table(C.byref(pb), 0, B, S))
    i = idx.cpu().numpy()
    assert np.array_equal(slots.cpu().numpy(), (i % size) * R + i // size)


@pytest.mark.parametrize("n_mb,mb", [(5, 1024), (3, 20000), (1, 700)])
def test_epoch_slot_table_and_launch_chain_agree(n_mb, mb, monkeypatch):
    """tsm_ppo_epoch, engine 1: the persistent launch (with and without a slot table) and the chain of dependent launches give
    the same parameters after an epoch (the last minibatch is longer: Batch.split merge_last)."""
    from tianshou_marl_b200 import _lib
    from tianshou_marl_b200.utils import ActorCriticNet

    P, L, S = _lib.ptr, _lib.lib(), _lib.stream_ptr()
    O, size = 18, 4
    total = n_mb * mb + (77 if n_mb > 1 else 0)
    R = (total + size - 1) // size
    B = size * R
    g = torch.Generator(device="cuda").manual_seed(n_mb)
    obs = torch.randn(B, O, device="cuda", generator=g)
    act = torch.randint(0, 5, (B,), device="cuda", generator=g, dtype=torch.int32)
    adv, ret = torch.randn(B, device="cuda", generator=g), torch.randn(B, device="cuda", generator=g)
    lp = torch.full((B,), float(np.log(0.2)), device="cuda")
    bounds = np.array([m * mb for m in range(n_mb)] + [total], dtype=np.int64)
    hb = bounds.ctypes.data_as(C.c_void_p)
    bounds_dev = torch.as_tensor(bounds, device="cuda")
    pc = _lib.TsmPpoCfg(eps_clip=0.2, dual_clip=0.0, vf_coef=0.5, ent_coef=0.01, value_clip=0, adv_norm=1)
    res = {}
    for variant in ("chain", "persistent", "persistent+slots"):
        if variant == "chain":
            monkeypatch.setenv("TSM_NO_PERSISTENT", "1")
        else:
            monkeypatch.delenv("TSM_NO_PERSISTENT", raising=False)
        net = ActorCriticNet(O, 64, 5, seed=9, engine="tc")
        slots = torch.full((B,), -1, dtype=torch.int32, device="cuda")
        pb = _lib.TsmPpoBatch(obs=P(obs), act=P(act), adv=P(adv), ret=P(ret), logp_old=P(lp), v_old=None, indices=None, size=size, n_rows=R,
                              start=0, len=size, perm_key=99, perm_n=B, records=None, slots=P(slots) if variant.endswith("slots") else None)
        rec = torch.empty(B, _lib.PPO_RECORD_BYTES, dtype=torch.uint8, device="cuda")
        _lib.check(L.tsm_ppo_pack_records(net.desc, C.byref(pb), P(rec), S))
        pb.records = P(rec)
        stats = torch.zeros(n_mb, 2, device="cuda")
        sums = torch.zeros(n_mb, 4, dtype=torch.float64, device="cuda")
        _lib.check(L.tsm_ppo_adv_stats(C.byref(pb), P(bounds_dev), n_mb, 1e-8, P(stats), S))
        _lib.check(L.tsm_ppo_epoch(net.desc, C.byref(pb), C.byref(pc), hb, n_mb, P(stats), 1e-3, 0.9, 0.999, 1e-8, 0.5, P(sums), S))
        torch.cuda.synchronize()
        res[variant] = (net.flat.detach().clone(), sums.cpu().numpy(), int(net.adam_step.item()), net.grads.abs().max().item())
    for variant in ("persistent", "persistent+slots"):
        assert res[variant][2] == n_mb and res[variant][3] == 0.0
        # same arithmetic, different summation order of the gradient reductions / the norm.  Adam's first steps move a
        # parameter by about lr * sign(g), so an entry whose gradient is rounding noise may differ by up to 2 * lr per step.
        # The four-group kernel accumulates the weight gradients of its groups in ONE TMEM accumulator, in the order the
        # groups' MMAs happen to issue: run-to-run rounding differences of 1e-8 in a parameter occasionally flip its fp16 image
        # entry (or a row across the ratio-clip edge) in a middle minibatch, after which ~6 % of the entries differ by up to
        # lr / 4 (measured on the (5, 1024) case: losses of the first two minibatches identical, 4e-5 relative afterwards).
        diff = (res[variant][0] - res["chain"][0]).abs()
        assert (diff < 2e-5).float().mean().item() > 0.90 and diff.max().item() < 2e-3 * n_mb + 1e-6, (variant, diff.max())
        assert np.allclose(res[variant][1][:1], res["chain"][1][:1], rtol=1e-6, atol=1e-6)  # the first minibatch: same parameters, same sums
        assert np.allclose(res[variant][1], res["chain"][1], rtol=1e-4, atol=1e-3)


# ---- the benched engine against the REFERENCE fixtures ----
# Tolerances = about 2x what tools/tc_error_budget.py measured on B200 (profiles/r02_tc_error_budget.json):
#   per-minibatch loss / clip / vf / ent: |d| <= 1e-3 + 1e-3 * |ref|          (measured <= 4.5e-4 abs, 3e-4 rel)
#   v_s, returns, adv, logp_old:          |d| <= 2e-3                          (measured <= 9.6e-4: fp16 operands)
#   parameters after the whole update:    RMS error <= 10 % of the RMS distance the reference moved them (measured 1.3 - 5 %),
#                                         and every entry within lr * steps (Adam moves an entry by at most ~lr per step; entries
#                                         whose gradient is rounding noise take +-lr steps of either sign: measured <= 0.48)
# The fp32 engine on the same C2-scale fixture: losses rtol 2e-5 / atol 3e-5, parameters RMS <= 0.5 % of the travel, max 5e-4.
TC_REF_TOL = dict(loss_rel=1e-3, loss_abs=1e-3, pre_abs=2e-3, param_rms_frac=0.10, param_abs_lr_steps=1.0)
FP32_LARGE_TOL = dict(loss_rel=2e-5, loss_abs=3e-5, pre_abs=4e-6, param_rms_frac=5e-3, param_abs=5e-4)


def _mk_case(c, engine):
    from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory, DiscreteActorPolicy
    from tianshou_marl_b200.data import Batch, VectorReplayBuffer
    from tianshou_marl_b200.utils import ActorCriticNet

    net = ActorCriticNet(c["O"], c["H"], c["nA"], engine=engine).load_arrays(c["W1"], c["b1"], c["W2"], c["b2"], c["W3"], c["b3"])
    pol = DiscreteActorPolicy(actor=net)
    k = c["cfg"]
    alg = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=c["lr"]), eps_clip=k["eps_clip"], dual_clip=k["dual_clip"] or None,
              value_clip=k["value_clip"], advantage_normalization=k["adv_norm"], vf_coef=k["vf_coef"], ent_coef=k["ent_coef"],
              max_grad_norm=c["max_norm"] or None, gamma=c["gamma"], gae_lambda=c["lam"], minibatch_order="numpy")
    buf = VectorReplayBuffer(c["R"] * c["T"], c["R"], index_type="torch")
    for t in range(c["T"]):
        buf.add(Batch(obs=c["obs"][t], act=c["act"][t], rew=c["rew"][t], terminated=c["term"][t], truncated=c["trunc"][t], obs_next=c["obs_next"][t]))
    pol.is_within_training_step = True
    return net, alg, buf


def _check_against_reference(c, net, alg, tol):
    seq = alg.last_sequences
    for k, ref in (("loss", "loss_seq"), ("clip", "clip_seq"), ("vf", "vf_seq"), ("ent", "ent_seq")):
        assert np.allclose(seq[k], c[ref], rtol=tol["loss_rel"], atol=tol["loss_abs"]), (k, np.abs(seq[k] - c[ref]).max())
    v = net.named_views()
    names = ("W1", "b1", "W2", "b2", "W3", "b3")
    err = np.concatenate([(v[k].cpu().numpy() - c["after_" + k]).reshape(-1) for k in names])
    trav = np.concatenate([(c[k] - c["after_" + k]).reshape(-1) for k in names])
    rms_frac = float(np.sqrt((err ** 2).mean() / (trav ** 2).mean()))
    assert rms_frac < tol["param_rms_frac"], rms_frac
    bound = tol["param_abs"] if "param_abs" in tol else tol["param_abs_lr_steps"] * c["lr"] * len(c["loss_seq"])
    assert np.abs(err).max() < bound, (np.abs(err).max(), bound)


@pytest.mark.parametrize("ci", [0, 1])
def test_update_tc_matches_reference(golden, ci):
    """PPO.update on the tcgen05 engine vs the fixtures recorded from the UNMODIFIED reference PPO (same minibatch index lists)."""
    from _ppo_golden import load_case

    c = load_case(golden("ppo.npz"), ci)
    net, alg, buf = _mk_case(c, "tc")
    np.random.seed(c["seed"])
    alg.update(buf, c["bs"], c["repeat"])
    _check_against_reference(c, net, alg, TC_REF_TOL)


@pytest.mark.parametrize("engine", ["fp32", "tc"])
@pytest.mark.parametrize("ci", [0, 1])
def test_update_large_matches_reference(golden, ci, engine):
    """C2-scale fixture (131 072 rows, minibatches of 4 096, 64 optimizer steps) recorded from the reference."""
    from _ppo_golden import load_large

    c = load_large(golden("ppo_large.npz"), ci)
    net, alg, buf = _mk_case(c, engine)
    np.random.seed(c["seed"])
    alg.update(buf, c["bs"], c["repeat"])
    em = lambda t: t.t().contiguous().view(-1).cpu().numpy()  # noqa: E731
    si = c["sample_idx"]
    tol = FP32_LARGE_TOL if engine == "fp32" else TC_REF_TOL
    for k, w in (("v_s", "v_s"), ("returns", "ret"), ("adv", "adv"), ("logp_old", "logp_old")):
        assert np.abs(em(alg._work[w])[si] - c[k]).max() < tol["pre_abs"], (k, np.abs(em(alg._work[w])[si] - c[k]).max())
    _check_against_reference(c, net, alg, tol)


@pytest.mark.parametrize("env", [{"TSM_TC_EPOCH_GROUPS": "2"}, {"TSM_TC_EPOCH_GROUPS": "2", "TSM_TC_EPOCH_HALVES": "1"}])
def test_two_group_epoch_kernels_still_agree(env):
    """The default persistent kernel keeps four tile groups per SM (mlp_tc_epoch4.cuh); the two-group kernels (two threads per row,
    and the round-1 shape with one) stay selectable.  The choice is read once per process, so they run in a child process:
    the chain-agreement and reference-fixture tests of this file under the given environment."""
    import os
    import subprocess
    import sys

    e = dict(os.environ)
    e.update(env)
    r = subprocess.run([sys.executable, "-m", "pytest", __file__, "-q", "-x", "-k", "launch_chain_agree or update_tc_matches_reference"],
                       env=e, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-1000:]
    assert " passed" in r.stdout
