"""Engine 2 (csrc/mlp_tcw.cu, the wide tcgen05 engine of the CTDE nets) through the C ABI against a plain torch fp32 restatement
of the same ops: input packing (bit-exact), forward, fused loss + backward (gradients), joint-clip Adam.  Shapes are the
reference's DecentralizedActor / CentralizedCritic (multiagent/ctde.py:346-414) at BASELINE.json configs[2]: 60 -> 128 -> 128 -> 5
and 600 -> 128 -> 128 -> 10.  Tolerances: fp16 operands / fp32 accumulation, stated per test."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _mods():
    from tianshou_marl_b200 import _lib
    from tianshou_marl_b200.utils import ActorCriticNet

    return _lib, ActorCriticNet


def _net(in_dim, n_act=0, heads=0, seed=0):
    _lib, Net = _mods()
    net = Net(in_dim, 128, n_act, with_value=False, n_value_only=heads, engine="tcw", seed=seed)
    g = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():
        for k in ("b1", "b2", "b3"):
            net.view(k).copy_(torch.randn(net.view(k).shape, generator=g) * 0.1)
    return net


def _torch_forward(net, x):
    v = net.named_views()
    h1 = torch.relu(x @ v["W1"].t() + v["b1"])
    h2 = torch.relu(h1 @ v["W2"].t() + v["b2"])
    return h2 @ v["W3"].t() + v["b3"]


@pytest.mark.parametrize("in_dim,n", [(60, 300), (600, 257), (18, 128), (54, 1)])
def test_pack_x_bit_exact(in_dim, n):
    _lib, _ = _mods()
    L = _lib.lib()
    KP = (in_dim + 1 + 15) // 16 * 16
    x = torch.randn(n + 7, in_dim, device="cuda")
    idx = torch.randperm(n + 7, device="cuda")[:n].contiguous()
    for use_idx in (False, True):
        nb = int(L.tsm_tcw_x_bytes(in_dim, n))
        tiles = (n + 127) // 128
        assert nb == tiles * (KP // 8) * 2048
        img = torch.full((nb,), 0xFF, dtype=torch.uint8, device="cuda")
        _lib.check(L.tsm_tcw_pack_x(_lib.ptr(x), in_dim, in_dim, _lib.ptr(idx) if use_idx else None, n, _lib.ptr(img), _lib.stream_ptr()))
        got = img.view(torch.float16).view(tiles, KP // 8, 128, 8).cpu().numpy()
        src = (x[idx] if use_idx else x[:n]).cpu().numpy()
        want = np.zeros((tiles * 128, KP), np.float16)
        want[:n, :in_dim] = src.astype(np.float16)
        want[:n, KP - 1] = 1.0
        want = want.reshape(tiles, 128, KP // 8, 8).transpose(0, 2, 1, 3)
        assert np.array_equal(got.view(np.uint16), want.view(np.uint16))


@pytest.mark.parametrize("in_dim,out", [(60, 5), (600, 10), (18, 5), (54, 3)])
def test_forward_raw(in_dim, out):
    """|out - torch fp32| < 6e-3 + 6e-3 |out| (11-bit operands, K up to 608, fp32 accumulation)."""
    n = 128 * 151 + 77  # more tiles than SMs: every CTA loops, the last tile is ragged
    net = _net(in_dim, heads=out)
    x = torch.randn(n, in_dim, device="cuda")
    got = net.forward(x, mode="raw")["out"]
    want = _torch_forward(net, x)
    assert got.shape == (n, out)
    err = (got - want).abs()
    assert bool((err < 6e-3 + 6e-3 * want.abs()).all()), float(err.max())


def test_forward_sample_eval_matches_own_logits():
    """Sampling / evaluation epilogue: logp and entropy consistent with the returned logits; actions follow the Philox stream of
    engines 0 / 1 (same key: seed, counter, row)."""
    _lib, Net = _mods()
    n = 1000
    net = _net(60, n_act=5)
    x = torch.randn(n, 60, device="cuda")
    o = net.forward(x, mode="sample", seed=11, counter=3, want_logits=True, want_entropy=True)
    ls = torch.log_softmax(o["logits"], -1)
    assert torch.allclose(o["logp"], ls.gather(1, o["act"].long()[:, None])[:, 0], atol=2e-3)
    assert torch.allclose(o["entropy"], -(ls.exp() * ls).sum(-1), atol=2e-3)
    e = net.forward(x, mode="eval", act=o["act"], want_logits=True)
    assert torch.equal(e["logp"], o["logp"])
    m = net.forward(x, mode="mode", want_logits=True)
    assert torch.equal(m["act"].long(), m["logits"].argmax(-1))
    # the categorical draw: u = u01(philox(row, counter, seed)), inverse CDF over softmax(logits) -- same stream as engine 0
    ref = Net(60, 128, 5, with_value=False, engine="fp32")
    with torch.no_grad():
        ref.flat.copy_(net.flat)
    r = ref.forward(x, mode="sample", seed=11, counter=3)
    assert float((r["act"] == o["act"]).float().mean()) > 0.99  # fp16 logits move a draw only when u sits on a CDF edge


def _train(net, ximg, n, loss, cfg, inv_n, gscale, stats=None, ids=None, sub=1, tile0=0, n_tiles=None):
    _lib, _ = _mods()
    tiles = _lib.TsmTcwTiles(ximg=_lib.ptr(ximg), n_rows=n, super_ids=_lib.ptr(ids), sub=sub, tile0=tile0,
                             n_tiles=(n + 127) // 128 if n_tiles is None else n_tiles)
    sums = torch.zeros(4, dtype=torch.float64, device="cuda")
    _lib.check(_lib.lib().tsm_tcw_train(net.desc, C.byref(tiles), C.byref(loss), C.byref(cfg), inv_n, gscale, _lib.ptr(stats), _lib.ptr(sums),
                                        _lib.stream_ptr()))
    return sums


def _emulated_grads(net, x, loss_fn):
    """The same gradient with the engine's operand rounding restated in torch: fp16 inputs, weights and stored activations
    (fp32 accumulation), ReLU masks taken from the ROUNDED activations.  A pre-activation within rounding of zero flips its mask
    against the fp32 net, which switches that unit's whole gradient contribution on or off; with this restatement the masks
    coincide and the comparison can be tight."""
    r = lambda t: t.half().float()  # noqa: E731
    v = {k: r(t) for k, t in net.named_views().items()}
    xh = r(x)
    h1 = r(torch.relu(xh @ v["W1"].t() + v["b1"]))
    h2 = r(torch.relu(h1 @ v["W2"].t() + v["b2"]))
    out = (h2 @ v["W3"].t() + v["b3"]).requires_grad_(True)
    loss_fn(out).backward()
    g = out.grad
    d2 = (g @ v["W3"]) * (h2 > 0)
    d1 = (d2 @ v["W2"]) * (h1 > 0)
    return {"W3": g.t() @ h2, "b3": g.sum(0), "W2": d2.t() @ h1, "b2": d2.sum(0), "W1": d1.t() @ xh, "b1": d1.sum(0)}


def _grad_close(net, want: dict, tol: float, rms: bool = False):
    """max |got - want| / max |want| per tensor, or (rms) the relative Frobenius error -- the stable measure where single rows
    switch on or off (comparisons with the fp32 net at small n)."""
    for k, g in want.items():
        off, shp = net._views[k]
        got = net.grads[off: off + g.numel()].view(shp)
        if rms:
            err = float((got - g).norm()) / (float(g.norm()) + 1e-12)
        else:
            err = float((got - g).abs().max()) / (float(g.abs().max()) + 1e-12)
        assert err < tol, (k, err, rms)


def _torch_grads(net, x, loss_fn):
    p = {k: v.clone().requires_grad_(True) for k, v in net.named_views().items()}
    h1 = torch.relu(x @ p["W1"].t() + p["b1"])
    h2 = torch.relu(h1 @ p["W2"].t() + p["b2"])
    out = h2 @ p["W3"].t() + p["b3"]
    loss_fn(out).backward()
    return {k: v.grad for k, v in p.items()}, out.detach()


@pytest.mark.parametrize("n,plain", [(128 * 3 + 50, 1), (128 * 300, 1), (128 * 3 + 50, 0), (128 * 300, 0)])
def test_actor_policy_gradients(n, plain):
    """Policy term + entropy bonus on 60 -> 128 -> 128 -> 5; plain: -logp * adv (a2c.py:247-288), else the PPO clip
    (ppo.py:181-211).  Against the fp32 net the gradients are within 12 % (n = 434) / 5 % (n = 38 400) of each tensor's scale
    (the bounds engine 1 is held to): 11-bit operands flip the ReLU mask of pre-activations within rounding of zero and move
    rows across the ratio-clip edge, each of which switches a whole contribution on or off.  Against the torch restatement WITH
    the engine's operand rounding (_emulated_grads) the plain gradient is within 1 %."""
    _lib, _ = _mods()
    net = _net(60, n_act=5, seed=3)
    x = torch.randn(n, 60, device="cuda")
    act = torch.randint(0, 5, (n,), device="cuda", dtype=torch.int32)
    adv = torch.randn(n, device="cuda")
    with torch.no_grad():
        lp0 = torch.log_softmax(_torch_forward(net, x), -1).gather(1, act.long()[:, None])[:, 0]
    logp_old = lp0 + 0.1 * torch.randn(n, device="cuda")
    mean, std = adv.mean(), adv.std()
    stats = torch.stack([mean, 1 / (std + 1e-8)]).float().contiguous()
    eps, ent_coef = 0.2, 0.01

    def loss_fn(out):
        ls = torch.log_softmax(out, -1)
        lp = ls.gather(1, act.long()[:, None])[:, 0]
        a = (adv - mean) / (std + 1e-8)
        ratio = (lp - logp_old).exp()
        clip = -(lp * a).mean() if plain else -torch.min(ratio * a, ratio.clamp(1 - eps, 1 + eps) * a).mean()
        ent = -(ls.exp() * ls).sum(-1).mean()
        return clip - ent_coef * ent

    want, _ = _torch_grads(net, x, loss_fn)
    ximg = net.pack_x(x)
    net.grads.zero_()
    loss = _lib.TsmTcwLoss(kind=0, act=_lib.ptr(act), adv=_lib.ptr(adv), logp_old=_lib.ptr(logp_old), target=None, v_old=None, weight=None, heads=0)
    cfg = _lib.TsmPpoCfg(eps_clip=eps, dual_clip=0.0, vf_coef=0.5, ent_coef=ent_coef, value_clip=0, adv_norm=1, plain_pg=plain, value_heads=0)
    gs = float(2 ** int(np.ceil(np.log2(n))))
    sums = _train(net, ximg, n, loss, cfg, 1.0 / n, gs, stats)
    assert float(sums[3]) == n
    _grad_close(net, want, 0.12 if n < 3000 else 0.05, rms=True)  # against the fp32 net: ReLU-mask / clip-edge flips of individual rows
    if plain:
        _grad_close(net, _emulated_grads(net, x, loss_fn), 0.01)  # same operand rounding: tight


@pytest.mark.parametrize("kind,n", [(1, 128 * 2 + 9), (1, 128 * 200), (2, 500)])
def test_critic_value_gradients(kind, n):
    """Centralized critic 600 -> 128 -> 128 -> 10 (streaming path): per-head value loss with a 0/1 mask and value clipping
    (kind 1, ppo.py:199-208), mean-of-heads TD loss (kind 2, ctde.py:153-166); tolerances as in test_actor_policy_gradients
    (kind 1 has the value-clip edge, so only kind 2 is compared with the rounding-aware restatement)."""
    _lib, _ = _mods()
    heads = 10
    net = _net(600, heads=heads, seed=5)
    x = torch.randn(n, 600, device="cuda")
    if kind == 1:
        target = torch.randn(n, heads, device="cuda")
        with torch.no_grad():
            v_old = _torch_forward(net, x) + 0.3 * torch.randn(n, heads, device="cuda")
        w = (torch.rand(n, heads, device="cuda") < 0.7).float()
        cnt = float(w.sum())
        eps = 0.2

        def loss_fn(out):
            vc = v_old + (out - v_old).clamp(-eps, eps)
            vf = torch.max((target - out) ** 2, (target - vc) ** 2)
            return 0.5 * (vf * w).sum() / cnt

        loss = _lib.TsmTcwLoss(kind=1, act=None, adv=None, logp_old=None, target=_lib.ptr(target), v_old=_lib.ptr(v_old), weight=_lib.ptr(w), heads=heads)
        cfg = _lib.TsmPpoCfg(eps_clip=eps, dual_clip=0.0, vf_coef=0.5, ent_coef=0.0, value_clip=1, adv_norm=0, plain_pg=0, value_heads=0)
        inv_n = 1.0 / cnt
    else:
        target = torch.randn(n, device="cuda")

        def loss_fn(out):
            return ((out.mean(-1) - target) ** 2).mean()

        loss = _lib.TsmTcwLoss(kind=2, act=None, adv=None, logp_old=None, target=_lib.ptr(target), v_old=None, weight=None, heads=heads)
        cfg = _lib.TsmPpoCfg(eps_clip=0.2, dual_clip=0.0, vf_coef=1.0, ent_coef=0.0, value_clip=0, adv_norm=0, plain_pg=0, value_heads=0)
        inv_n = 1.0 / n
    want, _ = _torch_grads(net, x, loss_fn)
    ximg = net.pack_x(x)
    net.grads.zero_()
    gs = float(2 ** int(np.ceil(np.log2(max(n, 2)))))
    sums = _train(net, ximg, n, loss, cfg, inv_n, gs)
    _grad_close(net, want, 0.12 if n < 3000 else 0.05, rms=True)
    if kind == 2:
        _grad_close(net, _emulated_grads(net, x, loss_fn), 0.01)
    if kind == 2:
        with torch.no_grad():
            assert abs(float(sums[1]) / n - float(loss_fn(_torch_forward(net, x)))) < 2e-3 * (1 + float(sums[1]) / n)


def test_tile_lists_select_minibatch():
    """A launch over a super-tile list touches exactly those tiles: gradient == torch gradient over the listed rows."""
    _lib, _ = _mods()
    A, worlds = 3, 128 * 6
    net = _net(18, n_act=5, seed=9)
    n = worlds * A  # rows; super-tile s = rows [s * 128 * A, (s + 1) * 128 * A) = A tiles
    x = torch.randn(n, 18, device="cuda")
    act = torch.randint(0, 5, (n,), device="cuda", dtype=torch.int32)
    adv = torch.randn(n, device="cuda")
    ids = torch.tensor([4, 1, 5, 0, 3, 2], dtype=torch.int32, device="cuda")
    # minibatch = logical super-tiles 1..2 of the list = physical super-tiles 1 and 5
    rows = torch.cat([torch.arange(s * 128 * A, (s + 1) * 128 * A, device="cuda") for s in (1, 5)])
    m = rows.numel()

    def loss_fn(out):
        lp = torch.log_softmax(out, -1).gather(1, act[rows].long()[:, None])[:, 0]
        return -(lp * adv[rows]).mean()

    want, _ = _torch_grads(net, x[rows], loss_fn)
    ximg = net.pack_x(x)
    net.grads.zero_()
    loss = _lib.TsmTcwLoss(kind=0, act=_lib.ptr(act), adv=_lib.ptr(adv), logp_old=None, target=None, v_old=None, weight=None, heads=0)
    cfg = _lib.TsmPpoCfg(eps_clip=0.2, dual_clip=0.0, vf_coef=0.0, ent_coef=0.0, value_clip=0, adv_norm=0, plain_pg=1, value_heads=0)
    sums = _train(net, ximg, n, loss, cfg, 1.0 / m, 1024.0, ids=ids, sub=A, tile0=1 * A, n_tiles=2 * A)
    assert float(sums[3]) == m
    _grad_close(net, want, 0.12, rms=True)
    _grad_close(net, _emulated_grads(net, x[rows], loss_fn), 0.01)


@pytest.mark.parametrize("max_norm", [0.0, 0.5])
def test_adam_joint_clip_matches_torch(max_norm):
    """Three optimizer steps over actor + critic with ONE gradient norm (clip_grad_norm_ over both nets' parameters) against
    torch.optim.Adam: parameters within 1e-6, fp16 image == pack of the new parameters."""
    _lib, _ = _mods()
    L = _lib.lib()
    actor, critic = _net(60, n_act=5, seed=1), _net(600, heads=10, seed=2)
    nets = (_lib.TsmNet * 2)()
    tp = [torch.nn.Parameter(n.flat.detach().clone()) for n in (actor, critic)]
    opt = torch.optim.Adam(tp, lr=1e-3)
    ws = torch.zeros(2048, dtype=torch.uint8, device="cuda")  # TSM_TCW_ADAM_WS_BYTES
    gn = torch.zeros(1, device="cuda")
    seq = 0
    for step in range(1, 4):
        for n, p in zip((actor, critic), tp):
            g = torch.randn(n.n_params, device="cuda") * (0.01 * step)
            n.grads.copy_(g)
            p.grad = g.clone() * 0.5  # grad_scale 0.5 below
        for i, n in enumerate((actor, critic)):
            n.desc  # refresh
            C.memmove(C.byref(nets, i * C.sizeof(_lib.TsmNet)), C.byref(n._desc), C.sizeof(_lib.TsmNet))
        _lib.check(L.tsm_tcw_adam(nets, 2, step, 1e-3, 0.9, 0.999, 1e-8, max_norm, 0.5, _lib.ptr(ws), seq, _lib.ptr(gn), _lib.stream_ptr()))
        if max_norm > 0:
            seq += 1
            want_norm = torch.nn.utils.clip_grad_norm_(tp, max_norm)
            assert abs(float(gn) - float(want_norm)) < 1e-5 * float(want_norm)
        opt.step()
        for n, p in zip((actor, critic), tp):
            assert float((n.flat.detach() - p.detach()).abs().max()) < 1e-6
            assert float(n.grads.abs().max()) == 0.0
            assert int(n.adam_step) == step
    for n in (actor, critic):
        img = n.tc_image.clone()
        n.tc_image.zero_()
        _lib.check(L.tsm_net_pack(n.desc, _lib.stream_ptr()))
        n._packed_version = n.flat._version
        assert torch.equal(img, n.tc_image)
