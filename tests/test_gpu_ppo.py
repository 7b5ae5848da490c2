"""K4/K5/K6 parity (GPU): csrc/mlp.cu through the C ABI and the PPO / ActorCriticNet classes vs
 (a) fixtures recorded from the UNMODIFIED reference PPO (tests/golden/ppo.npz), with numpy
     minibatch orders replayed bit-exactly, and
 (b) the torch-fp32 oracle (oracle/ppo.py) on larger random problems.
Tolerances (fp32, different summation order): forward values rtol 2e-5; GAE returns / adv rtol
2e-5 atol 2e-5; per-minibatch losses rtol 2e-4 atol 2e-5; parameters after the whole update
atol 5e-5 (Adam divides by sqrt(v)+1e-8, so early steps amplify 1e-7 gradient differences)."""
import numpy as np
import pytest
import torch

from _ppo_golden import load_case, load_reinforce

pytestmark = pytest.mark.gpu


def _mk(c, order="numpy", **ppo_kw):
    from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory, DiscreteActorPolicy
    from tianshou_marl_b200.data import Batch, VectorReplayBuffer
    from tianshou_marl_b200.utils import ActorCriticNet

    net = ActorCriticNet(c["O"], c["H"], c["nA"], softmax_output=c["probs"]).load_arrays(c["W1"], c["b1"], c["W2"], c["b2"], c["W3"], c["b3"])
    pol = DiscreteActorPolicy(actor=net)
    k = c["cfg"]
    alg = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=c["lr"]), eps_clip=k["eps_clip"], dual_clip=k["dual_clip"] or None,
              value_clip=k["value_clip"], advantage_normalization=k["adv_norm"], vf_coef=k["vf_coef"], ent_coef=k["ent_coef"],
              max_grad_norm=c["max_norm"] or None, gamma=c["gamma"], gae_lambda=c["lam"], minibatch_order=order, **ppo_kw)
    buf = VectorReplayBuffer(c["R"] * c["T"], c["R"], index_type="torch")
    for t in range(c["T"]):
        buf.add(Batch(obs=c["obs"][t], act=c["act"][t], rew=c["rew"][t], terminated=c["term"][t], truncated=c["trunc"][t],
                      obs_next=c["obs_next"][t]))
    return net, pol, alg, buf


@pytest.mark.parametrize("ci", [0, 1, 2, 3])
def test_forward_matches_reference(golden, ci):
    c = load_case(golden("ppo.npz"), ci)
    net, *_ = _mk(c)
    o = net.forward(torch.as_tensor(c["obs"][0], device="cuda"), mode="mode", want_logits=True)
    if c["probs"]:  # case 2: DiscreteActor(softmax_output=True) returns probabilities (utils/net/discrete.py:85-86)
        assert np.allclose(torch.softmax(o["logits"], -1).cpu().numpy(), c["fwd_actor_out"], rtol=2e-5, atol=2e-6)
        assert np.allclose(o["value"].cpu().numpy(), c["fwd_value"], rtol=2e-5, atol=2e-6)
        a = c["fwd_actor_out"].argmax(-1)
        assert np.array_equal(o["act"].cpu().numpy(), a)
        assert np.allclose(o["logp"].cpu().numpy(), np.log(c["fwd_actor_out"][np.arange(len(a)), a]), rtol=2e-5, atol=2e-6)
        return
    assert np.allclose(o["logits"].cpu().numpy(), c["fwd_actor_out"], rtol=2e-5, atol=2e-6)
    assert np.allclose(o["value"].cpu().numpy(), c["fwd_value"], rtol=2e-5, atol=2e-6)
    assert np.array_equal(o["act"].cpu().numpy(), c["fwd_actor_out"].argmax(-1))
    lp = torch.log_softmax(torch.as_tensor(c["fwd_actor_out"]), -1).numpy()
    assert np.allclose(o["logp"].cpu().numpy(), lp[np.arange(len(lp)), c["fwd_actor_out"].argmax(-1)], rtol=2e-5, atol=2e-6)


@pytest.mark.parametrize("ci", [0, 1, 2, 3])
def test_update_matches_reference(golden, ci):
    """Whole PPO.update vs the reference: v_s, returns, adv, logp_old, every minibatch loss, final parameters."""
    c = load_case(golden("ppo.npz"), ci)
    net, pol, alg, buf = _mk(c, "numpy")
    with pytest.raises(RuntimeError):
        alg.update(buf, c["bs"], c["repeat"])  # algorithm_base.py:610-615
    pol.is_within_training_step = True
    np.random.seed(c["seed"])
    stats = alg.update(buf, c["bs"], c["repeat"])
    em = lambda t: t.t().contiguous().view(-1).cpu().numpy()  # noqa: E731  time-major -> env-major
    assert np.allclose(em(alg._work["v_s"]), c["v_s"], rtol=2e-5, atol=2e-6)
    assert np.allclose(em(alg._work["ret"]), c["returns"], rtol=2e-5, atol=2e-5)
    assert np.allclose(em(alg._work["adv"]), c["adv"], rtol=2e-5, atol=2e-5)
    assert np.allclose(em(alg._work["logp_old"]), c["logp_old"], rtol=2e-5, atol=2e-6)
    seq = alg.last_sequences
    assert stats.gradient_steps == len(c["loss_seq"])
    for k, ref in (("loss", "loss_seq"), ("clip", "clip_seq"), ("vf", "vf_seq"), ("ent", "ent_seq")):
        assert np.allclose(seq[k], c[ref], rtol=2e-4, atol=2e-5), (k, np.abs(seq[k] - c[ref]).max())
    v = net.named_views()
    for k in ("W1", "b1", "W2", "b2", "W3", "b3"):
        err = np.abs(v[k].cpu().numpy() - c["after_" + k]).max()
        assert err < 5e-5, (k, err)
    assert int(net.adam_step.item()) == stats.gradient_steps
    assert float(net.grads.abs().max().item()) == 0.0  # zeroed by the Adam kernel


def test_single_minibatch_gradient_vs_oracle():
    """tsm_ppo_minibatch gradients vs torch autograd on a 3000-row minibatch (ragged tile), H=64 and H=128."""
    import ctypes as C

    from oracle import ppo as oppo
    from tianshou_marl_b200 import _lib
    from tianshou_marl_b200.utils import ActorCriticNet

    for (O, H, n) in ((18, 64, 3000), (60, 128, 1111), (7, 64, 200)):
        rng = np.random.default_rng(O)
        onet = oppo.make_net(O, H, 5, seed=O)
        for p in onet.p[1::2]:
            p.data.normal_(0, 0.1)
        net = ActorCriticNet(O, H, 5).load_arrays(*[p.detach().numpy() for p in onet.p])
        obs = rng.normal(size=(n, O)).astype(np.float32)
        act = rng.integers(0, 5, n)
        adv = rng.normal(size=n).astype(np.float32)
        ret = rng.normal(size=n).astype(np.float32)
        logp_old = (np.log(0.2) + 0.3 * rng.normal(size=n)).astype(np.float32)
        v_old = rng.normal(size=n).astype(np.float32)
        for cfg in (dict(adv_norm=True), dict(adv_norm=False, value_clip=True, dual_clip=2.0, eps_clip=0.1, vf_coef=0.3, ent_coef=0.02),
                    dict(adv_norm=True, probs=True, ent_coef=0.05), dict(adv_norm=True, probs=True, peaked=True, ent_coef=0.05)):
            T = lambda a: torch.from_numpy(a)  # noqa: E731
            if cfg.get("probs"):  # Categorical(probs=softmax(logits)), discrete.py:85-86; peaked: probabilities below 2^-23, where
                #                   torch's probs_to_logits clamps the log and stops the gradient
                if cfg.get("peaked"):
                    onet.p[4].data[:5] *= 60.0
                    with torch.no_grad():
                        pr = torch.softmax(onet.forward(T(obs))[0], -1)
                    assert float((pr < 1.19e-7).float().mean()) > 0.02
                net = ActorCriticNet(O, H, 5, softmax_output=True).load_arrays(*[p.detach().numpy() for p in onet.p])
            loss, cl, vf, en = oppo.minibatch_loss(onet, T(obs), T(act).float(), T(adv), T(ret), T(logp_old), T(v_old), cfg, bool(cfg.get("probs")))
            for p in onet.p:
                p.grad = None
            loss.backward()
            want = np.concatenate([p.grad.numpy().reshape(-1) for p in onet.p])
            d = lambda a, dt=torch.float32: torch.as_tensor(a, device="cuda").to(dt).contiguous()  # noqa: E731
            t = dict(obs=d(obs), act=d(act, torch.int32), adv=d(adv), ret=d(ret), lp=d(logp_old), vo=d(v_old))
            P = _lib.ptr
            pb = _lib.TsmPpoBatch(obs=P(t["obs"]), act=P(t["act"]), adv=P(t["adv"]), ret=P(t["ret"]), logp_old=P(t["lp"]), v_old=P(t["vo"]),
                                  indices=None, size=1, n_rows=n, start=0, len=1, perm_key=0, perm_n=0)
            pc = _lib.TsmPpoCfg(eps_clip=cfg.get("eps_clip", 0.2), dual_clip=cfg.get("dual_clip", 0.0), vf_coef=cfg.get("vf_coef", 0.5),
                                ent_coef=cfg.get("ent_coef", 0.01), value_clip=int(cfg.get("value_clip", False)), adv_norm=int(cfg["adv_norm"]))
            bounds = d(np.array([0, n]), torch.int64)
            stats = torch.zeros(2, device="cuda")
            sums = torch.zeros(4, dtype=torch.float64, device="cuda")
            L, S = _lib.lib(), _lib.stream_ptr()
            _lib.check(L.tsm_ppo_adv_stats(C.byref(pb), P(bounds), 1, 1e-8, P(stats), S))
            assert np.allclose(stats.cpu().numpy(), [adv.mean(), 1 / (adv.std(ddof=1) + 1e-8)], rtol=1e-5)
            net.grads.zero_()
            _lib.check(L.tsm_ppo_minibatch(net.desc, C.byref(pb), C.byref(pc), 0, n, P(stats), P(sums), S))
            got = net.grads.cpu().numpy()
            scale = np.abs(want).max()
            assert np.abs(got - want).max() < 2e-5 * max(scale, 1.0) + 1e-6, (O, H, cfg, np.abs(got - want).max(), scale)
            s = sums.cpu().numpy()
            assert np.allclose([s[0] / n, s[1] / n, s[2] / n], [float(cl), float(vf), float(en)], rtol=1e-4, atol=1e-5)


def test_feistel_permutation_is_a_bijection_and_matches_oracle():
    import ctypes as C

    from oracle.philox import feistel_permute
    from tianshou_marl_b200 import _lib
    from tianshou_marl_b200.utils import ActorCriticNet

    # drive the device permutation through tsm_ppo_adv_stats: with adv = position id, the mean over
    # the whole domain is invariant, and per-minibatch sums identify the permuted members
    for n in (1, 2, 5, 1000, 4097, 65536 + 17):
        R, T = n, 1
        adv = torch.arange(n, dtype=torch.float32, device="cuda")
        key = 0x1234ABCD5678EF01 + n
        pb = _lib.TsmPpoBatch(obs=None, act=None, adv=_lib.ptr(adv), ret=None, logp_old=None, v_old=None, indices=None,
                              size=1, n_rows=n, start=0, len=1, perm_key=key, perm_n=n)
        nb = min(n, 7)
        bounds = np.linspace(0, n, nb + 1).astype(np.int64)
        stats = torch.zeros(nb, 2, device="cuda")
        _lib.check(_lib.lib().tsm_ppo_adv_stats(C.byref(pb), _lib.ptr(torch.as_tensor(bounds, device="cuda")), nb, 0.0, _lib.ptr(stats),
                                                _lib.stream_ptr()))
        perm = feistel_permute(np.arange(n, dtype=np.uint64), n, key)
        assert np.array_equal(np.sort(perm), np.arange(n))
        want = [perm[bounds[i]:bounds[i + 1]].astype(np.float64).mean() for i in range(nb)]
        assert np.allclose(stats[:, 0].cpu().numpy(), want, rtol=1e-6)


def test_adam_matches_torch():
    from tianshou_marl_b200 import _lib
    from tianshou_marl_b200.utils import ActorCriticNet

    net = ActorCriticNet(18, 64, 5, seed=3)
    p = torch.nn.Parameter(net.flat.detach().clone())
    opt = torch.optim.Adam([p], lr=1e-3)
    g = torch.Generator(device="cuda").manual_seed(0)
    for it in range(5):
        grad = torch.randn(net.n_params, device="cuda", generator=g) * (10.0 if it == 2 else 0.01)
        net.grads.copy_(grad)
        gn = torch.zeros(1, device="cuda")
        _lib.check(_lib.lib().tsm_adam_step(net.desc, 1e-3, 0.9, 0.999, 1e-8, 0.5, 1.0, _lib.ptr(gn), _lib.stream_ptr()))
        p.grad = grad.clone()
        torch.nn.utils.clip_grad_norm_([p], 0.5)
        opt.step()
        assert torch.allclose(gn, grad.norm(), rtol=1e-5)
        assert torch.allclose(net.flat.detach(), p.detach(), rtol=1e-5, atol=1e-7), (it, (net.flat.detach() - p.detach()).abs().max())
    assert int(net.adam_step.item()) == 5


def test_device_order_update_runs_and_learns_value():
    """Throughput mode (Feistel minibatch order): not index-comparable with numpy, so check
    behaviour: every sample is visited once per epoch (gradient steps), losses finite, vf_loss falls."""
    from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory, DiscreteActorPolicy
    from tianshou_marl_b200.data import Batch, VectorReplayBuffer
    from tianshou_marl_b200.utils import ActorCriticNet

    R, T, O = 512, 32, 18
    net = ActorCriticNet(O, 64, 5, seed=1)
    pol = DiscreteActorPolicy(actor=net, seed=5)
    alg = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=3e-3), minibatch_order="device")
    buf = VectorReplayBuffer(R * T, R, index_type="torch")
    g = torch.Generator(device="cuda").manual_seed(0)
    obs = torch.randn(R, O, device="cuda", generator=g)
    for t in range(T):
        out = pol(Batch(obs=obs))
        nxt = obs + 0.1 * torch.randn(R, O, device="cuda", generator=g)
        buf.add(Batch(obs=obs, act=out.act, rew=obs[:, 0].contiguous(), terminated=torch.zeros(R, dtype=torch.bool, device="cuda"),
                      truncated=torch.full((R,), (t + 1) % 8 == 0, device="cuda"), obs_next=nxt))
        obs = nxt
    pol.is_within_training_step = True
    s = alg.update(buf, 2048, 6)
    assert s.gradient_steps == 6 * (R * T // 2048)
    seq = alg.last_sequences
    assert np.isfinite(seq["loss"]).all()
    assert seq["vf"][-1] < 0.7 * seq["vf"][0]


@pytest.mark.gpu
@pytest.mark.parametrize("engine", ["fp32", "tc"])
def test_reuse_rollout_policy_outputs(engine):
    """PPO(reuse_rollout_policy_outputs=True) takes critic(obs) / logp_old from the collect-time forward and evaluates the
    critic on obs_next only where an episode ended (a2c.py:120-127, ppo.py:157-161 recompute both).  Same parameters, same
    kernels: everything the update is fed with (values, logp_old, advantages, returns) must agree BIT FOR BIT with the
    recomputing path; the parameters after the update agree up to the order of the gradient atomics.  The stamp on the
    buffer goes stale once the parameters move."""
    from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory, DiscreteActorPolicy
    from tianshou_marl_b200.data import Collector, VectorReplayBuffer
    from tianshou_marl_b200.env import MPEVectorEnv
    from tianshou_marl_b200.utils import ActorCriticNet

    E, A, T = 96, 3, 60  # 60 steps: two truncations (max_cycles 25) inside the rollout
    res = []
    for reuse in (False, True):
        env = MPEVectorEnv("simple_spread_v3", n_envs=E, n_agents=A, seed=5)
        buf = VectorReplayBuffer(T * E * A, E * A, index_type="torch")
        net = ActorCriticNet(env.obs_dim, 64, 5, seed=3, engine=engine)
        pol = DiscreteActorPolicy(actor=net, seed=11)
        alg = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=1e-3), max_grad_norm=0.5, minibatch_order="device", seed=2,
                  reuse_rollout_policy_outputs=reuse)
        col = Collector(alg, env, buf)
        col.reset()
        pol.is_within_training_step = True
        col.collect(n_step=T * E * A)
        assert buf._policy_version == net.param_version
        alg.update(buf, 2048, 2)
        assert buf._policy_version != net.param_version  # parameters moved on: the stored outputs are stale now
        w = alg._work
        v_s = buf._store["value"] if reuse else w["v_s"]
        logp = buf._store["logp"] if reuse else w["logp_old"]
        res.append({"v_s": v_s.clone(), "v_n": w["v_n"].clone(), "logp": logp.clone(), "adv": w["adv"].clone(), "ret": w["ret"].clone(),
                    "p": net.flat.detach().clone(), "timed_passes": None})
        # a second update from the same (now stale) buffer must take the recomputing path and still work
        st2 = alg.update(buf, 2048, 1)
        assert np.isfinite(st2.loss.mean)
    for k in ("v_s", "v_n", "logp", "adv", "ret"):
        assert torch.equal(res[0][k], res[1][k]), k
    assert torch.allclose(res[0]["p"], res[1]["p"], atol=2e-5 if engine == "fp32" else 2e-3)


@pytest.mark.gpu
@pytest.mark.parametrize("engine", ["fp32", "tc"])
def test_dedup_next_values_is_bit_identical(engine):
    """PPO(dedup_next_values=True, the default) evaluates critic(obs_next) only where an episode ended and takes v_s[t + 1]
    elsewhere; the full second critic pass of a2c.py:120-127 must give the same bits (values, advantages, returns)."""
    from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory, DiscreteActorPolicy
    from tianshou_marl_b200.data import Collector, VectorReplayBuffer
    from tianshou_marl_b200.env import MPEVectorEnv
    from tianshou_marl_b200.utils import ActorCriticNet

    E, A, T = 96, 3, 60
    res = []
    for dedup in (False, True):
        env = MPEVectorEnv("simple_spread_v3", n_envs=E, n_agents=A, seed=5)
        buf = VectorReplayBuffer(T * E * A, E * A, index_type="torch")
        net = ActorCriticNet(env.obs_dim, 64, 5, seed=3, engine=engine)
        pol = DiscreteActorPolicy(actor=net, seed=11)
        alg = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=1e-3), minibatch_order="device", seed=2, dedup_next_values=dedup)
        col = Collector(alg, env, buf)
        col.reset()
        pol.is_within_training_step = True
        col.collect(n_step=T * E * A)
        assert buf._obs_chain is True
        alg.update(buf, 2048, 1)
        res.append({k: alg._work[k].clone() for k in ("v_s", "v_n", "adv", "ret")})
    for k in res[0]:
        assert torch.equal(res[0][k], res[1][k]), k


def test_reset_env_between_collects_breaks_the_obs_chain():
    """Collector.reset_env() while the lock-step buffer holds data: obs[t + 1] is a fresh reset observation although done[t] == 0,
    so critic(obs_next[t]) must be evaluated on obs_next itself there (a2c.py:124-127).  The buffer is flagged and the update
    with dedup_next_values=True equals the full second critic pass bit for bit."""
    from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory, DiscreteActorPolicy
    from tianshou_marl_b200.data import Collector, VectorReplayBuffer
    from tianshou_marl_b200.env import MPEVectorEnv
    from tianshou_marl_b200.utils import ActorCriticNet

    E, A, T = 64, 3, 20
    res = []
    for dedup in (False, True):
        env = MPEVectorEnv("simple_spread_v3", n_envs=E, n_agents=A, seed=5)
        buf = VectorReplayBuffer(T * E * A, E * A, index_type="torch")
        net = ActorCriticNet(env.obs_dim, 64, 5, seed=3)
        pol = DiscreteActorPolicy(actor=net, seed=11)
        alg = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=1e-3), minibatch_order="device", seed=2, dedup_next_values=dedup,
                  reuse_rollout_policy_outputs=dedup)
        col = Collector(alg, env, buf)
        col.reset()
        pol.is_within_training_step = True
        col.collect(n_step=7 * E * A)
        assert buf._obs_chain is True
        col.reset_env()
        assert buf._obs_chain is False and buf._policy_version is None
        col.collect(n_step=(T - 7) * E * A)
        assert buf._obs_chain is False
        st = buf._store
        assert not bool(st["done"][6].any()) and not torch.equal(st["obs"][7], st["obs_next"][6])  # the seam
        alg.update(buf, 1024, 1)
        res.append({k: alg._work[k].clone() for k in ("v_s", "v_n", "adv", "ret")})
    for k in res[0]:
        assert torch.equal(res[0][k], res[1][k]), k


def test_return_scaling_with_row_subset_is_refused():
    """ret_rms would be reduced over rows outside the group: refuse instead of contaminating the statistics."""
    from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory, DiscreteActorPolicy
    from tianshou_marl_b200.data import Batch, VectorReplayBuffer
    from tianshou_marl_b200.utils import ActorCriticNet

    R, T, O = 8, 4, 6
    net = ActorCriticNet(O, 64, 5, seed=1)
    pol = DiscreteActorPolicy(actor=net)
    alg = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(), return_scaling=True)
    buf = VectorReplayBuffer(R * T, R, index_type="torch")
    for t in range(T):
        buf.add(Batch(obs=np.zeros((R, O), np.float32), act=np.zeros(R, np.int64), rew=np.ones(R), terminated=np.zeros(R, bool),
                      truncated=np.zeros(R, bool), obs_next=np.zeros((R, O), np.float32)))
    pol.is_within_training_step = True
    with pytest.raises(NotImplementedError):
        alg.update(buf, 8, 1, rows=torch.tensor([0, 2, 4], device="cuda"))


@pytest.mark.parametrize("ci", [0, 1])
def test_recompute_advantage_matches_reference(golden, ci):
    """PPO(recompute_advantage=True) (ppo.py:174-178) vs fixtures recorded from the reference: every minibatch loss of every
    epoch (the later epochs use re-evaluated values / advantages) and the final parameters."""
    c = load_case(golden("ppo_recompute.npz"), ci)
    net, pol, alg, buf = _mk(c, "numpy", recompute_advantage=True)
    pol.is_within_training_step = True
    np.random.seed(c["seed"])
    stats = alg.update(buf, c["bs"], c["repeat"])
    seq = alg.last_sequences
    assert stats.gradient_steps == len(c["loss_seq"])
    for k, ref in (("loss", "loss_seq"), ("clip", "clip_seq"), ("vf", "vf_seq"), ("ent", "ent_seq")):
        assert np.allclose(seq[k], c[ref], rtol=2e-4, atol=2e-5), (k, np.abs(seq[k] - c[ref]).max())
    v = net.named_views()
    for k in ("W1", "b1", "W2", "b2", "W3", "b3"):
        err = np.abs(v[k].cpu().numpy() - c["after_" + k]).max()
        assert err < 5e-5, (k, err)


@pytest.mark.parametrize("ci", [0, 1])
def test_return_scaling_matches_reference(golden, ci):
    """PPO(return_scaling=True) (a2c.py:132-146) vs fixtures recorded from the reference with a warmed-up ret_rms: normalised
    returns, advantages, the running statistics after the update, every minibatch loss and the final parameters."""
    g = golden("ppo_retscale.npz")
    c = load_case(g, ci)
    pre = f"p{ci}_"
    net, pol, alg, buf = _mk(c, "numpy", return_scaling=True, recompute_advantage=(ci == 1))
    alg.ret_rms.mean, alg.ret_rms.var, alg.ret_rms.count = (float(x) for x in g[pre + "rms_before"])
    alg.ret_rms.count = int(alg.ret_rms.count)
    pol.is_within_training_step = True
    np.random.seed(c["seed"])
    stats = alg.update(buf, c["bs"], c["repeat"])
    em = lambda t: t.t().contiguous().view(-1).cpu().numpy()  # noqa: E731
    if ci == 0:  # (with recompute_advantage the work arrays hold the LAST epoch's values)
        assert np.allclose(em(alg._work["ret"]), c["returns"], rtol=2e-5, atol=2e-5)
        assert np.allclose(em(alg._work["adv"]), c["adv"], rtol=2e-5, atol=2e-5)
    assert np.allclose([alg.ret_rms.mean, alg.ret_rms.var, alg.ret_rms.count], g[pre + "rms_after"], rtol=1e-5)
    seq = alg.last_sequences
    assert stats.gradient_steps == len(c["loss_seq"])
    for k, ref in (("loss", "loss_seq"), ("clip", "clip_seq"), ("vf", "vf_seq"), ("ent", "ent_seq")):
        assert np.allclose(seq[k], c[ref], rtol=2e-4, atol=2e-5), (k, np.abs(seq[k] - c[ref]).max())
    v = net.named_views()
    for k in ("W1", "b1", "W2", "b2", "W3", "b3"):
        err = np.abs(v[k].cpu().numpy() - c["after_" + k]).max()
        assert err < 5e-5, (k, err)


@pytest.mark.parametrize("ci", [0, 1])
def test_a2c_matches_reference(golden, ci):
    """A2C.update (a2c.py:247-288; case 1 with return_scaling) vs fixtures recorded from the reference A2C."""
    from tianshou_marl_b200.algorithm import A2C, AdamOptimizerFactory, DiscreteActorPolicy
    from tianshou_marl_b200.data import Batch, VectorReplayBuffer
    from tianshou_marl_b200.utils import ActorCriticNet

    g = golden("a2c.npz")
    c = load_case(g, ci)
    pre = f"p{ci}_"
    net = ActorCriticNet(c["O"], c["H"], c["nA"]).load_arrays(c["W1"], c["b1"], c["W2"], c["b2"], c["W3"], c["b3"])
    pol = DiscreteActorPolicy(actor=net)
    alg = A2C(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=c["lr"]), vf_coef=c["cfg"]["vf_coef"], ent_coef=c["cfg"]["ent_coef"],
              max_grad_norm=c["max_norm"] or None, gamma=c["gamma"], gae_lambda=c["lam"], return_scaling=(ci == 1), minibatch_order="numpy")
    if ci == 1:
        alg.ret_rms.mean, alg.ret_rms.var = (float(x) for x in g[pre + "rms_before"][:2])
        alg.ret_rms.count = int(g[pre + "rms_before"][2])
    buf = VectorReplayBuffer(c["R"] * c["T"], c["R"], index_type="torch")
    for t in range(c["T"]):
        buf.add(Batch(obs=c["obs"][t], act=c["act"][t], rew=c["rew"][t], terminated=c["term"][t], truncated=c["trunc"][t],
                      obs_next=c["obs_next"][t]))
    pol.is_within_training_step = True
    np.random.seed(c["seed"])
    stats = alg.update(buf, c["bs"], c["repeat"])
    em = lambda t: t.t().contiguous().view(-1).cpu().numpy()  # noqa: E731
    assert np.allclose(em(alg._work["ret"]), c["returns"], rtol=2e-5, atol=2e-5)
    assert np.allclose(em(alg._work["adv"]), c["adv"], rtol=2e-5, atol=2e-5)
    seq = alg.last_sequences
    assert stats.gradient_steps == len(c["loss_seq"])
    for k, ref in (("loss", "loss_seq"), ("clip", "clip_seq"), ("vf", "vf_seq"), ("ent", "ent_seq")):
        assert np.allclose(seq[k], c[ref], rtol=2e-4, atol=2e-5), (k, np.abs(seq[k] - c[ref]).max())
    v = net.named_views()
    for k in ("W1", "b1", "W2", "b2", "W3", "b3"):
        err = np.abs(v[k].cpu().numpy() - c["after_" + k]).max()
        assert err < 5e-5, (k, err)


@pytest.mark.parametrize("ci", [0, 1])
def test_reinforce_matches_reference(golden, ci):
    """Reinforce.update (reinforce.py:249-382; case 1 with return_standardization and a warmed ret_rms) vs fixtures recorded from
    the reference: Monte-Carlo returns, running statistics, every minibatch loss, final parameters."""
    from tianshou_marl_b200.algorithm import AdamOptimizerFactory, DiscreteActorPolicy, Reinforce
    from tianshou_marl_b200.data import Batch, VectorReplayBuffer
    from tianshou_marl_b200.utils import ActorCriticNet

    c = load_reinforce(golden("reinforce.npz"), ci)
    net = ActorCriticNet(c["O"], c["H"], c["nA"], with_value=False).load_arrays(c["W1"], c["b1"], c["W2"], c["b2"], c["W3"], c["b3"])
    pol = DiscreteActorPolicy(actor=net)
    alg = Reinforce(policy=pol, gamma=c["gamma"], return_standardization=c["std"], optim=AdamOptimizerFactory(lr=c["lr"]), minibatch_order="numpy")
    alg.ret_rms.mean, alg.ret_rms.var = (float(x) for x in c["rms_before"][:2])
    alg.ret_rms.count = int(c["rms_before"][2])
    buf = VectorReplayBuffer(c["R"] * c["T"], c["R"], index_type="torch")
    for t in range(c["T"]):
        buf.add(Batch(obs=c["obs"][t], act=c["act"][t], rew=c["rew"][t], terminated=c["term"][t], truncated=c["trunc"][t],
                      obs_next=c["obs_next"][t]))
    pol.is_within_training_step = True
    np.random.seed(c["seed"])
    stats = alg.update(buf, c["bs"], c["repeat"])
    em = lambda t: t.t().contiguous().view(-1).cpu().numpy()  # noqa: E731
    assert np.allclose(em(alg._work["ret"]), c["returns"], rtol=2e-5, atol=2e-5)
    if c["std"]:
        assert np.allclose([alg.ret_rms.mean, alg.ret_rms.var, alg.ret_rms.count], c["rms_after"], rtol=1e-5)
    assert stats.gradient_steps == len(c["loss_seq"])
    assert np.allclose(alg.last_sequences["loss"], c["loss_seq"], rtol=2e-4, atol=2e-5), np.abs(alg.last_sequences["loss"] - c["loss_seq"]).max()
    v = net.named_views()
    for k in ("W1", "b1", "W2", "b2", "W3", "b3"):
        err = np.abs(v[k].cpu().numpy() - c["after_" + k]).max()
        assert err < 5e-5, (k, err)


_STATS_SNIPPET = r"""
import ctypes as C, sys, numpy as np, torch
from tianshou_marl_b200 import _lib
L, P, S = _lib.lib(), _lib.ptr, _lib.stream_ptr()
out = {}
for size, R, start, ln, mb in ((16, 5000, 0, 16, 8192), (8, 1237, 5, 6, 1000), (1, 70001, 0, 1, 70001), (4, 33, 1, 4, 7)):
    g = torch.Generator(device="cuda").manual_seed(size * 100 + R)
    adv = torch.randn(size, R, device="cuda", generator=g) * 3 + 0.5
    B = ln * R
    bounds = list(range(0, B, mb)) + [B]
    n_mb = len(bounds) - 1
    slots = torch.full((B,), -1, dtype=torch.int32, device="cuda")
    pb = _lib.TsmPpoBatch(obs=None, act=None, adv=P(adv), ret=None, logp_old=None, v_old=None, indices=None, size=size, n_rows=R,
                          start=start, len=ln, perm_key=0xC0FFEE1234 + R, perm_n=B, slots=P(slots))
    st = torch.zeros(n_mb, 2, device="cuda")
    _lib.check(L.tsm_ppo_adv_stats(C.byref(pb), P(torch.as_tensor(bounds, dtype=torch.int64, device="cuda")), n_mb, 1e-8, P(st), S))
    out[f"st_{R}"] = st.cpu().numpy()
    out[f"sl_{R}"] = slots.cpu().numpy()
np.savez(sys.argv[1], **out)
"""


@pytest.mark.gpu
def test_adv_stats_walk_kernel_is_bit_identical_to_the_plain_gather(tmp_path):
    """tsm_ppo_adv_stats over a keyed permutation runs adv_stats_walk_kernel (cycle walk with lane refill); TSM_ADV_STATS_PLAIN=1
    (read once per process) selects the plain per-element loop.  Same permutation, same accumulation order: equal bits, and the
    slot table written as a by-product is the same permutation of the view's slots."""
    import os
    import subprocess
    import sys

    res = []
    for k, env in enumerate(({}, {"TSM_ADV_STATS_PLAIN": "1"})):
        e = dict(os.environ)
        e.update(env)
        e["PYTHONPATH"] = os.path.dirname(os.path.dirname(os.path.abspath(__file__))) + os.pathsep + e.get("PYTHONPATH", "")
        f = str(tmp_path / f"s{k}.npz")
        r = subprocess.run([sys.executable, "-c", _STATS_SNIPPET, f], env=e, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        res.append(np.load(f))
    assert sorted(res[0].files) == sorted(res[1].files) and len(res[0].files) == 8
    for k in res[0].files:
        assert np.array_equal(res[0][k].view(np.uint32) if res[0][k].dtype == np.float32 else res[0][k],
                              res[1][k].view(np.uint32) if res[1][k].dtype == np.float32 else res[1][k]), k
    sl = res[0]["sl_1237"]  # view: 6 of 8 ring slots per row from slot 5 on
    t, row = sl // 1237, sl % 1237
    assert len(np.unique(sl)) == len(sl) == 6 * 1237 and set(np.unique(t)) == {5, 6, 7, 0, 1, 2} and row.max() == 1236
