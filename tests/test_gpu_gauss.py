"""Continuous-action PPO (diagonal Gaussian head) on the GPU against fixtures recorded from the reference's
ContinuousActorProbabilistic (unbounded, and tanh-bounded with max_action 1 and 2) + ProbabilisticActorPolicy + PPO
(tests/golden/ppo_gauss.npz <- make_golden_gauss.py).
fp32 engine: same tolerances as the discrete reference test (tests/test_gpu_ppo.py)."""
import numpy as np
import pytest
import torch

from _ppo_golden import load_gauss

pytestmark = pytest.mark.gpu


def _mk(c):
    from tianshou_marl_b200.algorithm import A2C, PPO, AdamOptimizerFactory, ContinuousActorPolicy
    from tianshou_marl_b200.data import Batch, VectorReplayBuffer
    from tianshou_marl_b200.utils import ActorCriticNet

    net = ActorCriticNet(c["O"], c["H"], c["nA"], gaussian=True, unbounded=c["mu_bound"] == 0.0, max_action=c["mu_bound"] or 1.0,
                         conditioned_sigma=c["cond_sigma"]).load_arrays(c["W1"], c["b1"], c["W2"], c["b2"], c["W3"], c["b3"])
    if not c["cond_sigma"]:
        with torch.no_grad():
            net.view("log_sigma").copy_(torch.as_tensor(c["log_sigma"]))
    pol = ContinuousActorPolicy(actor=net)
    k = c["cfg"]
    if c["a2c"]:  # case 7: recorded from the reference A2C
        alg = A2C(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=c["lr"]), vf_coef=k["vf_coef"], ent_coef=k["ent_coef"],
                  max_grad_norm=c["max_norm"] or None, gamma=c["gamma"], gae_lambda=c["lam"], minibatch_order="numpy")
    else:
        alg = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=c["lr"]), eps_clip=k["eps_clip"], dual_clip=k["dual_clip"] or None,
                  value_clip=k["value_clip"], advantage_normalization=k["adv_norm"], vf_coef=k["vf_coef"], ent_coef=k["ent_coef"],
                  max_grad_norm=c["max_norm"] or None, gamma=c["gamma"], gae_lambda=c["lam"], minibatch_order="numpy")
    buf = VectorReplayBuffer(c["R"] * c["T"], c["R"], index_type="torch")
    for t in range(c["T"]):
        buf.add(Batch(obs=c["obs"][t], act=c["act"][t], rew=c["rew"][t], terminated=c["term"][t], truncated=c["trunc"][t],
                      obs_next=c["obs_next"][t]))
    return net, pol, alg, buf


@pytest.mark.parametrize("ci", [0, 1, 2, 3, 4, 5, 6, 7])
def test_gaussian_update_matches_reference(golden, ci):
    """Whole PPO.update vs the reference: v_s, returns, adv, logp_old, every minibatch loss, final parameters incl. log sigma."""
    c = load_gauss(golden("ppo_gauss.npz"), ci)
    net, pol, alg, buf = _mk(c)
    # the buffer hands the float action vectors back (buffer_base.py:591-635)
    got = buf[np.arange(c["R"] * c["T"])]
    assert np.array_equal(np.asarray(got.act), c["flat"]["act"])
    raw = net.forward(torch.as_tensor(c["obs"][0], device="cuda"), mode="raw")["out"]
    o = raw.cpu().numpy()
    assert np.allclose(net.gaussian_mean(raw).cpu().numpy(), c["fwd_mu"], rtol=2e-5, atol=2e-6)  # cases 3, 4: max_action * tanh(head)
    assert np.allclose(o[:, net.value_index], c["fwd_value"], rtol=2e-5, atol=2e-6)
    if c["cond_sigma"]:  # cases 5, 6: per-row sigma = exp(clamp(head, -20, 2)), 7-18 % of the entries on the clamp
        assert np.allclose(net.gaussian_log_sigma(raw).exp().cpu().numpy(), c["fwd_sigma"], rtol=2e-5)
    pol.is_within_training_step = True
    np.random.seed(c["seed"])
    stats = alg.update(buf, c["bs"], c["repeat"])
    em = lambda t: t.t().contiguous().view(-1).cpu().numpy()  # noqa: E731  time-major -> env-major
    assert np.allclose(em(alg._work["v_s"]), c["v_s"], rtol=2e-5, atol=2e-6)
    assert np.allclose(em(alg._work["ret"]), c["returns"], rtol=2e-5, atol=2e-5)
    assert np.allclose(em(alg._work["adv"]), c["adv"], rtol=2e-5, atol=2e-5)
    assert c["a2c"] or np.allclose(em(alg._work["logp_old"]), c["logp_old"], rtol=2e-5, atol=2e-5)  # (A2C records no logp_old)
    seq = alg.last_sequences
    assert stats.gradient_steps == len(c["loss_seq"])
    for k, ref in (("loss", "loss_seq"), ("clip", "clip_seq"), ("vf", "vf_seq"), ("ent", "ent_seq")):
        assert np.allclose(seq[k], c[ref], rtol=2e-4, atol=2e-5), (k, np.abs(seq[k] - c[ref]).max())
    v = net.named_views()
    for k in c["param_names"]:
        err = np.abs(v[k].cpu().numpy() - c["after_" + k]).max()
        assert err < 5e-5, (k, err)
    if c["cond_sigma"]:
        A = c["nA"]
        assert np.abs(v["W3"].cpu().numpy()[A: 2 * A] - c["Wsig"]).max() > 1e-4  # the sigma head really moved
    else:
        assert np.abs(v["log_sigma"].cpu().numpy() - c["log_sigma"]).max() > 1e-4  # sigma really moved
    assert float(net.grads.abs().max().item()) == 0.0


def test_gaussian_policy_forward_and_map_action():
    """ContinuousActorPolicy.forward: act = mu + sigma * eps, logp = Independent(Normal).log_prob(act); the mean under
    deterministic_eval; map_action clips to [-1, 1] and scales into the Box (algorithm_base.py:255-288)."""
    from types import SimpleNamespace

    from tianshou_marl_b200.algorithm import ContinuousActorPolicy
    from tianshou_marl_b200.data import Batch
    from tianshou_marl_b200.utils import ActorCriticNet

    net = ActorCriticNet(9, 64, 3, gaussian=True, seed=2)
    with torch.no_grad():
        net.view("log_sigma").copy_(torch.tensor([-1.0, 0.0, 0.5]))
    space = SimpleNamespace(low=np.array([-2.0, 0.0, -1.0], np.float32), high=np.array([2.0, 4.0, 1.0], np.float32))
    pol = ContinuousActorPolicy(actor=net, action_space=space, seed=5)
    obs = torch.randn(4096, 9, device="cuda")
    out = pol(Batch(obs=obs), want_dist=True)
    mu, sigma = out.logits
    assert out.act.shape == (4096, 3) and torch.allclose(out.logp, out.dist.log_prob(out.act), atol=1e-5)
    z = ((out.act - mu) / sigma).cpu().numpy()
    assert abs(z.mean()) < 0.05 and abs(z.std() - 1.0) < 0.05
    assert torch.allclose(sigma[0], torch.tensor([-1.0, 0.0, 0.5], device="cuda").exp())
    pol.deterministic_eval = True
    assert torch.equal(pol(Batch(obs=obs)).act, mu)
    m = pol.map_action(out.act)
    lo, hi = torch.as_tensor(space.low, device="cuda"), torch.as_tensor(space.high, device="cuda")
    want = lo + (hi - lo) * (out.act.clamp(-1, 1) + 1.0) / 2.0
    assert torch.allclose(m, want, atol=1e-6)


def test_gaussian_net_rejected_by_tensor_core_engines():
    from tianshou_marl_b200.utils import ActorCriticNet

    with pytest.raises(ValueError):
        ActorCriticNet(18, 64, 3, gaussian=True, engine="tc")
