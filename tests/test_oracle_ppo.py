"""Pins oracle/ppo.py (torch-fp32 restatement) to fixtures recorded from the reference PPO
(tests/golden/ppo.npz <- tests/golden/make_golden_more.py).  Tolerances: fp32 reductions in a
different association order -> rtol 2e-5 on values, 1e-4 on losses, 2e-5 abs on parameters."""
import numpy as np
import pytest

from oracle import ppo as oppo
from _ppo_golden import load_case, load_gauss, load_reinforce


@pytest.mark.parametrize("ci", [0, 1, 2, 3])
def test_oracle_matches_reference_ppo(golden, ci):
    c = load_case(golden("ppo.npz"), ci)
    net = oppo.Net(c["W1"], c["b1"], c["W2"], c["b2"], c["W3"], c["b3"], c["nA"])
    f = c["flat"]
    # forward pin (actor output is probabilities when the reference ran softmax_output=True)
    import torch

    with torch.no_grad():
        logits, value = net.forward(torch.from_numpy(c["obs"][0]))
        a_out = torch.softmax(logits, -1) if c["probs"] else logits
    assert np.allclose(a_out.numpy(), c["fwd_actor_out"], rtol=1e-5, atol=1e-6)
    assert np.allclose(value.numpy(), c["fwd_value"], rtol=1e-5, atol=1e-6)
    v_s, ret, adv, logp_old = oppo.preprocess(net, f["obs"], f["obs_next"], f["act"], f["rew"], f["term"], f["trunc"], c["indices"],
                                              c["unfinished"], c["gamma"], c["lam"], c["probs"])
    assert np.allclose(v_s, c["v_s"], rtol=2e-5, atol=2e-6)
    assert np.allclose(ret, c["returns"], rtol=2e-5, atol=2e-5)
    assert np.allclose(adv, c["adv"], rtol=2e-5, atol=2e-5)
    assert np.allclose(logp_old, c["logp_old"], rtol=2e-5, atol=2e-6)
    log = oppo.update_with_orders(net, f["obs"], f["act"], c["adv"], c["returns"], c["logp_old"], c["v_s"], c["cfg"], c["orders"],
                                  c["bs"], c["lr"], c["max_norm"], c["probs"])
    assert len(log["loss"]) == len(c["loss_seq"])
    for k, ref in (("loss", "loss_seq"), ("clip", "clip_seq"), ("vf", "vf_seq"), ("ent", "ent_seq")):
        assert np.allclose(log[k], c[ref], rtol=1e-4, atol=1e-5), (k, np.abs(np.array(log[k]) - c[ref]).max())
    for k, p in zip(("W1", "b1", "W2", "b2", "W3", "b3"), net.p):
        assert np.allclose(p.detach().numpy(), c["after_" + k], rtol=0, atol=2e-5), (k, np.abs(p.detach().numpy() - c["after_" + k]).max())


@pytest.mark.parametrize("ci", [0, 1])
def test_oracle_matches_reference_ppo_recompute_advantage(golden, ci):
    """recompute_advantage=True (ppo.py:174-178): fixtures from the reference in tests/golden/ppo_recompute.npz."""
    c = load_case(golden("ppo_recompute.npz"), ci)
    net = oppo.Net(c["W1"], c["b1"], c["W2"], c["b2"], c["W3"], c["b3"], c["nA"])
    f = c["flat"]
    v_s, ret, adv, logp_old = oppo.preprocess(net, f["obs"], f["obs_next"], f["act"], f["rew"], f["term"], f["trunc"], c["indices"],
                                              c["unfinished"], c["gamma"], c["lam"], c["probs"])
    assert np.allclose(v_s, c["v_s"], rtol=2e-5, atol=2e-6) and np.allclose(adv, c["adv"], rtol=2e-5, atol=2e-5)
    rc = dict(obs_next=f["obs_next"], rew=f["rew"], term=f["term"], trunc=f["trunc"], indices=c["indices"], unfinished=c["unfinished"],
              gamma=c["gamma"], lam=c["lam"])
    log = oppo.update_with_orders(net, f["obs"], f["act"], c["adv"], c["returns"], c["logp_old"], c["v_s"], c["cfg"], c["orders"],
                                  c["bs"], c["lr"], c["max_norm"], c["probs"], recompute=rc)
    assert len(log["loss"]) == len(c["loss_seq"])
    for k, ref in (("loss", "loss_seq"), ("clip", "clip_seq"), ("vf", "vf_seq"), ("ent", "ent_seq")):
        assert np.allclose(log[k], c[ref], rtol=1e-4, atol=1e-5), (k, np.abs(np.array(log[k]) - c[ref]).max())
    for k, p in zip(("W1", "b1", "W2", "b2", "W3", "b3"), net.p):
        assert np.allclose(p.detach().numpy(), c["after_" + k], rtol=0, atol=2e-5), (k, np.abs(p.detach().numpy() - c["after_" + k]).max())
    # and without recomputation the later epochs must NOT match (the fixture really exercises the option)
    net2 = oppo.Net(c["W1"], c["b1"], c["W2"], c["b2"], c["W3"], c["b3"], c["nA"])
    log2 = oppo.update_with_orders(net2, f["obs"], f["act"], c["adv"], c["returns"], c["logp_old"], c["v_s"], c["cfg"], c["orders"],
                                   c["bs"], c["lr"], c["max_norm"], c["probs"])
    assert not np.allclose(log2["loss"], c["loss_seq"], rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("ci", [0, 1])
def test_oracle_matches_reference_ppo_return_scaling(golden, ci):
    """return_scaling=True (a2c.py:132-146, RunningMeanStd statistics.py:96-114): fixtures from the reference in
    tests/golden/ppo_retscale.npz (ret_rms warmed by one earlier preprocess; case 1 also recomputes advantages)."""
    g = golden("ppo_retscale.npz")
    c = load_case(g, ci)
    pre = f"p{ci}_"
    net = oppo.Net(c["W1"], c["b1"], c["W2"], c["b2"], c["W3"], c["b3"], c["nA"])
    f = c["flat"]
    rms = dict(zip(("mean", "var", "count"), (float(x) for x in g[pre + "rms_before"])))
    v_s, ret, adv, logp_old = oppo.preprocess(net, f["obs"], f["obs_next"], f["act"], f["rew"], f["term"], f["trunc"], c["indices"],
                                              c["unfinished"], c["gamma"], c["lam"], c["probs"], ret_rms=rms)
    assert np.allclose(v_s, c["v_s"], rtol=2e-5, atol=2e-6)
    assert np.allclose(ret, c["returns"], rtol=2e-5, atol=2e-5) and np.allclose(adv, c["adv"], rtol=2e-5, atol=2e-5)
    assert np.allclose([rms["mean"], rms["var"], rms["count"]], g[pre + "rms_after_preprocess"], rtol=1e-5)
    # scaling is really exercised: un-scaled returns differ
    _, ret_plain, _, _ = oppo.preprocess(net, f["obs"], f["obs_next"], f["act"], f["rew"], f["term"], f["trunc"], c["indices"],
                                         c["unfinished"], c["gamma"], c["lam"], c["probs"])
    assert not np.allclose(ret_plain, c["returns"], rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("ci", [0, 1])
def test_oracle_matches_reference_a2c(golden, ci):
    """A2C (a2c.py:247-288; case 1 with return_scaling): fixtures from the reference A2C in tests/golden/a2c.npz."""
    g = golden("a2c.npz")
    c = load_case(g, ci)
    pre = f"p{ci}_"
    net = oppo.Net(c["W1"], c["b1"], c["W2"], c["b2"], c["W3"], c["b3"], c["nA"])
    f = c["flat"]
    rms = dict(zip(("mean", "var", "count"), (float(x) for x in g[pre + "rms_before"]))) if ci == 1 else None
    v_s, ret, adv, _ = oppo.preprocess(net, f["obs"], f["obs_next"], f["act"], f["rew"], f["term"], f["trunc"], c["indices"],
                                       c["unfinished"], c["gamma"], c["lam"], c["probs"], ret_rms=rms)
    assert np.allclose(ret, c["returns"], rtol=2e-5, atol=2e-5) and np.allclose(adv, c["adv"], rtol=2e-5, atol=2e-5)
    cfg = dict(c["cfg"], a2c=True)
    log = oppo.update_with_orders(net, f["obs"], f["act"], c["adv"], c["returns"], np.zeros_like(c["adv"]), c["v_s"], cfg, c["orders"],
                                  c["bs"], c["lr"], c["max_norm"], c["probs"])
    for k, ref in (("loss", "loss_seq"), ("clip", "clip_seq"), ("vf", "vf_seq"), ("ent", "ent_seq")):
        assert np.allclose(log[k], c[ref], rtol=1e-4, atol=1e-5), (k, np.abs(np.array(log[k]) - c[ref]).max())
    for k, p in zip(("W1", "b1", "W2", "b2", "W3", "b3"), net.p):
        assert np.allclose(p.detach().numpy(), c["after_" + k], rtol=0, atol=2e-5), (k, np.abs(p.detach().numpy() - c["after_" + k]).max())


@pytest.mark.parametrize("ci", [0, 1])
def test_oracle_matches_reference_reinforce(golden, ci):
    """Reinforce (reinforce.py:249-382; case 1 with return_standardization): fixtures from the reference in reinforce.npz."""
    c = load_reinforce(golden("reinforce.npz"), ci)
    f = c["flat"]
    rms = dict(zip(("mean", "var", "count"), (float(x) for x in c["rms_before"])))
    ret = oppo.reinforce_returns(f["rew"], f["term"], f["trunc"], np.arange(c["R"] * c["T"]), c["unfinished"], c["gamma"], rms, c["std"])
    assert np.allclose(ret, c["returns"], rtol=1e-9, atol=1e-9)
    if c["std"]:
        assert np.allclose([rms["mean"], rms["var"], rms["count"]], c["rms_after"], rtol=1e-9)
    net = oppo.Net(c["W1"], c["b1"], c["W2"], c["b2"], c["W3"], c["b3"], c["nA"])
    losses = oppo.reinforce_update(net, f["obs"], f["act"], c["returns"], c["orders"], c["bs"], c["lr"])
    assert np.allclose(losses, c["loss_seq"], rtol=1e-4, atol=1e-5), np.abs(np.array(losses) - c["loss_seq"]).max()
    for k, p in zip(("W1", "b1", "W2", "b2", "W3", "b3"), net.p):
        assert np.allclose(p.detach().numpy(), c["after_" + k], rtol=0, atol=2e-5), k


@pytest.mark.parametrize("ci", [0, 1, 2, 3, 4, 5, 6, 7])
def test_oracle_matches_reference_ppo_gaussian(golden, ci):
    """Continuous actions: Independent(Normal(mu, exp(sigma_param)), 1) (tests/golden/ppo_gauss.npz, recorded from the reference's
    ContinuousActorProbabilistic + PPO; cases 3 and 4: unbounded=False, mu = max_action * tanh(head) with max_action 1 and 2)."""
    import torch

    c = load_gauss(golden("ppo_gauss.npz"), ci)
    net = oppo.Net(c["W1"], c["b1"], c["W2"], c["b2"], c["W3"], c["b3"], c["nA"], log_sigma=None if c["cond_sigma"] else c["log_sigma"],
                   mu_bound=c["mu_bound"], cond_sigma=c["cond_sigma"])
    f = c["flat"]
    with torch.no_grad():
        mu, value = net.forward(torch.from_numpy(c["obs"][0]))
    assert np.allclose(mu.numpy(), c["fwd_mu"], rtol=1e-5, atol=1e-6) and np.allclose(value.numpy(), c["fwd_value"], rtol=1e-5, atol=1e-6)
    if c["cond_sigma"]:  # cases 5, 6: sigma = exp(clamp(head, -20, 2)) per row; part of the rows sit on the clamp
        assert np.allclose(net._ls.exp().numpy(), c["fwd_sigma"], rtol=1e-5) and abs(float(c["fwd_sigma"].max()) - float(np.exp(2.0))) < 1e-5
    else:
        assert np.allclose(np.exp(c["log_sigma"])[None, :].repeat(len(c["fwd_sigma"]), 0), c["fwd_sigma"], rtol=1e-6)
    v_s, ret, adv, logp_old = oppo.preprocess(net, f["obs"], f["obs_next"], f["act"], f["rew"], f["term"], f["trunc"], c["indices"],
                                              c["unfinished"], c["gamma"], c["lam"])
    assert np.allclose(v_s, c["v_s"], rtol=2e-5, atol=2e-6) and np.allclose(ret, c["returns"], rtol=2e-5, atol=2e-5)
    assert np.allclose(adv, c["adv"], rtol=2e-5, atol=2e-5) and (c["a2c"] or np.allclose(logp_old, c["logp_old"], rtol=2e-5, atol=2e-5))
    cfg = dict(c["cfg"], a2c=True) if c["a2c"] else c["cfg"]  # case 7: A2C over the Gaussian actor (a2c.py:262-272)
    log = oppo.update_with_orders(net, f["obs"], f["act"], c["adv"], c["returns"], c["logp_old"], c["v_s"], cfg, c["orders"],
                                  c["bs"], c["lr"], c["max_norm"])
    assert len(log["loss"]) == len(c["loss_seq"])
    for k, ref in (("loss", "loss_seq"), ("clip", "clip_seq"), ("vf", "vf_seq"), ("ent", "ent_seq")):
        assert np.allclose(log[k], c[ref], rtol=1e-4, atol=1e-5), (k, np.abs(np.array(log[k]) - c[ref]).max())
    for k, p in zip(c["param_names"], net.p):
        assert np.allclose(p.detach().numpy(), c["after_" + k], rtol=0, atol=2e-5), (k, np.abs(p.detach().numpy() - c["after_" + k]).max())
    moved = (c["after_Wsig"] - c["Wsig"]) if c["cond_sigma"] else (c["after_log_sigma"] - c["log_sigma"])
    assert np.abs(moved).max() > 1e-4  # the fixture really moves sigma
