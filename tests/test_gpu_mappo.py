"""BASELINE.json configs[2] on the device (MAPPO = reference PPO arithmetic over the reference's CTDE nets; CTDEPolicy.learn at the
same shapes) against fixtures recorded from the UNMODIFIED reference (tests/golden/mappo.npz, ctde_c3.npz;
make_golden_mappo.py).  Engine "tcw" (csrc/mlp_tcw.cu): fp16 operands, fp32 accumulation -- tolerances stated below."""
import numpy as np
import pytest
import torch

from _mappo_golden import cases, em, net_arrays

pytestmark = pytest.mark.gpu
CASES = cases()
# fp16-operand engine against the fp32 reference (the budget engine 1 is held to against ppo.npz): forward outputs
# 6e-3 and preprocessed values / returns / advantages / logp_old 8e-3 absolute; per-step losses rtol 2e-2 + atol 4e-3; parameters
# after the update: RMS error below 10 % of the RMS distance the reference moved them (measured on B200: 5.0 - 6.1 %), no
# entry further than 2 x lr x steps from the reference.  The fixtures hold 2-3 optimizer steps from zero Adam state, where an
# update is lr * g / (|g| + eps) ~ lr * sign(g): an entry whose gradient is within rounding noise of zero moves by the full
# +-lr in either direction, so a sign flip costs 2 lr per step whatever the size of the gradient error.
TOL = dict(fwd=6e-3, pre=8e-3, loss_rel=2e-2, loss_abs=4e-3, rms_frac=0.10, lr_steps=2.0)


def _build(c):
    from tianshou_marl_b200.algorithm import AdamOptimizerFactory, DiscreteActorPolicy
    from tianshou_marl_b200.algorithm.ctde import CentralizedCritic, DecentralizedActor
    from tianshou_marl_b200.algorithm.mappo import MAPPO
    from tianshou_marl_b200.data import Batch, VectorReplayBuffer

    A, O, H, nA, T, R = c["A"], c["O"], c["H"], c["nA"], c["T"], c["R"]
    actor = DecentralizedActor(O, nA, H, engine="tcw").load_arrays(*net_arrays(c, "actor"))
    critic = CentralizedCritic(A * O, A, H, engine="tcw").load_arrays(*net_arrays(c, "critic"))
    pol = DiscreteActorPolicy(actor=actor)
    p = c["ppo"]
    alg = MAPPO(policy=pol, critic=critic, n_agents=A, optim=AdamOptimizerFactory(lr=c["lr"]), eps_clip=p["eps_clip"],
                dual_clip=p["dual_clip"] or None, value_clip=p["value_clip"], advantage_normalization=p["adv_norm"], vf_coef=p["vf_coef"],
                ent_coef=p["ent_coef"], max_grad_norm=c["max_grad_norm"] or None, gae_lambda=c["lam"], gamma=c["gamma"])
    buf = VectorReplayBuffer(R * T, R, index_type="torch")
    for t in range(T):
        buf.add(Batch(obs=c["obs"][t], act=c["act"][t], rew=c["rew"][t].astype(np.float32), terminated=c["term"][t], truncated=c["trunc"][t],
                      obs_next=c["obs_next"][t]))
    return actor, critic, pol, alg, buf


@pytest.mark.parametrize("ci", range(len(CASES)))
def test_forward_matches_reference(ci):
    c = CASES[ci]
    actor, critic, pol, alg, buf = _build(c)
    A, O = c["A"], c["O"]
    x = torch.as_tensor(c["obs"][0], device="cuda")
    lg = actor.forward(x, mode="raw")["out"].cpu().numpy()
    assert np.abs(lg - c["fwd_logits"]).max() < TOL["fwd"]
    v = critic.forward(x.view(-1, A * O), mode="raw")["out"].cpu().numpy()
    assert np.abs(v - c["fwd_values"]).max() < TOL["fwd"]


@pytest.mark.parametrize("ci", range(len(CASES)))
def test_update_matches_reference(ci):
    c = CASES[ci]
    actor, critic, pol, alg, buf = _build(c)
    T, R = c["T"], c["R"]
    pol.is_within_training_step = True
    stats = alg.update(buf, R * T, c["repeat"])
    assert stats.gradient_steps == c["repeat"]
    tm = lambda key: em(alg._work[key].cpu().numpy())  # noqa: E731  [T][R] -> buffer.sample(0) order
    for w, k in (("v_s", "v_s"), ("ret", "returns"), ("adv", "adv"), ("logp_old", "logp_old")):
        assert np.abs(tm(w) - c[k]).max() < TOL["pre"], (k, np.abs(tm(w) - c[k]).max())
    seq = alg.last_sequences
    for k, ref in (("loss", "loss_seq"), ("clip", "clip_seq"), ("vf", "vf_seq"), ("ent", "ent_seq")):
        assert np.allclose(seq[k], c[ref], rtol=TOL["loss_rel"], atol=TOL["loss_abs"]), (k, seq[k], c[ref])
    for which, net in (("actor", actor), ("critic", critic)):
        init = np.concatenate([a.reshape(-1) for a in net_arrays(c, which)])
        want = np.concatenate([a.reshape(-1) for a in net_arrays(c, which, "after")])
        got = net.flat.detach().cpu().numpy()
        travel = np.sqrt(np.mean((want - init) ** 2))
        assert np.sqrt(np.mean((got - want) ** 2)) < TOL["rms_frac"] * travel, (which, np.sqrt(np.mean((got - want) ** 2)), travel)
        assert np.abs(got - want).max() < TOL["lr_steps"] * c["lr"] * c["repeat"], which


def test_minibatches_of_super_tiles_cover_the_batch():
    """batch_size < batch: every super-tile is visited once per epoch, minibatch row counts add up, losses are finite."""
    from tianshou_marl_b200.algorithm import AdamOptimizerFactory, DiscreteActorPolicy
    from tianshou_marl_b200.algorithm.ctde import CentralizedCritic, DecentralizedActor
    from tianshou_marl_b200.algorithm.mappo import MAPPO
    from tianshou_marl_b200.data import Batch, VectorReplayBuffer

    A, O, E, T = 3, 18, 200, 7  # 1400 (step, world) pairs = 10.9 super-tiles: the last one is ragged
    R = E * A
    actor, critic = DecentralizedActor(O, 5, 128, engine="tcw", orthogonal=True), CentralizedCritic(A * O, A, 128, engine="tcw", orthogonal=True)
    pol = DiscreteActorPolicy(actor=actor)
    alg = MAPPO(policy=pol, critic=critic, n_agents=A, optim=AdamOptimizerFactory(lr=1e-3), max_grad_norm=0.5)
    buf = VectorReplayBuffer(R * T, R, index_type="torch")
    rng = np.random.default_rng(0)
    for t in range(T):
        buf.add(Batch(obs=rng.normal(size=(R, O)).astype(np.float32), act=rng.integers(0, 5, R), rew=rng.normal(size=R).astype(np.float32),
                      terminated=np.zeros(R, bool), truncated=np.full(R, t == 3), obs_next=rng.normal(size=(R, O)).astype(np.float32)))
    pol.is_within_training_step = True
    np.random.seed(1)
    p0 = actor.flat.detach().clone()
    stats = alg.update(buf, 3 * 128 * A, 2)  # 3 super-tiles per minibatch -> 4 minibatches per epoch
    assert stats.gradient_steps == 8
    assert all(np.isfinite(alg.last_sequences[k]).all() for k in ("loss", "clip", "vf", "ent"))
    assert float((actor.flat.detach() - p0).abs().max()) > 0
    assert int(actor.adam_step) == 8 and int(critic.adam_step) == 8


def test_ctde_learn_c3_shapes_matches_reference():
    """CTDEPolicy.learn (ctde.py:121-199) at A = 10, O = 60, H = 128 on engine "tcw" against ctde_c3.npz."""
    import os

    from tianshou_marl_b200.algorithm import AdamOptimizerFactory
    from tianshou_marl_b200.algorithm.ctde import CentralizedCritic, CTDEPolicy, DecentralizedActor, GlobalStateConstructor, load_linear3
    from tianshou_marl_b200.data import Batch

    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ctde_c3.npz"))
    g = lambda k: z["c0_" + k]  # noqa: E731
    A, O, H, nA, E = (int(g(k)) for k in ("A", "O", "H", "nA", "n_env"))
    actor = load_linear3(DecentralizedActor(O, nA, H, engine="tcw"), {k[len("c0_init_actor_"):]: z[k] for k in z.files if k.startswith("c0_init_actor_")})
    critic = load_linear3(CentralizedCritic(A * O, A, H, engine="tcw"), {k[len("c0_init_critic_"):]: z[k] for k in z.files if k.startswith("c0_init_critic_")})
    pol = CTDEPolicy(actor, critic, AdamOptimizerFactory(lr=float(g("lr_a"))), AdamOptimizerFactory(lr=float(g("lr_c"))), discount_factor=float(g("gamma")))
    obs, obs_next = torch.as_tensor(g("obs"), device="cuda"), torch.as_tensor(g("obs_next"), device="cuda")
    gsc = GlobalStateConstructor("concatenate", n_agents=A)
    g_rows, g_rows_next = gsc.build_rows(obs, A), gsc.build_rows(obs_next, A)
    assert np.array_equal(g_rows.view(E, A, -1)[:, 0].cpu().numpy(), g("global_concat"))
    fwd = pol.forward(Batch(obs=obs))
    assert np.abs(fwd.act.cpu().numpy() - g("fwd_act")).max() < TOL["fwd"]
    losses = []
    for _ in range(int(g("steps"))):
        r = pol.learn(Batch(obs=obs, act=g("act"), rew=g("rew"), obs_next=obs_next, terminated=g("term"), global_obs=g_rows, global_obs_next=g_rows_next))
        losses.append([r["actor_loss"], r["critic_loss"]])
    assert np.allclose(np.array(losses), g("losses"), rtol=2e-2, atol=4e-3), (losses, g("losses"))
    for which, net in (("actor", actor), ("critic", critic)):
        want = np.concatenate([z[f"c0_after_{which}_{k}"].reshape(-1) for k in ("fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias", "fc3.weight", "fc3.bias")])
        init = np.concatenate([z[f"c0_init_{which}_{k}"].reshape(-1) for k in ("fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias", "fc3.weight", "fc3.bias")])
        got = net.flat.detach().cpu().numpy()
        travel = np.sqrt(np.mean((want - init) ** 2))
        # n = 260 rows, three sign-like Adam steps: measured 17 % (critic: 76 800 of its 94 730 entries are fc1 weights whose
        # gradients are 260-term sums of products with N(0,1) inputs) -- see TOL above
        assert np.sqrt(np.mean((got - want) ** 2)) < 0.25 * travel, (which, np.sqrt(np.mean((got - want) ** 2)), travel)
