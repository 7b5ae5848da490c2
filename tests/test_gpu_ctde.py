"""CTDE row (a16) on the device against fixtures recorded from the unmodified reference (tests/golden/ctde.npz):
GlobalStateConstructor (bit-exact concatenate, mean to 1 ulp), DecentralizedActor forward, CTDEPolicy.learn losses and
parameters after several learn() calls.  Tolerances: fp32, different summation order -> losses rtol 2e-4, parameters 2e-5."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def load(golden, ci):
    g = golden("ctde.npz")
    pre = f"c{ci}_"
    return {k[len(pre):]: g[k] for k in g.files if k.startswith(pre)}


@pytest.mark.parametrize("ci", [0, 1, 2])
def test_ctde_matches_reference(golden, ci):
    from tianshou_marl_b200.algorithm import AdamOptimizerFactory
    from tianshou_marl_b200.algorithm.ctde import CTDEPolicy, CentralizedCritic, DecentralizedActor, GlobalStateConstructor, load_linear3
    from tianshou_marl_b200.data import Batch

    c = load(golden, ci)
    A, O, H, nA = int(c["A"]), int(c["O"]), int(c["H"]), int(c["nA"])
    use = bool(c["use_global"])
    dev = "cuda"
    obs, obs_next = torch.tensor(c["obs"], device=dev), torch.tensor(c["obs_next"], device=dev)
    gsc = GlobalStateConstructor("concatenate", obs_dim=O, n_agents=A)
    g_rows, g_rows_next = gsc.build_rows(obs), gsc.build_rows(obs_next)
    assert np.array_equal(g_rows.cpu().numpy(), np.repeat(c["global_concat"], A, axis=0))
    g_mean = GlobalStateConstructor("mean", n_agents=A).build_rows(obs)
    assert np.allclose(g_mean.cpu().numpy(), np.repeat(c["global_mean"], A, axis=0), rtol=0, atol=2e-7)
    # the reference's dict signature
    per_agent = {f"agent_{a}": obs.view(-1, A, O)[:, a].contiguous() for a in range(A)}
    assert np.array_equal(gsc.build(per_agent).cpu().numpy(), c["global_concat"])
    actor = load_linear3(DecentralizedActor(O, nA, H), {k[len("init_actor_"):]: v for k, v in c.items() if k.startswith("init_actor_")})
    critic = load_linear3(CentralizedCritic(A * O if use else O, A, H),
                          {k[len("init_critic_"):]: v for k, v in c.items() if k.startswith("init_critic_")})
    pol = CTDEPolicy(actor, critic, AdamOptimizerFactory(lr=float(c["lr_a"])), AdamOptimizerFactory(lr=float(c["lr_c"])),
                     enable_global_info=use, discount_factor=float(c["gamma"]))
    fwd = pol.forward(Batch(obs=obs))
    assert np.allclose(fwd.act.cpu().numpy(), c["fwd_act"], rtol=1e-5, atol=1e-5)
    for s in range(int(c["steps"])):
        b = Batch(obs=obs, act=torch.tensor(c["act"], device=dev), rew=torch.tensor(c["rew"], device=dev), obs_next=obs_next,
                  terminated=torch.tensor(c["term"], device=dev))
        if use:
            b.__dict__["global_obs"], b.__dict__["global_obs_next"] = g_rows, g_rows_next
        r = pol.learn(b)
        assert np.allclose([r["actor_loss"], r["critic_loss"]], c["losses"][s], rtol=2e-4, atol=2e-6), (s, r, c["losses"][s])
    for name, net in (("actor", actor), ("critic", critic)):
        v = net.named_views()
        for mine, ref in (("W1", "fc1.weight"), ("b1", "fc1.bias"), ("W2", "fc2.weight"), ("b2", "fc2.bias"), ("W3", "fc3.weight"), ("b3", "fc3.bias")):
            d = np.abs(v[mine].cpu().numpy() - c[f"after_{name}_{ref}"]).max()
            assert d < 2e-5, (name, mine, d)


def test_ctde_per_row_advantage_differs_from_broadcast():
    """broadcast_advantage=False (per-row advantage, the intended math) is a different update than the reference's literal one."""
    from tianshou_marl_b200.algorithm import AdamOptimizerFactory
    from tianshou_marl_b200.algorithm.ctde import CTDEPolicy, CentralizedCritic, DecentralizedActor
    from tianshou_marl_b200.data import Batch

    n, O, A = 96, 18, 3
    res = []
    for bc in (True, False):
        rng = np.random.default_rng(0)
        actor, critic = DecentralizedActor(O, 5, 64, seed=1, orthogonal=True), CentralizedCritic(O, A, 64, seed=2, orthogonal=True)
        pol = CTDEPolicy(actor, critic, AdamOptimizerFactory(lr=1e-3), AdamOptimizerFactory(lr=1e-3), enable_global_info=False,
                         broadcast_advantage=bc)
        b = Batch(obs=rng.normal(size=(n, O)).astype(np.float32), act=rng.integers(0, 5, n), rew=rng.normal(size=n).astype(np.float32),
                  obs_next=rng.normal(size=(n, O)).astype(np.float32), terminated=rng.random(n) < 0.1)
        r = pol.learn(b)
        assert np.isfinite(r["actor_loss"]) and np.isfinite(r["critic_loss"])
        res.append(actor.flat.detach().clone())
    assert not torch.equal(res[0], res[1])


@pytest.mark.gpu
def test_simultaneous_trainer_routes_joint_batch_to_ctde_policies():
    """SimultaneousTrainer.train_step (training_coordinator.py:292-339) over real CTDEPolicy objects: each agent's policy learns
    from its slice of the joint batch with the joint global state attached; result == calling learn() by hand."""
    import types

    from tianshou_marl_b200.algorithm import AdamOptimizerFactory
    from tianshou_marl_b200.algorithm.ctde import CTDEPolicy, CentralizedCritic, DecentralizedActor
    from tianshou_marl_b200.algorithm.training_coordinator import SimultaneousTrainer
    from tianshou_marl_b200.data import Batch

    n, O, A = 64, 18, 3
    rng = np.random.default_rng(1)
    g_obs = rng.normal(size=(n, A * O)).astype(np.float32)
    g_next = rng.normal(size=(n, A * O)).astype(np.float32)

    def mk():
        pols = {}
        for i in range(2):
            actor, critic = DecentralizedActor(O, 5, 64, seed=10 + i, orthogonal=True), CentralizedCritic(A * O, A, 64, seed=20 + i, orthogonal=True)
            pols[f"agent_{i}"] = CTDEPolicy(actor, critic, AdamOptimizerFactory(lr=1e-3), AdamOptimizerFactory(lr=1e-3), enable_global_info=True)
        return pols

    def agent_batch(i):
        r = np.random.default_rng(100 + i)
        return Batch(obs=r.normal(size=(n, O)).astype(np.float32), act=r.integers(0, 5, n), rew=r.normal(size=n).astype(np.float32),
                     obs_next=r.normal(size=(n, O)).astype(np.float32), terminated=r.random(n) < 0.1)

    pols = mk()
    joint = Batch(agent_0=agent_batch(0), agent_1=agent_batch(1))
    joint.__dict__["global_obs"], joint.__dict__["global_obs_next"] = g_obs, g_next
    pm = types.SimpleNamespace(mode="independent", agents=list(pols), policies=pols)
    out = SimultaneousTrainer(pm, agent_train_freq={"agent_1": 2}).train_step(joint)
    assert set(out) == {"agent_0"} and set(out["agent_0"]) == {"actor_loss", "critic_loss"}
    ref = mk()
    b0 = agent_batch(0)
    b0.__dict__["global_obs"], b0.__dict__["global_obs_next"] = g_obs, g_next
    want = ref["agent_0"].learn(b0)
    assert np.allclose([out["agent_0"]["actor_loss"], out["agent_0"]["critic_loss"]], [want["actor_loss"], want["critic_loss"]], rtol=1e-5)
