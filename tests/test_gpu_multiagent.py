"""FlexibleMultiAgentPolicyManager on the device (a15): static row routing, one fused forward per policy group,
per-group PPO updates over strided row sets (simple_tag: adversaries share one policy, the good agent has its own)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(E=64, T=8, engine="fp32"):
    from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory, DiscreteActorPolicy, FlexibleMultiAgentPolicyManager
    from tianshou_marl_b200.data import Collector, VectorReplayBuffer
    from tianshou_marl_b200.env import MPEVectorEnv
    from tianshou_marl_b200.utils import ActorCriticNet

    env = MPEVectorEnv("simple_tag_v3", n_envs=E, seed=3)
    A, O = env.n_agents, env.obs_dim
    algs = {}
    for name, seed in (("adv", 1), ("good", 2)):
        net = ActorCriticNet(O, 64, 5, seed=seed, engine=engine)
        pol = DiscreteActorPolicy(actor=net, seed=10 + seed)
        algs[name] = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=1e-3), max_grad_norm=0.5, minibatch_order="numpy", seed=seed)
    groups = {"adv": [a for a in env.agents if a.startswith("adversary")], "good": [a for a in env.agents if a.startswith("agent")]}
    man = FlexibleMultiAgentPolicyManager(algs, env, mode="grouped", agent_groups=groups)
    buf = VectorReplayBuffer(T * E * A, E * A, index_type="torch")
    return env, algs, man, buf, Collector, A, O


def test_forward_routes_rows_to_group_policies():
    env, algs, man, buf, _, A, O = _setup()
    obs, _ = env.reset()
    R = obs.shape[0]
    from tianshou_marl_b200.data import Batch

    for a in algs.values():
        a.policy.is_within_training_step = True
    out = man.forward(Batch(obs=obs))
    rows_adv = torch.tensor([r for r in range(R) if env.agents[r % A].startswith("adversary")], device="cuda")
    rows_good = torch.tensor([r for r in range(R) if env.agents[r % A].startswith("agent")], device="cuda")
    for name, rows in (("adv", rows_adv), ("good", rows_good)):
        pol = algs[name].policy
        ref = pol.actor.forward(obs, mode="sample", seed=pol.seed, counter=pol._counter - 1)  # whole batch through this net
        assert torch.equal(out.act[rows].to(torch.int32), ref["act"][rows])
        assert torch.equal(out.logp[rows], ref["logp"][rows])
        assert torch.equal(out.value[rows], ref["value"][rows])
    assert set(out.out) == set(env.agents) and out.out["adversary_0"] is out.out["adversary_2"] and out.out["agent_0"] is not out.out["adversary_0"]


def test_group_update_touches_only_its_rows():
    env, algs, man, buf, Collector, A, O = _setup()
    col = Collector(man, env, buf)
    col.reset()
    for a in algs.values():
        a.policy.is_within_training_step = True
    man.is_within_training_step = True
    E = env.n_envs
    col.collect(n_step=8 * E * A)
    before = {k: a.net.flat.detach().clone() for k, a in algs.items()}
    np.random.seed(0)
    stats = man.update(buf, 256, 2)
    assert set(stats) == {"adv", "good"} and all(np.isfinite(s.loss.mean) for s in stats.values())
    after = {k: a.net.flat.detach().clone() for k, a in algs.items()}
    assert all(not torch.equal(before[k], after[k]) for k in algs)
    # the same update again from the same start, with the OTHER group's rewards corrupted: identical result for "good"
    for k, a in algs.items():
        with torch.no_grad():
            a.net.flat.copy_(before[k])
        a.net.reset_optimizer()
    rows_adv = torch.tensor([r for r in range(E * A) if env.agents[r % A].startswith("adversary")], device="cuda")
    buf._store["rew"][:, rows_adv] += 100.0
    np.random.seed(0)
    man.update(buf, 256, 2)
    assert torch.equal(algs["good"].net.flat.detach(), after["good"])
    assert not torch.equal(algs["adv"].net.flat.detach(), after["adv"])


def test_shared_mode_equals_plain_ppo():
    from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory, DiscreteActorPolicy, FlexibleMultiAgentPolicyManager
    from tianshou_marl_b200.data import Collector, VectorReplayBuffer
    from tianshou_marl_b200.env import MPEVectorEnv
    from tianshou_marl_b200.utils import ActorCriticNet

    res = []
    for use_manager in (False, True):
        env = MPEVectorEnv("simple_spread_v3", n_envs=32, n_agents=3, seed=5)
        net = ActorCriticNet(env.obs_dim, 64, 5, seed=0)
        pol = DiscreteActorPolicy(actor=net, seed=4)
        alg = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=1e-3), minibatch_order="numpy", seed=0)
        buf = VectorReplayBuffer(8 * 96, 96, index_type="torch")
        pol.is_within_training_step = True
        if use_manager:
            man = FlexibleMultiAgentPolicyManager(alg, env, mode="shared")
            man.is_within_training_step = True
            col = Collector(man, env, buf)
        else:
            col = Collector(alg, env, buf)
        col.reset()
        col.collect(n_step=8 * 96)
        np.random.seed(1)
        if use_manager:
            man.update(buf, 128, 2)
        else:
            alg.update(buf, 128, 2)
        res.append((buf._store["act"].clone(), net.flat.detach().clone()))
    assert torch.equal(res[0][0], res[1][0])  # same actions through the manager's generic collect loop
    assert torch.allclose(res[0][1], res[1][1], atol=1e-6)
