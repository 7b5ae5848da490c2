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


def test_manager_forward_matches_reference_fixture():
    """tests/golden/marl.npz (make_golden_marl.py): the reference FlexibleMultiAgentPolicyManager.forward with copied weights and
    deterministic policies.  independent mode: holder.act, the agent keys of out / state and every agent's own out.act are the
    reference's.  shared / grouped: the reference's forward raises with real policies (recorded as found: `shared_error`,
    `grouped_error` -- MultiAgentPolicy.forward iterates self.policies, which FlexibleMultiAgentPolicyManager re-keys by group
    name, marl.py:141-150 / flexible_policy.py:93-105; _forward_shared drops batch.info); the build routes rows by the reference's
    own policy_map, checked against the mapped reference policies evaluated per agent (`*_routed_act`)."""
    import os

    from tianshou_marl_b200.algorithm import DiscreteActorPolicy, FlexibleMultiAgentPolicyManager
    from tianshou_marl_b200.data import Batch
    from tianshou_marl_b200.utils import ActorCriticNet

    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "marl.npz"))
    A, O, H, nA, E = (int(x) for x in z["dims"])
    agents = [str(a) for a in z["agents"]]

    class Env:
        pass

    env = Env()
    env.agents = agents
    obs = torch.as_tensor(z["obs"], device="cuda")
    assert int(z["independent_ok"]) == 1 and int(z["shared_ok"]) == 0 and int(z["grouped_ok"]) == 0  # as found in the reference
    for mode in ("independent", "shared", "grouped"):
        pof = z[f"{mode}_policy_of_agent"]
        pols = []
        for i in range(int(pof.max()) + 1):
            net = ActorCriticNet(O, H, nA, with_value=False).load_arrays(*[z[f"{mode}_p{i}_{k}"] for k in ("W1", "b1", "W2", "b2", "W3", "b3")])
            pols.append(DiscreteActorPolicy(actor=net, deterministic_eval=True))
        if mode == "independent":
            man = FlexibleMultiAgentPolicyManager([pols[int(i)] for i in pof], env, mode="independent")
        elif mode == "shared":
            man = FlexibleMultiAgentPolicyManager(pols[0], env, mode="shared")
        else:
            man = FlexibleMultiAgentPolicyManager({"adv": pols[0], "good": pols[1]}, env, mode="grouped",
                                                  agent_groups={"adv": agents[:3], "good": agents[3:]})
        assert [str(k) for k in z[f"{mode}_policies_keys"]] == list(man.policies.keys())
        h = man.forward(Batch(obs=obs))
        act = h.act.cpu().numpy()
        # logits within 1e-5 of the reference's; an argmax may only differ where the two best logits are that close
        want, lg = z[f"{mode}_routed_act"], z[f"{mode}_routed_logits"]
        top2 = np.sort(lg, axis=1)[:, -2:]
        clear = (top2[:, 1] - top2[:, 0]) > 1e-4
        assert np.array_equal(act[clear], want[clear]) and clear.mean() > 0.95
        assert list(h.out.keys()) == agents and list(h.state.keys()) == agents
        if mode == "independent":
            assert np.array_equal(z["independent_act"][clear], want[clear])  # the reference holder.act is the routed act
            assert [str(k) for k in z["independent_out_keys"]] == agents and [str(k) for k in z["independent_state_keys"]] == agents
            for ai, a in enumerate(agents):
                rows = np.arange(ai, E * A, A)
                got = h.out[a].act.cpu().numpy()
                assert got.shape == z[f"independent_out_{a}_act"].shape
                assert np.array_equal(got[clear[rows]], z[f"independent_out_{a}_act"][clear[rows]])


@pytest.mark.parametrize("engine", ["fp32", "tc"])
def test_fused_grouped_rollout_equals_generic_loop(engine, monkeypatch):
    """tsm_collect_lockstep_groups (one forward launch per policy group and step inside the C loop) writes the same buffer,
    bit for bit, as the per-step Python loop through FlexibleMultiAgentPolicyManager.forward."""
    res = {}
    for py in (True, False):
        if py:
            monkeypatch.setenv("TSM_COLLECT_PY", "1")
        else:
            monkeypatch.delenv("TSM_COLLECT_PY", raising=False)
        env, algs, man, buf, Collector, A, O = _setup(E=96, T=12, engine=engine)
        col = Collector(man, env, buf)
        col.reset()
        for a in algs.values():
            a.policy.is_within_training_step = True
        man.is_within_training_step = True
        st = col.collect(n_step=12 * 96 * A)
        res[py] = ({k: v.clone() for k, v in buf._store.items() if v is not None}, st.n_collected_episodes, [a.policy._counter for a in algs.values()])
    (sa, ea, ca), (sb, eb, cb) = res[True], res[False]
    assert ea == eb and ca == cb
    for k in sa:
        assert torch.equal(sa[k], sb[k]), k
