"""Golden fixture `marl.npz`: the UNMODIFIED reference FlexibleMultiAgentPolicyManager.forward (flexible_policy.py:189-231 ->
MultiAgentPolicy.forward, marl.py:108-185) with deterministic policies (DiscreteActorPolicy(deterministic_eval=True): act = argmax)
over rows row = env * A + agent carrying obs.agent_id.  Modes: independent (one policy per agent), shared (one policy, the
_forward_shared path), grouped (what the reference does there is recorded as found).  Per mode: weights of every policy, the
observations, holder.act, the agent keys of holder.out / holder.state and each agent's own out.act / out.logits.
Run in the build container:  cd tests/golden && python make_golden_marl.py"""
from __future__ import annotations

import numpy as np

from make_golden import save  # noqa: E402  (also sets sys.path for the reference + shims)


def gen() -> dict:
    import gymnasium as gym
    import torch
    from tianshou.algorithm.modelfree.reinforce import DiscreteActorPolicy
    from tianshou.algorithm.multiagent.flexible_policy import FlexibleMultiAgentPolicyManager
    from tianshou.data import Batch
    from tianshou.utils.net.common import Net
    from tianshou.utils.net.discrete import DiscreteActor

    A, O, H, nA, E = 4, 16, 64, 5, 6
    agents = ["adversary_0", "adversary_1", "adversary_2", "agent_0"]

    class Env:
        pass

    env = Env()
    env.agents = agents

    def mk(seed: int):
        torch.manual_seed(seed)
        actor = DiscreteActor(preprocess_net=Net(state_shape=(O,), hidden_sizes=[H, H]), action_shape=nA, softmax_output=False)
        pol = DiscreteActorPolicy(actor=actor, action_space=gym.spaces.Discrete(nA), deterministic_eval=True)
        pol.eval()
        return pol

    def weights(pol) -> list[np.ndarray]:
        sd = pol.actor.state_dict()
        pick = lambda suf: next(v for k, v in sd.items() if k.endswith(suf)).numpy().copy()  # noqa: E731
        return [pick("preprocess.model.model.0.weight"), pick("preprocess.model.model.0.bias"), pick("preprocess.model.model.2.weight"),
                pick("preprocess.model.model.2.bias"), pick("last.model.0.weight"), pick("last.model.0.bias")]

    rng = np.random.default_rng(9)
    obs = rng.normal(size=(E * A, O)).astype(np.float32)
    agent_id = np.array([agents[r % A] for r in range(E * A)])
    out: dict = {"obs": obs, "agents": np.array(agents), "dims": np.array([A, O, H, nA, E])}
    setups = {
        "independent": (lambda: [mk(10 + i) for i in range(A)], dict(mode="independent")),
        "shared": (lambda: mk(20), dict(mode="shared")),
        "grouped": (lambda: {"adv": mk(30), "good": mk(31)}, dict(mode="grouped", agent_groups={"adv": agents[:3], "good": agents[3:]})),
    }
    for name, (make, kw) in setups.items():
        pols = make()
        man = FlexibleMultiAgentPolicyManager(pols, env, **kw)
        man.eval()
        uniq = []
        for a in agents:
            p = man.policy_map[a]
            if not any(p is q for q in uniq):
                uniq.append(p)
        out[f"{name}_policy_of_agent"] = np.array([next(i for i, q in enumerate(uniq) if q is man.policy_map[a]) for a in agents])
        for i, p in enumerate(uniq):
            for k, w in zip(("W1", "b1", "W2", "b2", "W3", "b3"), weights(p)):
                out[f"{name}_p{i}_{k}"] = w
        out[f"{name}_policies_keys"] = np.array(list(man.policies.keys()))
        batch = Batch(obs=Batch(obs=obs, agent_id=agent_id), info=Batch())
        # what the manager's own policy_map says every row should get: the mapped reference policy evaluated on that agent's rows
        routed = np.zeros(E * A, np.int64)
        routed_logits = np.zeros((E * A, nA), np.float32)
        with torch.no_grad():
            for ai, a in enumerate(agents):
                rows = np.arange(ai, E * A, A)
                o = man.policy_map[a](Batch(obs=obs[rows], info=Batch()))
                routed[rows] = np.asarray(o.act)
                routed_logits[rows] = o.logits.numpy()
        out[f"{name}_routed_act"], out[f"{name}_routed_logits"] = routed, routed_logits
        try:
            with torch.no_grad():
                h = man.forward(batch)
            out[f"{name}_ok"] = np.array(1)
            out[f"{name}_act"] = np.asarray(h.act)
            out[f"{name}_out_keys"] = np.array(list(h.out.keys()))
            out[f"{name}_state_keys"] = np.array(list(h.state.keys()))
            for a in h.out.keys():
                o = h.out[a]
                if hasattr(o, "act"):
                    out[f"{name}_out_{a}_act"] = np.asarray(o.act)
                    out[f"{name}_out_{a}_logits"] = o.logits.detach().numpy().copy()
        except Exception as e:  # noqa: BLE001  (recorded as found)
            out[f"{name}_ok"] = np.array(0)
            out[f"{name}_error"] = np.array(f"{type(e).__name__}: {e}"[:300])
    return out


if __name__ == "__main__":
    r = gen()
    for k in sorted(r):
        if k.endswith("_ok") or k.endswith("_error") or k.endswith("_keys"):
            print(k, r[k])
    save("marl.npz", **r)
