"""FlexibleMultiAgentPolicyManager configuration rules (reference flexible_policy.py:108-186, 233-257); no GPU needed."""
import pytest

from tianshou_marl_b200.algorithm.multiagent import FlexibleMultiAgentPolicyManager as M


class Env:
    agents = ["adversary_0", "adversary_1", "adversary_2", "agent_0"]


class Pol:
    pass


def test_shared_single_list_dict():
    p = Pol()
    for arg in (p, [p, Pol()], {"x": p}):
        m = M(arg, Env(), mode="shared")
        assert all(q is p for q in m.policy_map.values()) and list(m.policies) == ["shared"]
        assert m.get_shared_parameters() and m.get_policy_groups() == {"shared": Env.agents}


def test_independent():
    ps = [Pol() for _ in range(4)]
    m = M(ps, Env(), mode="independent")
    assert [m.policy_map[a] for a in Env.agents] == ps and not m.get_shared_parameters()
    assert m.get_policy_groups() == {f"group_{i}": [a] for i, a in enumerate(Env.agents)}
    with pytest.raises(ValueError, match="Independent mode requires list or dict"):
        M(Pol(), Env(), mode="independent")
    with pytest.raises(ValueError, match="must match number of agents"):
        M(ps[:2], Env(), mode="independent")
    with pytest.raises(ValueError, match="Missing policies"):
        M({"agent_0": ps[0]}, Env(), mode="independent")


def test_grouped():
    adv, good = Pol(), Pol()
    groups = {"adv": ["adversary_0", "adversary_1", "adversary_2"], "good": ["agent_0"]}
    m = M({"adv": adv, "good": good}, Env(), mode="grouped", agent_groups=groups)
    assert m.policy_map["adversary_1"] is adv and m.policy_map["agent_0"] is good
    assert m.policies == {"adv": adv, "good": good} and m.get_shared_parameters()
    assert m.get_policy_groups() == {"group_0": groups["adv"], "group_1": groups["good"]}
    with pytest.raises(ValueError, match="agent_groups"):
        M({"adv": adv}, Env(), mode="grouped")
    with pytest.raises(ValueError, match="dict of policies mapped to group names"):
        M(adv, Env(), mode="grouped", agent_groups=groups)
    with pytest.raises(ValueError, match="No policy found for group good"):
        M({"adv": adv}, Env(), mode="grouped", agent_groups=groups)
    with pytest.raises(ValueError, match="not assigned to any group"):
        M({"adv": adv}, Env(), mode="grouped", agent_groups={"adv": groups["adv"]})


def test_custom():
    adv, good = Pol(), Pol()
    fn = lambda a: "adv" if a.startswith("adversary") else "good"  # noqa: E731
    m = M({"adv": adv, "good": good}, Env(), mode="custom", policy_mapping_fn=fn)
    assert m.policy_map["adversary_2"] is adv and m.policy_map["agent_0"] is good and len(m.policies) == 2
    with pytest.raises(ValueError, match="policy_mapping_fn"):
        M({"adv": adv}, Env(), mode="custom")
    with pytest.raises(ValueError, match="Policy good not found"):
        M({"adv": adv}, Env(), mode="custom", policy_mapping_fn=fn)
