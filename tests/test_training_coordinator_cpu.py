"""SimultaneousTrainer routing (reference training_coordinator.py:292-339) with stand-in policies: no device needed."""
import types

from tianshou_marl_b200.algorithm.training_coordinator import SimultaneousTrainer


class Pol:
    def __init__(self, name, log):
        self.name, self.log = name, log

    def train(self):
        self.log.append((self.name, "train"))

    def learn(self, b):
        self.log.append((self.name, "learn", b.tag, getattr(b, "global_obs", None), getattr(b, "global_obs_next", None)))
        return {"loss": len(self.log)}


class B(dict):
    __getattr__ = dict.get


def _batch(agents, with_global=True):
    b = B({a: types.SimpleNamespace(tag=a) for a in agents})
    if with_global:
        b["global_obs"], b["global_obs_next"] = "G", "G2"
    return b


def test_independent_mode_frequencies_and_global_state():
    log = []
    pm = types.SimpleNamespace(mode="independent", agents=["a0", "a1", "a2"], policies={a: Pol(a, log) for a in ("a0", "a1", "a2")})
    tr = SimultaneousTrainer(pm, agent_train_freq={"a1": 2})
    out1 = tr.train_step(_batch(["a0", "a1"]))            # a2 absent from the batch, a1 trains every 2nd step
    assert set(out1) == {"a0"} and tr.step_count == 1
    out2 = tr.train_step(_batch(["a0", "a1"]))
    assert set(out2) == {"a0", "a1"}
    learns = [e for e in log if e[1] == "learn"]
    assert learns == [("a0", "learn", "a0", "G", "G2"), ("a0", "learn", "a0", "G", "G2"), ("a1", "learn", "a1", "G", "G2")]
    assert log[0] == ("a0", "train")                      # policy.train() before learn (training_coordinator.py:335-337)


def test_shared_mode_trains_the_shared_policy_for_every_agent():
    log = []
    shared = Pol("shared", log)
    pm = types.SimpleNamespace(mode="shared", agents=["a0", "a1"], policies={"shared": shared})
    out = SimultaneousTrainer(pm).train_step(_batch(["a0", "a1"], with_global=False))
    assert set(out) == {"a0", "a1"}
    assert [e for e in log if e[1] == "learn"] == [("shared", "learn", "a0", None, None), ("shared", "learn", "a1", None, None)]
