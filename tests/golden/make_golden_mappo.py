"""Golden fixtures for BASELINE.json configs[2] shapes (A = 10 agents, obs 60, hidden 128) from the UNMODIFIED reference
(imported through oracle/shims).  Run in the build container:  cd tests/golden && python make_golden_mappo.py

`mappo.npz`: the reference PPO (modelfree/ppo.py:146-224: _preprocess_batch with critic(obs) / critic(obs_next) / GAE / logp_old,
_update_with_batch, Optimizer.step with one gradient norm over ActorCritic(actor, critic)) run over the reference's own CTDE
networks (multiagent/ctde.py:346-414): DecentralizedActor on the local observation of each agent row, CentralizedCritic on the
world's concatenated global state (GlobalStateConstructor("concatenate"), ctde.py:291-294), the row's value = the critic head of
its agent.  Rows are row = world * A + agent; observations are stored as Batch(obs=local, global_obs=concat, agent=index).
batch_size = the whole batch, so an epoch is one minibatch (its order does not matter) and `repeat` optimizer steps are recorded.

`ctde_c3.npz`: CTDEPolicy.learn (ctde.py:121-199) at the same shapes (make_golden_ctde.py's recorder with A = 10, O = 60, H = 128).
"""
from __future__ import annotations

import numpy as np

from make_golden import save  # noqa: E402  (also sets sys.path for the reference + shims)

CASES = [
    dict(E=16, A=10, T=12, O=60, H=128, nA=5, repeat=3, kw=dict(max_grad_norm=0.5, lr=1e-3)),
    dict(E=7, A=10, T=9, O=60, H=128, nA=5, repeat=2, kw=dict(value_clip=True, dual_clip=3.0, advantage_normalization=False, lr=5e-4, ent_coef=0.02)),
    dict(E=20, A=3, T=10, O=18, H=128, nA=5, repeat=2, kw=dict(max_grad_norm=1.0, eps_clip=0.1, vf_coef=0.25, lr=1e-3)),
]


def gen_mappo() -> dict:
    import gymnasium as gym
    import torch
    from tianshou.algorithm import PPO
    from tianshou.algorithm.modelfree.reinforce import DiscreteActorPolicy
    from tianshou.algorithm.multiagent.ctde import CentralizedCritic, DecentralizedActor, GlobalStateConstructor
    from tianshou.algorithm.optim import AdamOptimizerFactory
    from tianshou.data import Batch, VectorReplayBuffer
    from tianshou.data.stats import SequenceSummaryStats
    from tianshou.utils.torch_utils import policy_within_training_step, torch_train_mode

    class ActorOnLocalObs(torch.nn.Module):
        def __init__(self, net):
            super().__init__()
            self.net = net

        def forward(self, obs, state=None, info=None):
            return self.net(torch.as_tensor(np.asarray(obs.obs), dtype=torch.float32), state)

    class CriticOnGlobalState(torch.nn.Module):
        def __init__(self, net):
            super().__init__()
            self.net = net

        def forward(self, obs, **kw):
            g = torch.as_tensor(np.asarray(obs.global_obs), dtype=torch.float32)
            a = torch.as_tensor(np.asarray(obs.agent), dtype=torch.int64)
            return self.net(g).gather(1, a[:, None])

    out: dict = {"n_cases": np.array(len(CASES))}
    gsc = GlobalStateConstructor(mode="concatenate")
    for ci, c in enumerate(CASES):
        E, A, T, O, H, nA = c["E"], c["A"], c["T"], c["O"], c["H"], c["nA"]
        R = E * A
        torch.manual_seed(300 + ci)
        rng = np.random.default_rng(400 + ci)
        actor, critic = DecentralizedActor(O, nA, H), CentralizedCritic(A * O, A, H)
        kw = dict(c["kw"])
        lr = kw.pop("lr")
        policy = DiscreteActorPolicy(actor=ActorOnLocalObs(actor), action_space=gym.spaces.Discrete(nA))
        alg = PPO(policy=policy, critic=CriticOnGlobalState(critic), optim=AdamOptimizerFactory(lr=lr), **kw)
        pre = f"m{ci}_"
        for name, mod in (("actor", actor), ("critic", critic)):
            for k, v in mod.state_dict().items():
                out[pre + f"init_{name}_{k}"] = v.detach().numpy().copy()

        def wrap(local: np.ndarray) -> Batch:
            per_agent = {f"agent_{a}": torch.from_numpy(local.reshape(E, A, O)[:, a]) for a in range(A)}
            g = gsc.build(per_agent).numpy()  # [E, A * O]
            return Batch(obs=local, global_obs=np.repeat(g, A, axis=0), agent=np.tile(np.arange(A), E))

        buf = VectorReplayBuffer(R * T, R)
        obs = rng.normal(size=(R, O)).astype(np.float32)
        store = {k: [] for k in ("obs", "act", "rew", "term", "trunc", "obs_next")}
        for t in range(T):
            act = rng.integers(0, nA, R)
            rew = rng.normal(size=R).astype(np.float32).astype(np.float64)
            trunc_w = np.full(E, (t + 1) % 4 == 0) & (rng.random(E) < 0.8)   # whole worlds end together (MPE)
            term_w = (rng.random(E) < 0.06) & ~trunc_w
            trunc, term = np.repeat(trunc_w, A), np.repeat(term_w, A)
            obs_next = (obs + rng.normal(size=(R, O)) * 0.3).astype(np.float32)
            buf.add(Batch(obs=wrap(obs), act=act, rew=rew, terminated=term, truncated=trunc, obs_next=wrap(obs_next), info={}))
            for k, v in zip(store, (obs, act, rew, term, trunc, obs_next)):
                store[k].append(np.array(v))
            done = term | trunc
            obs = np.where(done[:, None], rng.normal(size=(R, O)).astype(np.float32), obs_next)
        for k, v in store.items():
            out[pre + k] = np.stack(v)  # [T][R]...
        with torch.no_grad():
            out[pre + "fwd_logits"] = actor(torch.from_numpy(store["obs"][0]))[0].numpy().copy()
            out[pre + "fwd_values"] = critic(torch.from_numpy(wrap(store["obs"][0]).global_obs[::A].copy())).numpy().copy()  # [E][A]
        seqs: list[list[float]] = []
        orig = SequenceSummaryStats.from_sequence.__func__

        def spy(cls, seq, _orig=orig, _seqs=seqs):
            _seqs.append([float(x) for x in seq])
            return _orig(cls, seq)

        SequenceSummaryStats.from_sequence = classmethod(spy)
        try:
            np.random.seed(11 + ci)
            with policy_within_training_step(alg.policy):
                batch, indices = buf.sample(0)
                batch = alg._preprocess_batch(batch, buf, indices)
                out[pre + "indices"] = np.asarray(indices)
                for k in ("v_s", "returns", "adv", "logp_old"):
                    out[pre + k] = batch[k].detach().cpu().numpy().copy()
                with torch_train_mode(alg):
                    stats = alg._update_with_batch(batch, R * T, c["repeat"])
        finally:
            SequenceSummaryStats.from_sequence = classmethod(orig)
        out[pre + "loss_seq"], out[pre + "clip_seq"], out[pre + "vf_seq"], out[pre + "ent_seq"] = (np.array(s) for s in seqs[:4])
        out[pre + "gradient_steps"] = np.array(stats.gradient_steps)
        for name, mod in (("actor", actor), ("critic", critic)):
            for k, v in mod.state_dict().items():
                out[pre + f"after_{name}_{k}"] = v.detach().numpy().copy()
        k2 = c["kw"]
        out[pre + "cfg"] = np.array([E, A, T, O, H, nA, c["repeat"]], np.int64)
        out[pre + "hyper"] = np.array([k2.get("eps_clip", 0.2), k2.get("dual_clip") or 0.0, k2.get("vf_coef", 0.5), k2.get("ent_coef", 0.01),
                                       float(k2.get("value_clip", False)), float(k2.get("advantage_normalization", True)),
                                       k2.get("max_grad_norm") or 0.0, k2["lr"], 0.99, 0.95])
    return out


def gen_ctde_c3() -> dict:
    """make_golden_ctde.gen_ctde's recorder at the configs[2] shapes."""
    import make_golden_ctde as m

    return m.gen_ctde([dict(A=10, O=60, H=128, nA=5, n_env=26, lr_a=1e-3, lr_c=2e-3, gamma=0.99, use_global=True, steps=3)])


if __name__ == "__main__":
    save("mappo.npz", **gen_mappo())
    save("ctde_c3.npz", **gen_ctde_c3())
