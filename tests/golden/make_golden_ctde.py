"""Golden fixtures for the CTDE row (a16) from the UNMODIFIED reference (imported through oracle/shims):
CTDEPolicy.learn (one-step actor-critic with a centralized critic), GlobalStateConstructor.build (concatenate / mean),
DecentralizedActor / CentralizedCritic.  Run in the build container:  cd tests/golden && python make_golden_ctde.py"""
from __future__ import annotations

import numpy as np

from make_golden import save  # noqa: E402  (also sets sys.path for the reference + shims)


def gen_ctde(cases: list | None = None) -> dict:
    import gymnasium as gym
    import torch
    from tianshou.algorithm.multiagent.ctde import CTDEPolicy, CentralizedCritic, DecentralizedActor, GlobalStateConstructor
    from tianshou.data import Batch

    out: dict = {}
    cases = cases or [dict(A=3, O=18, H=64, nA=5, n_env=40, lr_a=1e-3, lr_c=2e-3, gamma=0.99, use_global=True, steps=3),
             dict(A=3, O=18, H=128, nA=5, n_env=24, lr_a=5e-4, lr_c=1e-3, gamma=0.95, use_global=True, steps=2),
             dict(A=2, O=10, H=64, nA=4, n_env=33, lr_a=1e-3, lr_c=1e-3, gamma=0.9, use_global=False, steps=2)]
    out["n_cases"] = np.array(len(cases))
    for ci, c in enumerate(cases):
        A, O, H, nA, E = c["A"], c["O"], c["H"], c["nA"], c["n_env"]
        torch.manual_seed(100 + ci)
        rng = np.random.default_rng(200 + ci)
        actor = DecentralizedActor(O, nA, H)
        gdim = A * O if c["use_global"] else O
        critic = CentralizedCritic(gdim, A, H)
        oa = torch.optim.Adam(actor.parameters(), lr=c["lr_a"])
        oc = torch.optim.Adam(critic.parameters(), lr=c["lr_c"])
        pol = CTDEPolicy(actor=actor, critic=critic, optim_actor=oa, optim_critic=oc, observation_space=gym.spaces.Box(-1, 1, (O,)),
                         action_space=gym.spaces.Discrete(nA), enable_global_info=c["use_global"], discount_factor=c["gamma"])
        pre = f"c{ci}_"
        for k, v in c.items():
            out[pre + k] = np.array(v)
        for name, mod in (("actor", actor), ("critic", critic)):
            for k, v in mod.state_dict().items():
                out[pre + f"init_{name}_{k}"] = v.detach().numpy().copy()
        n = E * A  # rows = env * A + agent
        obs = rng.normal(size=(n, O)).astype(np.float32)
        obs_next = rng.normal(size=(n, O)).astype(np.float32)
        act = rng.integers(0, nA, n)
        rew = rng.normal(size=n).astype(np.float32)
        term = rng.random(n) < 0.2
        gsc = GlobalStateConstructor(mode="concatenate")
        gmean = GlobalStateConstructor(mode="mean")
        per_agent = {f"agent_{a}": torch.from_numpy(obs.reshape(E, A, O)[:, a]) for a in range(A)}
        per_agent_next = {f"agent_{a}": torch.from_numpy(obs_next.reshape(E, A, O)[:, a]) for a in range(A)}
        g_env = gsc.build(per_agent).numpy()            # [E, A*O]
        g_env_next = gsc.build(per_agent_next).numpy()
        out[pre + "global_concat"] = g_env
        out[pre + "global_mean"] = gmean.build(per_agent).numpy()
        g_rows, g_rows_next = np.repeat(g_env, A, axis=0), np.repeat(g_env_next, A, axis=0)  # every agent row sees its world's state
        for k, v in (("obs", obs), ("obs_next", obs_next), ("act", act), ("rew", rew), ("term", term)):
            out[pre + k] = v
        # forward (decentralized execution): the reference returns the actor's raw output as `act`
        fwd = pol.forward(Batch(obs=obs))
        out[pre + "fwd_act"] = fwd.act.detach().numpy().copy()
        losses = []
        for _ in range(c["steps"]):
            b = Batch(obs=obs, act=act, rew=rew, obs_next=obs_next, terminated=term)
            if c["use_global"]:
                b.global_obs, b.global_obs_next = g_rows, g_rows_next
            r = pol.learn(b)
            losses.append([r["actor_loss"], r["critic_loss"]])
        out[pre + "losses"] = np.array(losses)
        for name, mod in (("actor", actor), ("critic", critic)):
            for k, v in mod.state_dict().items():
                out[pre + f"after_{name}_{k}"] = v.detach().numpy().copy()
    return out


if __name__ == "__main__":
    save("ctde.npz", **gen_ctde())
    print("wrote ctde.npz")
