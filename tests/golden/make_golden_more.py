"""More golden fixtures from the UNMODIFIED reference (imported through oracle/shims):
`ppo` -- PPO._preprocess_batch + _update_with_batch + Optimizer.step on synthetic rollouts.
See make_golden.py for how to run.  Never runs on the GPU box."""
from __future__ import annotations

import numpy as np

from make_golden import save  # noqa: E402  (also sets sys.path for the reference + shims)


def _ref_ppo(O: int, H: int, nA: int, seed: int, softmax_probs: bool, **ppo_kw):
    import gymnasium as gym
    import torch
    from tianshou.algorithm import PPO
    from tianshou.algorithm.modelfree.reinforce import DiscreteActorPolicy
    from tianshou.algorithm.optim import AdamOptimizerFactory
    from tianshou.utils.net.common import ActorCritic, Net
    from tianshou.utils.net.discrete import DiscreteActor, DiscreteCritic

    torch.manual_seed(seed)
    net = Net(state_shape=(O,), hidden_sizes=[H, H])
    actor = DiscreteActor(preprocess_net=net, action_shape=nA, softmax_output=softmax_probs)
    critic = DiscreteCritic(preprocess_net=net)
    ac = ActorCritic(actor, critic)
    for m in ac.modules():  # test/discrete/test_ppo_discrete.py:101-105
        if isinstance(m, torch.nn.Linear):
            torch.nn.init.orthogonal_(m.weight)
            torch.nn.init.normal_(m.bias, std=0.1)  # non-zero so that bias paths are exercised
    if softmax_probs:
        policy = DiscreteActorPolicy(actor=actor, dist_fn=torch.distributions.Categorical, action_space=gym.spaces.Discrete(nA))
    else:
        policy = DiscreteActorPolicy(actor=actor, action_space=gym.spaces.Discrete(nA))
    lr = ppo_kw.pop("lr", 3e-4)
    if ppo_kw.pop("algo", "ppo") == "a2c":
        from tianshou.algorithm import A2C

        return A2C(policy=policy, critic=critic, optim=AdamOptimizerFactory(lr=lr), **ppo_kw), ac
    alg = PPO(policy=policy, critic=critic, optim=AdamOptimizerFactory(lr=lr), **ppo_kw)
    return alg, ac


def _flat_params(ac) -> dict[str, np.ndarray]:
    sd = {k: v.detach().cpu().numpy().copy() for k, v in ac.state_dict().items()}
    # shared trunk appears under actor.preprocess.* and critic.preprocess.* (same tensors)
    pick = lambda suffix, root: next(v for k, v in sd.items() if k.startswith(root) and k.endswith(suffix))  # noqa: E731
    return {
        "W1": pick("model.model.0.weight", "actor.preprocess"), "b1": pick("model.model.0.bias", "actor.preprocess"),
        "W2": pick("model.model.2.weight", "actor.preprocess"), "b2": pick("model.model.2.bias", "actor.preprocess"),
        "Wa": pick("last.model.0.weight", "actor"), "ba": pick("last.model.0.bias", "actor"),
        "Wv": pick("last.model.0.weight", "critic"), "bv": pick("last.model.0.bias", "critic"),
    }


PPO_CASES = [
    # R rows, T steps, O, H, nA, batch_size, repeat, probs-parameterisation, PPO kwargs
    dict(R=24, T=16, O=18, H=64, nA=5, bs=128, repeat=2, probs=False, kw=dict()),
    dict(R=12, T=25, O=18, H=64, nA=5, bs=64, repeat=3, probs=False,
         kw=dict(max_grad_norm=0.5, value_clip=True, dual_clip=3.0, advantage_normalization=False, lr=1e-3)),
    dict(R=8, T=13, O=16, H=64, nA=5, bs=40, repeat=2, probs=True, kw=dict(max_grad_norm=0.5, ent_coef=0.05, vf_coef=0.25)),
    dict(R=6, T=12, O=60, H=128, nA=5, bs=36, repeat=2, probs=False, kw=dict(max_grad_norm=1.0, eps_clip=0.1)),
]
# recompute_advantage (ppo.py:174-178): returns / advantages / v_s re-evaluated with the current parameters before every epoch after the first
PPO_RECOMPUTE_CASES = [
    dict(R=16, T=20, O=18, H=64, nA=5, bs=80, repeat=3, probs=False, kw=dict(recompute_advantage=True, max_grad_norm=0.5, lr=1e-3)),
    dict(R=10, T=15, O=18, H=64, nA=5, bs=50, repeat=2, probs=False,
         kw=dict(recompute_advantage=True, value_clip=True, advantage_normalization=False, lr=1e-3)),
]


# return_scaling (a2c.py:132-146): one un-recorded _preprocess_batch first, so that ret_rms holds non-trivial statistics
PPO_RETSCALE_CASES = [
    dict(R=16, T=20, O=18, H=64, nA=5, bs=80, repeat=2, probs=False, warm_rms=True, kw=dict(return_scaling=True, max_grad_norm=0.5, lr=1e-3)),
    dict(R=10, T=15, O=18, H=64, nA=5, bs=50, repeat=2, probs=False, warm_rms=True,
         kw=dict(return_scaling=True, recompute_advantage=True, value_clip=True, lr=1e-3)),
]


# A2C._update_with_batch (a2c.py:247-288)
A2C_CASES = [
    dict(R=16, T=20, O=18, H=64, nA=5, bs=80, repeat=2, probs=False, kw=dict(algo="a2c", max_grad_norm=0.5, lr=1e-3)),
    dict(R=10, T=15, O=18, H=64, nA=5, bs=50, repeat=1, probs=False, warm_rms=True,
         kw=dict(algo="a2c", return_scaling=True, vf_coef=0.25, ent_coef=0.02, lr=1e-3)),
]


def gen_ppo(cases=None, fname: str = "ppo.npz") -> None:
    import torch
    from tianshou.data import Batch, VectorReplayBuffer
    from tianshou.data.stats import SequenceSummaryStats
    from tianshou.utils.torch_utils import policy_within_training_step, torch_train_mode

    out: dict[str, np.ndarray] = {}
    cases = PPO_CASES if cases is None else cases
    for ci, c in enumerate(cases):
        R, T, O, H, nA = c["R"], c["T"], c["O"], c["H"], c["nA"]
        rng = np.random.default_rng(100 + ci)
        alg, ac = _ref_ppo(O, H, nA, seed=ci, softmax_probs=c["probs"], **dict(c["kw"]))
        pre = f"p{ci}_"
        for k, v in _flat_params(ac).items():
            out[pre + k] = v
        buf = VectorReplayBuffer(R * T, R)
        obs = rng.normal(size=(R, O)).astype(np.float32)
        store = {k: [] for k in ("obs", "act", "rew", "term", "trunc", "obs_next")}
        for t in range(T):
            act = rng.integers(0, nA, R)
            rew = rng.normal(size=R).astype(np.float32).astype(np.float64)  # f32-representable rewards
            trunc = np.full(R, (t + 1) % 5 == 0) & (rng.random(R) < 0.9)
            term = (rng.random(R) < 0.05) & ~trunc
            obs_next = (obs + rng.normal(size=(R, O)) * 0.3).astype(np.float32)
            buf.add(Batch(obs=obs, act=act, rew=rew, terminated=term, truncated=trunc, obs_next=obs_next, info={}))
            for k, v in zip(store, (obs, act, rew, term, trunc, obs_next)):
                store[k].append(np.array(v))
            done = term | trunc
            obs = np.where(done[:, None], rng.normal(size=(R, O)).astype(np.float32), obs_next)
        for k, v in store.items():
            out[pre + k] = np.stack(v)  # [T][R]...
        # one deterministic forward with the INITIAL weights for the policy-forward check
        with torch.no_grad():
            x = torch.from_numpy(store["obs"][0])
            logits, _ = alg.policy.actor(x)
            out[pre + "fwd_actor_out"] = logits.numpy().copy()
            out[pre + "fwd_value"] = alg.critic(x).flatten().numpy().copy()
        seqs: list[list[float]] = []
        orig = SequenceSummaryStats.from_sequence.__func__

        def spy(cls, seq, _orig=orig, _seqs=seqs):
            _seqs.append([float(x) for x in seq])
            return _orig(cls, seq)

        SequenceSummaryStats.from_sequence = classmethod(spy)
        try:
            np.random.seed(7 + ci)
            with policy_within_training_step(alg.policy):
                if c.get("warm_rms"):
                    b0, i0 = buf.sample(0)
                    alg._preprocess_batch(b0, buf, i0)
                    out[pre + "rms_before"] = np.array([float(alg.ret_rms.mean), float(alg.ret_rms.var), float(alg.ret_rms.count)])
                batch, indices = buf.sample(0)
                batch = alg._preprocess_batch(batch, buf, indices)
                if c.get("warm_rms"):
                    out[pre + "rms_after_preprocess"] = np.array([float(alg.ret_rms.mean), float(alg.ret_rms.var), float(alg.ret_rms.count)])
                out[pre + "indices"] = np.asarray(indices)
                for k in ("v_s", "returns", "adv", "logp_old"):
                    out[pre + k] = batch[k].detach().cpu().numpy().copy() if k in batch else np.zeros(len(indices), np.float32)  # (A2C: no logp_old)
                with torch_train_mode(alg):
                    stats = alg._update_with_batch(batch, c["bs"], c["repeat"])
        finally:
            SequenceSummaryStats.from_sequence = classmethod(orig)
        out[pre + "loss_seq"], out[pre + "clip_seq"], out[pre + "vf_seq"], out[pre + "ent_seq"] = (np.array(s) for s in seqs[:4])
        out[pre + "gradient_steps"] = np.array(stats.gradient_steps)
        for k, v in _flat_params(ac).items():
            out[pre + "after_" + k] = v
        if c.get("warm_rms"):
            out[pre + "rms_after"] = np.array([float(alg.ret_rms.mean), float(alg.ret_rms.var), float(alg.ret_rms.count)])
        kw = c["kw"]
        out[pre + "cfg"] = np.array([R, T, O, H, nA, c["bs"], c["repeat"], int(c["probs"]), 7 + ci], np.int64)
        out[pre + "hyper"] = np.array([kw.get("eps_clip", 0.2), kw.get("dual_clip") or 0.0, kw.get("vf_coef", 0.5), kw.get("ent_coef", 0.01),
                                       float(kw.get("value_clip", False)), float(kw.get("advantage_normalization", True)),
                                       kw.get("max_grad_norm") or 0.0, kw.get("lr", 3e-4), 0.99, 0.95])
    out["n_cases"] = np.array(len(cases))
    save(fname, **out)


GENS = {"ppo": gen_ppo, "ppo_recompute": lambda: gen_ppo(PPO_RECOMPUTE_CASES, "ppo_recompute.npz"),
        "ppo_retscale": lambda: gen_ppo(PPO_RETSCALE_CASES, "ppo_retscale.npz"), "a2c": lambda: gen_ppo(A2C_CASES, "a2c.npz")}


def gen_reinforce() -> None:
    """Reinforce (reinforce.py:249-382) with an actor-only net; case 1 with return_standardization and a warmed ret_rms."""
    import gymnasium as gym
    import torch
    from tianshou.algorithm import Reinforce
    from tianshou.algorithm.modelfree.reinforce import DiscreteActorPolicy
    from tianshou.algorithm.optim import AdamOptimizerFactory
    from tianshou.data import Batch, VectorReplayBuffer
    from tianshou.data.stats import SequenceSummaryStats
    from tianshou.utils.net.common import Net
    from tianshou.utils.net.discrete import DiscreteActor
    from tianshou.utils.torch_utils import policy_within_training_step, torch_train_mode

    out: dict[str, np.ndarray] = {}
    cases = [dict(R=16, T=20, O=18, H=64, nA=5, bs=80, repeat=2, std=False, lr=1e-3, gamma=0.99),
             dict(R=10, T=15, O=18, H=64, nA=5, bs=50, repeat=2, std=True, lr=1e-3, gamma=0.95)]
    for ci, c in enumerate(cases):
        R, T, O, H, nA = c["R"], c["T"], c["O"], c["H"], c["nA"]
        rng = np.random.default_rng(300 + ci)
        torch.manual_seed(50 + ci)
        actor = DiscreteActor(preprocess_net=Net(state_shape=(O,), hidden_sizes=[H, H]), action_shape=nA, softmax_output=False)
        for m in actor.modules():
            if isinstance(m, torch.nn.Linear):
                torch.nn.init.orthogonal_(m.weight)
                torch.nn.init.normal_(m.bias, std=0.1)
        policy = DiscreteActorPolicy(actor=actor, action_space=gym.spaces.Discrete(nA))
        alg = Reinforce(policy=policy, gamma=c["gamma"], return_standardization=c["std"], optim=AdamOptimizerFactory(lr=c["lr"]))
        pre = f"p{ci}_"

        def params(prefix):
            sd = {k: v.detach().cpu().numpy().copy() for k, v in actor.state_dict().items()}
            pick = lambda suf: next(v for k, v in sd.items() if k.endswith(suf))  # noqa: E731
            for name, suf in (("W1", "model.model.0.weight"), ("b1", "model.model.0.bias"), ("W2", "model.model.2.weight"),
                              ("b2", "model.model.2.bias"), ("W3", "last.model.0.weight"), ("b3", "last.model.0.bias")):
                out[pre + prefix + name] = pick(suf)

        params("")
        buf = VectorReplayBuffer(R * T, R)
        obs = rng.normal(size=(R, O)).astype(np.float32)
        store = {k: [] for k in ("obs", "act", "rew", "term", "trunc", "obs_next")}
        for t in range(T):
            act = rng.integers(0, nA, R)
            rew = rng.normal(size=R).astype(np.float32).astype(np.float64)
            trunc = np.full(R, (t + 1) % 5 == 0) & (rng.random(R) < 0.9)
            term = (rng.random(R) < 0.05) & ~trunc
            obs_next = (obs + rng.normal(size=(R, O)) * 0.3).astype(np.float32)
            buf.add(Batch(obs=obs, act=act, rew=rew, terminated=term, truncated=trunc, obs_next=obs_next, info={}))
            for k, v in zip(store, (obs, act, rew, term, trunc, obs_next)):
                store[k].append(np.array(v))
            obs = np.where((term | trunc)[:, None], rng.normal(size=(R, O)).astype(np.float32), obs_next)
        for k, v in store.items():
            out[pre + k] = np.stack(v)
        seqs: list[list[float]] = []
        orig = SequenceSummaryStats.from_sequence.__func__

        def spy(cls, seq, _orig=orig, _seqs=seqs):
            _seqs.append([float(x) for x in seq])
            return _orig(cls, seq)

        SequenceSummaryStats.from_sequence = classmethod(spy)
        try:
            np.random.seed(17 + ci)
            with policy_within_training_step(alg.policy):
                rms = alg.discounted_return_computation.ret_rms
                if c["std"]:
                    b0, i0 = buf.sample(0)
                    alg._preprocess_batch(b0, buf, i0)
                out[pre + "rms_before"] = np.array([float(rms.mean), float(rms.var), float(rms.count)])
                batch, indices = buf.sample(0)
                batch = alg._preprocess_batch(batch, buf, indices)
                out[pre + "returns"] = np.asarray(batch.returns, np.float64).copy()
                out[pre + "rms_after"] = np.array([float(rms.mean), float(rms.var), float(rms.count)])
                with torch_train_mode(alg):
                    alg._update_with_batch(batch, c["bs"], c["repeat"])
        finally:
            SequenceSummaryStats.from_sequence = classmethod(orig)
        out[pre + "loss_seq"] = np.array(seqs[0])
        params("after_")
        out[pre + "cfg"] = np.array([R, T, O, H, nA, c["bs"], c["repeat"], int(c["std"]), 17 + ci], np.int64)
        out[pre + "hyper"] = np.array([c["lr"], c["gamma"]])
    out["n_cases"] = np.array(len(cases))
    save("reinforce.npz", **out)


GENS["reinforce"] = gen_reinforce
