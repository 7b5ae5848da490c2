"""Golden fixture `ppo_gauss.npz`: the UNMODIFIED reference PPO with a continuous (diagonal Gaussian) actor --
ContinuousActorProbabilistic(conditioned_sigma=False; unbounded=True, and unbounded=False with max_action 1 and 2) + ContinuousCritic on a shared Net trunk,
dist_fn = Independent(Normal(mu, sigma), 1), as test/continuous/test_ppo.py:85-124 builds it -- on synthetic rollouts with
float actions.  Same recording points as make_golden_more.gen_ppo (PPO._preprocess_batch outputs, per-minibatch loss terms,
parameters after the update, sigma_param included).  Run in the build container:  cd tests/golden && python make_golden_gauss.py"""
from __future__ import annotations

import numpy as np

from make_golden import save  # noqa: E402  (also sets sys.path for the reference + shims)

CASES = [
    # R rows, T steps, O, H, action dimensions, batch_size, repeat, PPO kwargs
    dict(R=16, T=20, O=18, H=64, A=2, bs=80, repeat=2, kw=dict(max_grad_norm=0.5, lr=1e-3)),
    dict(R=10, T=15, O=11, H=64, A=3, bs=50, repeat=3,
         kw=dict(value_clip=True, dual_clip=3.0, advantage_normalization=False, ent_coef=0.02, vf_coef=0.25, lr=1e-3)),
    dict(R=8, T=12, O=24, H=128, A=5, bs=32, repeat=2, kw=dict(max_grad_norm=1.0, eps_clip=0.1, lr=3e-4)),
    # unbounded=False (the reference's default): mu = max_action * tanh(head output)
    dict(R=12, T=16, O=18, H=64, A=2, bs=64, repeat=2, bound=1.0, kw=dict(max_grad_norm=0.5, lr=1e-3)),
    dict(R=10, T=14, O=11, H=64, A=3, bs=48, repeat=3, bound=2.0, kw=dict(value_clip=True, dual_clip=2.0, ent_coef=0.02, lr=1e-3)),
    # conditioned_sigma=True: sigma = exp(clamp(Linear(features), -20, 2)); case 6 with the bounded mean as well
    dict(R=12, T=16, O=18, H=64, A=2, bs=64, repeat=2, cond=True, kw=dict(max_grad_norm=0.5, lr=1e-3)),
    dict(R=10, T=14, O=11, H=64, A=3, bs=48, repeat=3, cond=True, bound=1.5, kw=dict(value_clip=True, ent_coef=0.03, lr=1e-3)),
    # A2C (a2c.py:247-288) over the same continuous actor: the plain policy-gradient loss with the Gaussian log-density
    dict(R=12, T=16, O=18, H=64, A=2, bs=64, repeat=2, algo="a2c", kw=dict(max_grad_norm=0.5, vf_coef=0.25, ent_coef=0.02, lr=1e-3)),
]


def _ref(O: int, H: int, A: int, seed: int, bound: float = 0.0, cond: bool = False, algo: str = "ppo", **kw):
    import gymnasium as gym
    import torch
    from tianshou.algorithm import PPO
    from tianshou.algorithm.modelfree.reinforce import ProbabilisticActorPolicy
    from tianshou.algorithm.optim import AdamOptimizerFactory
    from tianshou.utils.net.common import ActorCritic, Net
    from tianshou.utils.net.continuous import ContinuousActorProbabilistic, ContinuousCritic
    from torch.distributions import Independent, Normal

    torch.manual_seed(seed)
    net = Net(state_shape=(O,), hidden_sizes=[H, H])
    actor = (ContinuousActorProbabilistic(preprocess_net=net, action_shape=(A,), unbounded=False, max_action=bound, conditioned_sigma=cond)
             if bound > 0 else ContinuousActorProbabilistic(preprocess_net=net, action_shape=(A,), unbounded=True, conditioned_sigma=cond))
    critic = ContinuousCritic(preprocess_net=net)
    ac = ActorCritic(actor, critic)
    for m in ac.modules():  # test/continuous/test_ppo.py:95-99 (orthogonal weights), biases non-zero so that their paths are exercised
        if isinstance(m, torch.nn.Linear):
            torch.nn.init.orthogonal_(m.weight)
            torch.nn.init.normal_(m.bias, std=0.1)
    with torch.no_grad():
        if cond:  # a wide spread of log sigma over the rows: part of them beyond the clamp at 2 (no gradient there)
            for m in actor.sigma.modules():
                if isinstance(m, torch.nn.Linear):
                    m.weight.mul_(3.0)
                    m.bias.add_(1.3)
        else:
            actor.sigma_param.copy_(torch.linspace(-0.5, 0.3, A).view(A, 1))  # distinct scales per action dimension

    def dist(loc_scale):  # test_ppo.py:104-106
        loc, scale = loc_scale
        return Independent(Normal(loc, scale), 1)

    space = gym.spaces.Box(low=-2.0, high=2.0, shape=(A,))
    policy = ProbabilisticActorPolicy(actor=actor, dist_fn=dist, action_space=space)
    lr = kw.pop("lr", 3e-4)
    if algo == "a2c":
        from tianshou.algorithm import A2C

        return A2C(policy=policy, critic=critic, optim=AdamOptimizerFactory(lr=lr), **kw), ac
    alg = PPO(policy=policy, critic=critic, optim=AdamOptimizerFactory(lr=lr), **kw)
    return alg, ac


def _flat(ac) -> dict[str, np.ndarray]:
    sd = {k: v.detach().cpu().numpy().copy() for k, v in ac.state_dict().items()}
    pick = lambda suffix, root: next(v for k, v in sd.items() if k.startswith(root) and k.endswith(suffix))  # noqa: E731
    return {
        "W1": pick("model.model.0.weight", "actor.preprocess"), "b1": pick("model.model.0.bias", "actor.preprocess"),
        "W2": pick("model.model.2.weight", "actor.preprocess"), "b2": pick("model.model.2.bias", "actor.preprocess"),
        "Wa": pick("mu.model.0.weight", "actor"), "ba": pick("mu.model.0.bias", "actor"),
        "Wv": pick("last.model.0.weight", "critic"), "bv": pick("last.model.0.bias", "critic"),
        **({"Ws": pick("sigma.model.0.weight", "actor"), "bs": pick("sigma.model.0.bias", "actor")} if any(k.startswith("actor.sigma.") for k in sd)
           else {"log_sigma": pick("sigma_param", "actor").reshape(-1)}),
    }


def gen() -> None:
    import torch
    from tianshou.data import Batch, VectorReplayBuffer
    from tianshou.data.stats import SequenceSummaryStats
    from tianshou.utils.torch_utils import policy_within_training_step, torch_train_mode

    out: dict[str, np.ndarray] = {}
    for ci, c in enumerate(CASES):
        R, T, O, H, A = c["R"], c["T"], c["O"], c["H"], c["A"]
        rng = np.random.default_rng(300 + ci)
        alg, ac = _ref(O, H, A, seed=40 + ci, bound=c.get("bound", 0.0), cond=c.get("cond", False), algo=c.get("algo", "ppo"), **dict(c["kw"]))
        pre = f"p{ci}_"
        for k, v in _flat(ac).items():
            out[pre + k] = v
        buf = VectorReplayBuffer(R * T, R)
        obs = rng.normal(size=(R, O)).astype(np.float32)
        store = {k: [] for k in ("obs", "act", "rew", "term", "trunc", "obs_next")}
        for t in range(T):
            act = rng.normal(size=(R, A)).astype(np.float32)  # stored (unmapped) actions, as the collector keeps them
            rew = rng.normal(size=R).astype(np.float32).astype(np.float64)
            trunc = np.full(R, (t + 1) % 5 == 0) & (rng.random(R) < 0.9)
            term = (rng.random(R) < 0.05) & ~trunc
            obs_next = (obs + rng.normal(size=(R, O)) * 0.3).astype(np.float32)
            buf.add(Batch(obs=obs, act=act, rew=rew, terminated=term, truncated=trunc, obs_next=obs_next, info={}))
            for k, v in zip(store, (obs, act, rew, term, trunc, obs_next)):
                store[k].append(np.array(v))
            done = term | trunc
            obs = np.where(done[:, None], rng.normal(size=(R, O)).astype(np.float32), obs_next)
        for k, v in store.items():
            out[pre + k] = np.stack(v)
        with torch.no_grad():
            x = torch.from_numpy(store["obs"][0])
            (mu, sigma), _ = alg.policy.actor(x)
            out[pre + "fwd_mu"] = mu.numpy().copy()
            out[pre + "fwd_sigma"] = sigma.numpy().copy()
            out[pre + "fwd_value"] = alg.critic(x).flatten().numpy().copy()
            if c.get("cond", False):  # how much of the batch sits beyond the clamp (those rows pass no gradient to the sigma head)
                feats, _ = alg.policy.actor.preprocess(torch.from_numpy(np.concatenate(store["obs"])))
                raw = alg.policy.actor.sigma(feats)
                out[pre + "ls_clamped_frac"] = np.array(float(((raw > 2.0) | (raw < -20.0)).float().mean()))
        seqs: list[list[float]] = []
        orig = SequenceSummaryStats.from_sequence.__func__

        def spy(cls, seq, _orig=orig, _seqs=seqs):
            _seqs.append([float(x) for x in seq])
            return _orig(cls, seq)

        SequenceSummaryStats.from_sequence = classmethod(spy)
        try:
            np.random.seed(70 + ci)
            with policy_within_training_step(alg.policy):
                batch, indices = buf.sample(0)
                batch = alg._preprocess_batch(batch, buf, indices)
                out[pre + "indices"] = np.asarray(indices)
                for k in ("v_s", "returns", "adv", "logp_old"):
                    out[pre + k] = batch[k].detach().cpu().numpy().copy() if k in batch else np.zeros(len(indices), np.float32)  # (A2C: no logp_old)
                with torch_train_mode(alg):
                    stats = alg._update_with_batch(batch, c["bs"], c["repeat"])
        finally:
            SequenceSummaryStats.from_sequence = classmethod(orig)
        out[pre + "loss_seq"], out[pre + "clip_seq"], out[pre + "vf_seq"], out[pre + "ent_seq"] = (np.array(s) for s in seqs[:4])
        out[pre + "gradient_steps"] = np.array(stats.gradient_steps)
        for k, v in _flat(ac).items():
            out[pre + "after_" + k] = v
        kw = c["kw"]
        out[pre + "cfg"] = np.array([R, T, O, H, A, c["bs"], c["repeat"], 0, 70 + ci], np.int64)
        out[pre + "mu_bound"] = np.array(c.get("bound", 0.0))
        out[pre + "cond_sigma"] = np.array(int(c.get("cond", False)))
        out[pre + "a2c"] = np.array(int(c.get("algo", "ppo") == "a2c"))
        out[pre + "hyper"] = np.array([kw.get("eps_clip", 0.2), kw.get("dual_clip") or 0.0, kw.get("vf_coef", 0.5), kw.get("ent_coef", 0.01),
                                       float(kw.get("value_clip", False)), float(kw.get("advantage_normalization", True)),
                                       kw.get("max_grad_norm") or 0.0, kw.get("lr", 3e-4), 0.99, 0.95])
    out["n_cases"] = np.array(len(CASES))
    save("ppo_gauss.npz", **out)


if __name__ == "__main__":
    gen()
