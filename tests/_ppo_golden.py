"""Shared loader for tests/golden/ppo.npz (recorded from the reference PPO)."""
import numpy as np


def load_case(g, ci):
    pre = f"p{ci}_"
    R, T, O, H, nA, bs, repeat, probs, seed = (int(v) for v in g[pre + "cfg"])
    eps_clip, dual_clip, vf_coef, ent_coef, value_clip, adv_norm, max_norm, lr, gamma, lam = (float(v) for v in g[pre + "hyper"])
    c = dict(R=R, T=T, O=O, H=H, nA=nA, bs=bs, repeat=repeat, probs=bool(probs), seed=seed, lr=lr, gamma=gamma, lam=lam,
             max_norm=max_norm,
             cfg=dict(eps_clip=eps_clip, dual_clip=dual_clip, vf_coef=vf_coef, ent_coef=ent_coef, value_clip=bool(value_clip),
                      adv_norm=bool(adv_norm)))
    for k in ("W1", "b1", "W2", "b2", "Wa", "ba", "Wv", "bv"):
        c[k] = g[pre + k]
        c["after_" + k] = g[pre + "after_" + k]
    for k in ("obs", "act", "rew", "term", "trunc", "obs_next", "indices", "v_s", "returns", "adv", "logp_old",
              "loss_seq", "clip_seq", "vf_seq", "ent_seq", "fwd_actor_out", "fwd_value"):
        c[k] = g[pre + k]
    c["W3"] = np.concatenate([c["Wa"], c["Wv"]], 0)
    c["b3"] = np.concatenate([c["ba"], c["bv"]], 0)
    c["after_W3"] = np.concatenate([c["after_Wa"], c["after_Wv"]], 0)
    c["after_b3"] = np.concatenate([c["after_ba"], c["after_bv"]], 0)
    B = R * T
    # buffer.sample(0) order of a fresh lock-step VectorReplayBuffer(R*T, R): env-major, index = r*T + t
    assert np.array_equal(c["indices"], np.arange(B))
    em = lambda a: np.swapaxes(a, 0, 1).reshape(B, *a.shape[2:])  # noqa: E731  [T][R] -> env-major flat
    c["flat"] = {k: em(c[k]) for k in ("obs", "act", "rew", "term", "trunc", "obs_next")}
    # unfinished_index: last written slot of each row unless done
    last = np.arange(R) * T + T - 1
    done = c["flat"]["term"] | c["flat"]["trunc"]
    c["unfinished"] = last[~done[last]]
    # np.random.permutation(B) per epoch with the recorded seed (batch.py:1219)
    rs = np.random.RandomState(seed)
    c["orders"] = [rs.permutation(B) for _ in range(repeat)]
    return c


def load_reinforce(g, ci):
    pre = f"p{ci}_"
    R, T, O, H, nA, bs, repeat, std, seed = (int(v) for v in g[pre + "cfg"])
    lr, gamma = (float(v) for v in g[pre + "hyper"])
    c = dict(R=R, T=T, O=O, H=H, nA=nA, bs=bs, repeat=repeat, std=bool(std), seed=seed, lr=lr, gamma=gamma)
    for k in ("W1", "b1", "W2", "b2", "W3", "b3", "after_W1", "after_b1", "after_W2", "after_b2", "after_W3", "after_b3", "obs", "act", "rew",
              "term", "trunc", "obs_next", "returns", "loss_seq", "rms_before", "rms_after"):
        c[k] = g[pre + k]
    B = R * T
    em = lambda a: np.swapaxes(a, 0, 1).reshape(B, *a.shape[2:])  # noqa: E731
    c["flat"] = {k: em(c[k]) for k in ("obs", "act", "rew", "term", "trunc")}
    last = np.arange(R) * T + T - 1
    done = c["flat"]["term"] | c["flat"]["trunc"]
    c["unfinished"] = last[~done[last]]
    rs = np.random.RandomState(seed)
    c["orders"] = [rs.permutation(B) for _ in range(repeat)]
    return c
