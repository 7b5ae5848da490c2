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


def load_gauss(g, ci):
    """tests/golden/ppo_gauss.npz (make_golden_gauss.py): load_case with float actions [T][R][A] and the log_sigma parameter."""
    pre = f"p{ci}_"
    R, T, O, H, A, bs, repeat, _, seed = (int(v) for v in g[pre + "cfg"])
    eps_clip, dual_clip, vf_coef, ent_coef, value_clip, adv_norm, max_norm, lr, gamma, lam = (float(v) for v in g[pre + "hyper"])
    c = dict(R=R, T=T, O=O, H=H, nA=A, bs=bs, repeat=repeat, probs=False, seed=seed, lr=lr, gamma=gamma, lam=lam, max_norm=max_norm,
             cfg=dict(eps_clip=eps_clip, dual_clip=dual_clip, vf_coef=vf_coef, ent_coef=ent_coef, value_clip=bool(value_clip),
                      adv_norm=bool(adv_norm)))
    c["mu_bound"] = float(g[pre + "mu_bound"]) if pre + "mu_bound" in g.files else 0.0  # > 0: unbounded=False, max_action
    c["cond_sigma"] = bool(int(g[pre + "cond_sigma"])) if pre + "cond_sigma" in g.files else False  # conditioned_sigma=True
    c["a2c"] = bool(int(g[pre + "a2c"])) if pre + "a2c" in g.files else False  # recorded from A2C instead of PPO (no logp_old)
    for k in ("W1", "b1", "W2", "b2", "Wa", "ba", "Wv", "bv") + (() if c["cond_sigma"] else ("log_sigma",)):
        c[k] = g[pre + k]
        c["after_" + k] = g[pre + "after_" + k]
    if c["cond_sigma"]:  # the sigma head (fixture keys Ws / bs; "bs" is the batch size in this dict)
        for a in ("", "after_"):
            c[a + "Wsig"], c[a + "bsig"] = g[pre + a + "Ws"], g[pre + a + "bs"]
    for k in ("obs", "act", "rew", "term", "trunc", "obs_next", "indices", "v_s", "returns", "adv", "logp_old",
              "loss_seq", "clip_seq", "vf_seq", "ent_seq", "fwd_mu", "fwd_sigma", "fwd_value"):
        c[k] = g[pre + k]
    for a in ("", "after_"):  # head rows: [mu | log sigma (conditioned_sigma) | value]
        mid_w, mid_b = ([c[a + "Wsig"]], [c[a + "bsig"]]) if c["cond_sigma"] else ([], [])
        c[a + "W3"] = np.concatenate([c[a + "Wa"]] + mid_w + [c[a + "Wv"]], 0)
        c[a + "b3"] = np.concatenate([c[a + "ba"]] + mid_b + [c[a + "bv"]], 0)
    c["param_names"] = ("W1", "b1", "W2", "b2", "W3", "b3") + (() if c["cond_sigma"] else ("log_sigma",))
    B = R * T
    assert np.array_equal(c["indices"], np.arange(B))
    em = lambda a: np.swapaxes(a, 0, 1).reshape(B, *a.shape[2:])  # noqa: E731  [T][R] -> env-major flat
    c["flat"] = {k: em(c[k]) for k in ("obs", "act", "rew", "term", "trunc", "obs_next")}
    last = np.arange(R) * T + T - 1
    done = c["flat"]["term"] | c["flat"]["trunc"]
    c["unfinished"] = last[~done[last]]
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


def large_inputs(seed, R, T, O, nA):
    """Synthetic rollout of the C2-scale fixtures (tests/golden/ppo_large.npz), regenerated from a frozen numpy RandomState
    stream by both the generator (make_golden_large.py) and the tests, so that only the reference's OUTPUTS are committed.
    MPE-like: every row truncates every 25 steps at its own phase, terminates with p = 0.01; returns [T][R] arrays."""
    rs = np.random.RandomState(seed)
    obs = rs.standard_normal((R, O)).astype(np.float32)
    phase = rs.randint(0, 25, R)
    out = {k: [] for k in ("obs", "act", "rew", "term", "trunc", "obs_next")}
    for t in range(T):
        act = rs.randint(0, nA, R)
        rew = rs.standard_normal(R).astype(np.float32)
        trunc = (t + phase) % 25 == 24
        term = (rs.random_sample(R) < 0.01) & ~trunc
        obs_next = (obs + 0.3 * rs.standard_normal((R, O))).astype(np.float32)
        for k, v in zip(out, (obs, act, rew, term, trunc, obs_next)):
            out[k].append(np.array(v))
        fresh = rs.standard_normal((R, O)).astype(np.float32)
        obs = np.where((term | trunc)[:, None], fresh, obs_next)
    out = {k: np.stack(v) for k, v in out.items()}
    out["checksum"] = np.array([float(out["obs"].astype(np.float64).sum()), float(out["act"].sum()), float(out["rew"].astype(np.float64).sum()),
                                float(out["term"].sum()), float(out["trunc"].sum()), float(out["obs_next"].astype(np.float64).sum())])
    return out


def load_large(g, ci):
    """One case of ppo_large.npz: reference outputs + the regenerated inputs (checksum-verified)."""
    pre = f"p{ci}_"
    R, T, O, H, nA, bs, repeat, seed, in_seed = (int(v) for v in g[pre + "cfg"])
    eps_clip, dual_clip, vf_coef, ent_coef, value_clip, adv_norm, max_norm, lr, gamma, lam = (float(v) for v in g[pre + "hyper"])
    c = dict(R=R, T=T, O=O, H=H, nA=nA, bs=bs, repeat=repeat, seed=seed, lr=lr, gamma=gamma, lam=lam, max_norm=max_norm,
             cfg=dict(eps_clip=eps_clip, dual_clip=dual_clip, vf_coef=vf_coef, ent_coef=ent_coef, value_clip=bool(value_clip), adv_norm=bool(adv_norm)))
    x = large_inputs(in_seed, R, T, O, nA)
    assert np.array_equal(x["checksum"], g[pre + "checksum"]), "regenerated inputs differ from the recorded ones"
    c.update({k: x[k] for k in ("obs", "act", "rew", "term", "trunc", "obs_next")})
    for k in ("W1", "b1", "W2", "b2", "Wa", "ba", "Wv", "bv"):
        c[k] = g[pre + k]
        c["after_" + k] = g[pre + "after_" + k]
    for k in ("loss_seq", "clip_seq", "vf_seq", "ent_seq", "sample_idx", "v_s", "returns", "adv", "logp_old"):
        c[k] = g[pre + k]
    c["W3"] = np.concatenate([c["Wa"], c["Wv"]], 0)
    c["b3"] = np.concatenate([c["ba"], c["bv"]], 0)
    c["after_W3"] = np.concatenate([c["after_Wa"], c["after_Wv"]], 0)
    c["after_b3"] = np.concatenate([c["after_ba"], c["after_bv"]], 0)
    return c
