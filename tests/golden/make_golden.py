"""Generate golden fixtures by running the UNMODIFIED reference in the build container.

Run (only where /root/reference exists; never on the GPU box):

    python tests/golden/make_golden.py [gae buffer ppo ppo_recompute ppo_retscale a2c reinforce ...]   (GENS lists the generators)
    python tests/golden/make_golden_{ctde,large,mappo,collector,marl}.py                                (one fixture file each)

The reference is imported from /root/reference through the import shims in oracle/shims/
(stand-ins for gymnasium, pettingzoo, overrides, sensai, h5py, deepdiff, matplotlib, which
are not installed here).  Outputs are small .npz files committed under tests/golden/; the
tests read only those files, so they run without the reference.
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, "/root/reference")
sys.path.insert(0, os.path.join(ROOT, "oracle", "shims"))
sys.path.insert(0, "/root/reference/test/base")  # env.py (MoveToRightEnv)


def save(name: str, **arrays) -> None:
    path = os.path.join(HERE, name)
    np.savez_compressed(path, **arrays)
    print(f"wrote {path}: {sorted(arrays)[:6]}... ({os.path.getsize(path)} B)")


# ----------------------------------------------------------------------------------------------
def gen_gae() -> None:
    """test/base/test_returns.py:23-158 vectors + random VectorReplayBuffer cases, all through
    the reference's Algorithm.compute_episodic_return (algorithm_base.py:651-717)."""
    from tianshou.algorithm import Algorithm
    from tianshou.data import Batch, ReplayBuffer, VectorReplayBuffer

    fn = Algorithm.compute_episodic_return
    out: dict[str, np.ndarray] = {}
    cases = [
        # (terminated, truncated, rew, v, gamma, lam, hand-computed answer in test_returns.py)
        ([1, 0, 0, 1, 0, 0, 0, 1.0], [0, 0, 0, 0, 0, 1, 0, 0], [0, 1, 2, 3, 4, 5, 6, 7.0], None, 0.1, 1.0,
         [0, 1.23, 2.3, 3, 4.5, 5, 6.7, 7]),
        ([0, 1, 0, 1, 0, 1, 0.0], [0, 0, 0, 0, 0, 0, 0.0], [7, 6, 1, 2, 3, 4, 5.0], None, 0.1, 1.0,
         [7.6, 6, 1.2, 2, 3.4, 4, 5]),
        ([0, 1, 0, 1, 0, 0, 1.0], [0, 0, 0, 0, 0, 0, 0], [7, 6, 1, 2, 3, 4, 5.0], None, 0.1, 1.0,
         [7.6, 6, 1.2, 2, 3.45, 4.5, 5]),
        ([0, 0, 0, 1.0, 0, 0, 0, 1, 0, 0, 0, 1], [0] * 12,
         [101, 102, 103.0, 200, 104, 105, 106, 201, 107, 108, 109, 202],
         [2.0, 3.0, 4, -1, 5.0, 6.0, 7, -2, 8.0, 9.0, 10, -3], 0.99, 0.95,
         [454.8344, 376.1143, 291.298, 200.0, 464.5610, 383.1085, 295.387, 201.0, 474.2876, 390.1027, 299.476, 202.0]),
        ([0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1], [0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0, 0],
         [101, 102, 103.0, 200, 104, 105, 106, 201, 107, 108, 109, 202],
         [2.0, 3.0, 4, -1, 5.0, 6.0, 7, -2, 8.0, 9.0, 10, -3], 0.99, 0.95,
         [454.0109, 375.2386, 290.3669, 199.01, 462.9138, 381.3571, 293.5248, 199.02, 474.2876, 390.1027, 299.476, 202.0]),
    ]
    for ci, (term, trunc, rew, v, gamma, lam, ans) in enumerate(cases):
        buf = ReplayBuffer(20)
        batch = Batch(terminated=np.array(term, float), truncated=np.array(trunc, float), rew=np.array(rew, float))
        for b in iter(batch):
            b.obs = b.act = 1
            buf.add(b)
        v_arr = None if v is None else np.array(v)
        ret, adv = fn(batch, buf, buf.sample_indices(0), v_arr, gamma=gamma, gae_lambda=lam)
        assert np.allclose(ret, ans), (ci, ret, ans)
        out[f"k{ci}_term"], out[f"k{ci}_trunc"], out[f"k{ci}_rew"] = np.array(term, float), np.array(trunc, float), np.array(rew, float)
        out[f"k{ci}_v"] = np.array([]) if v is None else np.array(v)
        out[f"k{ci}_gl"] = np.array([gamma, lam])
        out[f"k{ci}_ans"], out[f"k{ci}_ret"], out[f"k{ci}_adv"] = np.array(ans), ret, adv
    out["n_known"] = np.array(len(cases))

    # random rollouts in a VectorReplayBuffer, including wrap-around and ragged sub-buffers
    rng = np.random.default_rng(1234)
    specs = [(4, 8, 8, 1.0), (7, 16, 16, 1.0), (5, 12, 30, 0.7), (16, 25, 25, 1.0), (3, 128, 128, 1.0)]
    for ri, (nbuf, size, steps, p_add) in enumerate(specs):
        buf = VectorReplayBuffer(nbuf * size, nbuf)
        for t in range(steps):
            ids = np.nonzero(rng.random(nbuf) < p_add)[0] if p_add < 1 else np.arange(nbuf)
            if len(ids) == 0:
                continue
            n = len(ids)
            buf.add(
                Batch(obs=rng.normal(size=(n, 3)).astype(np.float32), act=rng.integers(0, 5, n),
                      rew=rng.normal(size=n), terminated=rng.random(n) < 0.05, truncated=rng.random(n) < 0.08,
                      obs_next=rng.normal(size=(n, 3)).astype(np.float32)),
                buffer_ids=ids,
            )
        batch, indices = buf.sample(0)
        B = len(indices)
        v_s = rng.normal(size=B).astype(np.float32)
        v_s_ = rng.normal(size=B).astype(np.float32)
        import torch

        ret, adv = fn(batch, buf, indices, torch.from_numpy(v_s_), torch.from_numpy(v_s), gamma=0.99, gae_lambda=0.95)
        pre = f"r{ri}_"
        out[pre + "shape"] = np.array([nbuf, size])
        out[pre + "indices"], out[pre + "rew"] = indices, batch.rew
        out[pre + "term"], out[pre + "trunc"] = batch.terminated, batch.truncated
        out[pre + "buf_term"], out[pre + "buf_done"] = buf.terminated, buf.done
        out[pre + "unfinished"], out[pre + "last_index"] = buf.unfinished_index(), buf.last_index
        out[pre + "lengths"] = buf._lengths
        out[pre + "v_s"], out[pre + "v_s_"], out[pre + "ret"], out[pre + "adv"] = v_s, v_s_, ret, adv
        # MC return-to-go variant: v_s_ None, lambda 1 (algorithm_base.py:704-706)
        ret1, adv1 = fn(batch, buf, indices, None, None, gamma=0.9, gae_lambda=1.0)
        out[pre + "ret_mc"] = ret1
    out["n_random"] = np.array(len(specs))

    from tianshou.algorithm.algorithm_base import episode_mc_return_to_go

    r = rng.normal(size=25)
    out["mc_rew"], out["mc_out"] = r, episode_mc_return_to_go(r, 0.97)
    save("gae.npz", **out)


# ----------------------------------------------------------------------------------------------
def gen_buffer() -> None:
    """Operation traces over the reference VectorReplayBuffer (manager.py / buffer_base.py)."""
    from tianshou.data import Batch, VectorReplayBuffer

    out: dict[str, np.ndarray] = {}
    rng = np.random.default_rng(42)
    configs = [(20, 4, 14, 0.8), (20, 4, 9, 1.0), (64, 8, 40, 0.6), (75, 5, 50, 1.0), (7, 3, 12, 0.9)]
    for ci, (total, nbuf, steps, p_add) in enumerate(configs):
        buf = VectorReplayBuffer(total, nbuf)
        size = buf.maxsize // nbuf
        pre = f"c{ci}_"
        out[pre + "cfg"] = np.array([total, nbuf, steps, size])
        for t in range(steps):
            ids = np.nonzero(rng.random(nbuf) < p_add)[0]
            if len(ids) == 0:
                ids = np.array([int(rng.integers(nbuf))])
            n = len(ids)
            term = rng.random(n) < 0.15
            trunc = rng.random(n) < 0.1
            rew = np.round(rng.normal(size=n), 3)
            obs = rng.integers(0, 100, size=(n, 2)).astype(np.float32)
            act = rng.integers(0, 5, n)
            ptr, ep_rew, ep_len, ep_idx = buf.add(
                Batch(obs=obs, act=act, rew=rew, terminated=term, truncated=trunc, obs_next=obs + 1), buffer_ids=ids
            )
            s = pre + f"t{t}_"
            out[s + "ids"], out[s + "term"], out[s + "trunc"], out[s + "rew"] = ids, term, trunc, rew
            out[s + "obs"], out[s + "act"] = obs, act
            out[s + "ptr"], out[s + "ep_rew"], out[s + "ep_len"], out[s + "ep_idx"] = ptr, ep_rew, ep_len, ep_idx
            out[s + "len"] = np.array(len(buf))
            out[s + "all"] = buf.sample_indices(0)
            out[s + "unfinished"] = buf.unfinished_index()
            out[s + "last_index"], out[s + "lengths"] = buf.last_index.copy(), buf._lengths.copy()
            every = np.arange(buf.maxsize)
            out[s + "prev"], out[s + "next"] = buf.prev(every), buf.next(every)
            if t == steps // 2 and ci in (1, 3):
                buf.reset(keep_statistics=(ci == 3))
                out[s + "reset"] = np.array([1, int(ci == 3)])
        out[pre + "done"], out[pre + "rew_store"] = buf.done.copy(), buf.rew.copy()
        out[pre + "obs_store"], out[pre + "act_store"] = buf.obs.copy(), buf.act.copy()
        out[pre + "obs_next_store"] = buf.obs_next.copy()
        # random sampling: manager RandomState(42) then per-child RandomState(42) (manager.py:215-229)
        out[pre + "sample7"] = buf.sample_indices(7)
        out[pre + "sample64"] = buf.sample_indices(64)
        out[pre + "sample_neg"] = buf.sample_indices(-1)
    out["n_cfg"] = np.array(len(configs))

    # test_buffer.py:1525-1549 get_buffer_indices pins
    buf = VectorReplayBuffer(10, 2)
    out["gbi_2_4"], out["gbi_4_2"], out["gbi_8_7"] = (buf.get_buffer_indices(2, 4), buf.get_buffer_indices(4, 2), buf.get_buffer_indices(8, 7))
    for bad in [(1, 6), (8, 1)]:
        try:
            buf.get_buffer_indices(*bad)
            raise AssertionError("expected ValueError")
        except ValueError:
            pass
    save("buffer.npz", **out)


GENS = {"gae": gen_gae, "buffer": gen_buffer}

if __name__ == "__main__":
    try:
        from make_golden_more import GENS as MORE  # ppo / a2c / reinforce fixtures

        GENS.update(MORE)
    except ImportError:
        pass
    which = sys.argv[1:] or list(GENS)
    for w in which:
        GENS[w]()
