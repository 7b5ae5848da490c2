"""Replays tests/golden/buffer.npz traces (recorded from the reference VectorReplayBuffer)
against any buffer implementation exposing the reference API."""
import numpy as np


def to_np(x):
    if hasattr(x, "detach"):
        x = x.detach().cpu().numpy()
    return np.asarray(x)


def replay_buffer_trace(g, ci, make_buffer, make_batch, check_store=True, check_random=True):
    pre = f"c{ci}_"
    total, nbuf, steps, size = (int(v) for v in g[pre + "cfg"])
    buf = make_buffer(total, nbuf)
    assert buf.maxsize == size * nbuf
    for t in range(steps):
        s = pre + f"t{t}_"
        if s + "ids" not in g:
            continue
        ids = g[s + "ids"]
        obs = g[s + "obs"]
        batch = make_batch(obs=obs, act=g[s + "act"], rew=g[s + "rew"], terminated=g[s + "term"],
                           truncated=g[s + "trunc"], obs_next=obs + 1)
        ptr, ep_rew, ep_len, ep_idx = buf.add(batch, buffer_ids=ids)
        assert np.array_equal(to_np(ptr), g[s + "ptr"]), (ci, t)
        assert np.array_equal(to_np(ep_len), g[s + "ep_len"]), (ci, t)
        assert np.array_equal(to_np(ep_idx), g[s + "ep_idx"]), (ci, t)
        assert np.allclose(to_np(ep_rew), g[s + "ep_rew"], rtol=1e-6, atol=1e-6), (ci, t)
        assert len(buf) == int(g[s + "len"]), (ci, t)
        assert np.array_equal(to_np(buf.sample_indices(0)), g[s + "all"]), (ci, t)
        assert np.array_equal(to_np(buf.unfinished_index()), g[s + "unfinished"]), (ci, t)
        assert np.array_equal(to_np(buf.last_index), g[s + "last_index"]), (ci, t)
        assert np.array_equal(to_np(buf._lengths), g[s + "lengths"]), (ci, t)
        every = np.arange(buf.maxsize)
        assert np.array_equal(to_np(buf.prev(every)), g[s + "prev"]), (ci, t)
        assert np.array_equal(to_np(buf.next(every)), g[s + "next"]), (ci, t)
        if s + "reset" in g:
            buf.reset(keep_statistics=bool(g[s + "reset"][1]))
    if check_store:
        written = np.zeros(buf.maxsize, bool)
        # only slots that are currently valid are comparable (the rest is allocator-defined)
        written[g[pre + f"t{steps - 1}_all"]] = True
        assert np.array_equal(to_np(buf.done)[written], g[pre + "done"][written])
        assert np.allclose(to_np(buf.rew)[written], g[pre + "rew_store"][written], rtol=1e-6)
        assert np.array_equal(to_np(buf.obs)[written], g[pre + "obs_store"][written])
        assert np.array_equal(to_np(buf.obs_next)[written], g[pre + "obs_next_store"][written])
        assert np.array_equal(to_np(buf.act)[written], g[pre + "act_store"][written])
    if check_random:
        assert np.array_equal(to_np(buf.sample_indices(7)), g[pre + "sample7"])
        assert np.array_equal(to_np(buf.sample_indices(64)), g[pre + "sample64"])
        assert len(to_np(buf.sample_indices(-1))) == 0
    return buf
