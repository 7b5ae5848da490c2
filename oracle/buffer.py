"""Oracle: VectorReplayBuffer index state machine.  Test infrastructure only.

numpy restatement of
  ReplayBuffer._update_state_pre_add   tianshou/data/buffer/buffer_base.py:355-410
  ReplayBuffer.sample_indices          tianshou/data/buffer/buffer_base.py:495-531
  ReplayBuffer.unfinished_index        tianshou/data/buffer/buffer_base.py:309-312
  ReplayBuffer.get_buffer_indices      tianshou/data/buffer/buffer_base.py:173-228
  ReplayBufferManager.add / sample_indices / prev / next / unfinished_index / reset
                                       tianshou/data/buffer/manager.py:67-229
  _prev_index / _next_index            tianshou/data/buffer/manager.py:306-358
  VectorReplayBuffer.__init__          tianshou/data/buffer/vecbuf.py:33-37
"""
from __future__ import annotations

import numpy as np


class OracleVectorReplayBuffer:
    def __init__(self, total_size: int, buffer_num: int, random_seed: int = 42) -> None:
        assert buffer_num > 0
        self.buffer_num = buffer_num
        self.size = int(np.ceil(total_size / buffer_num))  # vecbuf.py:35
        self.maxsize = self.size * buffer_num
        self._offset = np.arange(buffer_num, dtype=np.int64) * self.size  # manager.py:34-41
        self._extend_offset = np.concatenate([self._offset, [self.maxsize]])
        self._seed = random_seed
        self._random_state = np.random.RandomState(random_seed)  # manager's own (buffer_base.py:101)
        self._child_rs: dict[int, np.random.RandomState] = {}
        self.data: dict[str, np.ndarray] = {}
        self._init_state()

    # -- per-sub-buffer state, buffer_base.py:103-106 ------------------------------------
    def _init_state(self) -> None:
        n = self.buffer_num
        self.ins = np.zeros(n, np.int64)  # _insertion_idx
        self.sz = np.zeros(n, np.int64)  # _size
        self.ep_ret = np.zeros(n, np.float64)
        self.ep_len = np.zeros(n, np.int64)
        self.ep_start = np.zeros(n, np.int64)
        self.last_index = self._offset.copy()  # manager.py:72
        self._lengths = np.zeros(n, np.int64)

    def reset(self, keep_statistics: bool = False) -> None:
        """manager.py:70-75 + buffer_base.py:292-298."""
        self.last_index = self._offset.copy()
        self._lengths = np.zeros(self.buffer_num, np.int64)
        self.ins[:] = 0
        self.sz[:] = 0
        self.ep_start[:] = 0
        if not keep_statistics:
            self.ep_ret[:] = 0.0
            self.ep_len[:] = 0

    def __len__(self) -> int:
        return int(self._lengths.sum())  # manager.py:67-68

    @property
    def subbuffer_edges(self) -> np.ndarray:
        return self._extend_offset

    # -- add --------------------------------------------------------------------------------
    def add(self, batch: dict, buffer_ids=None):
        """manager.py:131-193.  `batch` is a dict of arrays with leading dim len(buffer_ids)."""
        term = np.asarray(batch["terminated"]).astype(bool)
        trunc = np.asarray(batch["truncated"]).astype(bool)
        done = np.logical_or(term, trunc)
        rew = np.asarray(batch["rew"]).astype(np.float64)
        if buffer_ids is None:
            buffer_ids = np.arange(self.buffer_num)
        buffer_ids = np.asarray(buffer_ids, dtype=np.int64)
        ptrs, ep_rets, ep_lens, ep_idxs = [], [], [], []
        for k, b in enumerate(buffer_ids):
            # buffer_base.py:373-410
            cur = self.ins[b]
            self.sz[b] = min(self.sz[b] + 1, self.size)
            self.ins[b] = (self.ins[b] + 1) % self.size
            self.ep_ret[b] += rew[k]
            self.ep_len[b] += 1
            if self.ep_start[b] > self.sz[b]:
                raise RuntimeError("MalformedBufferError")  # buffer_base.py:380-386
            if done[k]:
                er, el = self.ep_ret[b], self.ep_len[b]
            else:
                er, el = 0.0, 0
            es = self.ep_start[b]
            if done[k]:
                self.ep_ret[b], self.ep_len[b], self.ep_start[b] = 0.0, 0, self.ins[b]
            ptrs.append(cur + self._offset[b])
            ep_rets.append(er)
            ep_lens.append(el)
            ep_idxs.append(es + self._offset[b])
            self.last_index[b] = cur + self._offset[b]
            self._lengths[b] = self.sz[b]
        ptrs = np.array(ptrs, dtype=np.int64)
        store = dict(batch)
        store["terminated"], store["truncated"], store["done"], store["rew"] = term, trunc, done, rew
        for key, val in store.items():
            val = np.asarray(val)
            if key not in self.data:
                self.data[key] = np.zeros((self.maxsize, *val.shape[1:]), dtype=val.dtype)
            self.data[key][ptrs] = val
        return ptrs, np.array(ep_rets, np.float64), np.array(ep_lens, np.int64), np.array(ep_idxs, np.int64)

    # -- indices ----------------------------------------------------------------------------
    def _child_rng(self, b: int) -> np.random.RandomState:
        if b not in self._child_rs:
            self._child_rs[b] = np.random.RandomState(self._seed)  # every child: RandomState(42)
        return self._child_rs[b]

    def _child_sample_indices(self, b: int, batch_size: int) -> np.ndarray:
        """buffer_base.py:505-515 (stack_num == 1)."""
        if batch_size > 0:
            return self._child_rng(b).choice(self.sz[b], batch_size)
        if batch_size == 0:
            return np.concatenate([np.arange(self.ins[b], self.sz[b]), np.arange(self.ins[b])])
        return np.array([], int)

    def sample_indices(self, batch_size: int | None) -> np.ndarray:
        """manager.py:195-229."""
        if batch_size is not None and batch_size < 0:
            return np.array([], int)
        if batch_size == 0 or batch_size is None:
            sample_num = np.zeros(self.buffer_num, int)
        else:
            buffer_idx = self._random_state.choice(self.buffer_num, batch_size, p=self._lengths / self._lengths.sum())
            sample_num = np.bincount(buffer_idx, minlength=self.buffer_num)
            sample_num[sample_num == 0] = -1
        return np.concatenate(
            [self._child_sample_indices(b, int(k)) + self._offset[b] for b, k in enumerate(sample_num)]
        ).astype(np.int64)

    def unfinished_index(self) -> np.ndarray:
        """manager.py:85-91 + buffer_base.py:309-312."""
        out = []
        done = self.data.get("done")
        for b in range(self.buffer_num):
            if self.sz[b]:
                last = (self.ins[b] - 1) % self.sz[b]
                if not done[last + self._offset[b]]:
                    out.append(last + self._offset[b])
        return np.array(out, dtype=np.int64)

    def prev(self, index) -> np.ndarray:
        """manager.py:306-331, one index at a time (O(n) form of the masked loops)."""
        index = np.atleast_1d(np.asarray(index, dtype=np.int64)) % self.maxsize
        done = self.data["done"]
        out = np.zeros_like(index)
        for k, i in enumerate(index):
            b = i // self.size
            start = self._offset[b]
            cur_len = max(1, self._lengths[b])
            sub = (i - start - 1) % cur_len
            end_flag = bool(done[sub + start]) or (sub + start == self.last_index[b])
            out[k] = (sub + int(end_flag)) % cur_len + start
        return out

    def next(self, index) -> np.ndarray:
        """manager.py:334-358."""
        index = np.atleast_1d(np.asarray(index, dtype=np.int64)) % self.maxsize
        done = self.data["done"]
        out = np.zeros_like(index)
        for k, i in enumerate(index):
            b = i // self.size
            start = self._offset[b]
            cur_len = max(1, self._lengths[b])
            end_flag = bool(done[i]) or (i == self.last_index[b])
            out[k] = (i - start + 1 - int(end_flag)) % cur_len + start
        return out

    def get_buffer_indices(self, start: int, stop: int) -> np.ndarray:
        """buffer_base.py:173-228."""
        edges = self._extend_offset
        s_edge = np.searchsorted(edges, start, side="right") - 1
        e_edge = np.searchsorted(edges, stop - 1, side="right") - 1
        if s_edge != e_edge:
            raise ValueError("Start and stop indices must be within the same subbuffer.")
        if stop >= start:
            return np.arange(start, stop, dtype=int)
        upper = edges[s_edge + 1]
        lower = edges[s_edge]
        return np.concatenate((np.arange(start, upper, dtype=int), np.arange(lower, stop, dtype=int)))
