"""Batch: the dict-of-arrays container the hot path exchanges.

Host-side mirror of the part of tianshou/data/batch.py the path uses: attribute / key access,
row indexing, `split` (minibatch order, batch.py:1209-1226), `cat`, `stack`, `get_keys`, `pop`,
`update`, `to_torch` / `to_numpy`.  Values may be numpy arrays, torch tensors (any device) or
nested Batches.  It is plumbing: no arithmetic happens here.
"""
from __future__ import annotations

from collections.abc import Iterator, Sequence
from typing import Any

import numpy as np
import torch


def _convert(v: Any) -> Any:
    if isinstance(v, Batch | np.ndarray | torch.Tensor):
        return v
    if isinstance(v, dict):
        return Batch(v)
    if v is None:
        return None
    if isinstance(v, list | tuple):
        if len(v) > 0 and all(isinstance(e, dict | Batch) for e in v):
            return Batch.stack([Batch(e) if isinstance(e, dict) else e for e in v])
        if len(v) > 0 and all(isinstance(e, torch.Tensor) for e in v):
            return torch.stack(list(v))
        return np.asarray(v)
    if isinstance(v, int | float | bool | np.number | np.bool_):
        return np.asarray(v)
    return v  # opaque object (e.g. a distribution)


def _len(v: Any) -> int | None:
    if isinstance(v, Batch):
        return len(v) if len(v.__dict__) else None
    if isinstance(v, np.ndarray | torch.Tensor):
        return None if v.ndim == 0 else int(v.shape[0])
    return None


class Batch:
    def __init__(self, batch_dict: Any = None, copy: bool = False, **kwargs: Any) -> None:
        if batch_dict is not None:
            if isinstance(batch_dict, Batch):
                batch_dict = batch_dict.__dict__
            if isinstance(batch_dict, dict):
                for k, v in batch_dict.items():
                    self.__dict__[k] = _convert(v)
            elif isinstance(batch_dict, list | tuple) and len(batch_dict) > 0:
                st = Batch.stack([Batch(e) for e in batch_dict])
                self.__dict__.update(st.__dict__)
        for k, v in kwargs.items():
            self.__dict__[k] = _convert(v)
        if copy:
            for k, v in self.__dict__.items():
                if isinstance(v, np.ndarray):
                    self.__dict__[k] = v.copy()
                elif isinstance(v, torch.Tensor):
                    self.__dict__[k] = v.clone()
                elif isinstance(v, Batch):
                    self.__dict__[k] = Batch(v, copy=True)

    # -- mapping protocol ---------------------------------------------------------------------
    def __setattr__(self, key: str, value: Any) -> None:
        self.__dict__[key] = _convert(value)

    def __getattr__(self, key: str) -> Any:
        # only called when normal lookup fails
        raise AttributeError(key)

    def __contains__(self, key: str) -> bool:
        return key in self.__dict__

    def get_keys(self):
        return self.__dict__.keys()

    def keys(self):
        return self.__dict__.keys()

    def values(self):
        return self.__dict__.values()

    def items(self):
        return self.__dict__.items()

    def get(self, key: str, default: Any = None) -> Any:
        return self.__dict__.get(key, default)

    def pop(self, key: str, default: Any = None) -> Any:
        return self.__dict__.pop(key, default)

    def update(self, batch: Any = None, **kwargs: Any) -> None:
        if batch is not None:
            for k, v in (batch.items() if isinstance(batch, Batch | dict) else batch):
                self.__dict__[k] = _convert(v)
        for k, v in kwargs.items():
            self.__dict__[k] = _convert(v)

    def to_dict(self) -> dict:
        return {k: (v.to_dict() if isinstance(v, Batch) else v) for k, v in self.__dict__.items()}

    def is_empty(self) -> bool:
        return len(self.__dict__) == 0

    def __copy__(self) -> "Batch":
        b = Batch()
        b.__dict__.update(self.__dict__)
        return b

    # -- indexing -----------------------------------------------------------------------------
    def __getitem__(self, index: Any) -> Any:
        if isinstance(index, str):
            return self.__dict__[index]
        out = Batch()
        for k, v in self.__dict__.items():
            if isinstance(v, Batch):
                out.__dict__[k] = v[index] if len(v.__dict__) else Batch()
            elif isinstance(v, torch.Tensor):
                idx = index
                if isinstance(idx, np.ndarray):
                    idx = torch.as_tensor(idx, device=v.device)
                out.__dict__[k] = v[idx]
            elif isinstance(v, np.ndarray):
                idx = index
                if isinstance(idx, torch.Tensor):
                    idx = idx.detach().cpu().numpy()
                out.__dict__[k] = v[idx]
            else:
                out.__dict__[k] = v
        return out

    def __setitem__(self, index: Any, value: Any) -> None:
        if isinstance(index, str):
            self.__dict__[index] = _convert(value)
            return
        value = value if isinstance(value, Batch) else Batch(value)
        for k, v in value.items():
            dst = self.__dict__[k]
            if isinstance(dst, Batch):
                dst[index] = v
            elif isinstance(dst, torch.Tensor):
                idx = torch.as_tensor(index, device=dst.device) if isinstance(index, np.ndarray) else index
                dst[idx] = torch.as_tensor(v, device=dst.device, dtype=dst.dtype)
            else:
                idx = index.detach().cpu().numpy() if isinstance(index, torch.Tensor) else index
                dst[idx] = v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else v

    def __len__(self) -> int:
        lens = [n for n in (_len(v) for v in self.__dict__.values()) if n is not None]
        if not lens:
            if self.__dict__:
                raise TypeError("Object of type 'Batch' has no len()")
            return 0
        return min(lens)

    @property
    def shape(self) -> list[int]:
        return [len(self)] if self.__dict__ else []

    def __iter__(self) -> Iterator["Batch"]:
        for i in range(len(self)):
            yield self[i]

    def __repr__(self) -> str:
        inner = ", ".join(f"{k}: {tuple(v.shape) if hasattr(v, 'shape') else v!r}" for k, v in self.__dict__.items())
        return f"Batch({inner})"

    # -- conversion ---------------------------------------------------------------------------
    def to_torch(self, dtype: torch.dtype | None = None, device: Any = "cpu") -> "Batch":
        out = Batch()
        for k, v in self.__dict__.items():
            if isinstance(v, Batch):
                out.__dict__[k] = v.to_torch(dtype, device)
            elif isinstance(v, np.ndarray | torch.Tensor) and getattr(v, "dtype", None) != object:
                t = torch.as_tensor(v, device=device)
                out.__dict__[k] = t.to(dtype) if dtype is not None and t.is_floating_point() else t
            else:
                out.__dict__[k] = v
        return out

    def to_numpy(self) -> "Batch":
        out = Batch()
        for k, v in self.__dict__.items():
            if isinstance(v, Batch):
                out.__dict__[k] = v.to_numpy()
            elif isinstance(v, torch.Tensor):
                out.__dict__[k] = v.detach().cpu().numpy()
            else:
                out.__dict__[k] = v
        return out

    # -- split / cat / stack -------------------------------------------------------------------
    def split(self, size: int, shuffle: bool = True, merge_last: bool = False) -> Iterator["Batch"]:
        """tianshou/data/batch.py:1209-1226: np.random.permutation on the GLOBAL numpy RNG, then
        consecutive chunks; merge_last folds a short tail into the previous chunk."""
        for idx in split_indices(len(self), size, shuffle, merge_last):
            yield self[idx]

    @staticmethod
    def cat(batches: Sequence["Batch"]) -> "Batch":
        batches = [b for b in batches if isinstance(b, Batch) and not b.is_empty()]
        out = Batch()
        if not batches:
            return out
        for k in batches[0].keys():
            vals = [b[k] for b in batches]
            v0 = vals[0]
            if isinstance(v0, Batch):
                out.__dict__[k] = Batch.cat(vals)
            elif isinstance(v0, torch.Tensor):
                out.__dict__[k] = torch.cat(vals, dim=0)
            elif isinstance(v0, np.ndarray):
                out.__dict__[k] = np.concatenate(vals, axis=0)
            else:
                out.__dict__[k] = v0
        return out

    @staticmethod
    def stack(batches: Sequence["Batch"], axis: int = 0) -> "Batch":
        out = Batch()
        if not batches:
            return out
        for k in batches[0].keys():
            vals = [b[k] for b in batches]
            v0 = vals[0]
            if isinstance(v0, Batch):
                out.__dict__[k] = Batch.stack(vals, axis)
            elif isinstance(v0, torch.Tensor):
                out.__dict__[k] = torch.stack(vals, dim=axis)
            elif isinstance(v0, np.ndarray):
                out.__dict__[k] = np.stack(vals, axis=axis)
            else:
                out.__dict__[k] = np.asarray(vals)
        return out


def split_indices(length: int, size: int, shuffle: bool = True, merge_last: bool = False) -> Iterator[np.ndarray]:
    """Index chunks of Batch.split (batch.py:1209-1226), bit-exact with the reference when the
    global numpy RNG is seeded identically."""
    if size == -1:
        size = length
    assert size >= 1
    indices = np.random.permutation(length) if shuffle else np.arange(length)
    merge_last = merge_last and length % size > 0
    for idx in range(0, length, size):
        if merge_last and idx + size + size >= length:
            yield indices[idx:]
            break
        yield indices[idx : idx + size]
