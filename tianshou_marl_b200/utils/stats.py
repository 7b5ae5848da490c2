"""Stats dataclasses with the reference's field names (tianshou/data/stats.py:18-47,
tianshou/algorithm/modelfree/a2c.py:23-29, tianshou/data/collector.py:116-254)."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any

import numpy as np


@dataclass
class SequenceSummaryStats:
    mean: float
    std: float
    max: float
    min: float

    @classmethod
    def from_sequence(cls, sequence: Any) -> "SequenceSummaryStats":
        a = np.asarray(sequence, dtype=np.float64)
        if a.size == 0:
            return cls(mean=0.0, std=0.0, max=0.0, min=0.0)
        return cls(mean=float(a.mean()), std=float(a.std()), max=float(a.max()), min=float(a.min()))


@dataclass
class A2CTrainingStats:
    loss: SequenceSummaryStats
    actor_loss: SequenceSummaryStats
    vf_loss: SequenceSummaryStats
    ent_loss: SequenceSummaryStats
    gradient_steps: int = 0
    train_time: float = 0.0
    smoothed_loss: dict = field(default_factory=dict)


@dataclass
class CollectStats:
    """collector.py:116-149 (same field names; arrays are numpy)."""

    n_collected_episodes: int = 0
    n_collected_steps: int = 0
    collect_time: float = 0.0
    collect_speed: float = 0.0
    returns: np.ndarray = field(default_factory=lambda: np.array([], dtype=float))
    returns_stat: SequenceSummaryStats | None = None
    lens: np.ndarray = field(default_factory=lambda: np.array([], dtype=int))
    lens_stat: SequenceSummaryStats | None = None
    pred_dist_std_array: np.ndarray | None = None
    pred_dist_std_array_stat: dict | None = None

    def refresh_return_stats(self) -> None:
        self.returns_stat = SequenceSummaryStats.from_sequence(self.returns) if self.returns.size > 0 else None

    def refresh_len_stats(self) -> None:
        self.lens_stat = SequenceSummaryStats.from_sequence(self.lens) if self.lens.size > 0 else None

    def refresh_all_sequence_stats(self) -> None:
        self.refresh_return_stats()
        self.refresh_len_stats()

    def set_collect_time(self, t: float, update_collect_speed: bool = True) -> None:
        self.collect_time = t
        if update_collect_speed:
            self.collect_speed = self.n_collected_steps / t if t > 0 else 0.0


class RunningMeanStd:
    """Running mean / variance of a data stream (reference tianshou/utils/statistics.py:69-114; parallel-variance merge).
    The device path feeds it batch moments (`update_from_moments`); `update` takes an array like the reference."""

    def __init__(self, mean: float = 0.0, std: float = 1.0, clip_max: float | None = 10.0,
                 epsilon: float = float(np.finfo(np.float32).eps)) -> None:
        self.mean, self.var = mean, std  # (the reference initialises var with `std`, statistics.py:88)
        self.clip_max, self.count, self.eps = clip_max, 0, epsilon

    def norm(self, data_array: Any) -> Any:
        data_array = (data_array - self.mean) / np.sqrt(self.var + self.eps)
        if self.clip_max:
            data_array = np.clip(data_array, -self.clip_max, self.clip_max)
        return data_array

    def update_from_moments(self, batch_mean: float, batch_var: float, batch_count: int) -> None:
        delta = batch_mean - self.mean
        total_count = self.count + batch_count
        new_mean = self.mean + delta * batch_count / total_count
        m_2 = self.var * self.count + batch_var * batch_count + delta**2 * self.count * batch_count / total_count
        self.mean, self.var, self.count = new_mean, m_2 / total_count, total_count

    def update(self, data_array: np.ndarray) -> None:
        self.update_from_moments(np.mean(data_array, axis=0), np.var(data_array, axis=0), len(data_array))
