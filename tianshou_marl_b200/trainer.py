"""The on-policy training loop around the hot path.

Mirrors the call surface and loop semantics of the reference's `OnPolicyTrainer` / `OnPolicyTrainerParams`
(tianshou/trainer/trainer.py:81-330 parameters, :557-614 `execute_epoch`, :640-707 test step, :736-758 `run`,
:878-906 training step, :1079-1109 on-policy update + `reset_buffer(keep_statistics=True)`): per training step the
collector gathers `collection_step_num_env_steps` transitions (or episodes), the whole buffer goes to
`algorithm.update(buffer, batch_size, repeat)` and the buffer is cleared; an epoch ends after `epoch_num_steps`
environment steps and is followed by a test step when a test collector is given; training stops after `max_epochs` or
when `stop_fn(best score)` holds.  Everything that touches data runs in the device kernels behind `Collector` and
`PPO`; this module is host control flow only.  Loggers, checkpoints and progress bars are out of scope.
"""
from __future__ import annotations

import time
from dataclasses import dataclass, field
from typing import Any, Callable


@dataclass(kw_only=True)
class OnPolicyTrainerParams:
    """Field names as in the reference (trainer.py:81-330)."""

    train_collector: Any
    test_collector: Any = None
    max_epochs: int = 100
    epoch_num_steps: int = 30000
    collection_step_num_env_steps: int | None = 2048
    collection_step_num_episodes: int | None = None
    batch_size: int | None = 64
    update_step_num_repetitions: int = 1
    test_step_num_episodes: int = 1
    train_fn: Callable[[int, int], None] | None = None
    test_fn: Callable[[int, int | None], None] | None = None
    stop_fn: Callable[[float], bool] | None = None
    compute_score_fn: Callable[[Any], float] | None = None
    save_best_fn: Callable[[Any], None] | None = None
    test_in_train: bool = False
    verbose: bool = False

    def __post_init__(self) -> None:
        if (self.collection_step_num_env_steps is None) == (self.collection_step_num_episodes is None):
            raise ValueError("Exactly one of {collection_step_num_env_steps, collection_step_num_episodes} must be set")  # :277-280
        if self.test_collector is None:
            if self.stop_fn is not None:
                raise ValueError("stop_fn cannot be activated without test steps being enabled (test_collector being set)")  # :211-214
            if self.test_fn is not None:
                raise ValueError("test_fn is set while test steps are disabled (test_collector is None)")
            if self.save_best_fn is not None:
                raise ValueError("save_best_fn is set while test steps are disabled (test_collector is None)")
        elif self.test_step_num_episodes < 1:
            raise ValueError("test_step_num_episodes must be positive if test steps are enabled (test_collector not None)")
        if self.test_in_train and (self.test_collector is None or self.stop_fn is None):
            raise ValueError("test_in_train requires test_collector and stop_fn to be set")  # :281-282


@dataclass(kw_only=True)
class TimingStats:
    total_time: float = 0.0
    train_time_collect: float = 0.0
    train_time_update: float = 0.0
    test_time: float = 0.0


@dataclass(kw_only=True)
class InfoStats:
    """data/stats.py:83-104."""

    update_step: int
    best_score: float
    best_reward: float
    best_reward_std: float
    train_step: int
    train_episode: int
    test_step: int
    test_episode: int
    timing: TimingStats = field(default_factory=TimingStats)


@dataclass(kw_only=True)
class EpochStats:
    """data/stats.py:108-125."""

    epoch: int
    train_collect_stat: Any
    test_collect_stat: Any
    training_stat: Any
    info_stat: InfoStats


class OnPolicyTrainer:
    def __init__(self, algorithm: Any, params: OnPolicyTrainerParams) -> None:
        self.algorithm, self.params = algorithm, params
        self._epoch = self._env_step = self._current_update_step = 0
        self._stop_fn_flag = False
        self._best_score = self._best_reward = float("-inf")
        self._best_reward_std = 0.0
        self._best_epoch = 0
        self._timing = TimingStats()
        self._t0 = time.time()

    # -- trainer.py:418-470, :836-850 ---------------------------------------------------------------
    def reset(self, reset_collectors: bool = True, reset_collector_buffers: bool = False) -> None:
        p = self.params
        self._epoch = self._env_step = self._current_update_step = 0
        self._stop_fn_flag = False
        self._timing = TimingStats()
        self._t0 = time.time()
        if reset_collectors:
            p.train_collector.reset(reset_buffer=reset_collector_buffers)
            if p.test_collector is not None:
                p.test_collector.reset(reset_buffer=reset_collector_buffers)
        # without test steps the reference reports 0.0 for the best score / reward (trainer.py:367-368; tests/golden/trainer.npz "no_test")
        self._best_score = self._best_reward = float("-inf") if p.test_collector is not None else 0.0
        self._best_reward_std, self._best_epoch = 0.0, 0
        if p.test_collector is not None:  # initial test step fixes the "best" baseline (:460-468)
            self._test_step(force_update_best=True)

    def _score(self, stats: Any) -> float:
        if self.params.compute_score_fn is not None:
            return float(self.params.compute_score_fn(stats))
        return float(stats.returns_stat.mean)  # default score: mean test return (:367-375)

    def _policy(self) -> Any:
        return self.algorithm.policy

    # -- trainer.py:640-707 -------------------------------------------------------------------------
    def _test_step(self, force_update_best: bool = False) -> tuple[Any, bool]:
        p = self.params
        t = time.time()
        pol = self._policy()
        was = pol.is_within_training_step
        pol.is_within_training_step = False  # deterministic_eval applies outside training steps (reinforce.py:185-189)
        try:
            p.test_collector.reset(reset_stats=False)
            if p.test_fn:
                p.test_fn(self._epoch, self._env_step)
            stat = p.test_collector.collect(n_episode=p.test_step_num_episodes)
        finally:
            pol.is_within_training_step = was
        self._timing.test_time += time.time() - t
        score = self._score(stat)
        if score > self._best_score or force_update_best:
            self._best_score, self._best_epoch = score, self._epoch
            self._best_reward, self._best_reward_std = float(stat.returns_stat.mean), float(stat.returns_stat.std)
            if p.save_best_fn:
                p.save_best_fn(self.algorithm)
        if p.verbose:
            print(f"Epoch #{self._epoch}: test_reward: {stat.returns_stat.mean:.6f} ± {stat.returns_stat.std:.6f}, "
                  f"best_reward: {self._best_reward:.6f} ± {self._best_reward_std:.6f} in #{self._best_epoch}", flush=True)
        stop = bool(p.stop_fn(self._best_score)) if p.stop_fn is not None else False
        return stat, stop

    # -- trainer.py:878-906 + :1079-1109 -------------------------------------------------------------
    def _training_step(self) -> tuple[Any, Any, bool]:
        p = self.params
        pol = self._policy()
        was = pol.is_within_training_step
        pol.is_within_training_step = True
        try:
            if p.train_fn:
                p.train_fn(self._epoch, self._env_step)
            t = time.time()
            collect_stats = p.train_collector.collect(n_step=p.collection_step_num_env_steps, n_episode=p.collection_step_num_episodes)
            self._timing.train_time_collect += time.time() - t
            should_stop = False
            if p.test_in_train and collect_stats.n_collected_episodes > 0 and p.stop_fn(self._score(collect_stats)):  # :1000-1035
                _, should_stop = self._test_step()
            training_stats = None
            if not should_stop:
                t = time.time()
                training_stats = self.algorithm.update(buffer=p.train_collector.buffer, batch_size=p.batch_size,
                                                       repeat=p.update_step_num_repetitions)
                p.train_collector.reset_buffer(keep_statistics=True)  # :1104
                self._timing.train_time_update += time.time() - t
            return collect_stats, training_stats, should_stop
        finally:
            pol.is_within_training_step = was

    # -- trainer.py:557-614 ---------------------------------------------------------------------------
    def execute_epoch(self) -> EpochStats:
        p = self.params
        self._epoch += 1
        steps_done = 0
        collect_stats = training_stats = None
        while steps_done < p.epoch_num_steps and not self._stop_fn_flag:
            self._current_update_step += 1
            collect_stats, training_stats, self._stop_fn_flag = self._training_step()
            steps_done += collect_stats.n_collected_steps
            self._env_step += collect_stats.n_collected_steps
        test_stats = None
        if not self._stop_fn_flag and p.test_collector is not None:
            test_stats, self._stop_fn_flag = self._test_step()
        return EpochStats(epoch=self._epoch, train_collect_stat=collect_stats, test_collect_stat=test_stats,
                          training_stat=training_stats, info_stat=self._create_info_stats())

    def _create_info_stats(self) -> InfoStats:
        p = self.params
        self._timing.total_time = time.time() - self._t0
        tc = p.test_collector
        return InfoStats(update_step=self._current_update_step, best_score=self._best_score, best_reward=self._best_reward,
                         best_reward_std=self._best_reward_std, train_step=p.train_collector.collect_step,
                         train_episode=p.train_collector.collect_episode, test_step=tc.collect_step if tc is not None else 0,
                         test_episode=tc.collect_episode if tc is not None else 0, timing=self._timing)

    # -- trainer.py:736-758 ---------------------------------------------------------------------------
    def run(self, reset_collectors: bool = True, reset_collector_buffers: bool = False) -> InfoStats:
        self.reset(reset_collectors=reset_collectors, reset_collector_buffers=reset_collector_buffers)
        while self._epoch < self.params.max_epochs and not self._stop_fn_flag:
            self.execute_epoch()
        return self._create_info_stats()
