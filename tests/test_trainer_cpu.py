"""Host-side loop semantics of OnPolicyTrainer (reference tianshou/trainer/trainer.py:557-614, 640-707, 736-758,
878-906, 1079-1109) with stand-in collector / algorithm objects: no device needed."""
import types

import pytest

from tianshou_marl_b200.trainer import OnPolicyTrainer, OnPolicyTrainerParams


class FakeStats:
    def __init__(self, steps, episodes, mean=0.0):
        self.n_collected_steps, self.n_collected_episodes = steps, episodes
        self.returns_stat = types.SimpleNamespace(mean=mean, std=0.5)


class FakeCollector:
    def __init__(self, log, name, means=None):
        self.log, self.name, self.means = log, name, list(means or [])
        self.buffer = object()
        self.collect_step = self.collect_episode = 0

    def reset(self, reset_buffer=True, reset_stats=True, gym_reset_kwargs=None):
        self.log.append((self.name, "reset", reset_buffer, reset_stats))

    def reset_buffer(self, keep_statistics=False):
        self.log.append((self.name, "reset_buffer", keep_statistics))

    def collect(self, n_step=None, n_episode=None):
        self.log.append((self.name, "collect", n_step, n_episode, FakePolicy.current.is_within_training_step))
        steps = n_step if n_step is not None else 25 * n_episode
        self.collect_step += steps
        self.collect_episode += n_episode or 1
        return FakeStats(steps, n_episode or 1, self.means.pop(0) if self.means else 0.0)


class FakePolicy:
    current = None

    def __init__(self):
        self.is_within_training_step = False
        FakePolicy.current = self


class FakeAlg:
    def __init__(self, log):
        self.policy, self.log = FakePolicy(), log

    def update(self, buffer, batch_size, repeat):
        self.log.append(("alg", "update", batch_size, repeat, self.policy.is_within_training_step))
        return types.SimpleNamespace(train_time=0.0)


def test_param_validation_matches_reference():
    c = FakeCollector([], "train")
    with pytest.raises(ValueError):  # trainer.py:277-280
        OnPolicyTrainerParams(train_collector=c, collection_step_num_env_steps=10, collection_step_num_episodes=2)
    with pytest.raises(ValueError):
        OnPolicyTrainerParams(train_collector=c, collection_step_num_env_steps=None, collection_step_num_episodes=None)
    with pytest.raises(ValueError):  # :211-214
        OnPolicyTrainerParams(train_collector=c, stop_fn=lambda s: True)
    with pytest.raises(ValueError):  # :281-282
        OnPolicyTrainerParams(train_collector=c, test_collector=FakeCollector([], "test"), test_in_train=True)
    with pytest.raises(ValueError):  # :224-228
        OnPolicyTrainerParams(train_collector=c, test_collector=FakeCollector([], "test"), test_step_num_episodes=0)


def test_epoch_loop_collect_update_reset_order_and_counts():
    log = []
    train, test = FakeCollector(log, "train"), FakeCollector(log, "test", means=[-3.0, -2.0, -1.0])
    alg = FakeAlg(log)
    p = OnPolicyTrainerParams(train_collector=train, test_collector=test, max_epochs=2, epoch_num_steps=250,
                              collection_step_num_env_steps=100, batch_size=32, update_step_num_repetitions=4, test_step_num_episodes=3)
    info = OnPolicyTrainer(alg, p).run()
    # ceil(250 / 100) = 3 training steps per epoch, each collect -> update -> reset_buffer(keep_statistics=True)
    core = [e for e in log if e[1] in ("collect", "update", "reset_buffer") and e[0] != "test"]
    assert len(core) == 2 * 3 * 3
    for k in range(0, len(core), 3):
        assert core[k] == ("train", "collect", 100, None, True)        # inside a training step (trainer.py:888)
        assert core[k + 1] == ("alg", "update", 32, 4, True)
        assert core[k + 2] == ("train", "reset_buffer", True)            # :1104
    tests = [e for e in log if e[0] == "test" and e[1] == "collect"]
    assert tests == [("test", "collect", None, 3, False)] * 3            # initial test step + one per epoch, outside training steps
    assert info.update_step == 6 and info.train_step == 600 and info.test_episode == 9
    assert info.best_reward == -1.0 and info.best_score == -1.0
    assert alg.policy.is_within_training_step is False


def test_stop_fn_ends_training_after_the_test_step():
    log = []
    train, test = FakeCollector(log, "train"), FakeCollector(log, "test", means=[-5.0, 1.0, 2.0])
    p = OnPolicyTrainerParams(train_collector=train, test_collector=test, max_epochs=10, epoch_num_steps=100,
                              collection_step_num_env_steps=100, stop_fn=lambda s: s >= 0.5)
    tr = OnPolicyTrainer(FakeAlg(log), p)
    info = tr.run()
    assert tr._epoch == 1 and info.update_step == 1 and info.best_score == 1.0


def test_episode_mode_and_no_test_collector():
    log = []
    train = FakeCollector(log, "train")
    p = OnPolicyTrainerParams(train_collector=train, max_epochs=1, epoch_num_steps=100, collection_step_num_env_steps=None,
                              collection_step_num_episodes=2, batch_size=None)
    info = OnPolicyTrainer(FakeAlg(log), p).run()
    assert [e for e in log if e[1] == "collect"] == [("train", "collect", None, 2, True)] * 2  # 2 x 50 steps reach 100
    assert info.test_step == 0 and info.best_score == 0.0  # the reference initialises the best score to 0.0 (trainer.py:367; trainer.npz "no_test")
