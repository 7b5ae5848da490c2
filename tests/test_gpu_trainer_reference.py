"""OnPolicyTrainer against event traces recorded from the UNMODIFIED reference OnPolicyTrainer (tests/golden/trainer.npz,
make_golden_trainer.py; reference tianshou/trainer/trainer.py:557-1109): the order and arguments of train_fn / collect / update /
reset_buffer / test_fn / test collect / stop_fn / save_best_fn and the per-epoch counters (env_step, update steps, collector step
and episode counters, best reward) over five configurations (plain, stop_fn, test_in_train, episode-based collection, no test
collector).  The actor's last bias makes action 1 certain on both sides, so rollouts are deterministic."""
import json
import os

import numpy as np
import pytest
import torch

from _move_env import HostVectorEnv, MoveToRightEnv

pytestmark = pytest.mark.gpu
Z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "trainer.npz"))


@pytest.mark.parametrize("name", ["plain", "stop", "test_in_train", "episodes", "no_test"])
def test_trainer_trace_matches_reference(name):
    from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory, DiscreteActorPolicy
    from tianshou_marl_b200.data import Collector, VectorReplayBuffer
    from tianshou_marl_b200.trainer import OnPolicyTrainer, OnPolicyTrainerParams
    from tianshou_marl_b200.utils import ActorCriticNet

    c = json.loads(str(Z[name + "_cfg"]))
    want_events = json.loads(str(Z[name + "_events"]))
    net = ActorCriticNet(1, 64, 2, seed=0)
    with torch.no_grad():
        net.view("W3").zero_()
        net.view("b3").copy_(torch.tensor([-40.0, 40.0, 0.0]))  # always "right"; lr = 0 keeps it so
    policy = DiscreteActorPolicy(actor=net)
    alg = PPO(policy=policy, critic=net, optim=AdamOptimizerFactory(lr=0.0), minibatch_order="numpy")
    ev: list = []
    train_col = Collector(alg, HostVectorEnv([MoveToRightEnv(int(s)) for s in Z["train_sizes"]]), VectorReplayBuffer(400, len(Z["train_sizes"])))
    test_col = Collector(alg, HostVectorEnv([MoveToRightEnv(int(s)) for s in Z["test_sizes"]])) if c["test_eps"] else None

    def spy(obj, meth, tag, fmt):
        orig = getattr(obj, meth)

        def wrapped(*a, **k):
            r = orig(*a, **k)
            ev.append([tag] + fmt(a, k, r))
            return r

        setattr(obj, meth, wrapped)

    cfmt = lambda a, k, r: [k.get("n_step"), k.get("n_episode"), int(r.n_collected_steps), int(r.n_collected_episodes)]  # noqa: E731
    spy(train_col, "collect", "collect", cfmt)
    spy(train_col, "reset_buffer", "reset_buffer", lambda a, k, r: [bool(k.get("keep_statistics", a[0] if a else False))])
    spy(alg, "update", "update", lambda a, k, r: [k.get("batch_size"), k.get("repeat"), int(len(k["buffer"])), int(r.gradient_steps)])
    if test_col is not None:
        spy(test_col, "collect", "test_collect", cfmt)
    kw = dict(train_collector=train_col, test_collector=test_col, max_epochs=c["max_epochs"], epoch_num_steps=c["epoch_num_steps"],
              collection_step_num_env_steps=c["n_step"], collection_step_num_episodes=c["n_episode"], batch_size=c["batch_size"],
              update_step_num_repetitions=c["repeat"], test_in_train=c["test_in_train"],
              train_fn=lambda epoch, env_step: ev.append(["train_fn", int(epoch), int(env_step)]))
    if test_col is not None:
        kw.update(test_step_num_episodes=c["test_eps"], save_best_fn=lambda a: ev.append(["save_best_fn"]),
                  test_fn=lambda epoch, env_step: ev.append(["test_fn", int(epoch), None if env_step is None else int(env_step)]))
        if c["stop"] is not None:
            def stop_fn(score, _thr=c["stop"]):
                ev.append(["stop_fn", round(float(score), 6)])
                return score >= _thr

            kw["stop_fn"] = stop_fn
    tr = OnPolicyTrainer(alg, OnPolicyTrainerParams(**kw))
    tr.reset()
    ev.append(["reset_done"])
    epochs = []
    while tr._epoch < c["max_epochs"] and not tr._stop_fn_flag:
        es = tr.execute_epoch()
        i = es.info_stat
        row = [int(es.epoch), int(tr._env_step), int(i.update_step), int(i.train_step), int(i.train_episode), int(i.test_step), int(i.test_episode),
               float(i.best_reward), float(i.best_score), int(bool(tr._stop_fn_flag))]
        epochs.append(row)
        ev.append(["epoch_done"] + row[:7])
    assert len(ev) == len(want_events), (len(ev), len(want_events), ev[:12], want_events[:12])
    for k, (g, w) in enumerate(zip(ev, want_events)):
        assert g == w, (k, g, w)
    assert np.allclose(np.array(epochs, dtype=np.float64), Z[name + "_epochs"])
