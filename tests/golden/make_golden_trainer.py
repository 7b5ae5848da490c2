"""Golden fixture `trainer.npz`: event traces of the UNMODIFIED reference OnPolicyTrainer (tianshou/trainer/trainer.py:557-1109)
driving the reference Collector / PPO over DummyVectorEnv of the reference's MoveToRightEnv.  The actor's last bias makes action 1
("right") certain, so every rollout is deterministic and the trace -- the ORDER and ARGUMENTS of train_fn, collect, update,
reset_buffer, test_fn, test collect, stop_fn, save_best_fn, and the per-epoch counters (env_step, update steps, collector step /
episode counters, best reward) -- is a function of the trainer's control flow alone.
Run in the build container:  cd tests/golden && python make_golden_trainer.py"""
from __future__ import annotations

import json

import numpy as np

from make_golden import save  # noqa: E402  (also sets sys.path for the reference + shims + test/base)

SCENARIOS = {
    # test_in_train, stop threshold (None: no stop_fn), collect by episodes
    "plain": dict(max_epochs=3, epoch_num_steps=40, n_step=16, n_episode=None, test_eps=4, batch_size=8, repeat=2, stop=None, test_in_train=False),
    "stop": dict(max_epochs=5, epoch_num_steps=24, n_step=12, n_episode=None, test_eps=3, batch_size=8, repeat=1, stop=0.5, test_in_train=False),
    "test_in_train": dict(max_epochs=4, epoch_num_steps=24, n_step=12, n_episode=None, test_eps=3, batch_size=None, repeat=1, stop=0.5, test_in_train=True),
    "episodes": dict(max_epochs=2, epoch_num_steps=20, n_step=None, n_episode=5, test_eps=2, batch_size=16, repeat=1, stop=None, test_in_train=False),
    "no_test": dict(max_epochs=2, epoch_num_steps=32, n_step=16, n_episode=None, test_eps=0, batch_size=8, repeat=1, stop=None, test_in_train=False),
}
TRAIN_SIZES, TEST_SIZES = [2, 3, 4, 5], [3, 4]


def gen() -> dict:
    import gymnasium as gym
    import torch
    from env import MoveToRightEnv
    from tianshou.algorithm import PPO
    from tianshou.algorithm.modelfree.reinforce import DiscreteActorPolicy
    from tianshou.algorithm.optim import AdamOptimizerFactory
    from tianshou.data import Collector, VectorReplayBuffer
    from tianshou.env import DummyVectorEnv
    from tianshou.trainer import OnPolicyTrainerParams
    from tianshou.utils.net.common import Net
    from tianshou.utils.net.discrete import DiscreteActor, DiscreteCritic

    out: dict = {}
    for name, c in SCENARIOS.items():
        torch.manual_seed(0)
        net = Net(state_shape=(1,), hidden_sizes=[16, 16])
        actor = DiscreteActor(preprocess_net=net, action_shape=2, softmax_output=False)
        critic = DiscreteCritic(preprocess_net=net)
        with torch.no_grad():
            actor.last.model[0].weight.zero_()
            actor.last.model[0].bias.copy_(torch.tensor([-40.0, 40.0]))  # always "right"; lr = 0 keeps it so
        policy = DiscreteActorPolicy(actor=actor, action_space=gym.spaces.Discrete(2))
        alg = PPO(policy=policy, critic=critic, optim=AdamOptimizerFactory(lr=0.0))
        ev: list = []
        train_col = Collector(alg, DummyVectorEnv([lambda x=i: MoveToRightEnv(size=x, sleep=0) for i in TRAIN_SIZES]),
                              VectorReplayBuffer(total_size=400, buffer_num=len(TRAIN_SIZES)))
        test_col = None
        if c["test_eps"]:
            test_col = Collector(alg, DummyVectorEnv([lambda x=i: MoveToRightEnv(size=x, sleep=0) for i in TEST_SIZES]))

        def spy(obj, meth, tag, fmt):
            orig = getattr(obj, meth)

            def wrapped(*a, **k):
                r = orig(*a, **k)
                ev.append([tag] + fmt(a, k, r))
                return r

            setattr(obj, meth, wrapped)

        spy(train_col, "collect", "collect", lambda a, k, r: [k.get("n_step"), k.get("n_episode"), int(r.n_collected_steps), int(r.n_collected_episodes)])
        spy(train_col, "reset_buffer", "reset_buffer", lambda a, k, r: [bool(k.get("keep_statistics", a[0] if a else False))])
        spy(alg, "update", "update", lambda a, k, r: [k.get("batch_size"), k.get("repeat"), int(len(k["buffer"])) if k.get("buffer") is not None else -1,
                                                      int(r.gradient_steps) if hasattr(r, "gradient_steps") else -1])
        if test_col is not None:
            spy(test_col, "collect", "test_collect", lambda a, k, r: [k.get("n_step"), k.get("n_episode"), int(r.n_collected_steps), int(r.n_collected_episodes)])
        kw = dict(train_collector=train_col, test_collector=test_col, max_epochs=c["max_epochs"], epoch_num_steps=c["epoch_num_steps"],
                  collection_step_num_env_steps=c["n_step"], collection_step_num_episodes=c["n_episode"], batch_size=c["batch_size"],
                  update_step_num_repetitions=c["repeat"], test_in_train=c["test_in_train"], verbose=False, show_progress=False,
                  train_fn=lambda epoch, env_step: ev.append(["train_fn", int(epoch), int(env_step)]))
        if test_col is not None:
            kw.update(test_step_num_episodes=c["test_eps"], test_fn=lambda epoch, env_step: ev.append(["test_fn", int(epoch), None if env_step is None else int(env_step)]),
                      save_best_fn=lambda a: ev.append(["save_best_fn"]))
            if c["stop"] is not None:
                thr = c["stop"]

                def stop_fn(score, _thr=thr):
                    ev.append(["stop_fn", round(float(score), 6)])
                    return score >= _thr

                kw["stop_fn"] = stop_fn
        trainer = alg.create_trainer(OnPolicyTrainerParams(**kw))
        trainer.reset()
        ev.append(["reset_done"])
        epochs = []
        while trainer._epoch < c["max_epochs"] and not trainer._stop_fn_flag:
            es = trainer.execute_epoch()
            i = es.info_stat
            row = [int(es.epoch), int(trainer._env_step), int(i.update_step), int(i.train_step), int(i.train_episode), int(i.test_step), int(i.test_episode),
                   float(i.best_reward), float(i.best_score), int(bool(trainer._stop_fn_flag))]
            epochs.append(row)
            ev.append(["epoch_done"] + row[:7])
        out[name + "_events"] = np.array(json.dumps(ev))
        out[name + "_epochs"] = np.array(epochs, dtype=np.float64)
        out[name + "_cfg"] = np.array(json.dumps(c))
    out["train_sizes"], out["test_sizes"] = np.array(TRAIN_SIZES), np.array(TEST_SIZES)
    return out


if __name__ == "__main__":
    r = gen()
    for k in r:
        if k.endswith("_events"):
            print(k, str(r[k])[:600])
    save("trainer.npz", **r)
