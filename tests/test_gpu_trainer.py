"""OnPolicyTrainer over the real device classes (reference trainer.py:878-906, 1079-1109): a short PPO run on simple_spread."""
import numpy as np
import pytest
import torch


@pytest.mark.gpu
def test_on_policy_trainer_runs_ppo_on_simple_spread():
    from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory, DiscreteActorPolicy
    from tianshou_marl_b200.data import Collector, VectorReplayBuffer
    from tianshou_marl_b200.env import MPEVectorEnv
    from tianshou_marl_b200.trainer import OnPolicyTrainer, OnPolicyTrainerParams
    from tianshou_marl_b200.utils import ActorCriticNet

    E, A, T = 256, 3, 50
    R = E * A
    env = MPEVectorEnv("simple_spread_v3", n_envs=E, n_agents=A, seed=1)
    test_env = MPEVectorEnv("simple_spread_v3", n_envs=64, n_agents=A, seed=2)
    net = ActorCriticNet(env.obs_dim, 64, 5, seed=0, engine="tc")
    pol = DiscreteActorPolicy(actor=net, seed=3, deterministic_eval=True)
    alg = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=3e-4), max_grad_norm=0.5, minibatch_order="device", seed=4)
    train_c = Collector(alg, env, VectorReplayBuffer(T * R, R, index_type="torch"))
    test_c = Collector(alg, test_env, VectorReplayBuffer(30 * 64 * A, 64 * A, index_type="torch"))
    before = net.flat.detach().clone()
    seen = []
    p = OnPolicyTrainerParams(train_collector=train_c, test_collector=test_c, max_epochs=2, epoch_num_steps=2 * T * R,
                              collection_step_num_env_steps=T * R, batch_size=4096, update_step_num_repetitions=2,
                              test_step_num_episodes=64 * A, train_fn=lambda ep, st: seen.append((ep, st)))
    tr = OnPolicyTrainer(alg, p)
    info = tr.run(reset_collector_buffers=True)
    assert info.update_step == 4 and info.train_step == 4 * T * R and seen == [(1, 0), (1, T * R), (2, 2 * T * R), (2, 3 * T * R)]
    assert info.test_episode >= 3 * 64 * A and np.isfinite(info.best_reward) and info.best_score == info.best_reward
    assert len(train_c.buffer) == 0                                   # reset_buffer(keep_statistics=True) after every update
    assert not torch.equal(before, net.flat.detach())                 # the updates moved the parameters
    assert pol.is_within_training_step is False
    ep = tr.execute_epoch()                                            # one more epoch by hand
    assert ep.epoch == 3 and ep.training_stat.gradient_steps == 2 * (T * R // 4096) and ep.test_collect_stat.n_collected_episodes >= 64 * A
