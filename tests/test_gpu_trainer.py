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


def _train_simple_spread(engine: str, iters: int = 21, E: int = 4096) -> list[float]:
    from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory, DiscreteActorPolicy
    from tianshou_marl_b200.data import Collector, VectorReplayBuffer
    from tianshou_marl_b200.env import MPEVectorEnv
    from tianshou_marl_b200.utils import ActorCriticNet

    A, T = 3, 25
    R = E * A
    env = MPEVectorEnv("simple_spread_v3", n_envs=E, n_agents=A, seed=1)
    net = ActorCriticNet(env.obs_dim, 64, 5, seed=0, engine=engine)
    pol = DiscreteActorPolicy(actor=net, seed=3)
    alg = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=7e-4), max_grad_norm=0.5, minibatch_order="device", seed=4)
    col = Collector(pol, env, VectorReplayBuffer(T * R, R, index_type="torch"), episode_list=False)
    col.reset()
    pol.is_within_training_step = True
    curve = []
    for _ in range(iters):
        col.collect(n_step=T * R)
        s = col.last_episode_summary
        assert s["count"] == R  # every row finished exactly one 25-step episode
        curve.append(s["ret_sum"] / s["count"])
        alg.update(col.buffer, 16384, 4)
        col.reset_buffer(keep_statistics=True)
    return curve


@pytest.mark.gpu
def test_ppo_learns_simple_spread_on_both_engines():
    """The path does what it is for: 21 collect + update iterations (6.5 M agent-steps) lift the mean episode return of 12 288
    episodes per iteration from the random policy's -26.5 to about -19.7, and the tcgen05 engine (fp16 operands) learns the
    same curve as the exact fp32 engine (tools/learning_curve.py prints it)."""
    tc = _train_simple_spread("tc")
    fp = _train_simple_spread("fp32")
    for c in (tc, fp):
        assert -28.0 < c[0] < -25.0, c[0]           # random policy
        assert c[-1] > c[0] + 5.0, (c[0], c[-1])    # measured: +6.9 / +6.8
        assert c[-1] > c[10] > c[3] > c[0], c       # and it keeps improving
    assert abs(tc[0] - fp[0]) < 0.05                # same environments, same seeds: the first rollout differs by fp16 rounding only
    assert abs(tc[-1] - fp[-1]) < 0.6, (tc[-1], fp[-1])
