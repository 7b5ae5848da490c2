"""Checksums of a fused rollout (E = 8192, 30 steps): identical for every TSM_COLLECT_STREAMS setting."""
import sys, torch
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory, DiscreteActorPolicy
from tianshou_marl_b200.data import Collector, VectorReplayBuffer
from tianshou_marl_b200.env import MPEVectorEnv
from tianshou_marl_b200.utils import ActorCriticNet

E, A, T = 8192 + 37, 3, 30
env = MPEVectorEnv("simple_spread_v3", n_envs=E, n_agents=A, seed=5)
buf = VectorReplayBuffer(T * E * A, E * A, index_type="torch")
net = ActorCriticNet(env.obs_dim, 64, 5, seed=3, engine="tc")
pol = DiscreteActorPolicy(actor=net, seed=11)
alg = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=1e-3))
col = Collector(alg, env, buf); col.reset(); pol.is_within_training_step = True
cs = col.collect(n_step=T * E * A)
st = buf._store
print("CHK", " ".join(f"{k}={st[k].double().sum().item():.10e}" for k in ("obs", "obs_next", "act", "rew", "truncated", "logp", "value")),
      f"n_ep={cs.n_collected_episodes} ret={float(cs.returns_stat.mean) if cs.returns_stat else 0:.8f}")
