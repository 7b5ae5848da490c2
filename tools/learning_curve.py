"""Does the drop-in learn?  PPO (tcgen05 engine, device minibatch order) on simple_spread_v3 through Collector / PPO.update:
mean episode return of the training rollouts per iteration.  python tools/learning_curve.py [iterations] [envs] [engine]"""
import sys
import time

import torch

sys.path.insert(0, ".")
from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory, DiscreteActorPolicy  # noqa: E402
from tianshou_marl_b200.data import Collector, VectorReplayBuffer  # noqa: E402
from tianshou_marl_b200.env import MPEVectorEnv  # noqa: E402
from tianshou_marl_b200.utils import ActorCriticNet  # noqa: E402

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 40
E = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
engine = sys.argv[3] if len(sys.argv) > 3 else "tc"
A, T = 3, 25
R = E * A
env = MPEVectorEnv("simple_spread_v3", n_envs=E, n_agents=A, seed=1)
net = ActorCriticNet(env.obs_dim, 64, 5, seed=0, engine=engine)
pol = DiscreteActorPolicy(actor=net, seed=3)
alg = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=7e-4), max_grad_norm=0.5, ent_coef=0.01, gae_lambda=0.95, gamma=0.99,
          minibatch_order="device", seed=4)
col = Collector(pol, env, VectorReplayBuffer(T * R, R, index_type="torch"), episode_list=False)
col.reset()
pol.is_within_training_step = True
t0 = time.time()
for it in range(iters):
    st = col.collect(n_step=T * R)
    s = col.last_episode_summary
    mean_ret = s["ret_sum"] / max(s["count"], 1)
    ts = alg.update(col.buffer, 16384, 4)
    col.reset_buffer(keep_statistics=True)
    if it % 5 == 0 or it == iters - 1:
        print(f"iter {it:3d}: episodes {int(s['count']):6d} mean return {mean_ret:8.3f}  loss {ts.loss.mean:8.4f} vf {ts.vf_loss.mean:8.3f} ent {ts.ent_loss.mean:6.3f}", flush=True)
torch.cuda.synchronize()
print(f"{iters} iterations, {iters * T * R / 1e6:.1f} M agent-steps in {time.time() - t0:.1f} s")
