"""Two-or-more-rank check of the fused gradient exchange (run under torchrun, one rank per GPU):
  * parameters stay bit-identical across ranks after updates through tsm_ppo_epoch_dist (peer memory),
  * they match the per-minibatch NCCL all-reduce path from the same start (tolerance: fp32 atomics order).
Launched by tests/test_gpu_dist.py; prints DIST_OK on rank 0."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main() -> None:
    from tianshou_marl_b200.algorithm import PPO, AdamOptimizerFactory, DiscreteActorPolicy
    from tianshou_marl_b200.data import Collector, VectorReplayBuffer
    from tianshou_marl_b200.env import MPEVectorEnv
    from tianshou_marl_b200.utils import ActorCriticNet

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    E, A, T = 256, 3, 16
    res = {}
    for engine in ("fp32", "tc"):
        for sync in ("nccl", "peer"):
            env = MPEVectorEnv("simple_spread_v3", n_envs=E, n_agents=A, seed=100 + rank)  # a different shard per rank
            buf = VectorReplayBuffer(T * E * A, E * A, index_type="torch")
            net = ActorCriticNet(env.obs_dim, 64, 5, seed=0, engine=engine)  # the same init on every rank
            pol = DiscreteActorPolicy(actor=net, seed=7 + rank)
            alg = PPO(policy=pol, critic=net, optim=AdamOptimizerFactory(lr=1e-3), max_grad_norm=0.5, minibatch_order="device", seed=3,
                      grad_sync=sync)
            col = Collector(alg, env, buf)
            col.reset()
            pol.is_within_training_step = True
            for _ in range(2):
                col.collect(n_step=T * E * A)
                st = alg.update(buf, 1024, 2)
                col.reset_buffer(keep_statistics=True)
            flat = net.flat.detach().clone()
            gathered = [torch.empty_like(flat) for _ in range(world)]
            dist.all_gather(gathered, flat)
            for q in range(1, world):
                assert torch.equal(gathered[0], gathered[q]), (engine, sync, "parameters differ across ranks", (gathered[0] - gathered[q]).abs().max())
            assert np.isfinite(st.loss.mean)
            res[(engine, sync)] = flat
        a, b = res[(engine, "nccl")], res[(engine, "peer")]
        tol = 2e-5 if engine == "fp32" else 2e-3
        assert torch.allclose(a, b, atol=tol), (engine, (a - b).abs().max().item())
        if rank == 0:
            print(engine, "peer vs nccl max diff", (a - b).abs().max().item(), flush=True)
    dist.barrier()
    if rank == 0:
        print("DIST_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
