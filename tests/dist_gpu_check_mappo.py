"""Two-or-more-rank check of engine 2's gradient exchange (run under torchrun, one rank per GPU): MAPPO updates with the
exchange inside the optimizer kernel (NVLink peer loads, tsm_tcw_adam_peers) keep the parameters bit-identical across ranks and
match the per-minibatch NCCL all-reduce path from the same start.  Launched by tests/test_gpu_dist.py; prints DIST_OK on rank 0."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main() -> None:
    from tianshou_marl_b200.algorithm import AdamOptimizerFactory, DiscreteActorPolicy, MAPPO
    from tianshou_marl_b200.algorithm.ctde import CentralizedCritic, DecentralizedActor
    from tianshou_marl_b200.data import Collector, VectorReplayBuffer
    from tianshou_marl_b200.env import MPEVectorEnv

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    E, A, T = 256, 4, 16
    res = {}
    for sync in ("nccl", "peer"):
        env = MPEVectorEnv("simple_spread_v3", n_envs=E, n_agents=A, seed=100 + rank)  # a different shard per rank
        O = env.obs_dim
        buf = VectorReplayBuffer(T * E * A, E * A, index_type="torch")
        actor = DecentralizedActor(O, 5, 128, engine="tcw", orthogonal=True, seed=0)  # the same init on every rank
        critic = CentralizedCritic(A * O, A, 128, engine="tcw", orthogonal=True, seed=1)
        pol = DiscreteActorPolicy(actor=actor, seed=7 + rank)
        alg = MAPPO(policy=pol, critic=critic, n_agents=A, optim=AdamOptimizerFactory(lr=1e-3), max_grad_norm=0.5, grad_sync=sync)
        col = Collector(alg, env, buf)
        col.reset()
        pol.is_within_training_step = True
        np.random.seed(5)  # the same super-tile order on every rank and for both paths
        for _ in range(2):
            col.collect(n_step=T * E * A)
            st = alg.update(buf, 8 * 128 * A, 2)  # 4 minibatches per epoch
            col.reset_buffer(keep_statistics=True)
        assert (alg._peers is not None) == (sync == "peer")
        flat = torch.cat([actor.flat.detach(), critic.flat.detach()]).clone()
        gathered = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat)
        for q in range(1, world):
            assert torch.equal(gathered[0], gathered[q]), (sync, "parameters differ across ranks", (gathered[0] - gathered[q]).abs().max())
        assert np.isfinite(st.loss.mean)
        res[sync] = flat
    a, b = res["nccl"], res["peer"]
    assert torch.allclose(a, b, atol=2e-3), (a - b).abs().max().item()  # fp32 atomics order + sign-like early Adam steps
    if rank == 0:
        print("mappo peer vs nccl max diff", (a - b).abs().max().item(), flush=True)
    dist.barrier()
    if rank == 0:
        print("DIST_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
