"""Data-parallel plumbing of the PPO update: one process per GPU, environments sharded by rank, and ONE
exchange -- the flat parameter gradient of each minibatch.

`PeerGradients` owns that exchange for the fused path: every rank's receive area ([2 parities][world][n]) and flag pad
live in torch symmetric memory, so each rank holds device pointers to all of them (NVLink peer stores), and the optimizer
kernel (csrc/adam.cuh: push my gradient to every rank, rank barrier, sum the received slots in rank order, clip, Adam) does
gradient exchange and optimizer step in one launch (tsm_ppo_epoch_dist).  torch.distributed is used for the rendezvous only.  `shard_seed`, `grad_scale` and
`allreduce_mean_` are the host-side pieces that also run on CPU ranks (gloo) in tests/test_dist_cpu.py."""
from __future__ import annotations

import ctypes as C
from typing import Any

import torch
import torch.distributed as dist

from .. import _lib

FLAG_WORDS = 64  # uint32 slots behind the gradient blocks (TSM_MAX_RANKS = 8 used)


def world_info() -> tuple[int, int]:
    """(rank, world_size); (0, 1) when torch.distributed is not initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_seed(base_seed: int, rank: int) -> int:
    """Environment / sampling seeds differ per rank (independent shards); parameter init seeds must not."""
    return int(base_seed) + 1000003 * int(rank)


def grad_scale(world: int) -> float:
    """Each rank's minibatch loss is a mean over its own rows; the global gradient is the mean of the rank gradients."""
    return 1.0 / float(world)


def allreduce_mean_(flat: torch.Tensor) -> torch.Tensor:
    """Reference exchange (any backend, also gloo on CPU): sum over ranks, then scale.  The fused CUDA path computes the
    same quantity inside the optimizer kernel; this is the fallback and the CPU-testable statement of it."""
    rank, world = world_info()
    if world > 1:
        dist.all_reduce(flat)
        flat.mul_(grad_scale(world))
    return flat


class PeerGradients:
    """Symmetric-memory receive areas [2][world][n_stride] + flag pad of every rank, as a tsm_peers struct."""

    def __init__(self, n_params: int, device: torch.device) -> None:
        import torch.distributed._symmetric_memory as symm_mem

        self.rank, self.world = world_info()
        if self.world > 8:
            raise _lib.TsmError("the fused gradient exchange covers one node (<= 8 ranks)")
        self.n_params = int(n_params)
        self.n_stride = (self.n_params + 3) // 4 * 4
        n_pad = 2 * self.world * self.n_stride
        n_ll = 2 * n_pad  # the persistent epoch kernel's receive area: one (value, step) pair per element
        self.block = symm_mem.empty(n_pad + FLAG_WORDS + n_ll, dtype=torch.float32, device=device)
        self.block.zero_()
        self.handle = symm_mem.rendezvous(self.block, group=dist.group.WORLD.group_name)
        self.status = torch.zeros(1, dtype=torch.int32, device=device)
        self.seq = 0  # optimizer steps taken through this exchange
        p = _lib.TsmPeers()
        p.world, p.rank = self.world, self.rank
        for q in range(self.world):
            base = int(self.handle.buffer_ptrs[q])
            p.recv[q] = base
            p.flags[q] = base + 4 * n_pad
            p.ll[q] = base + 4 * (n_pad + FLAG_WORDS)
        p.status = self.status.data_ptr()
        self.desc = p
        torch.cuda.synchronize(device)
        dist.barrier()  # every rank has zeroed its block before anyone's kernels touch it

    def byref(self) -> Any:
        return C.byref(self.desc)

    def check(self) -> None:
        if int(self.status.item()) != 0:
            raise _lib.TsmError("cross-rank gradient barrier timed out (a peer rank did not reach the optimizer step)")
