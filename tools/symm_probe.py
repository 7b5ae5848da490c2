"""Probe: torch symmetric memory gives every rank device pointers to its peers' buffers (NVLink P2P)."""
import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
t = symm_mem.empty(8192, dtype=torch.float32, device="cuda")
hdl = symm_mem.rendezvous(t, group=dist.group.WORLD.group_name)
t.fill_(rank + 1)
dist.barrier(); torch.cuda.synchronize()
print(rank, "ptrs", [hex(p) for p in hdl.buffer_ptrs], "signal", [hex(p) for p in hdl.signal_pad_ptrs], "pad size", getattr(hdl, "signal_pad_size", None), flush=True)
peer = hdl.get_buffer((rank + 1) % world, (8192,), torch.float32)
print(rank, "peer value", float(peer[0]), flush=True)
dist.barrier()
dist.destroy_process_group()
