"""Probe: does torch's symmetric memory (peer pointers over NVLink) work on this box?  torchrun --nproc-per-node 2"""
import os
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
t = symm_mem.empty((1024, 56), dtype=torch.float64, device=f"cuda:{local}")
hdl = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
print(rank, "buffer_ptrs", [hex(p) for p in hdl.buffer_ptrs], "signal_pad_ptrs", [hex(p) for p in hdl.signal_pad_ptrs],
      "multicast", hex(hdl.multicast_ptr) if hdl.multicast_ptr else None, flush=True)
t.fill_(rank + 1)
hdl.barrier()
peer = hdl.get_buffer((rank + 1) % world, (1024, 56), torch.float64)
print(rank, "peer value", peer[0, 0].item(), flush=True)
hdl.barrier()
dist.destroy_process_group()
