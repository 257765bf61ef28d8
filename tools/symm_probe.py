"""2-rank probe: does torch symmetric memory (CUDA backend) rendezvous on this box, and do peer stores land?"""
import os
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as sm

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
t = sm.empty(1 << 20, dtype=torch.int32, device=dev)
t.zero_()
hdl = sm.rendezvous(t, dist.group.WORLD.group_name)
print(rank, "backend", sm.get_backend(dev) if hasattr(sm, "get_backend") else None, "ptrs", [hex(p) for p in hdl.buffer_ptrs], flush=True)
peer = (rank + 1) % world
remote = hdl.get_buffer(peer, (1 << 20,), torch.int32)
hdl.barrier(channel=0)
remote[rank * 16:(rank + 1) * 16] = rank + 100
hdl.barrier(channel=0)
torch.cuda.synchronize()
src = (rank - 1) % world
print(rank, "got", t[src * 16:src * 16 + 2].tolist(), flush=True)
# bandwidth of a plain peer copy
big = sm.empty(64 << 20, dtype=torch.uint8, device=dev)
h2 = sm.rendezvous(big, dist.group.WORLD.group_name)
rb = h2.get_buffer(peer, (64 << 20,), torch.uint8)
loc = torch.empty(64 << 20, dtype=torch.uint8, device=dev)
h2.barrier(channel=0)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for _ in range(3):
    rb.copy_(loc)
e0.record()
for _ in range(10):
    rb.copy_(loc)
e1.record()
torch.cuda.synchronize()
print(rank, "peer write GB/s", 10 * 64 / 1024 / (e0.elapsed_time(e1) * 1e-3), flush=True)
h2.barrier(channel=0)
dist.destroy_process_group()
