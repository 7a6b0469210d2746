"""Probe: torch symmetric memory on this box (peer pointers, barrier, capture in a CUDA graph)."""
import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
t = symm_mem.empty(1 << 20, dtype=torch.float32, device=dev)
hdl = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
print(rank, "ptrs", [hex(p) for p in hdl.buffer_ptrs], "signal", [hex(p) for p in hdl.signal_pad_ptrs][:2], flush=True)
t.fill_(rank + 1)
hdl.barrier()
peer = hdl.get_buffer((rank + 1) % world, (1 << 20,), torch.float32)
print(rank, "peer value", peer[:4].tolist(), flush=True)
hdl.barrier()
# graph capture of fill + barrier + peer read
s = torch.cuda.Stream()
out = torch.zeros(4, device=dev)
with torch.cuda.stream(s):
    for _ in range(2):
        t.add_(1.0); hdl.barrier(); out.copy_(peer[:4]); hdl.barrier()
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    t.add_(1.0); hdl.barrier(); out.copy_(peer[:4]); hdl.barrier()
for _ in range(3):
    g.replay()
torch.cuda.synchronize()
print(rank, "after graph", out.tolist(), flush=True)
dist.barrier(); torch.cuda.synchronize(); os._exit(0)
