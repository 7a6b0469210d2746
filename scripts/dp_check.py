"""2+ GPU check (torchrun): the fused NVLink reduce-Adam-broadcast step equals NCCL allreduce + per-rank Adam.
Prints max |param difference| between the two routes after a few steps, and between ranks."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
import nerf_kbs_b200 as nk
from nerf_kbs_b200.trainer import Trainer
from bench import make_batches
R, STEPS = 1024, int(os.environ.get("STEPS", "8"))
batches = [{k: v.to(dev) for k, v in b.items()} for b in make_batches(4, R, 100 + rank, False, False)]
def run(fused, graph):
    os.environ["NKBS_DP_FUSED"] = "1" if fused else "0"
    torch.manual_seed(0)
    m = nk.NerfactoModel(nk.ModelConfig()).to(dev).train()
    torch.cuda.manual_seed(7 + rank)
    tr = Trainer(m, R, use_graph=graph)
    losses = []
    for i in range(STEPS):
        b = batches[i % len(batches)]
        losses.append(float(tr.train_step(nk.RayBundle(b["origins"], b["directions"], b["pixel_area"], b["camera_indices"]), b)))
    torch.cuda.synchronize()
    return [p.detach().clone() for p in m.parameters()], losses
ref, l0 = run(False, False)
for graph in (False, True):
    got, l1 = run(True, graph)
    d = max(float((a - b).abs().max()) for a, b in zip(got, ref))
    scale = max(float(b.abs().max()) for b in ref)
    # ranks must hold identical parameters
    flat = torch.cat([p.reshape(-1) for p in got])
    other = flat.clone()
    dist.broadcast(other, src=0)
    dr = float((flat - other).abs().max())
    if rank == 0:
        print(f"graph={graph}: fused vs nccl max|dp| = {d:.3e} (max|p| {scale:.3e}); rank spread {dr:.3e}; losses {l1[-1]:.6f} vs {l0[-1]:.6f}", flush=True)
    else:
        print(f"  rank {rank}: spread vs rank 0 = {dr:.3e}", flush=True)
dist.barrier(); torch.cuda.synchronize(); os._exit(0)
