"""2+ GPU check (torchrun): the fused NVLink reduce-Adam-broadcast step equals NCCL allreduce + per-rank Adam
(nerf_kbs_b200/dp_check.py).  Prints one JSON line per mode (eager, graph, graph + deterministic accumulation) on rank 0."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
from nerf_kbs_b200.dp_check import run_dp_check
for graph, det in ((False, False), (True, False), (True, True)):
    r = run_dp_check(dev, rank, world, rays=1024, steps=int(os.environ.get("STEPS", "8")), graph=graph, deterministic=det)
    if rank == 0:
        print("DP_CHECK " + json.dumps(r), flush=True)
dist.barrier(); torch.cuda.synchronize(); os._exit(0)
