"""Correctness check of the data-parallel step on the GPUs of a running job (world size > 1): the fused
reduce-scatter + Adam + all-gather over NVLink peer memory (dp_optim.ShardedAdam -> nkbs_dp_adam_step) against the
NCCL route (sum-allreduce of the flat gradient slices + local Adam), same parameters, batches and jitter.
Called by bench.py outside its timed regions (``"dp_check"`` in the JSON line), by scripts/dp_check.py and by
tests/test_gpu_trainer.py::test_fused_dp_step_equals_nccl_route.  Follows the DDP setup nerfstudio's launcher
gives the reference (ref:dataparser/datamanager_z.py:225-231; optimizers ref:train.py:85-98)."""
from __future__ import annotations

import os
from typing import Dict

import torch
import torch.distributed as dist


def _batches(n, R, seed, dev):
    out = []
    for i in range(n):
        g = torch.Generator().manual_seed(seed * 1000 + i)
        o = torch.rand(R, 3, generator=g) - 0.5
        d = torch.randn(R, 3, generator=g)
        d = d / d.norm(dim=-1, keepdim=True)
        b = {"origins": o, "directions": d, "camera_indices": torch.randint(0, 115, (R, 1), generator=g),
             "pixel_area": torch.full((R, 1), 1e-6), "image": torch.rand(R, 3, generator=g)}
        out.append({k: v.to(dev) for k, v in b.items()})
    return out


def run_dp_check(dev, rank: int, world: int, rays: int = 1024, steps: int = 8, graph: bool = True,
                 deterministic: bool = False) -> Dict:
    """Returns {"max_param_diff", "max_param", "rank_spread", "loss_fused", "loss_nccl", "ok"} (identical on every
    rank).  ``rank_spread`` = max |parameter difference| between rank 0 and any rank under the fused route: must be
    exactly 0 (every rank receives the owner's stores).  The two routes sum the ranks' gradients in different orders,
    and Adam (eps 1e-15) turns an entry whose gradient is rounding noise into a step of about lr in either direction,
    so parameters are compared against ``steps * lr`` and the losses at 1e-4 relative.

    ``deterministic``: both routes run with fixed-point gradient accumulation (Trainer(deterministic=True)), the IEEE
    Adam kernel on the NCCL route (the fused kernel's arithmetic) and peer loads instead of the in-switch reduction.
    Every rank's local sums are then order-independent, and with two ranks the cross-rank sum is one commutative
    addition on either route: the parameters must agree bit for bit (``max_param_diff`` == 0); with more ranks NCCL's
    reduction order is its own and the loose bar applies."""
    from . import NerfactoModel, ModelConfig, RayBundle
    from .trainer import Trainer
    batches = _batches(4, rays, 100 + rank, dev)
    keys = ("NKBS_DP_FUSED", "NKBS_ADAM_IEEE", "NKBS_DP_MULTIMEM")
    prev = {k: os.environ.get(k) for k in keys}
    if deterministic:
        os.environ["NKBS_ADAM_IEEE"] = "1"
        os.environ["NKBS_DP_MULTIMEM"] = "0"

    def run(fused: bool, use_graph: bool):
        os.environ["NKBS_DP_FUSED"] = "1" if fused else "0"
        torch.manual_seed(0)
        m = NerfactoModel(ModelConfig()).to(dev).train()
        torch.cuda.manual_seed(7 + rank)
        tr = Trainer(m, rays, use_graph=use_graph, graph_warmup=2, deterministic=deterministic)
        losses = []
        for i in range(steps):
            b = batches[i % len(batches)]
            losses.append(float(tr.train_step(RayBundle(b["origins"], b["directions"], b["pixel_area"], b["camera_indices"]), b)))
        tr.flush()
        torch.cuda.synchronize()
        flat = torch.cat([p.detach().reshape(-1) for p in m.parameters()]).clone()
        tr.close()
        return flat, losses

    try:
        ref, l_nccl = run(False, False)
        got, l_fused = run(True, graph)
    finally:
        for k, v in prev.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    other = got.clone()
    dist.broadcast(other, src=0)
    stats = torch.tensor([float((got - ref).abs().max()), float(ref.abs().max()), float((got - other).abs().max()),
                          l_fused[-1], l_nccl[-1]], device=dev, dtype=torch.float64)
    mx = stats.clone()
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    dist.broadcast(stats, src=0)
    d, scale, spread = float(mx[0]), float(mx[1]), float(mx[2])
    lf, ln = float(stats[3]), float(stats[4])
    ok = spread == 0.0 and abs(lf - ln) <= 1e-4 * abs(ln) and d <= 1.5 * steps * 1e-3
    if deterministic and world == 2:
        ok = ok and d == 0.0
    return {"max_param_diff": d, "max_param": scale, "rank_spread": spread, "loss_fused": lf, "loss_nccl": ln,
            "rays_per_rank": rays, "steps": steps, "graph": bool(graph), "deterministic": bool(deterministic), "ok": bool(ok),
            "what": "fused NVLink reduce-scatter+Adam+all-gather vs NCCL allreduce + local Adam, same seeds"}
