"""Data-parallel optimizer step over NVLink peer memory (include/nkbs.h: nkbs_dp_adam_step).

Replaces "NCCL allreduce of the gradients, then Adam on every rank" (ref:dataparser/datamanager_z.py:
225-231, ref:train.py:85-98) by one kernel per parameter group: parameters and gradients of all groups live in two
flat SYMMETRIC buffers (same layout on every GPU, peer-mapped through torch's symmetric memory), every rank owns a
1/world slice of each group, reads that slice of the gradient from all peers, applies Adam with moments only it
keeps, and stores the new parameters into every peer's buffer.  Two device-side cross-rank barriers bracket the
kernels; everything is capturable in the step's CUDA graph.  torch.optim.Adam semantics per group (a group without
gradients in a step is skipped, its step count does not advance)."""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Sequence

import torch
import torch.distributed as dist

from ._lib import call, stream


def _round_up(n: int, m: int) -> int:
    return (n + m - 1) // m * m


def flat_layout(group_numels: Dict[str, Sequence[int]], world: int, rank: int) -> Dict[str, Dict]:
    """Offsets of every tensor in the flat buffers and this rank's slice of every group: tensors are 4-element
    aligned, every group is padded to a multiple of 4 * world elements and cut into ``world`` equal slices."""
    layout, off = {}, 0
    for name, numels in group_numels.items():
        start, offs = off, []
        for n in numels:
            offs.append(off)
            off += _round_up(n, 4)
        off = start + _round_up(off - start, 4 * world)
        shard = (off - start) // world
        layout[name] = {"start": start, "len": off - start, "offs": offs, "lo": start + rank * shard, "n": shard}
    layout["__total__"] = off
    return layout


class ShardedAdam:
    def __init__(self, groups: Dict[str, Sequence[torch.nn.Parameter]], lr: float = 1e-3, betas=(0.9, 0.999),
                 eps: float = 1e-15, weight_decay: float = 0.0, process_group=None):
        import torch.distributed._symmetric_memory as symm_mem
        pg = process_group if process_group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(pg), dist.get_rank(pg)
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.groups: Dict[str, List[torch.nn.Parameter]] = {k: [p for p in v if p.requires_grad] for k, v in groups.items()}
        first = next(p for ps in self.groups.values() for p in ps)
        dev = first.device
        if dev.type != "cuda":
            raise RuntimeError("ShardedAdam runs on CUDA parameters only (no CPU fallback)")
        for ps in self.groups.values():
            for p in ps:
                if p.dtype != torch.float32 or not p.is_contiguous():
                    raise ValueError("ShardedAdam takes contiguous fp32 parameters")
        self.layout = flat_layout({k: [p.numel() for p in ps] for k, ps in self.groups.items()}, self.world, self.rank)
        self.total = self.layout.pop("__total__")
        self.P = symm_mem.empty(self.total, dtype=torch.float32, device=dev)
        self.G = symm_mem.empty(self.total, dtype=torch.float32, device=dev)
        self.P.zero_()
        self.G.zero_()
        self.hP = symm_mem.rendezvous(self.P, pg.group_name)
        self.hG = symm_mem.rendezvous(self.G, pg.group_name)
        self._p_ptrs = (C.c_void_p * self.world)(*[int(x) for x in self.hP.buffer_ptrs])
        self._g_ptrs = (C.c_void_p * self.world)(*[int(x) for x in self.hG.buffer_ptrs])
        # parameters move into the symmetric buffer (their storage is swapped, the Parameter objects stay)
        self.grad_view: Dict[int, torch.Tensor] = {}
        with torch.no_grad():
            for name, ps in self.groups.items():
                for p, o in zip(ps, self.layout[name]["offs"]):
                    view = self.P[o:o + p.numel()].view(p.shape)
                    view.copy_(p.data)
                    p.data = view
                    self.grad_view[id(p)] = self.G[o:o + p.numel()].view(p.shape)
                    if p.numel() >= (1 << 22):
                        # large tables: the backward kernels scatter straight into the symmetric gradient buffer
                        p._nkbs_grad_sink = self.grad_view[id(p)]
        self.exp_avg = {k: torch.zeros(v["n"], device=dev) for k, v in self.layout.items()}
        self.exp_avg_sq = {k: torch.zeros(v["n"], device=dev) for k, v in self.layout.items()}
        self.step_t = {k: torch.zeros((), device=dev, dtype=torch.float32) for k in self.layout}
        torch.cuda.synchronize(dev)
        self.hP.barrier()
        torch.cuda.synchronize(dev)

    def zero_grad(self, set_to_none: bool = True) -> None:
        for ps in self.groups.values():
            for p in ps:
                p.grad = None

    def active_groups(self) -> List[str]:
        """Groups whose parameters received gradients in this backward (all of a group or none)."""
        out = []
        for name, ps in self.groups.items():
            have = [p.grad is not None for p in ps]
            if any(have):
                if not all(have):
                    missing = sum(1 for h in have if not h)
                    raise RuntimeError(f"ShardedAdam: {missing} parameters of group '{name}' have no gradient "
                                       "(torch.optim.Adam would skip them individually; shard-wise Adam cannot)")
                out.append(name)
        return out

    @torch.no_grad()
    def step(self) -> None:
        active = self.active_groups()
        if not active:
            return
        src, dst = [], []
        for name in active:
            for p in self.groups[name]:
                gv = self.grad_view[id(p)]
                if p.grad.data_ptr() != gv.data_ptr():
                    src.append(p.grad.reshape(gv.shape).float())
                    dst.append(gv)
        if dst:
            torch._foreach_copy_(dst, src)
        self.hG.barrier()  # every rank's gradients are complete and visible
        for name in active:
            lay = self.layout[name]
            self.step_t[name].add_(1.0)
            call("nkbs_dp_adam_step", self.world, self.rank, self._g_ptrs, self._p_ptrs, lay["lo"], lay["n"],
                 self.exp_avg[name].data_ptr(), self.exp_avg_sq[name].data_ptr(), float(self.lr), float(self.betas[0]),
                 float(self.betas[1]), float(self.eps), float(self.weight_decay), self.step_t[name].data_ptr(), stream())
        self.hP.barrier()  # every rank's parameters are complete before anyone runs the next forward
