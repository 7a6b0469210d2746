"""Data-parallel optimizer step over NVLink peer memory (include/nkbs.h: nkbs_dp_adam_step).

Replaces "NCCL allreduce of the gradients, then Adam on every rank" (ref:dataparser/datamanager_z.py:
225-231, ref:train.py:85-98) by one kernel per parameter group: parameters and gradients of all groups live in two
flat SYMMETRIC buffers (same layout on every GPU, peer-mapped through torch's symmetric memory), every rank owns a
1/world slice of each group, reads that slice of the gradient from all peers, applies Adam with moments only it
keeps, and stores the new parameters into every peer's buffer.  Two device-side cross-rank barriers bracket the
kernels; everything is capturable in the step's CUDA graph.  torch.optim.Adam semantics per group (a group that is
not active in a step is skipped, its step count does not advance).  The flat buffers are a ``ParamArena`` (optim.py):
the backward kernels accumulate into ``G`` directly, and the step leaves ``G`` zeroed (memset nodes after the closing
barrier) for the next backward pass."""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Sequence

import torch
import torch.distributed as dist

from ._lib import call, stream


from .optim import ParamArena, flat_layout  # noqa: F401  (flat_layout: re-exported for the CPU layout test)


class ShardedAdam:
    """Optimizer of the fused data-parallel step.  ``arena``: a ``ParamArena`` built with ``world`` / ``rank`` of the
    process group and ``alloc=ShardedAdam.symmetric_alloc(...)`` so that ``P`` and ``G`` are peer-mapped."""

    @staticmethod
    def make_arena(groups: Dict[str, Sequence[torch.nn.Parameter]], process_group=None, deterministic: bool = False) -> ParamArena:
        import torch.distributed._symmetric_memory as symm_mem
        pg = process_group if process_group is not None else dist.group.WORLD
        dev = next(p for ps in groups.values() for p in ps).device
        return ParamArena(groups, world=dist.get_world_size(pg), rank=dist.get_rank(pg),
                          alloc=lambda n: symm_mem.empty(n, dtype=torch.float32, device=dev), deterministic=deterministic)

    def __init__(self, arena: ParamArena, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-15,
                 weight_decay: float = 0.0, process_group=None):
        import torch.distributed._symmetric_memory as symm_mem
        pg = process_group if process_group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(pg), dist.get_rank(pg)
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.arena = arena
        self.layout = arena.layout
        dev = arena.device
        self.hP = symm_mem.rendezvous(arena.P, pg.group_name)
        self.hG = symm_mem.rendezvous(arena.G, pg.group_name)
        self._p_ptrs = (C.c_void_p * self.world)(*[int(x) for x in self.hP.buffer_ptrs])
        self._g_ptrs = (C.c_void_p * self.world)(*[int(x) for x in self.hG.buffer_ptrs])
        # NVSwitch multicast addresses of the two buffers (0 where the fabric / driver has no NVLS): in-switch reduction
        # of the gradients and replicated parameter stores (csrc/adam.cu).  Used from 4 ranks on: measured on the
        # B200 box, 4096-ray step, peer loads / stores vs multimem: N=2 0.814 / 0.886 ms, N=4 0.867 / 0.840 ms,
        # N=8 0.905 / 0.841 ms (profiles/r02_dp_multimem.txt) — per GPU the switch route moves 2/N of the parameter
        # bytes instead of 2 (N-1)/N, but one multimem round trip is slower than a peer access.
        # NKBS_DP_MULTIMEM=0 / 1 forces the choice.
        import os
        g_mc, p_mc = int(getattr(self.hG, "multicast_ptr", 0) or 0), int(getattr(self.hP, "multicast_ptr", 0) or 0)
        want = os.environ.get("NKBS_DP_MULTIMEM", "auto")
        self.multimem = bool(g_mc and p_mc) and (want == "1" or (want not in ("0", "1") and self.world >= 4))
        self._g_mc, self._p_mc = (g_mc, p_mc) if self.multimem else (None, None)
        self.exp_avg = {k: torch.zeros(v["n"], device=dev) for k, v in self.layout.items()}
        self.exp_avg_sq = {k: torch.zeros(v["n"], device=dev) for k, v in self.layout.items()}
        self.step_t = {k: torch.zeros((), device=dev, dtype=torch.float32) for k in self.layout}
        self._counters = {k: torch.zeros((), device=dev, dtype=torch.int32) for k in self.layout}
        torch.cuda.synchronize(dev)
        self.hP.barrier()
        torch.cuda.synchronize(dev)

    @torch.no_grad()
    def step(self, active: Optional[Sequence[str]] = None) -> None:
        """Gradient mean over ranks + Adam + parameter broadcast for every group in ``active`` (default: all).  The
        local gradients of those groups are zero afterwards (ready for the next backward pass to accumulate into)."""
        active = list(active) if active is not None else list(self.layout)
        if not active:
            return
        # deterministic mode: every rank's local sums are order-independent; the cross-rank sum is taken in rank order
        # by the peer-load route (the in-switch multimem reduction has no documented order)
        self.arena.flush_fixed(active)
        self.hG.barrier()  # every rank's gradients are complete and visible
        for name in active:
            lay = self.layout[name]
            call("nkbs_dp_adam_step", self.world, self.rank, self._g_ptrs, self._p_ptrs, lay["lo"], lay["n"],
                 self.exp_avg[name].data_ptr(), self.exp_avg_sq[name].data_ptr(), float(self.lr), float(self.betas[0]),
                 float(self.betas[1]), float(self.eps), float(self.weight_decay), self.step_t[name].data_ptr(),
                 self._counters[name].data_ptr(), self._g_mc, self._p_mc, stream())
        self.hP.barrier()  # every rank's parameters are complete (and every peer has read this rank's gradients)
        for name in active:
            lay = self.layout[name]
            call("nkbs_zero_async", self.arena.G.data_ptr() + 4 * lay["start"], 4 * lay["len"], stream())
