"""Adam over the nkbs multi-tensor kernel (include/nkbs.h: nkbs_adam_step).

Stands in for the ``AdamOptimizerConfig(lr=1e-3, eps=1e-15)`` optimizers of the reference's two
parameter groups (ref:train.py:85-98, groups ref:nerfacto.py:359-363).  torch.optim.Adam semantics
(amsgrad off, no scheduler: the reference sets ``"scheduler": None``).  The step count lives on the
device so that a captured CUDA graph advances it on every replay."""
from __future__ import annotations

import ctypes as C
from typing import Dict, Iterable, List, Optional, Sequence

import torch

from ._lib import call, stream


class Adam:
    def __init__(self, params: Iterable[torch.nn.Parameter], lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-15,
                 weight_decay: float = 0.0):
        self.params = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("Adam got no parameters")
        dev = self.params[0].device
        if dev.type != "cuda":
            raise RuntimeError("nkbs Adam runs on CUDA parameters only (no CPU fallback)")
        for p in self.params:
            if p.dtype != torch.float32 or not p.is_contiguous():
                raise ValueError("nkbs Adam takes contiguous fp32 parameters")
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.exp_avg = [torch.zeros_like(p) for p in self.params]
        self.exp_avg_sq = [torch.zeros_like(p) for p in self.params]
        self.step_t = torch.zeros((), device=dev, dtype=torch.float32)

    def zero_grad(self, set_to_none: bool = True) -> None:
        for p in self.params:
            if p.grad is not None:
                if set_to_none:
                    p.grad = None
                else:
                    p.grad.zero_()

    @torch.no_grad()
    def step(self, inv_scale: Optional[torch.Tensor] = None, zero_grads: bool = False) -> None:
        """One update of every parameter that has a gradient.  A call in which no parameter has one (the proposal
        group on a gated step) is a no-op and does NOT advance the step count — torch.optim.Adam keeps
        the count per parameter and skips parameters whose ``grad`` is None."""
        ps, gs, ms, vs, ns, keep = [], [], [], [], [], []
        for p, m, v in zip(self.params, self.exp_avg, self.exp_avg_sq):
            g = p.grad
            if g is None:
                continue
            if g.dtype != torch.float32 or not g.is_contiguous():
                g = g.float().contiguous()
            ps.append(p.data_ptr()), gs.append(g.data_ptr()), ms.append(m.data_ptr()), vs.append(v.data_ptr())
            ns.append(p.numel())
            keep.append(g)
        n = len(ps)
        if n == 0:
            return
        self.step_t.add_(1.0)
        arr = lambda xs: (C.c_void_p * n)(*xs)
        call("nkbs_adam_step", n, arr(ps), arr(gs), arr(ms), arr(vs), (C.c_int64 * n)(*ns), float(self.lr),
             float(self.betas[0]), float(self.betas[1]), float(self.eps), float(self.weight_decay),
             self.step_t.data_ptr(), inv_scale.data_ptr() if inv_scale is not None else None, int(zero_grads), None,
             stream())

    def state_dict(self):
        return {"step": self.step_t.clone(), "exp_avg": [m.clone() for m in self.exp_avg],
                "exp_avg_sq": [v.clone() for v in self.exp_avg_sq]}

    def load_state_dict(self, sd):
        self.step_t.copy_(sd["step"])
        for a, b in zip(self.exp_avg, sd["exp_avg"]):
            a.copy_(b)
        for a, b in zip(self.exp_avg_sq, sd["exp_avg_sq"]):
            a.copy_(b)


# ----------------------------------------------------------------------------------------------
# flat parameter / gradient arena of the step driver
# ----------------------------------------------------------------------------------------------
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


class ParamArena:
    """Parameters and gradients of all groups in two flat fp32 buffers ``P`` and ``G`` (group after group, every
    tensor 16-byte aligned).  The Parameter objects stay; their storage moves into ``P``, ``p.grad`` becomes the
    matching view of ``G`` for good, and ``p._nkbs_grad_sink`` tells the nkbs backward kernels to ACCUMULATE their
    parameter gradients straight into ``G`` (ops.py) instead of returning fresh zero-filled tensors to autograd: no
    fill launches, no AccumulateGrad copies, and the optimizer is one launch per group that re-zeroes ``G`` as it
    reads it.  ``alloc(numel)`` lets the data-parallel step place the buffers in symmetric (peer-mapped) memory.

    ``deterministic``: the backward kernels add their partial sums to a 64-bit fixed-point shadow ``F`` of ``G``
    (include/nkbs.h: nkbs_fixed_point_register; 2^log2_scale units per 1.0) instead of issuing float atomics, and
    ``flush_fixed`` (called by the optimizers before they read ``G``) folds it into ``G``: the gradients, and with them
    the whole training run, no longer depend on the arrival order of warps and CTAs.  8 more bytes per parameter; until
    the flush ``p.grad`` does not hold the step's gradient.  ``log2_scale`` = 56: partial sums are kept to 1.4e-17
    (70 times below Adam's eps of 1e-15) and a gradient element must stay below 128 in magnitude."""

    def __init__(self, groups: Dict[str, Sequence[torch.nn.Parameter]], world: int = 1, rank: int = 0, alloc=None,
                 deterministic: bool = False, log2_scale: int = 56):
        self.groups: Dict[str, List[torch.nn.Parameter]] = {k: [p for p in v if p.requires_grad] for k, v in groups.items()}
        first = next(p for ps in self.groups.values() for p in ps)
        dev = first.device
        if dev.type != "cuda":
            raise RuntimeError("nkbs optimizers run on CUDA parameters only (no CPU fallback)")
        for ps in self.groups.values():
            for p in ps:
                if p.dtype != torch.float32 or not p.is_contiguous():
                    raise ValueError("nkbs optimizers take contiguous fp32 parameters")
        self.device = dev
        self.layout = flat_layout({k: [p.numel() for p in ps] for k, ps in self.groups.items()}, world, rank)
        self.total = self.layout.pop("__total__")
        alloc = alloc if alloc is not None else (lambda n: torch.empty(n, dtype=torch.float32, device=dev))
        self.P, self.G = alloc(self.total), alloc(self.total)
        self.P.zero_()
        self.G.zero_()
        self.F = None
        if deterministic:
            self.F = torch.zeros(self.total, dtype=torch.int64, device=dev)
            call("nkbs_fixed_point_register", self.G.data_ptr(), self.total, self.F.data_ptr(), int(log2_scale))
            from . import ops
            ops._DET["count"] += 1  # gradient buffers outside the arena get a shadow per backward call (ops._temp_shadow)
            ops._DET["log2_scale"] = int(log2_scale)
        self.grad_view: Dict[int, torch.Tensor] = {}
        with torch.no_grad():
            for name, ps in self.groups.items():
                for p, o in zip(ps, self.layout[name]["offs"]):
                    view = self.P[o:o + p.numel()].view(p.shape)
                    view.copy_(p.data)
                    p.data = view
                    gv = self.G[o:o + p.numel()].view(p.shape)
                    self.grad_view[id(p)] = gv
                    p._nkbs_grad_sink = gv
                    p.grad = gv

    def group_slice(self, name: str, buf: torch.Tensor) -> torch.Tensor:
        lay = self.layout[name]
        return buf[lay["start"]:lay["start"] + lay["len"]]

    def flush_fixed(self, names: Optional[Sequence[str]] = None) -> None:
        """Deterministic mode: fold the fixed-point shadow of the groups ``names`` (default all) into ``G``."""
        if self.F is None:
            return
        for name in (names if names is not None else list(self.layout)):
            lay = self.layout[name]
            call("nkbs_fixed_point_flush", self.G.data_ptr() + lay["start"] * 4, lay["len"], stream())

    def __del__(self):
        # a registered buffer must not outlive its memory: later allocations could land inside the stale range
        try:
            if getattr(self, "F", None) is not None:
                self.release()
        except Exception:
            pass

    def release(self) -> None:
        """Detach the parameters from the arena's gradient buffer (they keep their storage in ``P``)."""
        if self.F is not None:
            call("nkbs_fixed_point_register", self.G.data_ptr(), self.total, None, 0)
            self.F = None
            from . import ops
            ops._DET["count"] -= 1
        for ps in self.groups.values():
            for p in ps:
                if hasattr(p, "_nkbs_grad_sink"):
                    del p._nkbs_grad_sink
                p.grad = None


class ArenaAdam:
    """torch.optim.Adam semantics per parameter group over a ``ParamArena``: one ``nkbs_adam_step`` launch per
    active group over the group's flat slice; the kernel writes zeros back over the gradients it has read.  A group
    that is not active in a step (the proposal networks on a gated step) is skipped and its step count
    does not advance, like torch.optim.Adam skips parameters whose ``grad`` is None."""

    def __init__(self, arena: ParamArena, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-15,
                 weight_decay: float = 0.0):
        self.arena = arena
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        dev = arena.device
        self.exp_avg = {k: torch.zeros(v["len"], device=dev) for k, v in arena.layout.items()}
        self.exp_avg_sq = {k: torch.zeros(v["len"], device=dev) for k, v in arena.layout.items()}
        self.step_t = {k: torch.zeros((), device=dev, dtype=torch.float32) for k in arena.layout}
        # arrival counters of the kernels' last-CTA step-count store (nkbs_adam_step: step_counter)
        self._counters = {k: torch.zeros((), device=dev, dtype=torch.int32) for k in arena.layout}

    @torch.no_grad()
    def step(self, active: Optional[Sequence[str]] = None, inv_scale: Optional[torch.Tensor] = None,
             zero_grads: bool = True) -> None:
        a = self.arena
        a.flush_fixed(active)
        for name in (active if active is not None else list(a.layout)):
            lay = a.layout[name]
            one = lambda x: (C.c_void_p * 1)(x)
            off = lay["start"] * 4
            call("nkbs_adam_step", 1, one(a.P.data_ptr() + off), one(a.G.data_ptr() + off),
                 one(self.exp_avg[name].data_ptr()), one(self.exp_avg_sq[name].data_ptr()), (C.c_int64 * 1)(lay["len"]),
                 float(self.lr), float(self.betas[0]), float(self.betas[1]), float(self.eps), float(self.weight_decay),
                 self.step_t[name].data_ptr(), inv_scale.data_ptr() if inv_scale is not None else None, int(zero_grads),
                 self._counters[name].data_ptr(), stream())

    def moments_of(self, p: torch.nn.Parameter):
        """(exp_avg, exp_avg_sq) views shaped like ``p`` (checkpointing, tests)."""
        a = self.arena
        for name, ps in a.groups.items():
            for q, o in zip(ps, a.layout[name]["offs"]):
                if q is p:
                    lo = o - a.layout[name]["start"]
                    return (self.exp_avg[name][lo:lo + p.numel()].view(p.shape),
                            self.exp_avg_sq[name][lo:lo + p.numel()].view(p.shape))
        raise KeyError("parameter is not in the arena")
