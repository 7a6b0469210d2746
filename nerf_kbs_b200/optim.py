"""Adam over the nkbs multi-tensor kernel (include/nkbs.h: nkbs_adam_step).

Stands in for the ``AdamOptimizerConfig(lr=1e-3, eps=1e-15)`` optimizers of the reference's two
parameter groups (ref:train.py:85-98, groups ref:nerfacto.py:359-363).  torch.optim.Adam semantics
(amsgrad off, no scheduler: the reference sets ``"scheduler": None``).  The step count lives on the
device so that a captured CUDA graph advances it on every replay."""
from __future__ import annotations

import ctypes as C
from typing import Iterable, Optional

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
    def step(self, inv_scale: Optional[torch.Tensor] = None) -> None:
        """One update of every parameter that has a gradient."""
        self.step_t.add_(1.0)
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
        arr = lambda xs: (C.c_void_p * n)(*xs)
        call("nkbs_adam_step", n, arr(ps), arr(gs), arr(ms), arr(vs), (C.c_int64 * n)(*ns), float(self.lr),
             float(self.betas[0]), float(self.betas[1]), float(self.eps), float(self.weight_decay),
             self.step_t.data_ptr(), inv_scale.data_ptr() if inv_scale is not None else None, stream())

    def state_dict(self):
        return {"step": self.step_t.clone(), "exp_avg": [m.clone() for m in self.exp_avg],
                "exp_avg_sq": [v.clone() for v in self.exp_avg_sq]}

    def load_state_dict(self, sd):
        self.step_t.copy_(sd["step"])
        for a, b in zip(self.exp_avg, sd["exp_avg"]):
            a.copy_(b)
        for a, b in zip(self.exp_avg_sq, sd["exp_avg_sq"]):
            a.copy_(b)
