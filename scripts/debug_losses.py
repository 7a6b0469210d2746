import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import nerfstudio_torch as O
import nerf_kbs_b200 as nk
from nerf_kbs_b200 import ops
from nerf_kbs_b200.losses import ray_samples_to_sdist
from tests.test_gpu_model import _pair, _batch
from tests.util import make_jitter, to_nk_bundle
DEV = "cuda"
R = 64
om, nm = _pair("nerfacto")
om.train(); nm.train()
rb = O.synthetic_ray_bundle(R, seed=2)
for step in (0, 1, 2):
    nm.set_step(step)
    jit = make_jitter(R, seed=step)
    out = nm(to_nk_bundle(rb), jitter=[j.to(DEV) for j in jit])
    wl, rsl = out["weights_list"], out["ray_samples_list"]
    print("step", step, "requires_grad", [w.requires_grad for w in wl])
    ws = [w[..., 0].detach().clone().requires_grad_() for w in wl]
    sd = [ray_samples_to_sdist(rs).detach() for rs in rsl]
    _, inter, dist, _ = ops.ray_losses(w_final=ws[-1], sdist_final=sd[-1], w_props=ws[:-1], sdist_props=sd[:-1])
    inter.backward()
    print("  kernel inter", inter.item(), [w.grad.abs().sum().item() if w.grad is not None else None for w in ws])
    ws2 = [w[..., 0].detach().clone().requires_grad_() for w in wl]
    class RS:
        def __init__(s, e): s.spacing_starts = e[..., :-1, None]; s.spacing_ends = e[..., 1:, None]
    want = O.interlevel_loss([w[..., None] for w in ws2], [RS(e) for e in sd])
    want.backward()
    print("  torch  inter", want.item(), [w.grad.abs().sum().item() if w.grad is not None else None for w in ws2])
    nm.after_step(step)
