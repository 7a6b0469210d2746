"""Sweep dedup_levels of nkbs_encode_bwd on ray-ordered samples (CUDA events)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import nerf_kbs_b200 as nk
from nerf_kbs_b200 import ops
from nerf_kbs_b200._lib import call, ptr, host_floats, stream
R = int(os.environ.get("R", "4096"))
torch.manual_seed(0)
m = nk.NerfactoModel(nk.ModelConfig()).cuda().train()
g = torch.Generator().manual_seed(1)
o = torch.rand(R, 3, generator=g) - 0.5
d = torch.randn(R, 3, generator=g); d = d / d.norm(dim=-1, keepdim=True)
rb = nk.RayBundle(o.cuda(), d.cuda(), torch.ones(R, 1).cuda(), torch.zeros(R, 1, dtype=torch.long).cuda())
rb = m.collider(rb)
with torch.no_grad():
    rs, wl, rsl = m.proposal_sampler(rb, density_fns=m.density_fns)
levels = [("prop0", rsl[0], m.proposal_networks[0].encoding, m.proposal_networks[0]._norm),
          ("prop1", rsl[1], m.proposal_networks[1].encoding, m.proposal_networks[1]._norm),
          ("main", rs, m.field.mlp_base_grid, m.field._norm)]
for name, s, enc, norm in levels:
    fr = s.frustums
    feats, sel, xn, _ = ops.encode_raw(2, enc.grid, norm, enc.hash_table.detach(), origins=fr.ray_origins, directions=fr.ray_directions,
                                       starts=fr.starts[..., 0].contiguous(), ends=fr.ends[..., 0].contiguous())
    N = xn.shape[0]
    dfeat = torch.randn(N, enc.grid.out_dim, device="cuda")
    L = enc.grid.L
    res = []
    for dl in ([L] if os.environ.get('DL_ONLY') else range(0, L + 1)):
        dt = torch.zeros_like(enc.hash_table)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for it in range(3):
            if it == 1: e0.record()
            call("nkbs_encode_bwd", ptr(xn), ptr(dfeat), N, host_floats(enc.grid.scalings), L, enc.grid.log2_T, 2, dl, ptr(dt), stream())
        e1.record(); torch.cuda.synchronize()
        res.append(e0.elapsed_time(e1) / 2)
    print(name, "N", N, " ".join(f"{dl}:{t*1e3:.0f}us" for dl, t in enumerate(res)) if not os.environ.get('DL_ONLY') else f"dedup all: {res[0]*1e3:.0f}us")
