"""Fused proposal density kernels vs the encode + MLP route at the 65 536-ray proposal sizes (CUDA events)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import nerf_kbs_b200 as nk
from nerf_kbs_b200 import ops
R = int(os.environ.get("R", "65536"))
torch.manual_seed(0)
# MODEL=semantic_nerfw: the default HashMLPDensityField proposals (8 levels, 2^18, hidden 64)
if os.environ.get("MODEL", "nerfacto") == "semantic_nerfw":
    m = nk.SemanticNerfWModel(nk.ModelConfig(use_semantic=True)).cuda().train()
else:
    m = nk.NerfactoModel(nk.ModelConfig()).cuda().train()
g = torch.Generator().manual_seed(1)
o = torch.rand(R, 3, generator=g) - 0.5
d = torch.randn(R, 3, generator=g); d = d / d.norm(dim=-1, keepdim=True)
rb = m.collider(nk.RayBundle(o.cuda(), d.cuda(), torch.ones(R, 1).cuda(), torch.zeros(R, 1, dtype=torch.long).cuda()))
with torch.no_grad():
    rs, wl, rsl = m.proposal_sampler(rb, density_fns=m.density_fns)
def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
for i, (s, net) in enumerate(zip(rsl, m.proposal_networks)):
    N = s.frustums.starts.numel()
    dy = torch.rand(s.frustums.starts.shape, device="cuda")
    for fused in (True, False):
        ops.set_fused_density(fused)
        out = {}
        def fwd():
            out["d"], _ = net.get_density(s)
        def bwd():
            for p in net.parameters(): p.grad = None
            out["d"].backward(dy, retain_graph=True)
        tf = timeit(fwd); tb = timeit(bwd)
        print(f"prop{i} N={N} fused={fused}: fwd {tf:.3f} ms ({N/tf/1e6:.2f} Gsamp/s)  bwd {tb:.3f} ms ({N/tb/1e6:.2f} Gsamp/s)")
ops.set_fused_density(True)
