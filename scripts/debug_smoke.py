import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import nerf_kbs_b200 as nk
from nerf_kbs_b200 import ops
from oracle import nerfstudio_torch as O
from tests.util import copy_model, make_jitter, rel_err, to_nk_bundle, field_grad_pairs, density_grad_pairs
def run(prec, fhead, fdens, R=256):
    ops.set_mlp_precision(prec); ops.set_fused_head(fhead); ops.set_fused_density(fdens)
    torch.manual_seed(0)
    cfg = dict(use_semantic=True)
    om, nm = O.NerfactoModel(O.ModelConfig(**cfg)), nk.NerfactoModel(nk.ModelConfig(**cfg)).to("cuda:0")
    with torch.no_grad():
        for p in om.parameters():
            if p.dim() == 2 and p.shape[1] == 2:
                p.mul_(300.0)
    copy_model(nm, om)
    om.train(), nm.train()
    rb = O.synthetic_ray_bundle(R, seed=1)
    jit = make_jitter(R)
    g = torch.Generator().manual_seed(2)
    batch = {"image": torch.rand(R, 3, generator=g), "semantics_label": torch.randint(0, 19, (R,), generator=g)}
    lo = sum(om.get_loss_dict(om(rb, jitter=jit), batch).values())
    out = nm(to_nk_bundle(rb, "cuda:0"), jitter=[j.cuda() for j in jit])
    ln = sum(nm.get_loss_dict(out, {k: v.cuda() for k, v in batch.items()}).values())
    lo.backward(); ln.backward(); torch.cuda.synchronize()
    errs = {n: rel_err(a, b) for n, a, b in field_grad_pairs(nm.field, om.field)}
    for i, (a, b) in enumerate(zip(nm.proposal_networks, om.proposal_networks)):
        for n, x, y in density_grad_pairs(a, b): errs[f"prop{i}.{n}"] = rel_err(x, y)
    print(f"prec={prec} fused_head={fhead} fused_density={fdens}: loss {ln.item():.6f} vs {lo.item():.6f}")
    print("   ", {k: f"{v:.1e}" for k, v in errs.items()})
for prec, fh, fd in [("fp32", True, True), ("tc", False, False), ("tc", True, False), ("tc", False, True), ("tc", True, True)]:
    run(prec, fh, fd)
