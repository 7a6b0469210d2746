"""Debug helper (GPU): per-parameter gradient error of the nkbs model vs the fp32 and an fp64 oracle."""
import copy, sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import nerfstudio_torch as O
import nerf_kbs_b200 as nk
from tests.util import make_jitter, rel_err, copy_model, to_nk_bundle, field_grad_pairs, density_grad_pairs
torch.manual_seed(0)
cfg = dict(use_semantic=True)
om = O.NerfactoModel(O.ModelConfig(**cfg)); nm = nk.NerfactoModel(nk.ModelConfig(**cfg)).cuda()
with torch.no_grad():
    for p in om.parameters():
        if p.dim() == 2 and p.shape[1] == 2: p.mul_(300.0)
copy_model(nm, om); om.train(); nm.train()
R = 384
rb = O.synthetic_ray_bundle(R, seed=2); jit = make_jitter(R, seed=3)
g = torch.Generator().manual_seed(5)
batch = {"image": torch.rand(R, 3, generator=g), "semantics_label": torch.randint(0, 19, (R,), generator=g)}
out = om(rb, jitter=jit); lo = om.get_loss_dict(out, batch); sum(lo.values()).backward()
outn = nm(to_nk_bundle(rb), jitter=[j.cuda() for j in jit]); ln = nm.get_loss_dict(outn, {k: v.cuda() for k, v in batch.items()}); sum(ln.values()).backward()
print({k: (lo[k].item(), ln[k].item()) for k in lo})
for name, a, b in field_grad_pairs(nm.field, om.field):
    print(f"field {name:14s} rel {rel_err(a, b):.3e}  max {b.abs().max().item():.3e}")
for i, (a, b) in enumerate(zip(nm.proposal_networks, om.proposal_networks)):
    for name, x, y in density_grad_pairs(a, b):
        print(f"prop{i} {name:14s} rel {rel_err(x, y):.3e}  max {y.abs().max().item():.3e}")
# fp64 truth for the final level, using the oracle's fp32 samples
om64 = copy.deepcopy(om).double(); om64.zero_grad()
rs = out["ray_samples_list"][-1]; fr = rs.frustums
d = lambda t: t.double() if t is not None and t.is_floating_point() else t
rs64 = O.RaySamples(O.Frustums(d(fr.origins), d(fr.directions), d(fr.starts), d(fr.ends), None), rs.camera_indices, d(rs.deltas), d(rs.spacing_starts), d(rs.spacing_ends))
torch.set_default_dtype(torch.float64)
fo = om64.field(rs64); w = rs64.get_weights(fo[O.FieldHeadNames.DENSITY])
rgb = om64.renderer_rgb(rgb=fo[O.FieldHeadNames.RGB], weights=w)
sem = om64.renderer_semantics(fo[O.FieldHeadNames.SEMANTICS], weights=w.detach())
loss = torch.nn.functional.mse_loss(batch["image"].double(), rgb) + 0.002 * O.distortion_loss([w], [rs64]) + 0.05 * torch.nn.functional.cross_entropy(sem, batch["semantics_label"], ignore_index=19)
loss.backward()
torch.set_default_dtype(torch.float32)
print("final-level field grads vs fp64 truth:   [oracle fp32]   [nkbs]")
for (name, a, b), (_, _, c) in zip(field_grad_pairs(nm.field, om.field), field_grad_pairs(nm.field, om64.field)):
    print(f"  {name:14s} {rel_err(b, c):.3e}   {rel_err(a, c):.3e}")
# positions of the final samples: are they identical?
sb_o = O.ray_samples_to_sdist(out["ray_samples_list"][-1]); sb_n = outn["ray_samples_list"][-1].spacing_bins.cpu()
print("final bins identical frac", (sb_o == sb_n).float().mean().item(), "max diff", (sb_o - sb_n).abs().max().item())
