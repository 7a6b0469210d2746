"""Mint the golden vectors under tests/golden/ from the CPU oracle (oracle/nerfstudio_torch.py).

The reference ships no fixtures for this path and nerfstudio cannot be imported here, so these
vectors pin the ORACLE: a later change to the oracle, to torch's CPU kernels or to numpy that alters a bit of
the integer / bin outputs, or moves a float output by more than 1e-6, fails tests/test_golden.py.
Run:  python scripts/make_golden.py     (CPU only; records torch / numpy versions in the file)"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import nerfstudio_torch as O  # noqa: E402


def build():
    out = {"torch_version": np.array(torch.__version__), "numpy_version": np.array(np.__version__)}
    # --- hash encoding: indices + features, main grid and a proposal grid ---
    for tag, (L, max_res, log2) in {"main": (16, 2048, 19), "prop": (5, 128, 17)}.items():
        torch.manual_seed(11)
        enc = O.HashEncoding(num_levels=L, min_res=16, max_res=max_res, log2_hashmap_size=log2)
        g = torch.Generator().manual_seed(12)
        x = torch.rand(256, 3, generator=g)
        x[:8] = torch.randint(0, 17, (8, 3), generator=g).float() / 16.0
        hs, _ = enc.corner_indices(x)
        idx = torch.stack(hs, -1)
        rows = torch.unique(idx)  # only the touched table rows are stored
        out[f"enc_{tag}_scalings"] = enc.scalings.numpy()
        out[f"enc_{tag}_x"] = x.numpy()
        out[f"enc_{tag}_idx"] = idx.numpy()
        out[f"enc_{tag}_rows"] = rows.numpy()
        out[f"enc_{tag}_row_values"] = enc.hash_table[rows].detach().numpy()
        out[f"enc_{tag}_feats"] = enc(x).detach().numpy()
    # --- samplers: initial bins (bit-exact) and one PDF level ---
    rb = O.NearFarCollider(0.001, 1000.0)(O.synthetic_ray_bundle(96, seed=21))
    g = torch.Generator().manual_seed(22)
    t_rand = torch.rand(96, 1, generator=g)
    init = O.UniformLinDispPiecewiseSampler(single_jitter=True).train()
    rs0 = init(rb, num_samples=256, t_rand=t_rand)
    w0 = torch.rand(96, 256, 1, generator=g) ** 5
    rand1 = torch.rand(96, 1, generator=g)
    rs1 = O.PDFSampler(include_original=False, single_jitter=True).train()(rb, rs0, w0, num_samples=96, rand=rand1)
    out["smp_origins"] = rb.origins.numpy()
    out["smp_directions"] = rb.directions.numpy()
    out["smp_t_rand"] = t_rand.numpy()
    out["smp_sbins0"] = O.ray_samples_to_sdist(rs0).numpy()
    out["smp_ebins0"] = torch.cat([rs0.frustums.starts[..., 0], rs0.frustums.ends[..., -1:, 0]], -1).numpy()
    out["smp_w0"] = w0[..., 0].numpy()
    out["smp_rand1"] = rand1.numpy()
    out["smp_sbins1"] = O.ray_samples_to_sdist(rs1).numpy()
    out["smp_inds1"] = rs1._dbg_inds.numpy()
    # --- compositing ---
    R, S, C = 64, 48, 19
    dens = torch.rand(R, S, 1, generator=g) * 40
    rgb = torch.rand(R, S, 3, generator=g)
    sem = torch.randn(R, S, C, generator=g)
    rs2 = O.PDFSampler(include_original=False, single_jitter=True).train()(
        rb, rs1, torch.rand(96, 96, 1, generator=g), num_samples=S, rand=torch.rand(96, 1, generator=g))
    st, en = rs2.frustums.starts[:R], rs2.frustums.ends[:R]
    sub = O.RaySamples(O.Frustums(rs2.frustums.origins[:R], rs2.frustums.directions[:R], st, en), None, en - st,
                       rs2.spacing_starts[:R], rs2.spacing_ends[:R])
    w = sub.get_weights(dens)
    out["cmp_density"], out["cmp_rgb"], out["cmp_sem"] = dens.numpy(), rgb.numpy(), sem.numpy()
    out["cmp_starts"], out["cmp_ends"] = st.numpy(), en.numpy()
    out["cmp_weights"] = w.numpy()
    out["cmp_rgb_out"] = O.RGBRenderer("last_sample").train()(rgb=rgb, weights=w).numpy()
    out["cmp_acc"] = O.AccumulationRenderer()(weights=w).numpy()
    out["cmp_depth_med"] = O.DepthRenderer("median")(weights=w, ray_samples=sub).numpy()
    out["cmp_depth_exp"] = O.DepthRenderer("expected")(weights=w, ray_samples=sub).numpy()
    out["cmp_sem_out"] = O.SemanticRenderer()(sem, weights=w).numpy()
    # --- whole nerfacto forward, eval mode (no jitter): seeds only, outputs stored ---
    torch.manual_seed(31)
    m = O.NerfactoModel(O.ModelConfig(use_semantic=True, log2_hashmap_size=14,
                                      proposal_net_args_list=({"hidden_dim": 16, "log2_hashmap_size": 12, "num_levels": 5, "max_res": 128, "use_linear": False},
                                                              {"hidden_dim": 16, "log2_hashmap_size": 12, "num_levels": 5, "max_res": 256, "use_linear": False}))).eval()
    with torch.no_grad():
        for p in m.parameters():
            if p.dim() == 2 and p.shape[1] == 2:
                p.mul_(300.0)
        o = m(O.synthetic_ray_bundle(48, seed=32))
    for k in ("rgb", "accumulation", "depth", "expected_depth", "semantics"):
        out[f"model_{k}"] = o[k].numpy()
    # --- proposal losses (N1): interlevel + distortion over three histogram levels, values and d/dw ---
    g = torch.Generator().manual_seed(41)
    edges, ws = [], []
    for S in (64, 24, 12):
        e = torch.sort(torch.rand(40, S + 1, generator=g), dim=-1).values
        w = torch.rand(40, S, generator=g) + 0.05
        w = (w / w.sum(-1, keepdim=True) * torch.rand(40, 1, generator=g)).requires_grad_()
        edges.append(e)
        ws.append(w)

    class _RS:  # the losses read only the spacing-space bin edges
        def __init__(self, sd):
            self.spacing_starts, self.spacing_ends = sd[..., :-1, None], sd[..., 1:, None]

    rsl = [_RS(e) for e in edges]
    inter = O.interlevel_loss([w[..., None] for w in ws], rsl)
    dist = O.distortion_loss([w[..., None] for w in ws], rsl)
    (inter + 0.5 * dist).backward()
    out["loss_interlevel"], out["loss_distortion"] = inter.detach().numpy(), dist.detach().numpy()
    for i, (e, w) in enumerate(zip(edges, ws)):
        out[f"loss_sdist{i}"], out[f"loss_w{i}"], out[f"loss_dw{i}"] = e.numpy(), w.detach().numpy(), w.grad.numpy()
    # --- ray generation, masked compaction, label palette (N3 / N4) ---
    g = torch.Generator().manual_seed(51)
    q, _ = torch.linalg.qr(torch.randn(3, 3, 3, generator=g))
    c2w = torch.cat([q, torch.rand(3, 3, 1, generator=g) - 0.5], -1)
    intr = torch.tensor([[721.5, 721.5, 621.0, 187.5]]).repeat(3, 1)
    idx = torch.stack([torch.randint(0, 3, (300,), generator=g), torch.randint(0, 375, (300,), generator=g),
                       torch.randint(0, 1242, (300,), generator=g)], -1)
    rbg = O.generate_rays(idx, c2w, intr)
    out["ray_c2w"], out["ray_intrinsics"], out["ray_idx"] = c2w.numpy(), intr.numpy(), idx.numpy()
    out["ray_origins"], out["ray_directions"] = rbg.origins.numpy(), rbg.directions.numpy()
    out["ray_pixel_area"], out["ray_dnorm"] = rbg.pixel_area.numpy(), rbg.metadata["directions_norm"].numpy()
    masks = (torch.rand(3, 500, generator=g) < 0.6).to(torch.uint8) * 255
    ci, cp = torch.randint(0, 3, (200,), generator=g), torch.randint(0, 500, (200,), generator=g)
    wi, wp, wn = O.compact_masked(ci, cp, masks, 150)
    out["msk_masks"], out["msk_img"], out["msk_pix"] = masks.numpy(), ci.numpy(), cp.numpy()
    out["msk_out_img"], out["msk_out_pix"], out["msk_count"] = wi.numpy(), wp.numpy(), np.array(wn)
    palette = [(128, 64, 128), (244, 35, 232), (70, 70, 70), (102, 102, 156), (190, 153, 153), (0, 0, 142)]
    pix = torch.tensor(palette, dtype=torch.uint8)[torch.randint(0, 6, (400,), generator=g)]
    pix[::9] = torch.randint(0, 256, (len(pix[::9]), 3), generator=g).to(torch.uint8)
    logits = torch.randn(400, 6, generator=g)
    lab, col = O.semantic_colormap(logits, palette)
    out["pal_palette"], out["pal_pixels"] = np.array(palette, dtype=np.uint8), pix.numpy()
    out["pal_labels"] = O.color_to_label(pix, palette, 6).numpy()
    out["pal_logits"], out["pal_argmax"], out["pal_colours"] = logits.numpy(), lab.numpy(), col.numpy()
    return out


if __name__ == "__main__":
    vec = build()
    path = os.path.join(ROOT, "tests", "golden", "oracle_vectors.npz")
    np.savez_compressed(path, **vec)
    print(path, os.path.getsize(path), "bytes")
