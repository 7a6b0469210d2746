"""GPU parity of ray generation, masked sampling, the semantic palette kernels and the chunked full-frame render
driver against the oracle restatements."""
import pytest
import torch

from oracle import nerfstudio_torch as O
import nerf_kbs_b200 as nk
from nerf_kbs_b200 import data as D
from tests.util import assert_close, copy_model

pytestmark = pytest.mark.gpu
DEV = "cuda"
PALETTE = [(128, 64, 128), (244, 35, 232), (70, 70, 70), (102, 102, 156), (190, 153, 153), (153, 153, 153),
           (250, 170, 30), (220, 220, 0), (107, 142, 35), (152, 251, 152), (70, 130, 180), (220, 20, 60), (255, 0, 0),
           (0, 0, 142), (0, 0, 70), (0, 60, 100), (0, 80, 100), (0, 0, 230), (119, 11, 32)]


def _cameras(n=5, H=375, W=1242, seed=0):
    g = torch.Generator().manual_seed(seed)
    q, _ = torch.linalg.qr(torch.randn(n, 3, 3, generator=g))
    t = torch.rand(n, 3, 1, generator=g) - 0.5
    c2w = torch.cat([q, t], -1)
    f = torch.full((n,), 721.5)
    return nk.Cameras(c2w, f, f.clone(), torch.full((n,), W / 2.0), torch.full((n,), H / 2.0), H, W)


def test_generate_rays_matches_oracle():
    cams = _cameras()
    g = torch.Generator().manual_seed(1)
    N = 20000
    idx = torch.stack([torch.randint(0, len(cams), (N,), generator=g), torch.randint(0, cams.height, (N,), generator=g),
                       torch.randint(0, cams.width, (N,), generator=g)], -1)
    want = O.generate_rays(idx, cams.camera_to_worlds, cams.intrinsics)
    got = nk.RayGenerator(cams).to(DEV)(idx.to(DEV))
    assert torch.equal(got.origins.cpu(), want.origins)
    assert torch.equal(got.camera_indices.cpu(), want.camera_indices)
    assert_close(got.directions, want.directions, 1e-6, "directions")
    assert_close(got.metadata["directions_norm"], want.metadata["directions_norm"], 1e-6, "directions_norm")
    # a product of two differences of unit vectors 1/721 apart: one fp32 ulp of a direction component is 6e-8 * 721 =
    # 4e-5 of a difference, so two correct fp32 evaluations (the host's vector ISA decides the oracle's rounding) differ
    # by a few 1e-4 of the area; the bar sits above that noise, and the directions themselves are held to 1e-6 above
    assert_close(got.pixel_area, want.pixel_area, 1e-3, "pixel_area")
    assert float((got.directions.norm(dim=-1) - 1).abs().max()) < 1e-6


def test_full_image_indices_are_row_major():
    idx = D.image_ray_indices(3, 4, 5, DEV)
    assert idx.shape == (20, 3) and idx[:, 0].eq(3).all()
    assert idx[:, 1].tolist() == [r for r in range(4) for _ in range(5)]
    assert idx[:, 2].tolist() == list(range(5)) * 4


@pytest.mark.parametrize("keep,R", [(0.8, 8192), (0.3, 4096), (0.0, 64), (1.0, 1000)])
def test_masked_compaction_matches_oracle(keep, R):
    g = torch.Generator().manual_seed(R)
    n_img, n_pix = 7, 1242 * 375
    masks = (torch.rand(n_img, n_pix, generator=g) < keep).to(torch.uint8) * 255
    M = int(R * 1.5) + 32
    img = torch.randint(0, n_img, (M,), generator=g)
    pix = torch.randint(0, n_pix, (M,), generator=g)
    wi, wp, wn = O.compact_masked(img, pix, masks, R)
    gi, gp, gn = D.compact_masked(img.to(DEV), pix.to(DEV), masks.to(DEV), R)
    assert int(gn) == wn
    if wn > 0:
        assert torch.equal(gi.cpu(), wi) and torch.equal(gp.cpu(), wp)
        assert masks[gi.cpu(), gp.cpu()].ne(0).all()


@pytest.mark.parametrize("keep,R,M", [(0.8, 200000, 300001), (0.02, 50000, 1200000), (0.0, 64, 70000), (1.0, 1000, 40000)])
def test_masked_compaction_long_candidate_lists(keep, R, M, monkeypatch):
    """The multi-CTA route (per-CTA counts, scan, ordered scatter, cyclic fill) for candidate lists beyond one CTA's
    reach — pixel chunks like the reference's (ref:dataparser/datamanager_z.py:49) — against the oracle and against the
    one-CTA kernel."""
    g = torch.Generator().manual_seed(R)
    n_img, n_pix = 3, 1242 * 375
    masks = (torch.rand(n_img, n_pix, generator=g) < keep).to(torch.uint8)
    img = torch.randint(0, n_img, (M,), generator=g)
    pix = torch.randint(0, n_pix, (M,), generator=g)
    wi, wp, wn = O.compact_masked(img, pix, masks, R)
    assert M > D.COMPACT_ONE_CTA_MAX
    gi, gp, gn = D.compact_masked(img.to(DEV), pix.to(DEV), masks.to(DEV), R)
    assert int(gn) == wn
    if wn > 0:
        assert torch.equal(gi.cpu(), wi) and torch.equal(gp.cpu(), wp)
    monkeypatch.setattr(D, "COMPACT_ONE_CTA_MAX", 1 << 22)  # force the one-CTA kernel on the same list
    oi, op, on = D.compact_masked(img.to(DEV), pix.to(DEV), masks.to(DEV), R)
    assert int(on) == int(gn) and (wn == 0 or (torch.equal(oi, gi) and torch.equal(op, gp)))


def test_palette_round_trip_and_colormap():
    pal = nk.SemanticPalette(PALETTE, DEV)
    g = torch.Generator().manual_seed(2)
    labels = torch.randint(0, len(PALETTE), (5000,), generator=g)
    rgb = torch.tensor(PALETTE, dtype=torch.uint8)[labels]
    rgb[::11] = torch.tensor([1, 2, 3], dtype=torch.uint8)  # colours outside the palette -> default label
    want = O.color_to_label(rgb, PALETTE, len(PALETTE))
    got = pal.color_to_label(rgb.to(DEV))
    assert torch.equal(got.cpu(), want)
    logits = torch.randn(5000, len(PALETTE), generator=g)
    logits[5, 3] = logits[5, 9] = 10.0  # tie: the first maximum wins
    wl, wc = O.semantic_colormap(logits, PALETTE)
    gl, gc = pal.colormap(logits.to(DEV))
    assert torch.equal(gl.cpu(), wl) and torch.equal(gc.cpu(), wc)


def test_render_camera_matches_oracle_chunks():
    """A small frame rendered in uneven chunks equals the oracle's eval forward on the same rays."""
    torch.manual_seed(0)
    cfg = dict(use_semantic=True)
    om, nm = O.NerfactoModel(O.ModelConfig(**cfg)), nk.NerfactoModel(nk.ModelConfig(**cfg)).to(DEV)
    with torch.no_grad():
        for p in om.parameters():
            if p.dim() == 2 and p.shape[1] == 2:
                p.mul_(300.0)
    copy_model(nm, om)
    om.eval(), nm.eval()
    cams = _cameras(n=3, H=12, W=20)
    gen = nk.RayGenerator(cams).to(DEV)
    pal = nk.SemanticPalette(PALETTE, DEV)
    prev = nk.ops.get_mlp_precision()
    nk.ops.set_mlp_precision("fp32")
    try:
        out = nk.render_camera(nm, gen, 1, num_rays_per_chunk=64, palette=pal)
    finally:
        nk.ops.set_mlp_precision(prev)
    idx = D.image_ray_indices(1, 12, 20, "cpu")
    with torch.no_grad():
        want = om(O.generate_rays(idx, cams.camera_to_worlds, cams.intrinsics))
    assert out["rgb"].shape == (12, 20, 3) and out["semantics_colormap"].dtype == torch.uint8
    assert_close(out["rgb"].reshape(-1, 3), want["rgb"], 1e-4, "rgb")
    assert_close(out["accumulation"].reshape(-1, 1), want["accumulation"], 1e-4, "accumulation")
    assert_close(out["depth"].reshape(-1, 1), want["depth"], 1e-3, "median depth")
    assert_close(out["semantics"].reshape(-1, 19), want["semantics"], 1e-4, "semantics")
    assert float(out["rgb"].min()) >= 0.0 and float(out["rgb"].max()) <= 1.0


def test_colormaps_full_kitti_frame_match_oracle():
    """N4: accumulation / depth colour maps of a full 1242x375 frame against the oracle (torch restatement of
    nerfstudio utils.colormaps), same colour tables in both arms: bit-exact table indices, hence equal colours."""
    H, W = 375, 1242
    g = torch.Generator().manual_seed(3)
    acc = torch.rand(H, W, 1, generator=g)
    acc[0, :7, 0] = torch.tensor([0.0, 1.0, 0.5, 1.0 / 255.0, 254.999 / 255.0, 0.999999, 2.0 / 255.0])
    depth = torch.rand(H, W, 1, generator=g) * 80.0 + 0.5
    depth[1, :3, 0] = torch.tensor([0.05, 1000.0, 3.0])  # sets the frame's own range
    for name in ("viridis", "turbo"):
        lut = nk.colormap_table(name)
        assert lut.shape == (256, 3) and float(lut.min()) >= 0.0 and float(lut.max()) <= 1.0
        got = nk.apply_colormap(acc.to(DEV), cmap=name)
        assert got.shape == (H, W, 3)
        assert torch.equal(got.cpu(), O.apply_colormap(acc, lut))
    lut = nk.colormap_table("turbo")
    for kw in (dict(), dict(near_plane=0.5), dict(far_plane=60.0), dict(near_plane=1.0, far_plane=50.0)):
        for use_acc in (True, False):
            want = O.apply_depth_colormap(depth, lut, accumulation=acc if use_acc else None, **kw)
            got = nk.apply_depth_colormap(depth.to(DEV), accumulation=acc.to(DEV) if use_acc else None, **kw)
            assert torch.equal(got.cpu(), want), (kw, use_acc, (got.cpu() - want).abs().max())
    # a caller-supplied table (what matplotlib would hand over) is used as given
    custom = torch.rand(256, 3, generator=g)
    assert torch.equal(nk.apply_colormap(acc.to(DEV), cmap=custom).cpu(), O.apply_colormap(acc, custom))


def test_render_camera_emits_colormapped_images():
    torch.manual_seed(0)
    nm = nk.NerfactoModel(nk.ModelConfig(use_semantic=True)).to(DEV).eval()
    cams = _cameras(n=2, H=10, W=16)
    gen = nk.RayGenerator(cams).to(DEV)
    out = nk.render_camera(nm, gen, 0, num_rays_per_chunk=50, palette=nk.SemanticPalette(PALETTE, DEV))
    lut_t, lut_v = nk.colormap_table("turbo"), nk.colormap_table("viridis")
    assert torch.equal(out["accumulation_colormap"].cpu(), O.apply_colormap(out["accumulation"].cpu(), lut_v))
    for k in ("depth", "prop_depth_0", "prop_depth_1"):
        want = O.apply_depth_colormap(out[k].cpu(), lut_t, accumulation=out["accumulation"].cpu())
        assert torch.equal(out[k + "_colormap"].cpu(), want), k


def test_composite_backward_at_max_samples():
    """S = 512 needs more than 48 KB of dynamic shared memory in the backward kernel."""
    g = torch.Generator().manual_seed(0)
    R, S = 33, 512
    dens = (torch.rand(R, S, generator=g) * 3).to(DEV).requires_grad_(True)
    rgb = torch.rand(R, S, 3, generator=g).to(DEV).requires_grad_(True)
    bins = torch.sort(torch.rand(R, S + 1, generator=g) * 4, dim=-1).values
    out = nk.ops.composite(density=dens, rgb=rgb, starts=bins[:, :-1].to(DEV), ends=bins[:, 1:].to(DEV),
                           want=("rgb", "acc"))
    (out["rgb"].sum() + out["accumulation"].sum()).backward()
    d2 = dens.detach().cpu().double().requires_grad_(True)
    c2 = rgb.detach().cpu().double().requires_grad_(True)
    delta = (bins[:, 1:] - bins[:, :-1]).double()
    ds = d2 * delta
    T = torch.exp(-torch.cat([torch.zeros(R, 1, dtype=torch.float64), torch.cumsum(ds, -1)[:, :-1]], -1))
    w = (1 - torch.exp(-ds)) * T
    acc = w.sum(-1)
    col = (w[..., None] * c2).sum(1) + c2[:, -1] * (1 - acc[:, None])
    (col.sum() + acc.sum()).backward()
    assert_close(dens.grad, d2.grad.float(), 2e-4, "d density at S=512")
    assert_close(rgb.grad, c2.grad.float(), 2e-4, "d rgb at S=512")
