"""GPU parity of the PDF sampler, compositing kernels and the operator classes against the CPU oracle."""
import pytest
import torch

from oracle import c_oracle as CO
from oracle import nerfstudio_torch as O
import nerf_kbs_b200 as nk
from nerf_kbs_b200 import ops
from tests.util import assert_close, rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda"
FP32_TOL = 1e-4


def _level0(R, S, seed):
    rb = O.NearFarCollider(0.001, 1000.0)(O.synthetic_ray_bundle(R, seed=seed))
    samp = O.UniformLinDispPiecewiseSampler(single_jitter=True)
    samp.train()
    g = torch.Generator().manual_seed(seed)
    rs = samp(rb, num_samples=S, t_rand=torch.rand(R, 1, generator=g))
    return rb, rs, g


def _weights_like(R, S, g, kind):
    if kind == "peaky":
        w = torch.zeros(R, S)
        idx = torch.randint(0, S, (R, 3), generator=g)
        w.scatter_(1, idx, torch.rand(R, 3, generator=g))
        w = w / w.sum(-1, keepdim=True).clamp_min(1e-6) * torch.rand(R, 1, generator=g)
    elif kind == "zero":
        w = torch.zeros(R, S)
    elif kind == "onehot":
        w = torch.zeros(R, S)
        w[:, S // 3] = 1.0
    else:
        w = torch.rand(R, S, generator=g) ** 4
        w = w / w.sum(-1, keepdim=True) * torch.rand(R, 1, generator=g)
    return w[..., None]


@pytest.mark.parametrize("S,ns", [(256, 96), (96, 48), (64, 31)])
@pytest.mark.parametrize("kind", ["smooth", "peaky", "zero", "onehot"])
@pytest.mark.parametrize("train", [True, False])
def test_pdf_sampler_bins(S, ns, kind, train):
    R = 513
    rb, rs, g = _level0(R, S, seed=S + ns)
    w = _weights_like(R, S, g, kind)
    pdf = O.PDFSampler(include_original=False, single_jitter=True)
    pdf.train(train)
    rand = torch.rand(R, 1, generator=g)
    out = pdf(rb, rs, w, num_samples=ns, rand=rand)
    want_s = torch.cat([out.spacing_starts[..., 0], out.spacing_ends[..., -1:, 0]], -1)
    want_e = torch.cat([out.frustums.starts[..., 0], out.frustums.ends[..., -1:, 0]], -1)
    sb_in = torch.cat([rs.spacing_starts[..., 0], rs.spacing_ends[..., -1:, 0]], -1).contiguous()
    sb, eb, inds = ops.pdf_sample(w[..., 0].to(DEV), sb_in.to(DEV), rb.nears.to(DEV), rb.fars.to(DEV), ns,
                                  anneal=1.0, rand=rand.to(DEV) if train else None, want_indices=True)
    # (1) bit-exact against the C restatement, which fixes the one order torch leaves open (the normaliser)
    c_sb, c_eb, c_inds = CO.pdf_sample(w[..., 0], sb_in, rb.nears, rb.fars, ns, rand=rand if train else None)
    assert torch.equal(inds.cpu(), c_inds), "searchsorted indices differ from the C oracle"
    assert torch.equal(sb.cpu(), c_sb), "spacing bins differ from the C oracle"
    assert torch.equal(eb.cpu(), c_eb), "euclidean bins differ from the C oracle"
    # (2) against the torch oracle: torch.sum's vectorised order can move the normaliser by an ulp, so rays
    # whose normaliser agrees must be bit-identical and the rest within a couple of ulps
    wsum_t = (w[..., 0] + 0.01).sum(-1)
    same = wsum_t == (w[..., 0] + 0.01).double().sum(-1).float()
    assert torch.equal(sb.cpu()[same], want_s[same])
    assert torch.equal(inds.cpu()[same], out._dbg_inds[same])
    assert_close(sb, want_s, 2e-6, "spacing bins")
    assert (sb[:, 1:] >= sb[:, :-1]).all(), "bins not monotone"
    assert_close(eb, want_e, 1e-3, "euclidean bins")  # 1/(2-2s) amplifies an ulp of s near the far plane


def test_pdf_sampler_anneal():
    R, S, ns = 300, 256, 96
    rb, rs, g = _level0(R, S, seed=9)
    w = _weights_like(R, S, g, "smooth")
    pdf = O.PDFSampler(include_original=False, single_jitter=True)
    pdf.train()
    rand = torch.rand(R, 1, generator=g)
    out = pdf(rb, rs, torch.pow(w, 0.37), num_samples=ns, rand=rand)
    want_s = torch.cat([out.spacing_starts[..., 0], out.spacing_ends[..., -1:, 0]], -1)
    sb_in = torch.cat([rs.spacing_starts[..., 0], rs.spacing_ends[..., -1:, 0]], -1).contiguous()
    anneal_dev = torch.tensor([0.37], device=DEV)
    for kw in (dict(anneal=0.37), dict(anneal=1.0, anneal_dev=anneal_dev)):
        sb, _ = ops.pdf_sample(w[..., 0].to(DEV), sb_in.to(DEV), rb.nears.to(DEV), rb.fars.to(DEV), ns,
                               rand=rand.to(DEV), **kw)
        assert_close(sb, want_s, 1e-5, "annealed bins")


def _final_level(R, S, C, seed):
    rb, rs, g = _level0(R, S, seed)
    density = (torch.rand(R, S, 1, generator=g) * 40.0) ** 2 * (torch.rand(R, S, 1, generator=g) > 0.3)
    rgb = torch.rand(R, S, 3, generator=g)
    sem = torch.randn(R, S, C, generator=g)
    return rb, rs, density, rgb, sem, g


@pytest.mark.parametrize("S", [48, 96, 256, 37])
def test_get_weights(S):
    R = 257
    rb, rs, density, _, _, g = _final_level(R, S, 3, seed=S)
    density.requires_grad_(True)
    want = rs.get_weights(density)
    gw = torch.randn(R, S, 1, generator=g)
    want.backward(gw)
    nrb = nk.RayBundle(origins=rb.origins.to(DEV), directions=rb.directions.to(DEV), nears=rb.nears.to(DEV), fars=rb.fars.to(DEV))
    nrs = nrb.get_ray_samples(rs.frustums.starts.to(DEV), rs.frustums.ends.to(DEV))
    d = density.detach().to(DEV).requires_grad_(True)
    got = nrs.get_weights(d)
    assert got.shape == want.shape
    assert_close(got, want, FP32_TOL, "weights")
    assert (got >= 0).all() and (got.sum(-2) <= 1 + 1e-5).all()
    got.backward(gw.to(DEV))
    assert_close(d.grad, density.grad, FP32_TOL, "d_density")


def test_weights_known_answer():
    """Constant density sigma over unit-length bins: w_i = (1-e^-sigma) e^(-i sigma)."""
    S = 48
    st = torch.arange(S, dtype=torch.float32)[None, :].repeat(3, 1).to(DEV)
    en = st + 1
    dens = torch.full((3, S), 0.25, device=DEV)
    w = ops.composite(density=dens, starts=st, ends=en, want=())["weights"][..., 0]
    i = torch.arange(S, dtype=torch.float64)
    want = (1 - torch.exp(torch.tensor(-0.25, dtype=torch.float64))) * torch.exp(-0.25 * i)
    assert torch.allclose(w[0].double().cpu(), want, rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("bg", ["last_sample", "black", "white"])
@pytest.mark.parametrize("train", [True, False])
def test_renderers_match_oracle(bg, train):
    R, S, C = 301, 48, 19
    rb, rs, density, rgb, sem, g = _final_level(R, S, C, seed=11)
    if not train:
        rgb[0, 0, 0] = float("nan")
    w = rs.get_weights(density)
    o = {"rgb": O.RGBRenderer(bg), "acc": O.AccumulationRenderer(), "med": O.DepthRenderer("median"),
         "exp": O.DepthRenderer("expected"), "sem": O.SemanticRenderer(), "unc": O.UncertaintyRenderer()}
    n = {"rgb": nk.RGBRenderer(bg), "acc": nk.AccumulationRenderer(), "med": nk.DepthRenderer("median"),
         "exp": nk.DepthRenderer("expected"), "sem": nk.SemanticRenderer(), "unc": nk.UncertaintyRenderer()}
    for m in list(o.values()) + list(n.values()):
        m.train(train)
    nrb = nk.RayBundle(origins=rb.origins.to(DEV), directions=rb.directions.to(DEV))
    nrs = nrb.get_ray_samples(rs.frustums.starts.to(DEV), rs.frustums.ends.to(DEV))
    wg = w.to(DEV)
    assert_close(n["rgb"](rgb=rgb.to(DEV), weights=wg), o["rgb"](rgb=rgb, weights=w), FP32_TOL, "rgb")
    assert_close(n["acc"](weights=wg), o["acc"](weights=w), FP32_TOL, "accumulation")
    assert_close(n["exp"](weights=wg, ray_samples=nrs), o["exp"](weights=w, ray_samples=rs), FP32_TOL, "expected depth")
    assert_close(n["sem"](sem.to(DEV), weights=wg), o["sem"](sem, weights=w), FP32_TOL, "semantics")
    assert_close(n["unc"](sem[..., :1].to(DEV), wg), o["unc"](sem[..., :1], w), FP32_TOL, "uncertainty")
    got_med, want_med = n["med"](weights=wg, ray_samples=nrs), o["med"](weights=w, ray_samples=rs)
    assert (got_med.cpu() == want_med).float().mean() > 0.995
    assert got_med.shape == want_med.shape == (R, 1)


def test_composite_fused_backward_matches_autograd():
    R, S, C = 129, 48, 19
    rb, rs, density, rgb, sem, g = _final_level(R, S, C, seed=21)
    for t in (density, rgb, sem):
        t.requires_grad_(True)
    for pass_sem in (False, True):
        for t in (density, rgb, sem):
            t.grad = None
        w = rs.get_weights(density)
        out_rgb = O.RGBRenderer("last_sample").train()(rgb=rgb, weights=w)
        acc = O.AccumulationRenderer()(weights=w)
        dexp = O.DepthRenderer("expected")(weights=w, ray_samples=rs)
        osem = O.SemanticRenderer()(sem, weights=w if pass_sem else w.detach())
        gs = [torch.randn(R, 3, generator=g), torch.randn(R, 1, generator=g), torch.randn(R, 1, generator=g),
              torch.randn(R, C, generator=g), torch.randn(R, S, 1, generator=g)]
        loss = (out_rgb * gs[0]).sum() + (acc * gs[1]).sum() + (dexp * gs[2]).sum() + (osem * gs[3]).sum() + (w * gs[4]).sum()
        loss.backward()
        d = density.detach().to(DEV).requires_grad_(True)
        c = rgb.detach().to(DEV).requires_grad_(True)
        s = sem.detach().to(DEV).requires_grad_(True)
        out = ops.composite(density=d, rgb=c, sem=s, starts=rs.frustums.starts.to(DEV), ends=rs.frustums.ends.to(DEV),
                            sem_weight_grad=pass_sem)
        assert_close(out["rgb"], out_rgb, FP32_TOL, "rgb")
        assert_close(out["expected_depth"], dexp, FP32_TOL, "expected depth")
        assert_close(out["semantics"], osem, FP32_TOL, "semantics")
        gd = [t.to(DEV) for t in gs]
        l2 = ((out["rgb"] * gd[0]).sum() + (out["accumulation"] * gd[1]).sum() + (out["expected_depth"] * gd[2]).sum()
              + (out["semantics"] * gd[3]).sum() + (out["weights"] * gd[4]).sum())
        l2.backward()
        assert_close(d.grad, density.grad, FP32_TOL, f"d_density pass_sem={pass_sem}")
        assert_close(c.grad, rgb.grad, FP32_TOL, "d_rgb")
        assert_close(s.grad, sem.grad, FP32_TOL, "d_sem")


def test_empty_batches():
    e = torch.empty(0, 48, device=DEV)
    out = ops.composite(density=e, starts=e, ends=e, rgb=torch.empty(0, 48, 3, device=DEV))
    assert out["rgb"].shape == (0, 3) and out["weights"].shape == (0, 48, 1)
    y = ops.mlp(torch.empty(0, 10, device=DEV), [torch.randn(16, 10, device=DEV), torch.randn(1, 16, device=DEV)],
                [torch.randn(16, device=DEV), torch.randn(1, device=DEV)], 1)
    assert y.shape == (0, 1)
