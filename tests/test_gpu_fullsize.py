"""GPU: the kernels at BASELINE.json's full sizes (65 536 rays, 48 / 96 / 256 samples, 16-level 2^19 grid), where the
CPU oracle would take minutes, held to properties that do not depend on the size:

* chunk invariance — a batch gives bit-identical per-sample results whether it is launched whole (one slice of the
  level range, many waves) or in small chunks (sliced sub-wave grids, csrc/encode.cu level_slices);
* linearity of the encoding in the table;
* a checksum of the gradient scatter: the trilinear weights of a sample sum to one on every level, so the column sums of
  the table gradient over a level's rows equal the column sums of the incoming feature gradient;
* closed forms of compositing (accumulation = 1 - exp(-sum delta*sigma)) and sortedness / range of the resampled bins.
"""
import math

import pytest
import torch

import nerf_kbs_b200 as nk
from nerf_kbs_b200 import ops

pytestmark = pytest.mark.gpu
DEV = "cuda"
R_FULL = 65536


def _main_grid():
    enc = nk.HashEncoding(num_levels=16, min_res=16, max_res=2048, log2_hashmap_size=19)
    return enc.grid


def _points(n, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.rand(n, 3, generator=g).to(DEV)


def test_encode_full_size_chunk_invariance_and_linearity():
    grid = _main_grid()
    N = R_FULL * 48
    g = torch.Generator().manual_seed(0)
    t1 = ((torch.rand(grid.L * 2 ** grid.log2_T, 2, generator=g) * 2 - 1) * 1e-3).to(DEV)
    t2 = ((torch.rand(grid.L * 2 ** grid.log2_T, 2, generator=g) * 2 - 1) * 1e-3).to(DEV)
    x = _points(N, 1)
    whole, _, _, _ = ops.encode_raw(0, grid, None, t1, x=x)
    assert whole.shape == (N, 32) and torch.isfinite(whole).all()
    # chunks of 4096 x 48 samples take the sliced grid; the whole batch takes the unsliced one
    step = 4096 * 48
    for lo in (0, 5 * step, N - step):
        part, _, _, _ = ops.encode_raw(0, grid, None, t1, x=x[lo:lo + step].contiguous())
        assert torch.equal(part, whole[lo:lo + step]), f"chunk at {lo} differs from the whole batch"
    # linear in the table (fp32: three roundings apart)
    a = 0.75
    mixed, _, _, _ = ops.encode_raw(0, grid, None, a * t1 + t2, x=x[:step].contiguous())
    second, _, _, _ = ops.encode_raw(0, grid, None, t2, x=x[:step].contiguous())
    want = a * whole[:step] + second
    assert (mixed - want).abs().max() <= 1e-5 * want.abs().max()
    # values stay inside the hull of the table entries (convex combination of 8 rows)
    assert whole.abs().max() <= t1.abs().max() * (1 + 1e-6)


def test_encode_backward_full_size_checksum_and_chunk_sum():
    from nerf_kbs_b200._lib import call, host_floats, ptr, stream
    grid = _main_grid()
    L, T = grid.L, 2 ** grid.log2_T
    N = R_FULL * 48
    x = _points(N, 2)
    g = torch.Generator().manual_seed(3)
    dfeat = torch.randn(N, 2 * L, generator=g).to(DEV)

    def scatter(xs, ds):
        dt = torch.zeros(L * T, 2, device=DEV)
        call("nkbs_encode_bwd", ptr(xs), ptr(ds), xs.shape[0], host_floats(grid.scalings), L, grid.log2_T, 2, L, ptr(dt), stream())
        return dt

    whole = scatter(x, dfeat)
    # checksum: per level and feature, sum over the level's rows == sum over samples of the incoming gradient
    got = whole.double().reshape(L, T, 2).sum(1)                       # [L, 2]
    want = dfeat.double().reshape(N, L, 2).sum(0)                      # [L, 2]
    scale = dfeat.double().abs().reshape(N, L, 2).sum(0)               # magnitude the rounding error scales with
    assert ((got - want).abs() <= 1e-5 * scale).all(), (got - want).abs().max().item()
    # the whole batch equals the sum of its halves (different grids: unsliced, both many waves) up to summation order
    h = N // 2
    halves = scatter(x[:h].contiguous(), dfeat[:h].contiguous()) + scatter(x[h:].contiguous(), dfeat[h:].contiguous())
    assert (whole - halves).abs().max() <= 1e-4 * whole.abs().max()
    # and a small (sliced) chunk equals the same chunk scattered with every level in one slice's worth of threads
    step = 4096 * 48
    small = scatter(x[:step].contiguous(), dfeat[:step].contiguous())
    got = small.double().reshape(L, T, 2).sum(1)
    want = dfeat[:step].double().reshape(step, L, 2).sum(0)
    scale = dfeat[:step].double().abs().reshape(step, L, 2).sum(0)
    assert ((got - want).abs() <= 1e-5 * scale).all()


@pytest.mark.parametrize("S", [48, 96, 256])
def test_composite_full_size_closed_forms(S):
    R = R_FULL
    g = torch.Generator().manual_seed(S)
    edges = torch.sort(torch.rand(R, S + 1, generator=g) * 4 + 0.05, dim=-1).values.to(DEV)
    density = (torch.rand(R, S, generator=g) * 3).to(DEV)
    rgb = torch.rand(R, S, 3, generator=g).to(DEV)
    starts, ends = edges[:, :-1], edges[:, 1:]
    out = ops.composite(density=density, rgb=rgb, starts=starts, ends=ends, want=("rgb", "acc", "depth_exp", "depth_med"))
    w = out["weights"][..., 0]
    assert torch.isfinite(w).all() and (w >= 0).all()
    acc = out["accumulation"][:, 0]
    # accumulation = 1 - T_end = 1 - exp(-sum delta*sigma)
    tau = ((ends - starts).double() * density.double()).sum(-1)
    want = 1.0 - torch.exp(-tau)
    assert (acc.double() - want).abs().max() <= 2e-5
    assert (w.double().sum(-1) - acc.double()).abs().max() <= 1e-5
    # rgb with the last sample as background is a convex combination of the samples' colours
    assert (out["rgb"] >= rgb.min(dim=1).values - 1e-5).all() and (out["rgb"] <= rgb.max(dim=1).values + 1e-5).all()
    # median depth is one of the interval mid points and lies inside the ray's span
    med = out["depth"][:, 0]
    assert (med >= starts[:, 0]).all() and (med <= ends[:, -1]).all()
    # chunk invariance (bit-exact: one warp per ray either way)
    part = ops.composite(density=density[:4096].contiguous(), rgb=rgb[:4096].contiguous(), starts=starts[:4096], ends=ends[:4096],
                         want=("rgb", "acc", "depth_med"))
    assert torch.equal(part["weights"], out["weights"][:4096]) and torch.equal(part["rgb"], out["rgb"][:4096])
    assert torch.equal(part["depth"], out["depth"][:4096])


def test_samplers_full_size_sorted_and_in_range():
    R = R_FULL
    nears, fars = torch.full((R, 1), 0.001, device=DEV), torch.full((R, 1), 1000.0, device=DEV)
    g = torch.Generator().manual_seed(5)
    t_rand = torch.rand(R, 1, generator=g).to(DEV)
    sb, eb = ops.initial_sampler(nears, fars, 256, t_rand)
    assert sb.shape == (R, 257)
    assert (sb[:, 1:] >= sb[:, :-1]).all() and (eb[:, 1:] >= eb[:, :-1]).all()
    assert sb.min() >= 0 and sb.max() <= 1 and eb.min() >= 0.001 * (1 - 1e-6) and eb.max() <= 1000.0 * (1 + 1e-3)
    prev_s, prev_n = sb, 256
    for ns in (96, 48):
        w = torch.rand(R, prev_n, generator=g).to(DEV) ** 4          # peaky histograms
        w[::7] = 0.0                                                  # empty histograms: the padding keeps them valid
        rand = torch.rand(R, 1, generator=g).to(DEV)
        ob, oe = ops.pdf_sample(w, prev_s, nears, fars, ns, anneal=0.6, rand=rand)
        assert ob.shape == (R, ns + 1) and torch.isfinite(ob).all() and torch.isfinite(oe).all()
        assert (ob[:, 1:] >= ob[:, :-1]).all(), "spacing bins must be sorted"
        assert (oe[:, 1:] >= oe[:, :-1]).all(), "euclidean bins must be sorted"
        assert ob.min() >= prev_s.min() - 1e-6 and ob.max() <= prev_s.max() + 1e-6
        # chunk invariance, bit-exact
        pb, pe = ops.pdf_sample(w[:2048].contiguous(), prev_s[:2048].contiguous(), nears[:2048], fars[:2048], ns, anneal=0.6,
                                rand=rand[:2048].contiguous())
        assert torch.equal(pb, ob[:2048]) and torch.equal(pe, oe[:2048])
        prev_s, prev_n = ob, ns


def test_proposal_density_full_size_chunk_invariance():
    """Fused proposal field (encode + MLP + trunc_exp) on 65 536 x 96 samples: the whole batch equals its chunks."""
    torch.manual_seed(0)
    m = nk.NerfactoModel(nk.ModelConfig()).to(DEV).eval()
    R = R_FULL
    g = torch.Generator().manual_seed(9)
    o = (torch.rand(R, 3, generator=g) - 0.5).to(DEV)
    d = torch.randn(R, 3, generator=g)
    d = (d / d.norm(dim=-1, keepdim=True)).to(DEV)
    rb = m.collider(nk.RayBundle(o, d, torch.ones(R, 1, device=DEV), torch.zeros(R, 1, dtype=torch.long, device=DEV)))
    sb, eb = ops.initial_sampler(rb.nears, rb.fars, 96, None)
    with torch.no_grad():
        rs = rb.samples_from_bins(sb, eb)
        whole, _ = m.proposal_networks[1].get_density(rs)
        assert whole.shape[:2] == (R, 96) and torch.isfinite(whole).all() and (whole >= 0).all()
        k = 4096
        rb2 = nk.RayBundle(o[:k].contiguous(), d[:k].contiguous(), torch.ones(k, 1, device=DEV),
                           torch.zeros(k, 1, dtype=torch.long, device=DEV), nears=rb.nears[:k], fars=rb.fars[:k])
        part, _ = m.proposal_networks[1].get_density(rb2.samples_from_bins(sb[:k].contiguous(), eb[:k].contiguous()))
        assert torch.equal(part, whole[:k])
