"""GPU: one proposal level as ONE launch (nkbs_proposal_level_fwd: [initial sampler] + density field + get_weights +
median depth + PDF resampling) against the route of four stand-alone launches it fuses — the arithmetic is the same,
so bins, weights and depths must be bit-identical — and against the CPU oracle's ProposalNetworkSampler
(ref:nerfacto.py:325-332,397; ref:semantic_nerfw.py:132-137,187)."""
import pytest
import torch

from oracle import nerfstudio_torch as O
import nerf_kbs_b200 as nk
from nerf_kbs_b200 import ops
from tests.util import assert_close, copy_model, make_jitter, rel_err, to_nk_bundle

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _models(kind):
    torch.manual_seed(0)
    if kind == "nerfacto":
        om, nm = O.NerfactoModel(O.ModelConfig()), nk.NerfactoModel(nk.ModelConfig())
    else:
        cfg = dict(use_semantic=True)
        om, nm = O.SemanticNerfWModel(O.ModelConfig(**cfg)), nk.SemanticNerfWModel(nk.ModelConfig(**cfg))
    with torch.no_grad():
        for p in om.parameters():
            if p.dim() == 2 and p.shape[1] == 2:
                p.mul_(300.0)
    nm = nm.to(DEV)
    copy_model(nm, om)
    return om, nm


def _sample(nm, rb, jit, fused):
    ops.set_fused_level(fused)
    try:
        for p in nm.parameters():
            p.grad = None
        rs, wl, rsl = nm.proposal_sampler(nm.collider(to_nk_bundle(rb, DEV)), density_fns=nm.density_fns,
                                          jitter=None if jit is None else [j.to(DEV) for j in jit])
        grads = None
        if torch.is_grad_enabled() and wl[0].requires_grad:
            sum((w ** 2).sum() * (i + 1) for i, w in enumerate(wl)).backward()
            grads = [p.grad.detach().clone() for net in nm.proposal_networks for p in net.parameters()]
        return rs, wl, rsl, grads
    finally:
        ops.set_fused_level(False)


@pytest.mark.parametrize("kind", ["nerfacto", "semantic_nerfw"])
@pytest.mark.parametrize("train", [True, False])
def test_fused_level_equals_unfused_route(kind, train):
    R = 515  # not a multiple of the rays per CTA
    om, nm = _models(kind)
    nm.train(train)
    rb = O.synthetic_ray_bundle(R, seed=3)
    jit = make_jitter(R, seed=4) if train else None
    a = _sample(nm, rb, jit, fused=False)
    b = _sample(nm, rb, jit, fused=True)
    assert torch.equal(a[0].spacing_bins, b[0].spacing_bins) and torch.equal(a[0].euclid_bins, b[0].euclid_bins)
    for lvl in range(2):
        assert torch.equal(a[1][lvl], b[1][lvl]), f"weights of level {lvl}"
        assert torch.equal(a[2][lvl].spacing_bins, b[2][lvl].spacing_bins), f"spacing bins of level {lvl}"
        assert torch.equal(a[2][lvl].euclid_bins, b[2][lvl].euclid_bins), f"euclidean bins of level {lvl}"
        assert torch.equal(a[1][lvl]._nkbs_depth_med, b[1][lvl]._nkbs_depth_med), f"median depth of level {lvl}"
        assert b[1][lvl].shape == (R, (256, 96)[lvl], 1)
    if train:
        for ga, gb in zip(a[3], b[3]):
            assert rel_err(gb, ga) < 2e-5  # same kernels behind both backward routes; atomics order only


@pytest.mark.parametrize("kind", ["nerfacto", "semantic_nerfw"])
def test_fused_level_matches_oracle_sampler(kind):
    R = 300
    om, nm = _models(kind)
    om.train(), nm.train()
    rb = O.synthetic_ray_bundle(R, seed=5)
    jit = make_jitter(R, seed=6)
    want = om.proposal_sampler(om.collider(rb), om.density_fns, jitter=jit)
    got = _sample(nm, rb, jit, fused=True)
    # level 0: the initial sampler's bins, bit-exact
    assert torch.equal(got[2][0].spacing_bins.cpu(), O.ray_samples_to_sdist(want[2][0]))
    for lvl in range(2):
        assert_close(got[1][lvl], want[1][lvl], 1e-4, f"weights level {lvl}")
    for lvl, rs in ((1, got[2][1]), (2, got[0])):
        ref = want[2][1] if lvl == 1 else want[0]
        assert_close(rs.spacing_bins, O.ray_samples_to_sdist(ref), 1e-4, f"spacing bins level {lvl}")


def _sample_resample(nm, rb, jit, fused):
    ops.set_fused_level(False)
    ops.set_fused_resample(fused)
    try:
        for p in nm.parameters():
            p.grad = None
        rs, wl, rsl = nm.proposal_sampler(nm.collider(to_nk_bundle(rb, DEV)), density_fns=nm.density_fns,
                                          jitter=None if jit is None else [j.to(DEV) for j in jit])
        grads = None
        if torch.is_grad_enabled() and wl[0].requires_grad:
            sum((w ** 2).sum() * (i + 1) for i, w in enumerate(wl)).backward()
            grads = [p.grad.detach().clone() for net in nm.proposal_networks for p in net.parameters()]
        return rs, wl, rsl, grads
    finally:
        ops.set_fused_resample(True)


@pytest.mark.parametrize("kind", ["nerfacto", "semantic_nerfw"])
@pytest.mark.parametrize("train", [True, False])
def test_weights_resample_launch_equals_two_launches(kind, train):
    """nkbs_weights_resample_fwd (get_weights + median depth + the next level's PDF sampler in one launch, the default
    route of ProposalNetworkSampler) against nkbs_composite_fwd followed by nkbs_pdf_sample: bit-identical."""
    R = 517
    om, nm = _models(kind)
    nm.train(train)
    rb = O.synthetic_ray_bundle(R, seed=7)
    jit = make_jitter(R, seed=8) if train else None
    a = _sample_resample(nm, rb, jit, fused=False)
    b = _sample_resample(nm, rb, jit, fused=True)
    assert torch.equal(a[0].spacing_bins, b[0].spacing_bins) and torch.equal(a[0].euclid_bins, b[0].euclid_bins)
    for lvl in range(2):
        assert torch.equal(a[1][lvl], b[1][lvl]), f"weights of level {lvl}"
        assert torch.equal(a[2][lvl].spacing_bins, b[2][lvl].spacing_bins), f"spacing bins of level {lvl}"
        assert torch.equal(a[1][lvl]._nkbs_depth_med, b[1][lvl]._nkbs_depth_med), f"median depth of level {lvl}"
    if train:
        for ga, gb in zip(a[3], b[3]):
            assert rel_err(gb, ga) < 2e-5
