"""GPU parity of the one-launch per-ray losses (nkbs_ray_losses, csrc/losses.cu) against the oracle's
interlevel / distortion losses and torch's MSE / cross-entropy, values and gradients."""
import pytest
import torch

from oracle import nerfstudio_torch as O
from nerf_kbs_b200 import ops
from tests.util import assert_close

pytestmark = pytest.mark.gpu
DEV = "cuda"


class _RS:
    """Minimal stand-in for RaySamples: only the spacing-space bin edges matter to the losses."""

    def __init__(self, sdist):
        self.spacing_starts = sdist[..., :-1, None]
        self.spacing_ends = sdist[..., 1:, None]


def _hist(R, S, g):
    """Random sorted bin edges in [0,1] and positive weights of sum < 1.  Weights stay well above the loss's 1e-7
    regulariser: clip(w - w_outer)^2 / (w + 1e-7) is ill-conditioned in fp32 for w ~ 1e-7 (both sides then differ
    by their cumsum order), which is not what this test is about."""
    edges = torch.sort(torch.rand(R, S + 1, generator=g), dim=-1).values
    w = torch.rand(R, S, generator=g) + 0.05
    w = w / w.sum(-1, keepdim=True) * torch.rand(R, 1, generator=g)
    return edges, w


@pytest.mark.parametrize("R,levels", [(257, (256, 96, 48)), (5, (64, 48)), (33, (7, 5, 3)), (1, (1, 1))])
def test_interlevel_and_distortion_match_oracle(R, levels):
    g = torch.Generator().manual_seed(R)
    hs = [_hist(R, S, g) for S in levels]
    ws = [w.clone().requires_grad_() for _, w in hs]
    rsl = [_RS(e) for e, _ in hs]
    want_i = O.interlevel_loss([w[..., None] for w in ws], rsl)
    want_d = O.distortion_loss([w[..., None] for w in ws], rsl)
    (want_i + 0.5 * want_d).backward()
    gw = [w.detach().to(DEV).requires_grad_() for _, w in hs]
    _, inter, dist, _ = ops.ray_losses(w_final=gw[-1], sdist_final=hs[-1][0].to(DEV), w_props=gw[:-1],
                                       sdist_props=[e.to(DEV) for e, _ in hs[:-1]])
    assert_close(inter, want_i, 1e-4, "interlevel")
    assert_close(dist, want_d, 1e-4, "distortion")
    (inter + 0.5 * dist).backward()
    for lvl, (a, b) in enumerate(zip(gw, ws)):
        assert_close(a.grad, b.grad, 1e-4, f"d/dw level {lvl}")


def test_interlevel_known_answer():
    """One ray, proposal = one bin [0,1] of weight 0.5; final = two bins of weights 0.4 and 0.7: only the second
    exceeds the envelope: loss = mean(0, (0.7-0.5)^2/(0.7+1e-7))."""
    sd_p = torch.tensor([[0.0, 1.0]], device=DEV)
    w_p = torch.tensor([[0.5]], device=DEV, requires_grad=True)
    sd_f = torch.tensor([[0.0, 0.5, 1.0]], device=DEV)
    w_f = torch.tensor([[0.4, 0.7]], device=DEV)
    _, inter, _, _ = ops.ray_losses(w_final=w_f, sdist_final=sd_f, w_props=[w_p], sdist_props=[sd_p], want_distortion=False)
    want = 0.5 * (0.2 ** 2) / (0.7 + 1e-7)
    assert abs(inter.item() - want) < 1e-7
    inter.backward()
    assert abs(w_p.grad.item() - (-2 * 0.2 / (0.7 + 1e-7) / 2)) < 1e-6


@pytest.mark.parametrize("C,ignore", [(19, 19), (19, -100), (40, 7)])
def test_pixel_losses_match_torch(C, ignore):
    R = 1000
    g = torch.Generator().manual_seed(C)
    rgb = torch.rand(R, 3, generator=g).requires_grad_()
    img = torch.rand(R, 3, generator=g)
    logits = (torch.randn(R, C, generator=g) * 3).requires_grad_()
    labels = torch.randint(0, C, (R,), generator=g)
    if ignore >= 0:
        labels[::7] = ignore
    want_m = torch.nn.functional.mse_loss(img, rgb)
    want_c = torch.nn.functional.cross_entropy(logits, labels, ignore_index=ignore)
    (want_m + 0.05 * want_c).backward()
    grgb, glog = rgb.detach().to(DEV).requires_grad_(), logits.detach().to(DEV).requires_grad_()
    mse, _, _, ce = ops.ray_losses(rgb=grgb, image=img.to(DEV), logits=glog, labels=labels.to(DEV), ignore_index=ignore)
    assert_close(mse, want_m, 1e-5, "mse")
    assert_close(ce, want_c, 1e-5, "cross-entropy")
    (mse + 0.05 * ce).backward()
    assert_close(grgb.grad, rgb.grad, 1e-5, "d/drgb")
    assert_close(glog.grad, logits.grad, 1e-5, "d/dlogits")


def test_all_labels_ignored_gives_zero():
    logits = torch.randn(8, 5, device=DEV, requires_grad=True)
    labels = torch.full((8,), 5, device=DEV)
    _, _, _, ce = ops.ray_losses(logits=logits, labels=labels, ignore_index=5)
    ce.backward()
    assert ce.item() == 0.0 and float(logits.grad.abs().max()) == 0.0
