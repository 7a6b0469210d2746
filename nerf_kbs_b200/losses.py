"""Losses adjacent to the hot path: interlevel (proposal) loss, distortion loss, colour MSE
and semantic cross-entropy as called at ref:nerfacto.py:468,483-515 and ref:semantic_nerfw.py:253-279.  All of them
run in one ray-major CUDA kernel (csrc/losses.cu, ``ops.ray_losses``) that also writes the gradients."""
from __future__ import annotations

import torch

from . import ops


def ray_samples_to_sdist(ray_samples):
    if getattr(ray_samples, "spacing_bins", None) is not None:
        return ray_samples.spacing_bins
    return torch.cat([ray_samples.spacing_starts[..., 0], ray_samples.spacing_ends[..., -1:, 0]], dim=-1)


def _w(weights):
    # squeeze, not [..., 0]: the backward of a select builds a zero tensor and copies the gradient into it (two
    # launches per level), the backward of a squeeze is a view
    return weights.squeeze(-1) if (weights.dim() == 3 and weights.shape[-1] == 1) else (weights[..., 0] if weights.dim() == 3 else weights)


def interlevel_loss(weights_list, ray_samples_list):
    """Sum over proposal levels of mean(lossfun_outer(final histogram, proposal histogram))."""
    _, inter, _, _ = ops.ray_losses(
        w_final=_w(weights_list[-1]).detach(), sdist_final=ray_samples_to_sdist(ray_samples_list[-1]).detach(),
        w_props=[_w(w) for w in weights_list[:-1]], sdist_props=[ray_samples_to_sdist(rs) for rs in ray_samples_list[:-1]],
        want_distortion=False)
    return inter


def distortion_loss(weights_list, ray_samples_list):
    """mean over rays of lossfun_distortion(final bins, final weights)."""
    _, _, dist, _ = ops.ray_losses(w_final=_w(weights_list[-1]), sdist_final=ray_samples_to_sdist(ray_samples_list[-1]))
    return dist


def training_losses(weights_list, ray_samples_list, rgb=None, image=None, logits=None, labels=None, ignore_index=-100):
    """Every per-ray term of a training step in ONE launch: (rgb_mse, interlevel, distortion, cross_entropy)."""
    return ops.ray_losses(
        w_final=_w(weights_list[-1]), sdist_final=ray_samples_to_sdist(ray_samples_list[-1]),
        w_props=[_w(w) for w in weights_list[:-1]], sdist_props=[ray_samples_to_sdist(rs) for rs in ray_samples_list[:-1]],
        rgb=rgb, image=image, logits=logits, labels=labels, ignore_index=ignore_index)


def training_loss_total(weights_list, ray_samples_list, rgb, image, term_weights, unit_seed=True):
    """(total, (mse, interlevel, distortion), the same three weighted): the weighted total as the root of the backward
    pass, gradients weighted inside the kernel (``ops.ray_loss_total``).  Used by the step driver; the loss dict route stays ``training_losses``."""
    return ops.ray_loss_total(
        _w(weights_list[-1]), ray_samples_to_sdist(ray_samples_list[-1]),
        [_w(w) for w in weights_list[:-1]], [ray_samples_to_sdist(rs) for rs in ray_samples_list[:-1]],
        rgb, image, term_weights, unit_seed=unit_seed)


def pixel_losses(rgb=None, image=None, logits=None, labels=None, ignore_index=-100):
    """Colour MSE and semantic cross-entropy only (evaluation: no weights lists)."""
    mse, _, _, ce = ops.ray_losses(rgb=rgb, image=image, logits=logits, labels=labels, ignore_index=ignore_index)
    return mse, ce
