"""Full-frame inference driver: an image's rays are generated on the device, pushed through
the model in chunks of ``eval_num_rays_per_chunk`` (1 << 16 in the reference, ref:train.py:77), and the semantic
logits are turned into a label / colour image by one kernel (ref:nerfacto.py:574-578; semantic_nerfw.py:346-351).
With several ranks the chunks are dealt round-robin (rank-strided like ref:dataparser/datamanager_z.py:163-165) and
summed into rank-identical full-size buffers (each pixel is written by exactly one rank)."""
from __future__ import annotations

from typing import Dict, Optional

import torch
import torch.distributed as dist

from .data import RayGenerator, SemanticPalette, image_ray_indices
from .rays import RayBundle

RENDER_KEYS = ("rgb", "depth", "accumulation", "expected_depth", "semantics", "prop_depth_0", "prop_depth_1")


def _slice(rb: RayBundle, a: int, b: int) -> RayBundle:
    cut = lambda t: t[a:b] if t is not None else None
    return RayBundle(cut(rb.origins), cut(rb.directions), cut(rb.pixel_area), cut(rb.camera_indices), cut(rb.nears),
                     cut(rb.fars), {k: cut(v) for k, v in (rb.metadata or {}).items()})


@torch.no_grad()
def get_outputs_for_camera_ray_bundle(model, ray_bundle: RayBundle, num_rays_per_chunk: int = 1 << 16,
                                      rank: int = 0, world_size: int = 1, group=None, keys=RENDER_KEYS
                                      ) -> Dict[str, torch.Tensor]:
    """nerfstudio Model.get_outputs_for_camera_ray_bundle: chunked eval forward over all rays of ``ray_bundle``;
    returns flat [N, ...] tensors (the caller reshapes to the image)."""
    if model.training:
        raise RuntimeError("rendering runs in eval mode (model.eval())")
    N = len(ray_bundle)
    out: Dict[str, torch.Tensor] = {}
    for ci, a in enumerate(range(0, N, num_rays_per_chunk)):
        if ci % world_size != rank:
            continue
        b = min(a + num_rays_per_chunk, N)
        res = model(_slice(ray_bundle, a, b))
        for k in keys:
            v = res.get(k)
            if v is None:
                continue
            if k not in out:
                out[k] = torch.zeros((N,) + tuple(v.shape[1:]), device=v.device, dtype=v.dtype)
            out[k][a:b] = v
    if world_size > 1:
        for k in sorted(out):
            dist.all_reduce(out[k], group=group)
    return out


@torch.no_grad()
def render_camera(model, ray_generator: RayGenerator, camera_index: int, num_rays_per_chunk: int = 1 << 16,
                  palette: Optional[SemanticPalette] = None, rank: int = 0, world_size: int = 1, group=None,
                  colormaps: bool = True) -> Dict[str, torch.Tensor]:
    """Render one full frame: rgb [H,W,3], depth / accumulation [H,W,1], semantics logits [H,W,C]; with a palette,
    ``semantics_label`` [H,W] int64 and ``semantics_colormap`` [H,W,3] uint8; with ``colormaps``, the colour-mapped
    ``accumulation_colormap`` and ``depth_colormap`` / ``prop_depth_i_colormap`` [H,W,3] float images."""
    cams = ray_generator.cameras
    H, W = cams.height, cams.width
    idx = image_ray_indices(camera_index, H, W, ray_generator.c2w.device)
    rb = ray_generator(idx)
    flat = get_outputs_for_camera_ray_bundle(model, rb, num_rays_per_chunk, rank, world_size, group)
    out = {k: v.reshape(H, W, *v.shape[1:]) for k, v in flat.items()}
    if palette is not None and "semantics" in out:
        out["semantics_label"], out["semantics_colormap"] = palette.colormap(out["semantics"])
    if colormaps:
        # the images of get_image_metrics_and_images (ref:nerfacto.py:574-578,608-613), on the device
        from .colormaps import apply_colormap, apply_depth_colormap
        out["accumulation_colormap"] = apply_colormap(out["accumulation"])
        for k in ("depth", "prop_depth_0", "prop_depth_1"):
            if k in out:
                out[k + "_colormap"] = apply_depth_colormap(out[k], accumulation=out["accumulation"])
    return out
