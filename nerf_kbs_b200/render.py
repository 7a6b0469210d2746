"""Full-frame inference driver: an image's rays are generated on the device, pushed through
the model in chunks of ``eval_num_rays_per_chunk`` (1 << 16 in the reference, ref:train.py:77), and the semantic
logits are turned into a label / colour image by one kernel (ref:nerfacto.py:574-578; semantic_nerfw.py:346-351).
With several ranks every rank renders one contiguous block of the frame's rays (the reference strides whole eval images
over ranks, ref:dataparser/datamanager_z.py:163-165; a single frame is split here) and one all-gather of the packed
outputs hands every rank the whole frame."""
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


def rank_ray_range(num_rays: int, rank: int, world_size: int) -> range:
    """Rays of a frame that ``rank`` renders: ``world_size`` contiguous blocks of ceil(N / world) rays (the last one
    shorter), so that the gathered per-rank results are the frame in ray order."""
    per = (num_rays + world_size - 1) // world_size
    return range(min(rank * per, num_rays), min((rank + 1) * per, num_rays))


@torch.no_grad()
def get_outputs_for_camera_ray_bundle(model, ray_bundle: RayBundle, num_rays_per_chunk: int = 1 << 16,
                                      rank: int = 0, world_size: int = 1, group=None, keys=RENDER_KEYS
                                      ) -> Dict[str, torch.Tensor]:
    """nerfstudio Model.get_outputs_for_camera_ray_bundle: chunked eval forward over all rays of ``ray_bundle``;
    returns flat [N, ...] tensors (the caller reshapes to the image).  With several ranks every rank renders its
    block of the rays (``rank_ray_range``) in chunks of at most ``num_rays_per_chunk`` and ONE all-gather of the packed
    outputs gives every rank the whole frame (N x 27 floats for rgb + depths + accumulation + 19 class logits)."""
    if model.training:
        raise RuntimeError("rendering runs in eval mode (model.eval())")
    N = len(ray_bundle)
    mine = rank_ray_range(N, rank, world_size)
    out: Dict[str, torch.Tensor] = {}
    for a in range(mine.start, mine.stop, num_rays_per_chunk):
        b = min(a + num_rays_per_chunk, mine.stop)
        res = model(_slice(ray_bundle, a, b))
        for k in keys:
            v = res.get(k)
            if v is None:
                continue
            if k not in out:
                out[k] = torch.empty((len(mine),) + tuple(v.shape[1:]), device=v.device, dtype=v.dtype)
            out[k][a - mine.start:b - mine.start] = v
    if world_size == 1:
        return out
    # pack every output of this rank's block into one [per, C_total] float buffer, all-gather, slice per key
    per = (N + world_size - 1) // world_size
    if per * (world_size - 1) >= N:
        raise ValueError(f"{N} rays cannot be split over {world_size} ranks (a rank would render nothing)")
    names = [k for k in keys if k in out]
    widths = [int(out[k][0].numel()) for k in names]
    dev = out[names[0]].device
    packed = torch.zeros((per, sum(widths)), device=dev, dtype=torch.float32)
    c = 0
    for k, w in zip(names, widths):
        packed[:len(mine), c:c + w] = out[k].reshape(len(mine), w)
        c += w
    full = torch.empty((world_size * per, sum(widths)), device=dev, dtype=torch.float32)
    dist.all_gather_into_tensor(full, packed, group=group)
    res, c = {}, 0
    for k, w in zip(names, widths):
        res[k] = full[:N, c:c + w].reshape((N,) + tuple(out[k].shape[1:])).to(out[k].dtype)
        c += w
    return res


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
