"""Scalar colour maps of the rendered frame on the device — nerfstudio ``utils.colormaps.apply_colormap`` /
``apply_depth_colormap`` as the reference calls them (ref:nerfacto.py:574-578,608-613; ref:semantic_nerfw.py:305):
same names, arguments and defaults.  One kernel maps the image through a 256-entry colour table
(``nkbs_apply_colormap`` / ``nkbs_apply_depth_colormap``, csrc/raygen.cu); the depth range, when not given, is
reduced on the device (the reference's ``float(torch.min(depth))`` is a host synchronisation per image).

The colour TABLES are data: nerfstudio takes them from matplotlib (``cm.get_cmap(name).colors``).  matplotlib is
not installed in this image, so ``colormap_table`` uses it when importable and otherwise evaluates the published
polynomial fits of the two maps the reference uses ("viridis", "turbo"); any [256,3] table can be passed in."""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

from ._lib import call, ptr, require_cuda, stream

_POLY = {
    # 6th-order fit of matplotlib's viridis (M. Zucker's public-domain fit), coefficients c0..c6 per channel
    "viridis": ((0.2777273272234177, 0.005407344544966578, 0.3340998053353061),
                (0.1050930431085774, 1.404613529898575, 1.384590162594685),
                (-0.3308618287255563, 0.214847559468213, 0.09509516302823659),
                (-4.634230498983486, -5.799100973351585, -19.33244095627987),
                (6.228269936347081, 14.17993336680509, 56.69055260068105),
                (4.776384997670288, -13.74514537774601, -65.35303263337234),
                (-5.435455855934631, 4.645852612178535, 26.3124352495832)),
    # 5th-order fit of turbo (A. Mikhailov, Google, Apache-2.0 blog post), coefficients of 1, x, .., x^5
    "turbo": ((0.13572138, 0.09140261, 0.10667330), (4.61539260, 2.19418839, 12.64194608),
              (-42.66032258, 4.84296658, -60.58204836), (132.13108234, -14.18503333, 110.36276771),
              (-152.94239396, 4.27729857, -89.90310912), (59.28637943, 2.82956604, 27.34824973)),
}
_TABLES: Dict[Tuple[str, str], torch.Tensor] = {}


def colormap_table(cmap: str = "viridis") -> torch.Tensor:
    """[256,3] float32 CPU colour table of ``cmap``: matplotlib's when it is installed, else the polynomial fit."""
    try:
        from matplotlib import colormaps as _mpl  # type: ignore
        return torch.tensor(_mpl[cmap].colors, dtype=torch.float32)
    except Exception:
        pass
    if cmap not in _POLY:
        raise ValueError(f"no colour table for {cmap!r} (matplotlib is not installed; built in: {sorted(_POLY)})")
    x = torch.linspace(0.0, 1.0, 256, dtype=torch.float64)[:, None]
    acc = torch.zeros(256, 3, dtype=torch.float64)
    for c in reversed(_POLY[cmap]):
        acc = acc * x + torch.tensor(c, dtype=torch.float64)
    return acc.clamp(0.0, 1.0).float()


def _device_table(cmap, device) -> torch.Tensor:
    if isinstance(cmap, torch.Tensor):
        if cmap.shape != (256, 3):
            raise ValueError("a colour table is [256,3]")
        return cmap.to(device=device, dtype=torch.float32).contiguous()
    key = (cmap, str(device))
    if key not in _TABLES:
        _TABLES[key] = colormap_table(cmap).to(device)
    return _TABLES[key]


@torch.no_grad()
def apply_colormap(image: torch.Tensor, cmap="viridis") -> torch.Tensor:
    """image [..., 1] in [0,1] -> rgb [..., 3] (``colormap[(image * 255).long()[..., 0]]``)."""
    require_cuda(image)
    x = image.float().reshape(-1).contiguous()
    out = torch.empty(x.shape[0], 3, device=x.device, dtype=torch.float32)
    call("nkbs_apply_colormap", x.shape[0], ptr(x), ptr(_device_table(cmap, x.device)), ptr(out), stream())
    return out.reshape(*image.shape[:-1], 3)


@torch.no_grad()
def apply_depth_colormap(depth: torch.Tensor, accumulation: Optional[torch.Tensor] = None,
                         near_plane: Optional[float] = None, far_plane: Optional[float] = None, cmap="turbo") -> torch.Tensor:
    """depth [..., 1] -> rgb [..., 3]: normalised to [0,1] by (near_plane, far_plane) — the image's own min / max
    where not given (``near_plane or float(torch.min(depth))``: a zero plane counts as not given, as in nerfstudio)
    — mapped through ``cmap`` and blended over white by ``accumulation``."""
    require_cuda(depth, accumulation)
    d = depth.float().reshape(-1).contiguous()
    acc = accumulation.float().reshape(-1).contiguous() if accumulation is not None else None
    out = torch.empty(d.shape[0], 3, device=d.device, dtype=torch.float32)
    have_near, have_far = bool(near_plane), bool(far_plane)
    ws = None if (have_near and have_far) else torch.empty(2, device=d.device, dtype=torch.int32)
    call("nkbs_apply_depth_colormap", d.shape[0], ptr(d), ptr(acc), ptr(_device_table(cmap, d.device)), int(have_near),
         float(near_plane or 0.0), int(have_far), float(far_plane or 0.0), ptr(ws), ptr(out), stream())
    return out.reshape(*depth.shape[:-1], 3)
