"""Device-side neighbours of the ray-batch path: pinhole ray generation from (camera, row, col)
indices, mask-aware pixel sampling and the semantic palette, as the reference's datamanager / dataset drive them
(ref:dataparser/datamanager_z.py:96,252-278; ref:dataparser/dataset_z.py:180-191; ref:nerfacto.py:508-512)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Sequence, Tuple

import torch

from ._lib import call, ptr, require_cuda, stream
from .rays import RayBundle


@dataclass
class Cameras:
    """Undistorted perspective cameras: ``camera_to_worlds`` [N,3,4], per-camera fx, fy, cx, cy, one image size."""
    camera_to_worlds: torch.Tensor
    fx: torch.Tensor
    fy: torch.Tensor
    cx: torch.Tensor
    cy: torch.Tensor
    height: int
    width: int

    def __len__(self) -> int:
        return self.camera_to_worlds.shape[0]

    def to(self, device) -> "Cameras":
        return Cameras(self.camera_to_worlds.to(device), self.fx.to(device), self.fy.to(device), self.cx.to(device),
                       self.cy.to(device), self.height, self.width)

    @property
    def intrinsics(self) -> torch.Tensor:
        return torch.stack([self.fx, self.fy, self.cx, self.cy], dim=-1).float().contiguous()


class RayGenerator(torch.nn.Module):
    """nerfstudio RayGenerator: ``ray_indices`` [N,3] (camera, row, col) -> RayBundle (one kernel)."""

    def __init__(self, cameras: Cameras):
        super().__init__()
        self.cameras = cameras
        self.register_buffer("c2w", cameras.camera_to_worlds.float().contiguous(), persistent=False)
        self.register_buffer("intrinsics", cameras.intrinsics, persistent=False)

    def forward(self, ray_indices: torch.Tensor) -> RayBundle:
        require_cuda(ray_indices, self.c2w)
        idx = ray_indices.to(torch.int64).contiguous()
        N, dev = idx.shape[0], idx.device
        o = torch.empty(N, 3, device=dev)
        d = torch.empty(N, 3, device=dev)
        dn = torch.empty(N, 1, device=dev)
        pa = torch.empty(N, 1, device=dev)
        cam = torch.empty(N, 1, device=dev, dtype=torch.int64)
        call("nkbs_generate_rays", N, ptr(idx), ptr(self.c2w), ptr(self.intrinsics), self.c2w.shape[0], ptr(o), ptr(d),
             ptr(dn), ptr(pa), ptr(cam), stream())
        return RayBundle(origins=o, directions=d, pixel_area=pa, camera_indices=cam, metadata={"directions_norm": dn})


def image_ray_indices(camera_index: int, height: int, width: int, device, rows: Optional[range] = None) -> torch.Tensor:
    """(camera, row, col) of every pixel of one image (or of a range of its rows), row-major."""
    rows = range(height) if rows is None else rows
    ys = torch.arange(rows.start, rows.stop, device=device, dtype=torch.int64)
    xs = torch.arange(width, device=device, dtype=torch.int64)
    yy, xx = torch.meshgrid(ys, xs, indexing="ij")
    cam = torch.full_like(yy, int(camera_index))
    return torch.stack([cam, yy, xx], dim=-1).reshape(-1, 3)


def sample_masked_pixels(masks: torch.Tensor, num_rays: int, oversample: float = 1.5, generator=None
                         ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """Draw (image, pixel) candidates uniformly and keep those whose motion mask is non-zero, compacted in order to
    exactly ``num_rays`` (ref:dataparser/dataset_z.py:180-191).  ``masks`` uint8 [num_images, H*W] on the device.
    Returns (image_indices [R], pixel_indices [R], n_kept device int32)."""
    require_cuda(masks)
    n_img, n_pix = masks.shape
    M = int(num_rays * oversample) + 32
    dev = masks.device
    img = torch.randint(0, n_img, (M,), device=dev, generator=generator)
    pix = (torch.rand(M, device=dev, generator=generator) * n_pix).floor().long().clamp_(max=n_pix - 1)
    return compact_masked(img, pix, masks, num_rays)


COMPACT_ONE_CTA_MAX = 1 << 15  # candidates one 1024-thread CTA still walks faster than four launches


def compact_masked(image_indices, pixel_indices, masks, num_rays: int):
    require_cuda(masks, image_indices, pixel_indices)
    dev = masks.device
    out_i = torch.zeros(num_rays, device=dev, dtype=torch.int64)
    out_p = torch.zeros(num_rays, device=dev, dtype=torch.int64)
    kept = torch.zeros(1, device=dev, dtype=torch.int32)
    M = image_indices.shape[0]
    # long candidate lists (pixel chunks) take the multi-CTA route; per-step lists one CTA
    ws = torch.empty((M + 1023) // 1024 + 1, device=dev, dtype=torch.int32) if M > COMPACT_ONE_CTA_MAX else None
    call("nkbs_compact_masked", M, ptr(image_indices.contiguous()), ptr(pixel_indices.contiguous()),
         ptr(masks.contiguous()), masks.shape[1], num_rays, ptr(out_i), ptr(out_p), ptr(kept), ptr(ws), stream())
    return out_i, out_p, kept


class SemanticPalette:
    """Class colours <-> labels on the device (KITTI-STEP-like palettes, ref:process_data/motion_mask_util.py:226-246)."""

    def __init__(self, colors: Sequence[Sequence[int]], device):
        if not 1 <= len(colors) <= 256:
            raise ValueError("1..256 classes")
        keys = [int(r) | (int(g) << 8) | (int(b) << 16) for r, g, b in colors]
        buf = torch.zeros(257, dtype=torch.int32)
        buf[0] = len(keys)
        buf[1:1 + len(keys)] = torch.tensor(keys, dtype=torch.int32)
        self.colors = [tuple(int(v) for v in c) for c in colors]
        self.table = buf.to(device)
        if self.table.device.type != "cuda":
            raise RuntimeError("nkbs operators run on CUDA tensors only (no CPU fallback)")

    def __len__(self) -> int:
        return len(self.colors)

    def color_to_label(self, rgb_u8: torch.Tensor, default_label: Optional[int] = None) -> torch.Tensor:
        """uint8 [..., 3] semantic image -> int64 labels; unknown colours map to ``default_label`` (= num classes)."""
        require_cuda(rgb_u8)
        flat = rgb_u8.to(torch.uint8).reshape(-1, 3).contiguous()
        out = torch.empty(flat.shape[0], device=flat.device, dtype=torch.int64)
        call("nkbs_color_to_label", flat.shape[0], ptr(flat), ptr(self.table),
             len(self) if default_label is None else int(default_label), ptr(out), stream())
        return out.reshape(rgb_u8.shape[:-1])

    def colormap(self, logits: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        """logits [..., C] -> (argmax labels int64 [...], palette colours uint8 [..., 3])."""
        require_cuda(logits)
        flat = logits.float().reshape(-1, logits.shape[-1]).contiguous()
        lab = torch.empty(flat.shape[0], device=flat.device, dtype=torch.int64)
        rgb = torch.empty(flat.shape[0], 3, device=flat.device, dtype=torch.uint8)
        call("nkbs_semantic_colormap", flat.shape[0], flat.shape[1], ptr(flat), ptr(self.table), ptr(lab), ptr(rgb), stream())
        return lab.reshape(logits.shape[:-1]), rgb.reshape(*logits.shape[:-1], 3)
