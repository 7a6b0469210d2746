"""Ray containers with the attribute names the reference touches (nerfstudio cameras/rays.py):
``RayBundle`` (ref:nerfacto.py:396, datamanager_z.py:252-278), ``RaySamples``/``Frustums``
(ref:nerfacto.py:397-399, semantic_nerfw.py:187-193).  nerfstudio is not installed here, so these are
stand-alone; the operator classes only duck-type on these attributes."""
from __future__ import annotations

from dataclasses import dataclass, field as dc_field
from typing import Callable, Dict, Optional

import torch

from . import ops


@dataclass
class Frustums:
    origins: torch.Tensor      # [..., 3]
    directions: torch.Tensor   # [..., 3]
    starts: torch.Tensor       # [..., 1]
    ends: torch.Tensor         # [..., 1]
    pixel_area: Optional[torch.Tensor] = None
    # per-ray views kept so kernels can read [R,3] instead of the broadcast [R,S,3]
    ray_origins: Optional[torch.Tensor] = None
    ray_directions: Optional[torch.Tensor] = None

    def get_positions(self) -> torch.Tensor:
        if self.starts is None:
            return self.origins
        return self.origins + self.directions * (self.starts + self.ends) / 2

    @property
    def shape(self):
        return self.starts.shape[:-1] if self.starts is not None else self.origins.shape[:-1]


@dataclass
class RaySamples:
    frustums: Frustums
    camera_indices: Optional[torch.Tensor] = None
    deltas_: Optional[torch.Tensor] = None  # see ``deltas``: materialised on first use (no kernel reads it)
    spacing_starts: Optional[torch.Tensor] = None
    spacing_ends: Optional[torch.Tensor] = None
    spacing_to_euclidean_fn: Optional[Callable] = None
    metadata: Optional[Dict[str, torch.Tensor]] = None
    # contiguous [R,S+1] bin edges both domains + per-ray camera index, for the kernels
    spacing_bins: Optional[torch.Tensor] = None
    euclid_bins: Optional[torch.Tensor] = None
    ray_camera_indices: Optional[torch.Tensor] = None
    ray_nears: Optional[torch.Tensor] = None
    ray_fars: Optional[torch.Tensor] = None

    @property
    def shape(self):
        return self.frustums.shape

    @property
    def deltas(self) -> Optional[torch.Tensor]:
        """``ends - starts`` [..., S, 1].  The compositing kernels read the bin edges themselves, so the tensor is only
        built when somebody asks for it (one launch per level saved on the training step)."""
        if self.deltas_ is None and self.frustums is not None and self.frustums.starts is not None:
            self.deltas_ = self.frustums.ends - self.frustums.starts
        return self.deltas_

    @deltas.setter
    def deltas(self, value: Optional[torch.Tensor]) -> None:
        self.deltas_ = value

    def get_weights(self, densities: torch.Tensor) -> torch.Tensor:
        """alpha-compositing weights [..., S, 1] (ref:nerfacto.py:399; semantic_nerfw.py:192-193)."""
        S = densities.shape[-2]
        d = densities.reshape(-1, S)
        st = self.frustums.starts.reshape(-1, S)
        en = self.frustums.ends.reshape(-1, S)
        # the median depth of the level comes out of the same launch (the models ask for it as prop_depth_i)
        out = ops.composite(density=d, starts=st, ends=en, want=("depth_med",))
        w = out["weights"].reshape(densities.shape)
        w._nkbs_depth_med = out["depth"]
        return w


@dataclass
class RayBundle:
    origins: torch.Tensor
    directions: torch.Tensor
    pixel_area: Optional[torch.Tensor] = None
    camera_indices: Optional[torch.Tensor] = None
    nears: Optional[torch.Tensor] = None
    fars: Optional[torch.Tensor] = None
    metadata: Dict[str, torch.Tensor] = dc_field(default_factory=dict)

    def __len__(self):
        return self.origins.shape[0]

    def to(self, device):
        def mv(t):
            return t.to(device) if t is not None else None
        return RayBundle(mv(self.origins), mv(self.directions), mv(self.pixel_area), mv(self.camera_indices),
                         mv(self.nears), mv(self.fars), {k: mv(v) for k, v in (self.metadata or {}).items()})

    def get_ray_samples(self, bin_starts, bin_ends, spacing_starts=None, spacing_ends=None,
                        spacing_to_euclidean_fn=None) -> RaySamples:
        R, S = bin_starts.shape[0], bin_starts.shape[-2]
        cam = self.camera_indices[:, None, :].expand(R, S, 1) if self.camera_indices is not None else None
        pa = self.pixel_area[:, None, :].expand(R, S, 1) if self.pixel_area is not None else None
        fr = Frustums(origins=self.origins[:, None, :].expand(R, S, 3),
                      directions=self.directions[:, None, :].expand(R, S, 3),
                      starts=bin_starts, ends=bin_ends, pixel_area=pa,
                      ray_origins=self.origins, ray_directions=self.directions)
        return RaySamples(frustums=fr, camera_indices=cam, deltas_=None,
                          spacing_starts=spacing_starts, spacing_ends=spacing_ends,
                          spacing_to_euclidean_fn=spacing_to_euclidean_fn, metadata=self.metadata,
                          ray_camera_indices=self.camera_indices[:, 0] if self.camera_indices is not None else None,
                          ray_nears=self.nears, ray_fars=self.fars)

    def samples_from_bins(self, spacing_bins, euclid_bins, spacing_to_euclidean_fn=None) -> RaySamples:
        """RaySamples over contiguous [R,S+1] bin-edge arrays produced by the sampler kernels."""
        rs = self.get_ray_samples(
            bin_starts=euclid_bins[:, :-1, None], bin_ends=euclid_bins[:, 1:, None],
            spacing_starts=spacing_bins[:, :-1, None], spacing_ends=spacing_bins[:, 1:, None],
            spacing_to_euclidean_fn=spacing_to_euclidean_fn)
        rs.spacing_bins, rs.euclid_bins = spacing_bins, euclid_bins
        return rs
