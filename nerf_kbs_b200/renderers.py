"""Drop-in renderers and collider (nerfstudio model_components/renderers.py, scene_colliders.py) built at
ref:nerfacto.py:335-343 / ref:semantic_nerfw.py:129,140-144 and called at ref:nerfacto.py:404-449 /
ref:semantic_nerfw.py:194-229.  Each is one launch of the ray-major compositing kernel."""
from __future__ import annotations

import torch
from torch import nn

from . import ops
from .rays import RayBundle, RaySamples


class NearFarCollider(nn.Module):
    def __init__(self, near_plane: float, far_plane: float, **kwargs):
        super().__init__()
        self.near_plane, self.far_plane = near_plane, far_plane
        self._planes = {}

    def set_nears_and_fars(self, ray_bundle: RayBundle) -> RayBundle:
        # the two [R,1] constants are built once per (rays, device) and handed out read-only afterwards: four launches
        # fewer on every step (the sampler kernels only read them).  Entries are never evicted while somebody else
        # still holds them (a captured CUDA graph would keep reading the freed memory); the step driver owns its
        # planes itself (trainer.Trainer._planes) and does not come through here.
        o = ray_bundle.origins
        key = (tuple(o.shape[:-1]), o.device, o.dtype, float(self.near_plane), float(self.far_plane))
        planes = self._planes.get(key)
        if planes is None:
            if len(self._planes) > 8:
                import sys
                # getrefcount: the dict's tuple + the argument = 2 when nobody else holds the pair
                for k in [k for k, v in self._planes.items() if sys.getrefcount(v[0]) <= 2 and sys.getrefcount(v[1]) <= 2]:
                    del self._planes[k]
            planes = (torch.full_like(o[..., 0:1], self.near_plane), torch.full_like(o[..., 0:1], self.far_plane))
            self._planes[key] = planes
        ray_bundle.nears, ray_bundle.fars = planes
        return ray_bundle

    def forward(self, ray_bundle: RayBundle) -> RayBundle:
        if ray_bundle.nears is not None and ray_bundle.fars is not None:
            return ray_bundle
        return self.set_nears_and_fars(ray_bundle)


class RGBRenderer(nn.Module):
    def __init__(self, background_color="random"):
        super().__init__()
        if background_color not in ops.BG_MODES:
            raise NotImplementedError(f"background_color={background_color!r}: nkbs supports {sorted(ops.BG_MODES)} "
                                      "(the reference uses 'last_sample')")
        self.background_color = background_color

    def forward(self, rgb, weights, ray_indices=None, num_rays=None):
        if ray_indices is not None:
            raise NotImplementedError("packed samples are not on the reference path")
        out = ops.composite(weights=weights, rgb=rgb, background=self.background_color, eval_mode=not self.training,
                            want=("rgb",))
        return out["rgb"]


class AccumulationRenderer(nn.Module):
    def forward(self, weights, ray_indices=None, num_rays=None):
        if ray_indices is not None:
            raise NotImplementedError("packed samples are not on the reference path")
        return ops.composite(weights=weights, want=("acc",))["accumulation"]


class DepthRenderer(nn.Module):
    def __init__(self, method="median"):
        super().__init__()
        if method not in ("median", "expected"):
            raise NotImplementedError(f"Method {method} not implemented")
        self.method = method

    def forward(self, weights, ray_samples: RaySamples, ray_indices=None, num_rays=None):
        if ray_indices is not None:
            raise NotImplementedError("packed samples are not on the reference path")
        st, en = ray_samples.frustums.starts, ray_samples.frustums.ends
        if self.method == "median":
            return ops.composite(weights=weights, starts=st, ends=en, want=("depth_med",))["depth"]
        return ops.composite(weights=weights, starts=st, ends=en, want=("depth_exp",))["expected_depth"]


class SemanticRenderer(nn.Module):
    def forward(self, semantics, weights):
        return ops.composite(weights=weights, sem=semantics, sem_weight_grad=True, want=("sem",))["semantics"]


class UncertaintyRenderer(nn.Module):
    def forward(self, betas, weights):
        return ops.composite(weights=weights, sem=betas, sem_weight_grad=True, want=("sem",))["semantics"]
