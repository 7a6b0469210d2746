"""Drop-in field operators: same class names, constructor kwargs, defaults, parameter layouts and
return values as the nerfstudio 0.3.x fields the reference builds at ref:nerfacto.py:275-311 and
ref:semantic_nerfw.py:102-126 — computed by the nkbs sm_100a kernels."""
from __future__ import annotations

from enum import Enum
from typing import Dict, Optional, Tuple

import numpy as np
import torch
from torch import nn

from . import ops
from .rays import Frustums, RaySamples


class FieldHeadNames(Enum):
    RGB = "rgb"
    SH = "sh"
    DENSITY = "density"
    NORMALS = "normals"
    PRED_NORMALS = "pred_normals"
    UNCERTAINTY = "uncertainty"
    TRANSIENT_RGB = "transient_rgb"
    TRANSIENT_DENSITY = "transient_density"
    SEMANTICS = "semantics"


class SceneContraction(nn.Module):
    """Marker + torch expression of the L-inf contraction (ref:nerfacto.py:272).  Inside the fields the
    contraction runs fused in the encode kernel; calling the module directly evaluates it with torch
    ops on whatever device the input lives on."""

    def __init__(self, order=float("inf")):
        super().__init__()
        self.order = order

    def forward(self, positions):
        mag = torch.linalg.norm(positions, ord=self.order, dim=-1)[..., None]
        return torch.where(mag < 1, positions, (2 - (1 / mag)) * (positions / mag))


def hash_scalings(num_levels: int, min_res: int, max_res: int) -> torch.Tensor:
    """floor(min_res * g**l) evaluated as nerfstudio does (float32 via Tensor.__rpow__)."""
    levels = torch.arange(num_levels)
    growth = np.exp((np.log(max_res) - np.log(min_res)) / (num_levels - 1)) if num_levels > 1 else 1
    return torch.floor(min_res * growth**levels)


class HashEncoding(nn.Module):
    """Multiresolution hash grid with the torch path's table layout [L*T, F] and init."""

    def __init__(self, num_levels=16, min_res=16, max_res=1024, log2_hashmap_size=19, features_per_level=2,
                 hash_init_scale=0.001, implementation="nkbs"):
        super().__init__()
        if features_per_level != 2:
            raise ValueError("nkbs HashEncoding supports features_per_level=2 (the reference's value)")
        self.num_levels = num_levels
        self.features_per_level = features_per_level
        self.log2_hashmap_size = log2_hashmap_size
        self.hash_table_size = 2**log2_hashmap_size
        self.scalings = hash_scalings(num_levels, min_res, max_res)
        self.hash_offset = torch.arange(num_levels) * self.hash_table_size
        table = torch.rand(size=(self.hash_table_size * num_levels, features_per_level)) * 2 - 1
        table *= hash_init_scale
        self.hash_table = nn.Parameter(table)
        self.grid = ops.GridSpec(tuple(float(s) for s in self.scalings.tolist()), log2_hashmap_size, features_per_level)

    def get_out_dim(self) -> int:
        return self.num_levels * self.features_per_level

    def forward(self, in_tensor):
        """in_tensor [..., 3] in [0,1] -> [..., L*F]."""
        lead = in_tensor.shape[:-1]
        return ops.hash_encode(self.hash_table, self.grid, in_tensor).reshape(*lead, self.get_out_dim())


class MLP(nn.Module):
    """nn.Linear stack with ReLU between layers (nerfstudio field_components.mlp.MLP)."""

    def __init__(self, in_dim, num_layers, layer_width, out_dim=None, activation=nn.ReLU(), out_activation=None,
                 implementation="nkbs"):
        super().__init__()
        self.in_dim = in_dim
        self.out_dim = out_dim if out_dim is not None else layer_width
        if not (activation is None or isinstance(activation, nn.ReLU)):
            raise ValueError("nkbs MLP supports ReLU hidden activations")
        if not (out_activation is None or isinstance(out_activation, nn.Sigmoid)):
            raise ValueError("nkbs MLP supports None or Sigmoid output activations")
        if not 1 <= num_layers <= 3:
            raise ValueError("nkbs MLP supports 1..3 linear layers")
        layers = []
        if num_layers == 1:
            layers.append(nn.Linear(in_dim, self.out_dim))
        else:
            for i in range(num_layers - 1):
                layers.append(nn.Linear(in_dim if i == 0 else layer_width, layer_width))
            layers.append(nn.Linear(layer_width, self.out_dim))
        self.layers = nn.ModuleList(layers)
        self.activation = activation
        self.out_activation = out_activation

    def get_out_dim(self) -> int:
        return self.out_dim

    @property
    def relu_mask(self) -> int:
        n = len(self.layers)
        return ((1 << (n - 1)) - 1) if self.activation is not None else 0

    def weights_biases(self):
        return [l.weight for l in self.layers], [l.bias for l in self.layers]

    def forward(self, in_tensor):
        lead = in_tensor.shape[:-1]
        W, b = self.weights_biases()
        act = ops.ACT_SIGMOID if self.out_activation is not None else ops.ACT_NONE
        y = ops.mlp(in_tensor.reshape(-1, self.in_dim), W, b, self.relu_mask, act)
        return y.reshape(*lead, self.out_dim)


class Embedding(nn.Module):
    """nerfstudio field_components.embedding.Embedding (parameter name ``embedding.weight``)."""

    def __init__(self, in_dim: int, out_dim: int):
        super().__init__()
        self.in_dim, self.out_dim = in_dim, out_dim
        self.embedding = nn.Embedding(in_dim, out_dim)

    def mean(self, dim=0):
        return self.embedding.weight.mean(dim)

    def forward(self, in_tensor):
        return self.embedding(in_tensor)


class FieldHead(nn.Module):
    """Linear head (parameter name ``net``) as in nerfstudio field_heads."""

    def __init__(self, in_dim: int, out_dim: int):
        super().__init__()
        self.net = nn.Linear(in_dim, out_dim)


def _norm_spec(aabb: torch.Tensor, spatial_distortion) -> ops.NormSpec:
    if spatial_distortion is not None and not isinstance(spatial_distortion, SceneContraction):
        raise ValueError("nkbs fields support SceneContraction(order=inf) or no distortion")
    if spatial_distortion is not None and spatial_distortion.order != float("inf"):
        raise ValueError("nkbs fields support the L-inf SceneContraction the reference uses")
    return ops.NormSpec(spatial_distortion is not None, tuple(float(v) for v in aabb.reshape(-1).tolist()))


def _frustum_inputs(ray_samples: RaySamples):
    """(origins [R,3], directions [R,3], starts [R,S], ends [R,S]) if the samples came from a ray bundle,
    else None (then positions are materialised)."""
    fr = ray_samples.frustums
    if fr.ray_origins is not None and fr.starts.dim() == 3:
        return fr.ray_origins, fr.ray_directions, fr.starts[..., 0], fr.ends[..., 0]
    return None


class _Field(nn.Module):
    def density_fn(self, positions: torch.Tensor) -> torch.Tensor:
        """positions [..., 3] -> density [..., 1] (nerfstudio Field.density_fn)."""
        # nerfstudio wraps the points as zero-length frustums (starts = ends = 0, directions = 1) so that
        # get_positions() returns them unchanged; here they go straight to the encode kernel.
        rs = RaySamples(frustums=Frustums(origins=positions, directions=None, starts=None, ends=None))
        rs._positions = positions
        density, _ = self.get_density(rs)
        return density

    def _encode(self, ray_samples: RaySamples, encoding: HashEncoding, norm: ops.NormSpec):
        pos = getattr(ray_samples, "_positions", None)
        fi = _frustum_inputs(ray_samples) if pos is None else None
        if fi is not None:
            o, d, st, en = fi
            return ops.contract_encode(encoding.hash_table, encoding.grid, norm, origins=o, directions=d, starts=st,
                                       ends=en)
        if pos is None:
            pos = ray_samples.frustums.get_positions()
        return ops.contract_encode(encoding.hash_table, encoding.grid, norm, positions=pos)


class HashMLPDensityField(_Field):
    """Proposal density field (ref:nerfacto.py:299-311; defaults used at ref:semantic_nerfw.py:119-126)."""

    def __init__(self, aabb, num_layers: int = 2, hidden_dim: int = 64, spatial_distortion=None,
                 use_linear: bool = False, num_levels: int = 8, max_res: int = 1024, base_res: int = 16,
                 log2_hashmap_size: int = 18, features_per_level: int = 2, implementation: str = "nkbs"):
        super().__init__()
        self.register_buffer("aabb", torch.as_tensor(aabb, dtype=torch.float32))
        self.spatial_distortion = spatial_distortion
        self.use_linear = use_linear
        self.encoding = HashEncoding(num_levels=num_levels, min_res=base_res, max_res=max_res,
                                     log2_hashmap_size=log2_hashmap_size, features_per_level=features_per_level)
        if not use_linear:
            network = MLP(in_dim=self.encoding.get_out_dim(), num_layers=num_layers, layer_width=hidden_dim, out_dim=1,
                          activation=nn.ReLU(), out_activation=None)
            self.mlp_base = nn.Sequential(self.encoding, network)
        else:
            self.linear = nn.Linear(self.encoding.get_out_dim(), 1)
        self._norm = _norm_spec(self.aabb, spatial_distortion)

    def _fused_ok(self) -> bool:
        if self.use_linear or not ops.fused_density_enabled():
            return False
        net = self.mlp_base[1]
        shape = (self.encoding.grid.L, net.layers[0].out_features)
        if len(net.layers) != 2 or shape not in ops.FUSED_DENSITY_SHAPES:
            return False
        if torch.is_grad_enabled() and self.encoding.hash_table.requires_grad:
            # other shapes: fused forward + backward through the MLP / encode kernels (ops.HYBRID_DENSITY)
            return shape in ops.FUSED_DENSITY_TRAIN_SHAPES or ops.HYBRID_DENSITY
        return True

    def get_density(self, ray_samples: RaySamples) -> Tuple[torch.Tensor, None]:
        if self._fused_ok():
            # one kernel: ray point -> contraction -> hash grid -> MLP -> trunc_exp * selector (csrc/density.cu)
            (W1, W2), (b1, b2) = self.mlp_base[1].weights_biases()
            enc = self.encoding
            pos = getattr(ray_samples, "_positions", None)
            fi = _frustum_inputs(ray_samples) if pos is None else None
            if fi is not None:
                o, d, st, en = fi
                density = ops.fused_density(enc.hash_table, enc.grid, self._norm, W1, b1, W2, b2, origins=o,
                                            directions=d, starts=st, ends=en)
            else:
                if pos is None:
                    pos = ray_samples.frustums.get_positions()
                density = ops.fused_density(enc.hash_table, enc.grid, self._norm, W1, b1, W2, b2, positions=pos)
            return density.reshape(*ray_samples.frustums.shape, 1), None
        feats, sel = self._encode(ray_samples, self.encoding, self._norm)
        if not self.use_linear:
            W, b = self.mlp_base[1].weights_biases()
            relu = self.mlp_base[1].relu_mask
        else:
            W, b, relu = [self.linear.weight], [self.linear.bias], 0
        density = ops.mlp(feats, W, b, relu, ops.ACT_DENSITY, selector=sel)
        return density.reshape(*ray_samples.frustums.shape, 1), None

    def get_outputs(self, ray_samples, density_embedding=None) -> dict:
        return {}


class NerfactoField(_Field):
    """Main field (ref:nerfacto.py:275-290; semantic_nerfw.py:102-114): hash grid -> base MLP -> density +
    15-d geometry feature; SH(dir) + geo + appearance -> colour MLP; semantic and transient heads."""

    def __init__(self, aabb, num_images: int, num_layers: int = 2, hidden_dim: int = 64, geo_feat_dim: int = 15,
                 num_levels: int = 16, base_res: int = 16, max_res: int = 2048, log2_hashmap_size: int = 19,
                 num_layers_color: int = 3, num_layers_transient: int = 2, features_per_level: int = 2,
                 hidden_dim_color: int = 64, hidden_dim_transient: int = 64, appearance_embedding_dim: int = 32,
                 transient_embedding_dim: int = 16, use_transient_embedding: bool = False,
                 use_semantics: bool = False, num_semantic_classes: int = 100,
                 pass_semantic_gradients: bool = False, use_pred_normals: bool = False,
                 use_average_appearance_embedding: bool = False, spatial_distortion=None,
                 implementation: str = "nkbs"):
        super().__init__()
        if use_pred_normals:
            raise NotImplementedError("predicted normals are outside the nkbs hot path (SURVEY a19)")
        self.register_buffer("aabb", torch.as_tensor(aabb, dtype=torch.float32))
        self.geo_feat_dim = geo_feat_dim
        self.spatial_distortion = spatial_distortion
        self.num_images = num_images
        self.appearance_embedding_dim = appearance_embedding_dim
        self.embedding_appearance = Embedding(num_images, appearance_embedding_dim)
        self.use_average_appearance_embedding = use_average_appearance_embedding
        self.use_transient_embedding = use_transient_embedding
        self.use_semantics = use_semantics
        self.use_pred_normals = use_pred_normals
        self.pass_semantic_gradients = pass_semantic_gradients
        self.mlp_base_grid = HashEncoding(num_levels=num_levels, min_res=base_res, max_res=max_res,
                                          log2_hashmap_size=log2_hashmap_size, features_per_level=features_per_level)
        self.mlp_base_mlp = MLP(in_dim=self.mlp_base_grid.get_out_dim(), num_layers=num_layers, layer_width=hidden_dim,
                                out_dim=1 + geo_feat_dim, activation=nn.ReLU(), out_activation=None)
        self.mlp_base = nn.Sequential(self.mlp_base_grid, self.mlp_base_mlp)
        if use_transient_embedding:
            self.transient_embedding_dim = transient_embedding_dim
            self.embedding_transient = Embedding(num_images, transient_embedding_dim)
            self.mlp_transient = MLP(in_dim=geo_feat_dim + transient_embedding_dim, num_layers=num_layers_transient,
                                     layer_width=hidden_dim_transient, out_dim=hidden_dim_transient,
                                     activation=nn.ReLU(), out_activation=None)
            self.field_head_transient_uncertainty = FieldHead(hidden_dim_transient, 1)
            self.field_head_transient_rgb = FieldHead(hidden_dim_transient, 3)
            self.field_head_transient_density = FieldHead(hidden_dim_transient, 1)
        if use_semantics:
            self.mlp_semantics = MLP(in_dim=geo_feat_dim, num_layers=2, layer_width=64, out_dim=hidden_dim_transient,
                                     activation=nn.ReLU(), out_activation=None)
            self.field_head_semantics = FieldHead(hidden_dim_transient, num_semantic_classes)
        self.mlp_head = MLP(in_dim=16 + geo_feat_dim + appearance_embedding_dim, num_layers=num_layers_color,
                            layer_width=hidden_dim_color, out_dim=3, activation=nn.ReLU(), out_activation=nn.Sigmoid())
        self._norm = _norm_spec(self.aabb, spatial_distortion)

    # -- density -------------------------------------------------------------------------------
    def get_density(self, ray_samples: RaySamples):
        feats, sel = self._encode(ray_samples, self.mlp_base_grid, self._norm)
        W, b = self.mlp_base_mlp.weights_biases()
        # split_bwd: x is the hash grid's output — its backward overlaps the scatter with the MLP (ops._BWD_SPLIT)
        h, d0 = ops.mlp(feats, W, b, self.mlp_base_mlp.relu_mask, ops.ACT_DENSITY, selector=sel, col0=True, split_bwd=True)
        h3 = h.reshape(*ray_samples.frustums.shape, -1)
        # torch.split(h3, [1, geo]) of the reference, without a split node in the backward pass (ops._Mlp, col0)
        density = d0.reshape(*ray_samples.frustums.shape, 1)
        base_mlp_out = h3[..., 1:]
        # the head kernels read the geometry columns of the base MLP's output in place (no slice copy)
        base_mlp_out._nkbs_src = (h, 1)
        return density, base_mlp_out

    # -- heads ---------------------------------------------------------------------------------
    def _per_ray(self, ray_samples: RaySamples):
        """(S, directions, dir_per_ray, cam, cam_per_ray) in the cheapest form the samples allow."""
        fr = ray_samples.frustums
        shape = fr.directions.shape[:-1]
        S = shape[-1] if len(shape) >= 2 else 1
        if fr.ray_directions is not None:
            dirs, dpr = fr.ray_directions, True
        else:
            dirs, dpr = fr.directions.reshape(-1, 3), False
        cam, cpr = None, False
        if ray_samples.ray_camera_indices is not None:
            cam, cpr = ray_samples.ray_camera_indices, True
        elif ray_samples.camera_indices is not None:
            cam, cpr = ray_samples.camera_indices.reshape(-1), False
        return S, dirs, dpr, cam, cpr

    def get_outputs(self, ray_samples: RaySamples, density_embedding: Optional[torch.Tensor] = None
                    ) -> Dict[FieldHeadNames, torch.Tensor]:
        assert density_embedding is not None
        outputs = {}
        if ray_samples.camera_indices is None and ray_samples.ray_camera_indices is None:
            raise AttributeError("Camera indices are not provided.")
        S, dirs, dpr, cam, cpr = self._per_ray(ray_samples)
        outputs_shape = ray_samples.frustums.directions.shape[:-1]
        geo = density_embedding.reshape(-1, self.geo_feat_dim)

        # appearance row: gathered per camera in training; mean (or zeros) in eval (ref:nerfacto.py:202,286)
        emb_const = None
        if not self.training:
            if self.use_average_appearance_embedding:
                emb_const = self.embedding_appearance.mean(dim=0).detach().contiguous()
            else:
                emb_const = torch.zeros(self.appearance_embedding_dim, device=geo.device)

        # tensor-core path: the head kernels assemble their input rows themselves (csrc/mlp_tc.cu, IN = 1 / 2)
        src, col = getattr(density_embedding, "_nkbs_src", (None, 0))
        fused_in = (ops.get_mlp_precision() == "tc" and ops.fused_head_enabled() and self.geo_feat_dim == 15
                    and cam is not None)
        if fused_in and src is None:
            src, col = geo.contiguous(), 0

        if self.use_transient_embedding and self.training:
            W, b = self.mlp_transient.weights_biases()
            heads = (self.field_head_transient_uncertainty, self.field_head_transient_rgb,
                     self.field_head_transient_density)
            W = W + [torch.cat([h.net.weight for h in heads], dim=0)]
            b = b + [torch.cat([h.net.bias for h in heads], dim=0)]
            relu = self.mlp_transient.relu_mask  # no activation between mlp_transient's output and the heads
            if fused_in and self.transient_embedding_dim == 16 and len(W) == 3:
                x = ops.head_mlp(src, col, self.embedding_transient.embedding.weight, None, None, cam, S, 0, False, cpr,
                                 W, b, relu, ops.ACT_TRANSIENT)
            else:
                t_in = ops.assemble(geo, self.embedding_transient.embedding.weight, None, None, cam, S, 0, False, cpr)
                x = ops.mlp(t_in, W, b, relu, ops.ACT_TRANSIENT)
            x = x.reshape(*outputs_shape, 5)
            outputs[FieldHeadNames.UNCERTAINTY] = x[..., 0:1]
            outputs[FieldHeadNames.TRANSIENT_RGB] = x[..., 1:4]
            outputs[FieldHeadNames.TRANSIENT_DENSITY] = x[..., 4:5]

        if self.use_semantics:
            sem_in = geo if self.pass_semantic_gradients else geo.detach()
            W, b = self.mlp_semantics.weights_biases()
            W = W + [self.field_head_semantics.net.weight]
            b = b + [self.field_head_semantics.net.bias]
            if fused_in and len(W) == 3:
                # geometry columns read in place from the base MLP's output (no slice copy), rows stored coalesced
                sem_src = src if self.pass_semantic_gradients else src.detach()
                x = ops.head_mlp(sem_src, col, None, None, None, None, S, 0, False, False, W, b,
                                 self.mlp_semantics.relu_mask, ops.ACT_NONE)
            else:
                x = ops.mlp(sem_in, W, b, self.mlp_semantics.relu_mask, ops.ACT_NONE)
            outputs[FieldHeadNames.SEMANTICS] = x.reshape(*outputs_shape, -1)

        W, b = self.mlp_head.weights_biases()
        if fused_in and self.appearance_embedding_dim == 32 and len(W) == 3:
            rgb = ops.head_mlp(src, col, self.embedding_appearance.embedding.weight, emb_const, dirs, cam, S, 16, dpr,
                               cpr, W, b, self.mlp_head.relu_mask, ops.ACT_SIGMOID)
        else:
            h = ops.assemble(geo, self.embedding_appearance.embedding.weight, emb_const, dirs, cam, S, 16, dpr, cpr)
            rgb = ops.mlp(h, W, b, self.mlp_head.relu_mask, ops.ACT_SIGMOID)
        outputs[FieldHeadNames.RGB] = rgb.reshape(*outputs_shape, -1)
        return outputs

    def forward(self, ray_samples: RaySamples, compute_normals: bool = False) -> Dict[FieldHeadNames, torch.Tensor]:
        if compute_normals:
            raise NotImplementedError("analytic normals are outside the nkbs hot path (SURVEY a19)")
        density, density_embedding = self.get_density(ray_samples)
        field_outputs = self.get_outputs(ray_samples, density_embedding=density_embedding)
        field_outputs[FieldHeadNames.DENSITY] = density
        return field_outputs
