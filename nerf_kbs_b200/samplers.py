"""Drop-in samplers (nerfstudio model_components/ray_samplers.py) built at ref:nerfacto.py:325-332 and
ref:semantic_nerfw.py:132-137, called at ref:nerfacto.py:397 / ref:semantic_nerfw.py:187."""
from __future__ import annotations

import contextlib
import os

from typing import Callable, List, Optional, Tuple

import torch
from torch import nn

from . import ops
from .rays import RayBundle, RaySamples


def _spacing_fn(x):
    return torch.where(x < 1, x / 2, 1 - 1 / (2 * x))


def _spacing_fn_inv(x):
    return torch.where(x < 0.5, 2 * x, 1 / (2 - 2 * x))


def _make_s2e(nears, fars):
    """spacing -> euclidean map handed to RaySamples (only evaluated by foreign callers; the kernels
    compute it in-line)."""
    def fn(x):
        s_near, s_far = _spacing_fn(nears), _spacing_fn(fars)
        return _spacing_fn_inv(x * s_far + (1 - x) * s_near)
    return fn


class UniformLinDispPiecewiseSampler(nn.Module):
    """Initial sampler: uniform in distance up to 1, uniform in disparity beyond."""

    def __init__(self, num_samples: Optional[int] = None, train_stratified: bool = True, single_jitter: bool = False):
        super().__init__()
        self.num_samples = num_samples
        self.train_stratified = train_stratified
        self.single_jitter = single_jitter

    def generate_ray_samples(self, ray_bundle: RayBundle, num_samples: Optional[int] = None, t_rand=None) -> RaySamples:
        assert ray_bundle.nears is not None and ray_bundle.fars is not None
        num_samples = num_samples or self.num_samples
        R = ray_bundle.origins.shape[0]
        if self.train_stratified and self.training:
            if t_rand is None:
                n = 1 if self.single_jitter else num_samples + 1
                t_rand = torch.rand((R, n), dtype=torch.float32, device=ray_bundle.origins.device)
        else:
            t_rand = None
        sb, eb = ops.initial_sampler(ray_bundle.nears, ray_bundle.fars, num_samples, t_rand)
        return ray_bundle.samples_from_bins(sb, eb, _make_s2e(ray_bundle.nears, ray_bundle.fars))

    def forward(self, *args, **kwargs):
        return self.generate_ray_samples(*args, **kwargs)


class PDFSampler(nn.Module):
    """Inverse-CDF resampling along a weight histogram."""

    def __init__(self, num_samples: Optional[int] = None, train_stratified: bool = True, single_jitter: bool = False,
                 include_original: bool = True, histogram_padding: float = 0.01):
        super().__init__()
        if include_original:
            raise NotImplementedError("include_original=True is not on the reference path (ProposalNetworkSampler "
                                      "builds PDFSampler(include_original=False))")
        self.num_samples = num_samples
        self.train_stratified = train_stratified
        self.single_jitter = single_jitter
        self.histogram_padding = histogram_padding
        self.anneal_dev = None  # optional device scalar (CUDA-graph replay)

    def generate_ray_samples(self, ray_bundle: RayBundle, ray_samples: RaySamples, weights: torch.Tensor,
                             num_samples: Optional[int] = None, eps: float = 1e-5, rand=None,
                             anneal: float = 1.0) -> RaySamples:
        """``weights`` are the pre-anneal weights; ``anneal`` is applied inside the kernel
        (pow(weights, anneal), ref ProposalNetworkSampler)."""
        num_samples = num_samples or self.num_samples
        S = weights.shape[-2]
        w = weights.reshape(-1, S)
        R = w.shape[0]
        if ray_samples.spacing_bins is not None:
            sb = ray_samples.spacing_bins
        else:
            sb = torch.cat([ray_samples.spacing_starts[..., 0], ray_samples.spacing_ends[..., -1:, 0]], dim=-1)
        if self.train_stratified and self.training:
            if rand is None:
                n = 1 if self.single_jitter else num_samples + 1
                rand = torch.rand((R, n), device=w.device)
        else:
            rand = None
        nsb, neb = ops.pdf_sample(w, sb, ray_bundle.nears, ray_bundle.fars, num_samples, anneal=anneal, rand=rand,
                                  histogram_padding=self.histogram_padding, eps=eps, anneal_dev=self.anneal_dev)
        return ray_bundle.samples_from_bins(nsb, neb, ray_samples.spacing_to_euclidean_fn)

    def forward(self, *args, **kwargs):
        return self.generate_ray_samples(*args, **kwargs)


class ProposalNetworkSampler(nn.Module):
    """Hierarchical proposal sampling with the reference's update gating and weight annealing
    (ref:nerfacto.py:314-332,365-394)."""

    def __init__(self, num_nerf_samples_per_ray: int = 32, num_proposal_samples_per_ray: Tuple[int, ...] = (64,),
                 num_proposal_network_iterations: int = 2, single_jitter: bool = False,
                 update_sched: Callable = lambda x: 1, initial_sampler=None):
        super().__init__()
        self.num_nerf_samples_per_ray = num_nerf_samples_per_ray
        self.num_proposal_samples_per_ray = num_proposal_samples_per_ray
        self.num_proposal_network_iterations = num_proposal_network_iterations
        self.update_sched = update_sched
        if self.num_proposal_network_iterations < 1:
            raise ValueError("num_proposal_network_iterations must be >= 1")
        if initial_sampler is None:
            self.initial_sampler = UniformLinDispPiecewiseSampler(single_jitter=single_jitter)
        else:
            self.initial_sampler = initial_sampler
        self.pdf_sampler = PDFSampler(include_original=False, single_jitter=single_jitter)
        self._anneal = 1.0
        self._steps_since_update = 0
        self._step = 0
        self.force_updated = None  # Trainer sets this while it captures the two CUDA-graph variants
        # training proposal levels run on their own CUDA streams (see generate_ray_samples); NKBS_SIDE_STREAMS=0 disables
        self.side_streams = os.environ.get("NKBS_SIDE_STREAMS", "1") != "0"
        self._streams = {}

    def _side_stream(self, level: int, device) -> "torch.cuda.Stream":
        key = (level, device.index)
        if key not in self._streams:
            # Same (high) priority as the step driver's capture stream.  The backward of a proposal level and the main
            # field's backward become ready together and each of them fills the register file of the SMs it gets; with
            # unequal priorities the lower one is starved at every kernel boundary and the persistent tensor-core MLP
            # kernels start on fewer SMs (measured on the B200, same box, 4096-ray captured step: 0.814 ms with the
            # proposal streams one level below the step stream, 0.787 ms with equal priorities, 0.801 ms on one stream).
            self._streams[key] = torch.cuda.Stream(device, priority=int(os.environ.get("NKBS_SIDE_PRIO", "-1")))
        return self._streams[key]

    def set_anneal(self, anneal: float) -> None:
        self._anneal = anneal

    def step_cb(self, step):
        self._step = step
        self._steps_since_update += 1

    def will_update(self) -> bool:
        if self.force_updated is not None:
            return bool(self.force_updated)
        return self._steps_since_update > self.update_sched(self._step) or self._step < 10

    def _level_fusable(self, owner, i_level: int, jitter, stratified: bool) -> bool:
        """A proposal level runs as one launch (ops.proposal_level) when its field is a fused-shape
        HashMLPDensityField and the jitter is per ray (the reference's use_single_jitter) or absent (evaluation)."""
        if not ops.fused_level_enabled() or owner is None or not hasattr(owner, "_fused_ok") or not owner._fused_ok():
            return False
        if getattr(owner, "_norm", None) is None:
            return False
        if not isinstance(self.initial_sampler, UniformLinDispPiecewiseSampler) and i_level == 0:
            return False
        if stratified:
            if jitter is None or not (getattr(self.initial_sampler, "single_jitter", False) and self.pdf_sampler.single_jitter):
                return False
            if any(t is not None and t.shape[-1] != 1 for t in (jitter[i_level], jitter[i_level + 1])):
                return False
        return True

    def _resample_fusable(self, ray_samples, density, jitter, i_level: int, stratified: bool) -> bool:
        """get_weights of a level and the PDF resampling of the next run as one launch (ops.weights_resample) when the
        samples carry their [R,S+1] bin arrays and the jitter is per ray (use_single_jitter) or absent (evaluation)."""
        if not ops.fused_resample_enabled() or ray_samples.spacing_bins is None or ray_samples.euclid_bins is None:
            return False
        if density.dim() != 3 or density.shape[-1] != 1 or density.shape[1] > 512:
            return False
        if stratified:
            if jitter is None or not self.pdf_sampler.single_jitter:
                return False
            t = jitter[i_level + 1]
            if t is None or t.shape[-1] != 1:
                return False
        elif self.training and self.pdf_sampler.train_stratified:
            return False  # stratified training without injected / pre-drawn jitter: the sampler draws its own
        return True

    def generate_ray_samples(self, ray_bundle: Optional[RayBundle] = None, density_fns: Optional[List[Callable]] = None,
                             jitter=None) -> Tuple[RaySamples, List, List]:
        """``jitter`` (optional [t_rand0, rand1, rand2]) injects the random draws, for parity tests."""
        assert ray_bundle is not None and density_fns is not None
        weights_list, ray_samples_list = [], []
        n = self.num_proposal_network_iterations
        weights, ray_samples = None, None
        updated = self.will_update()
        if jitter is None and self.training and getattr(self.initial_sampler, "single_jitter", False) \
                and self.pdf_sampler.single_jitter and getattr(self.initial_sampler, "train_stratified", False) \
                and self.pdf_sampler.train_stratified:
            # one launch draws the per-ray jitter of every level (the samplers would draw [R,1] each)
            dev = ray_bundle.origins.device
            jitter = torch.rand((n + 1, ray_bundle.origins.shape[0], 1), dtype=torch.float32, device=dev).unbind(0)
        stratified = self.training and getattr(self.initial_sampler, "train_stratified", False) \
            and self.pdf_sampler.train_stratified
        next_bins = None  # (spacing, euclidean) bins of the coming level, produced by the previous level's fused pass
        for i_level in range(n + 1):
            is_prop = i_level < n
            num_samples = self.num_proposal_samples_per_ray[i_level] if is_prop else self.num_nerf_samples_per_ray
            j = None if jitter is None else jitter[i_level]
            fn = density_fns[i_level] if is_prop else None
            owner = getattr(fn, "__self__", None) if is_prop else None
            train_level = is_prop and updated and torch.is_grad_enabled()
            # A level that trains runs on its own CUDA stream.  The forward gains nothing (the next level waits
            # for these weights) but autograd replays each node on the stream of its forward, so the backward of
            # the proposal fields (atomics-bound table scatters) overlaps the main field's backward (tensor-core
            # MLPs + its own scatter) instead of queueing behind it.  Works the same under CUDA-graph capture.
            side = self._side_stream(i_level, ray_bundle.origins.device) if (train_level and self.side_streams) else None
            cur = torch.cuda.current_stream() if side is not None else None
            if is_prop and self._level_fusable(owner, i_level, jitter, stratified):
                # ONE launch: [initial sampler] -> field -> weights + median depth -> the next level's bins
                num_next = (self.num_proposal_samples_per_ray[i_level + 1] if i_level + 1 < n
                            else self.num_nerf_samples_per_ray)
                net = owner.mlp_base[1]
                (W1, W2), (b1, b2) = net.weights_biases()
                enc = owner.encoding
                if side is not None:
                    side.wait_stream(cur)
                with torch.set_grad_enabled(train_level), (torch.cuda.stream(side) if side is not None else contextlib.nullcontext()):
                    sb_in, eb_in = next_bins if next_bins is not None else (None, None)
                    w, dmed, sb, eb, nsb, neb = ops.proposal_level(
                        enc.hash_table, enc.grid, owner._norm, W1, b1, W2, b2, ray_bundle.origins, ray_bundle.directions,
                        ray_bundle.nears, ray_bundle.fars, num_samples, num_next, sbins=sb_in, ebins=eb_in,
                        t_rand=j if (i_level == 0 and stratified) else None,
                        rand=(jitter[i_level + 1] if stratified else None), anneal=float(self._anneal),
                        anneal_dev=self.pdf_sampler.anneal_dev, histogram_padding=self.pdf_sampler.histogram_padding)
                    weights = w[..., None]
                    weights._nkbs_depth_med = dmed[..., None]
                if side is not None:
                    cur.wait_stream(side)
                    for t in (w, dmed, sb, eb, nsb, neb):
                        t.record_stream(cur)
                ray_samples = ray_bundle.samples_from_bins(
                    sb, eb, _make_s2e(ray_bundle.nears, ray_bundle.fars) if i_level == 0 else ray_samples.spacing_to_euclidean_fn)
                next_bins = (nsb, neb)
                weights_list.append(weights)
                ray_samples_list.append(ray_samples)
                continue
            if next_bins is not None:
                ray_samples = ray_bundle.samples_from_bins(next_bins[0], next_bins[1], ray_samples.spacing_to_euclidean_fn)
                next_bins = None
            elif i_level == 0:
                ray_samples = self.initial_sampler(ray_bundle, num_samples=num_samples, t_rand=j)
            else:
                # pow(weights, anneal) happens inside the PDF kernel
                ray_samples = self.pdf_sampler(ray_bundle, ray_samples, weights, num_samples=num_samples, rand=j,
                                               anneal=float(self._anneal))
            if is_prop:
                if side is not None:
                    side.wait_stream(cur)
                with torch.set_grad_enabled(train_level), (torch.cuda.stream(side) if side is not None else contextlib.nullcontext()):
                    if owner is not None and hasattr(owner, "get_density") and getattr(owner, "_norm", None) is not None:
                        density, _ = owner.get_density(ray_samples)  # per-ray frustum path, no [R,S,3] tensor
                    else:
                        density = fn(ray_samples.frustums.get_positions())
                    if self._resample_fusable(ray_samples, density, jitter, i_level, stratified):
                        # weights + median depth of this level and the bins of the next one in ONE ray-major launch
                        num_next = (self.num_proposal_samples_per_ray[i_level + 1] if i_level + 1 < n
                                    else self.num_nerf_samples_per_ray)
                        w, dmed, nsb, neb = ops.weights_resample(
                            density, ray_samples.euclid_bins, ray_samples.spacing_bins, ray_bundle.nears, ray_bundle.fars,
                            num_next, rand=(jitter[i_level + 1] if stratified else None), anneal=float(self._anneal),
                            anneal_dev=self.pdf_sampler.anneal_dev, histogram_padding=self.pdf_sampler.histogram_padding)
                        weights = w[..., None]
                        weights._nkbs_depth_med = dmed[..., None]
                        next_bins = (nsb, neb)
                    else:
                        weights = ray_samples.get_weights(density)
                if side is not None:
                    cur.wait_stream(side)
                    for t in (weights, getattr(weights, "_nkbs_depth_med", None)) + (tuple(next_bins) if next_bins is not None else ()):
                        if t is not None:
                            t.record_stream(cur)
                weights_list.append(weights)
                ray_samples_list.append(ray_samples)
        if updated:
            self._steps_since_update = 0
        assert ray_samples is not None
        return ray_samples, weights_list, ray_samples_list

    def forward(self, *args, **kwargs):
        return self.generate_ray_samples(*args, **kwargs)
