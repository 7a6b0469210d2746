"""torch.autograd bindings of the nkbs C-ABI kernels (include/nkbs.h).

Every function here launches hand-written sm_100a kernels on the current CUDA stream.  PyTorch only
owns the buffers.  There is no CPU path: CPU tensors raise.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import call, host_floats, host_i32, host_ptrs, ptr, require_cuda, rows, stream

ACT_NONE, ACT_SIGMOID, ACT_DENSITY, ACT_TRANSIENT = 0, 1, 2, 3
BG_MODES = {"last_sample": 0, "black": 1, "white": 2}


def _zeros_like_many(tensors):
    """One zero-filled flat buffer carved into views shaped like ``tensors`` (one fill launch instead of one per
    tensor; the kernels accumulate into these gradient buffers)."""
    sizes = [(t.numel() + 3) // 4 * 4 for t in tensors]  # 16-byte aligned pieces
    flat = torch.zeros(sum(sizes), device=tensors[0].device, dtype=torch.float32)
    out, o = [], 0
    for t, n in zip(tensors, sizes):
        out.append(flat[o:o + t.numel()].view(t.shape))
        o += n
    return out, flat


def _sinks(params):
    """Gradient sinks of ``params`` (``p._nkbs_grad_sink``: the parameter's view of a step driver's flat gradient
    buffer, optim.ParamArena) if EVERY one has one, else None.  Collected outside ``Function.apply`` from the caller's
    own Parameter objects."""
    sinks = [getattr(p, "_nkbs_grad_sink", None) for p in params]
    return sinks if sinks and all(s is not None for s in sinks) else None


# Deterministic mode (optim.ParamArena(deterministic=True), include/nkbs.h: nkbs_fixed_point_register): gradient buffers
# that are NOT views of the arena — parameters the model feeds through torch ops first, like the concatenated transient
# heads — get a fixed-point shadow of their own for the duration of the backward call: registered here, folded into the
# float buffer and unregistered by ``_finish_temps`` right after the kernel launches.
_DET = {"count": 0, "log2_scale": 56}
_TEMPS: List = []


def deterministic_enabled() -> bool:
    return _DET["count"] > 0


def _temp_shadow(flat: torch.Tensor) -> None:
    if _DET["count"] > 0 and flat is not None and flat.numel() > 0:
        shadow = torch.zeros(flat.numel(), device=flat.device, dtype=torch.int64)
        call("nkbs_fixed_point_register", flat.data_ptr(), flat.numel(), shadow.data_ptr(), _DET["log2_scale"])
        _TEMPS.append((flat, shadow))


def _finish_temps() -> None:
    while _TEMPS:
        flat, _shadow = _TEMPS.pop()
        call("nkbs_fixed_point_flush", flat.data_ptr(), flat.numel(), stream())
        call("nkbs_fixed_point_register", flat.data_ptr(), flat.numel(), None, 0)


def _grad_buffers(sinks, like):
    """(buffers the backward kernel accumulates the parameter gradients into, what ``backward`` returns for them).
    With sinks the kernel adds into the step driver's buffer and autograd sees no gradient for the parameter (no
    fill launch, no AccumulateGrad); without, fresh zero-filled tensors are handed to autograd."""
    if sinks is not None:
        return list(sinks), [None] * len(sinks)
    zs, flat = _zeros_like_many(like)
    _temp_shadow(flat)
    return zs, zs


def _c(t: Optional[torch.Tensor], dtype=torch.float32) -> Optional[torch.Tensor]:
    if t is None:
        return None
    if t.dtype != dtype:
        t = t.to(dtype)
    return t.contiguous()


# ----------------------------------------------------------------------------------------------
# K1/K2 hash-grid encoding
# ----------------------------------------------------------------------------------------------
@dataclass(frozen=True)
class GridSpec:
    scalings: Tuple[float, ...]
    log2_T: int
    features_per_level: int = 2

    @property
    def L(self) -> int:
        return len(self.scalings)

    @property
    def out_dim(self) -> int:
        return self.L * self.features_per_level


@dataclass(frozen=True)
class NormSpec:
    contraction: bool
    aabb: Tuple[float, ...]  # 6 floats


def encode_raw(mode: int, grid: GridSpec, norm: Optional[NormSpec], table, x=None, origins=None, directions=None,
               starts=None, ends=None, want_indices=False):
    """Launch K1.  Returns (features [N,2L], selector [N], xn [N,3], indices or None)."""
    require_cuda(table, x, origins, starts)
    if mode == 2:
        R, S = starts.shape[0], starts.shape[1]
        dev = starts.device
    else:
        R, S = x.shape[0], 1
        dev = x.device
    N = R * S
    feats = torch.empty((N, grid.out_dim), device=dev, dtype=torch.float32)
    sel = torch.empty((N,), device=dev, dtype=torch.float32)
    xn = torch.empty((N, 3), device=dev, dtype=torch.float32)
    idx = torch.empty((N, grid.L, 8), device=dev, dtype=torch.int64) if want_indices else None
    aabb = host_floats(norm.aabb) if norm is not None else None
    p_st, p_en, bstride, _keep = rows(starts, ends)
    call("nkbs_encode_fwd", mode, ptr(x), ptr(origins), ptr(directions), p_st, p_en, bstride, R, S,
         int(norm.contraction) if norm is not None else 0, aabb, ptr(table), host_floats(grid.scalings), grid.L,
         grid.log2_T, grid.features_per_level, ptr(xn), ptr(sel), ptr(feats), ptr(idx), stream())
    return feats, sel, xn, idx


# Split backward of the main field's base MLP + hash grid (opt-in: NKBS_BWD_SPLIT = a fraction in (0, 1)): the MLP's
# backward runs as two launches over [0, k) and [k, N) samples, and the gradient scatter of the first part starts on a
# side stream as soon as ITS feature gradients exist, beside the MLP backward of the second part — the tensor-core
# kernel leaves most issue slots and all of L2's reduction bandwidth idle, the scatter is bound by exactly that
# bandwidth.  _Mlp.backward leaves (event after part one, k) here under the feature gradient's address and
# _HashEncode.backward picks it up; any other consumer of that gradient just sees a complete tensor on the stream.
# MEASURED on the B200 (scripts/gpu_ab_split.sh, same build back to back): 4096-ray step 0.753 ms without, 0.771-0.773 ms
# with (fractions 0.375 / 0.5 / 0.625); 65 536 rays 7.65 / 7.66 ms; 8192-ray semantic step 2.754 / 2.759 ms.  The
# proposal fields' backward on the sampler's side streams already fills what the tensor-core kernels leave idle, and
# the two extra launches and half-size grids cost more than the overlap returns.  Hence OFF by default.
_BWD_SPLIT = {"fraction": float(__import__("os").environ.get("NKBS_BWD_SPLIT", "0")), "min_rows": 1 << 16,
              "pending": {}, "streams": {}}


def set_bwd_split(fraction: float) -> None:
    _BWD_SPLIT["fraction"] = float(fraction)


def _split_stream(device) -> "torch.cuda.Stream":
    st = _BWD_SPLIT["streams"].get(device)
    if st is None:
        st = _BWD_SPLIT["streams"][device] = torch.cuda.Stream(device, priority=-1)
    return st


class _HashEncode(torch.autograd.Function):
    """features = HashEncoding(normalise(positions)); gradient w.r.t. the table only (sample positions
    are detached in the reference path: PDF bins are detached)."""

    @staticmethod
    def forward(ctx, table, mode, grid, norm, dedup_levels, x, origins, directions, starts, ends, sink=None):
        feats, sel, xn, _ = encode_raw(mode, grid, norm, table, x, origins, directions, starts, ends)
        # the non-differentiable outputs would otherwise receive materialised zero gradients ([N] and [N,3] fills)
        ctx.set_materialize_grads(False)
        ctx.sink = sink  # optional gradient sink (optim.ParamArena): the scatter accumulates into it
        ctx.grid = grid
        # run merging (segmented shuffle reduction, csrc/encode.cuh) costs a ballot on levels without duplicates and
        # never loses on B200 (scripts/encode_bench.py sweep): -1 = every level
        ctx.dedup = dedup_levels if dedup_levels >= 0 else grid.L
        ctx.table_shape = table.shape
        # Optional (set_ray_scatter(encode=True)): proposal-sized ray samples (64..256 per ray) take the ray-major
        # scatter in the backward pass.  Off by default: on the 8-level grids that reach this unfused route (the
        # semantic-NeRF-W proposals) it measured slower than the run-merged sample-major scatter (8192-ray step 2.676 ->
        # 2.714 ms), unlike on the 5-level grids of the fused field (see _density_backward).
        ctx.rays = None
        if mode == 2 and _RAY_SCATTER["encode"] and norm is not None and 64 <= starts.shape[1] <= 256:
            ctx.rays = (norm, starts.shape[0], starts.shape[1])
            ctx.save_for_backward(xn, origins, directions, starts, ends)
        else:
            ctx.save_for_backward(xn)
        ctx.mark_non_differentiable(sel, xn)
        return feats, sel, xn

    @staticmethod
    def backward(ctx, dfeat, _dsel, _dxn):
        xn = ctx.saved_tensors[0]
        g = ctx.grid
        if dfeat is None:
            return (None,) * 11
        (dtable,), (ret,) = _grad_buffers([ctx.sink] if ctx.sink is not None else None, [xn.new_empty(ctx.table_shape)])
        dfeat = _c(dfeat)
        if ctx.rays is not None:
            norm, R, S = ctx.rays
            _, origins, directions, starts, ends = ctx.saved_tensors
            p_st, p_en, bstride, _keep = rows(starts, ends)
            call("nkbs_encode_bwd_rays", R, S, ptr(origins), ptr(directions), p_st, p_en, bstride, int(norm.contraction),
                 host_floats(norm.aabb), ptr(dfeat), host_floats(g.scalings), g.L, g.log2_T, ptr(dtable), stream())
            _finish_temps()
            return (ret,) + (None,) * 10
        split = _BWD_SPLIT["pending"].pop(dfeat.data_ptr(), None)
        N = xn.shape[0]
        if split is not None and split[1] < N and not _TEMPS:
            ev, k = split
            cur = torch.cuda.current_stream(xn.device)
            side = _split_stream(xn.device)
            side.wait_event(ev)  # part one of the MLP backward: dfeat[:k] is complete
            W = dfeat.shape[1]
            with torch.cuda.stream(side):
                call("nkbs_encode_bwd", ptr(xn), ptr(dfeat), k, host_floats(g.scalings), g.L, g.log2_T,
                     g.features_per_level, ctx.dedup, ptr(dtable), stream())
            call("nkbs_encode_bwd", xn.data_ptr() + 12 * k, dfeat.data_ptr() + 4 * W * k, N - k, host_floats(g.scalings),
                 g.L, g.log2_T, g.features_per_level, ctx.dedup, ptr(dtable), stream())
            cur.wait_stream(side)
            return (ret,) + (None,) * 10
        call("nkbs_encode_bwd", ptr(xn), ptr(dfeat), N, host_floats(g.scalings), g.L, g.log2_T,
             g.features_per_level, ctx.dedup, ptr(dtable), stream())
        _finish_temps()
        return (ret,) + (None,) * 10


def hash_encode(table, grid: GridSpec, x_normalised):
    """HashEncoding.forward on positions already in [0,1] (mode 0)."""
    x = _c(x_normalised).reshape(-1, 3)
    feats, _, _ = _HashEncode.apply(table, 0, grid, None, -1, x, None, None, None, None)
    return feats


def contract_encode(table, grid: GridSpec, norm: NormSpec, positions=None, origins=None, directions=None,
                    starts=None, ends=None, dedup_levels=-1):
    """Front end + encoding.  Either raw ``positions`` [N,3] or per-ray ``origins/directions`` [R,3] with
    per-sample ``starts/ends`` [R,S].  Returns (features [N,2L], selector [N])."""
    sink = getattr(table, "_nkbs_grad_sink", None)
    if positions is not None:
        feats, sel, _ = _HashEncode.apply(table, 1, grid, norm, dedup_levels, _c(positions).reshape(-1, 3), None,
                                          None, None, None, sink)
    else:
        feats, sel, _ = _HashEncode.apply(table, 2, grid, norm, dedup_levels, None, _c(origins), _c(directions),
                                          starts.float(), ends.float(), sink)
    return feats, sel


# ----------------------------------------------------------------------------------------------
# K8 fused proposal density field (encode + two-layer MLP + trunc_exp * selector)
# ----------------------------------------------------------------------------------------------
FUSED_DENSITY_SHAPES = ((5, 16), (8, 64))  # (levels, hidden width) instantiated in csrc/density.cu
# Shapes whose TRAINING passes take the fused kernels too.  The hidden-64 field (semantic-NeRF-W's default proposals)
# does not: its fused backward holds 64 + 16 accumulator registers per lane next to the recomputed forward (255
# registers, one 8-warp CTA per SM) and is latency-bound — measured on the B200 at 8192 x 256 samples: fused forward +
# backward 0.37 + 1.47 ms, encode + tensor-core MLP route 0.52 + 0.65 ms (scripts/density_bench.py,
# profiles/r02_density_h64.txt).  Without gradients (evaluation, gated steps) every shape uses the fused forward.
FUSED_DENSITY_TRAIN_SHAPES = ((5, 16),) if __import__("os").environ.get("NKBS_FUSED_DENSITY_H64_BWD", "0") != "1" else FUSED_DENSITY_SHAPES
# The other shapes still take the fused FORWARD when they train (nkbs_density_fwd_save: the kernel also leaves the
# features, normalised positions and selectors) and run their backward through the MLP and encode kernels — measured at
# 8192 x 256 samples of the hidden-64 field: forward 0.52 ms (encode + MLP) -> 0.37 ms + the extra stores.
# NKBS_HYBRID_DENSITY=0: encode + MLP kernels in both directions.
HYBRID_DENSITY = __import__("os").environ.get("NKBS_HYBRID_DENSITY", "1") != "0"
# Shapes with a fused backward (the nerfacto proposal fields) on ray samples: the forward also stores features and
# selectors (44 B/sample) and the backward reads them back (nkbs_density_bwd_saved) instead of recomputing positions
# and the 8 L table gathers per sample; the ray-major scatter follows as before.  NKBS_SAVED_DENSITY_BWD=0: recompute.
SAVED_DENSITY_BWD = __import__("os").environ.get("NKBS_SAVED_DENSITY_BWD", "1") != "0"
_FUSED_DENSITY = {"enabled": True}


def set_fused_density(enabled: bool) -> None:
    """Route HashMLPDensityField through the one-kernel path (default) or through encode + mlp."""
    _FUSED_DENSITY["enabled"] = bool(enabled)


def fused_density_enabled() -> bool:
    return _FUSED_DENSITY["enabled"]


# Backward of the fused proposal field for samples that come as rays: the kernel writes the feature gradients and a
# ray-major kernel scatters them (nkbs_encode_bwd_rays: consecutive samples of a ray that share a cell are summed in
# registers).  NKBS_RAY_SCATTER=0 keeps the sample-major scatter inside nkbs_density_bwd.
_RAY_SCATTER = {"enabled": __import__("os").environ.get("NKBS_RAY_SCATTER", "1") != "0", "encode": False}


def set_ray_scatter(enabled: bool, encode: bool = False) -> None:
    """``enabled``: backward of the fused proposal field; ``encode``: backward of contract_encode on ray samples."""
    _RAY_SCATTER["enabled"] = bool(enabled)
    _RAY_SCATTER["encode"] = bool(encode)


def _density_backward(mode, x, origins, directions, p_st, p_en, bstride, R, S, norm, table, grid, W1, b1, W2, b2,
                      ddensity, dedup, dtable, dW1, db1, dW2, db2):
    """nkbs_density_bwd, with the table-gradient scatter done ray-major when the samples come as rays of <= 256."""
    head = (mode, ptr(x), ptr(origins), ptr(directions), p_st, p_en, bstride, R, S, int(norm.contraction),
            host_floats(norm.aabb), ptr(table), host_floats(grid.scalings), grid.L, grid.log2_T, ptr(W1), ptr(b1), ptr(W2),
            ptr(b2), W1.shape[0], ptr(ddensity), dedup)
    if _RAY_SCATTER["enabled"] and mode == 2 and S <= 256 and R > 0:
        dfeat = torch.empty((R * S, grid.out_dim), device=table.device, dtype=torch.float32)
        call("nkbs_density_bwd", *head, None, ptr(dW1), ptr(db1), ptr(dW2), ptr(db2), ptr(dfeat), stream())
        call("nkbs_encode_bwd_rays", R, S, ptr(origins), ptr(directions), p_st, p_en, bstride, int(norm.contraction),
             host_floats(norm.aabb), ptr(dfeat), host_floats(grid.scalings), grid.L, grid.log2_T, ptr(dtable), stream())
        _finish_temps()
        return
    call("nkbs_density_bwd", *head, ptr(dtable), ptr(dW1), ptr(db1), ptr(dW2), ptr(db2), None, stream())
    _finish_temps()


class _FusedDensity(torch.autograd.Function):
    @staticmethod
    def forward(ctx, table, W1, b1, W2, b2, grid, norm, dedup_levels, x, origins, directions, starts, ends, sinks=None):
        ctx.sinks = sinks
        table, W1, b1, W2, b2 = _c(table), _c(W1), _c(b1), _c(W2), _c(b2)
        require_cuda(table, W1, x, origins, starts)
        if x is not None:
            mode, R, S, dev = 1, x.shape[0], 1, x.device
        else:
            mode, R, S, dev = 2, starts.shape[0], starts.shape[1], starts.device
        H = W1.shape[0]
        density = torch.empty((R * S,), device=dev, dtype=torch.float32)
        aabb = host_floats(norm.aabb)
        p_st, p_en, bstride, _keep = rows(starts, ends)
        head = (mode, ptr(x), ptr(origins), ptr(directions), p_st, p_en, bstride, R, S, int(norm.contraction), aabb,
                ptr(table), host_floats(grid.scalings), grid.L, grid.log2_T, ptr(W1), ptr(b1), ptr(W2), ptr(b2), H, ptr(density))
        ctx.hybrid = (grid.L, H) not in FUSED_DENSITY_TRAIN_SHAPES and any(ctx.needs_input_grad[:5])
        if ctx.hybrid:
            # fused forward, backward through the MLP and encode kernels (see HYBRID_DENSITY)
            N = R * S
            feat = torch.empty((N, grid.out_dim), device=dev, dtype=torch.float32)
            xn = torch.empty((N, 3), device=dev, dtype=torch.float32)
            sel = torch.empty((N,), device=dev, dtype=torch.float32)
            call("nkbs_density_fwd_save", *head, ptr(feat), ptr(xn), ptr(sel), stream())
            ctx.cfg = (grid, dedup_levels, _MLP_CFG["precision"] == "tc", _MLP_CFG["bwd_fmt"])
            ctx.save_for_backward(table, W1, b1, W2, b2, feat, xn, sel)
            return density
        ctx.saved_feat = (SAVED_DENSITY_BWD and any(ctx.needs_input_grad[:5]) and (grid.L, H) == (5, 16) and mode == 2
                          and _RAY_SCATTER["enabled"] and S <= 256 and R > 0)
        ctx.cfg = (mode, R, S, H, grid, norm, dedup_levels if dedup_levels >= 0 else grid.L)
        if ctx.saved_feat:
            feat = torch.empty((R * S, grid.out_dim), device=dev, dtype=torch.float32)
            sel = torch.empty((R * S,), device=dev, dtype=torch.float32)
            call("nkbs_density_fwd_save", *head, ptr(feat), None, ptr(sel), stream())
            ctx.save_for_backward(table, W1, b1, W2, b2, feat, sel, origins, directions, starts, ends)
            return density
        call("nkbs_density_fwd", *head, stream())
        # proposal rays carry 96-256 samples: neighbouring samples share cells on every level of these grids
        ctx.save_for_backward(table, W1, b1, W2, b2, x, origins, directions, starts, ends)
        return density

    @staticmethod
    def backward(ctx, ddensity):
        if ctx.hybrid:
            grid, dedup, tc, bwd_fmt = ctx.cfg
            table, W1, b1, W2, b2, feat, xn, sel = ctx.saved_tensors
            ddensity = _c(ddensity).reshape(-1, 1)
            N = feat.shape[0]
            (dtable, dW1, db1, dW2, db2), ret = _grad_buffers(ctx.sinks, [table, W1, b1, W2, b2])
            dfeat = torch.empty_like(feat)
            dims, Ws, bs, dWs, dbs = [grid.out_dim, W1.shape[0], 1], [W1, W2], [b1, b2], [dW1, dW2], [db1, db2]
            if tc:
                call("nkbs_mlp_tc_bwd", ptr(feat), ptr(ddensity), N, 2, host_i32(dims), host_ptrs(Ws), host_ptrs(bs), 1,
                     ACT_DENSITY, bwd_fmt, ptr(sel), None, ptr(dfeat), host_ptrs(dWs), host_ptrs(dbs), stream())
            else:
                call("nkbs_mlp_bwd", ptr(feat), ptr(ddensity), N, 2, host_i32(dims), host_ptrs(Ws), host_ptrs(bs), 1,
                     ACT_DENSITY, ptr(sel), None, ptr(dfeat), host_ptrs(dWs), host_ptrs(dbs), stream())
            call("nkbs_encode_bwd", ptr(xn), ptr(dfeat), N, host_floats(grid.scalings), grid.L, grid.log2_T,
                 grid.features_per_level, dedup, ptr(dtable), stream())
            _finish_temps()
            return (*ret,) + (None,) * 9
        mode, R, S, H, grid, norm, dedup = ctx.cfg
        if ctx.saved_feat:
            table, W1, b1, W2, b2, feat, sel, origins, directions, starts, ends = ctx.saved_tensors
            ddensity = _c(ddensity)
            (dtable, dW1, db1, dW2, db2), ret = _grad_buffers(ctx.sinks, [table, W1, b1, W2, b2])
            p_st, p_en, bstride, _keep = rows(starts, ends)
            dfeat = torch.empty_like(feat)
            call("nkbs_density_bwd_saved", R * S, grid.L, H, ptr(feat), ptr(sel), ptr(ddensity), ptr(W1), ptr(b1), ptr(W2),
                 ptr(b2), ptr(dW1), ptr(db1), ptr(dW2), ptr(db2), ptr(dfeat), stream())
            call("nkbs_encode_bwd_rays", R, S, ptr(origins), ptr(directions), p_st, p_en, bstride, int(norm.contraction),
                 host_floats(norm.aabb), ptr(dfeat), host_floats(grid.scalings), grid.L, grid.log2_T, ptr(dtable), stream())
            _finish_temps()
            return (*ret,) + (None,) * 9
        table, W1, b1, W2, b2, x, origins, directions, starts, ends = ctx.saved_tensors
        ddensity = _c(ddensity)
        (dtable, dW1, db1, dW2, db2), ret = _grad_buffers(ctx.sinks, [table, W1, b1, W2, b2])
        p_st, p_en, bstride, _keep = rows(starts, ends)
        _density_backward(mode, x, origins, directions, p_st, p_en, bstride, R, S, norm, table, grid, W1, b1, W2, b2,
                          ddensity, dedup, dtable, dW1, db1, dW2, db2)
        return (*ret,) + (None,) * 9


def fused_density(table, grid: GridSpec, norm: NormSpec, W1, b1, W2, b2, positions=None, origins=None, directions=None,
                  starts=None, ends=None, dedup_levels=-1):
    """density [N] of the proposal field.  Either raw ``positions`` [N,3] or per-ray ``origins/directions`` [R,3]
    with per-sample ``starts/ends`` [R,S].  Gradients: table and the four MLP tensors (sample positions are
    detached in the reference path)."""
    if (grid.L, W1.shape[0]) not in FUSED_DENSITY_SHAPES or W1.shape[1] != grid.out_dim or W2.shape[0] != 1:
        raise ValueError(f"no fused density kernel for L={grid.L}, MLP {tuple(W1.shape)} -> {tuple(W2.shape)}")
    sinks = _sinks([table, W1, b1, W2, b2])
    if positions is not None:
        return _FusedDensity.apply(table, W1, b1, W2, b2, grid, norm, dedup_levels, _c(positions).reshape(-1, 3), None,
                                   None, None, None, sinks)
    return _FusedDensity.apply(table, W1, b1, W2, b2, grid, norm, dedup_levels, None, _c(origins), _c(directions),
                               starts.float(), ends.float(), sinks)


# ----------------------------------------------------------------------------------------------
# K3 fused MLP (fp32 CUDA-core path)
# ----------------------------------------------------------------------------------------------
# "fp32": CUDA-core kernels (parity 1e-4).  "tc": tcgen05 tensor cores, 16-bit operands, fp32 accumulate
# (parity 1e-2; what the reference's autocast training does, ref:train.py:50).
FMT_FP16, FMT_BF16 = 0, 1
_MLP_CFG = {"precision": "tc", "fwd_fmt": FMT_FP16, "bwd_fmt": FMT_BF16}


def set_mlp_precision(precision: str, fwd_fmt: int = FMT_FP16, bwd_fmt: int = FMT_BF16) -> None:
    """Select the MLP kernels: ``"fp32"`` (CUDA cores) or ``"tc"`` (tensor cores; operand formats per pass)."""
    if precision not in ("fp32", "tc"):
        raise ValueError("precision must be 'fp32' or 'tc'")
    _MLP_CFG.update(precision=precision, fwd_fmt=int(fwd_fmt), bwd_fmt=int(bwd_fmt))


def get_mlp_precision() -> str:
    return _MLP_CFG["precision"]


class _Mlp(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, selector, relu_mask, out_act, col0, split, sinks, *params):
        ctx.sinks = sinks
        ctx.split = bool(split)  # see _BWD_SPLIT  # gradient sinks in the order W_0..W_{n-1}, b_0..b_{n-1}, or None
        nl = len(params) // 2
        Ws = [_c(p) for p in params[0::2]]
        bs = [_c(p) for p in params[1::2]]
        x = _c(x)
        require_cuda(x, *Ws)
        ctx.set_materialize_grads(False)
        dims = [Ws[0].shape[1]] + [w.shape[0] for w in Ws]
        N = x.shape[0]
        y = torch.empty((N, dims[-1]), device=x.device, dtype=torch.float32)
        # column 0 as its own contiguous output, written by the same launch (no copy kernel)
        y0 = torch.empty((N,), device=x.device, dtype=torch.float32) if col0 else None
        tc = _MLP_CFG["precision"] == "tc"
        if tc:
            call("nkbs_mlp_tc_fwd", ptr(x), N, nl, host_i32(dims), host_ptrs(Ws), host_ptrs(bs), relu_mask, out_act,
                 _MLP_CFG["fwd_fmt"], ptr(selector), ptr(y), ptr(y0), stream())
        else:
            call("nkbs_mlp_fwd", ptr(x), N, nl, host_i32(dims), host_ptrs(Ws), host_ptrs(bs), relu_mask, out_act,
                 ptr(selector), ptr(y), ptr(y0), stream())
        ctx.cfg = (nl, dims, relu_mask, out_act, tc, _MLP_CFG["bwd_fmt"])
        ctx.save_for_backward(x, selector, *Ws, *bs)
        if col0:
            # column 0 also leaves as its own contiguous output (the density of the base MLP): its gradient comes back
            # to THIS node and is added into column 0 of dy, instead of a split node that zero-fills, concatenates and
            # adds three [N,16] tensors
            return y, y0
        return y

    @staticmethod
    def backward(ctx, dy, dy0=None):
        nl, dims, relu_mask, out_act, tc, bwd_fmt = ctx.cfg
        saved = ctx.saved_tensors
        x, selector = saved[0], saved[1]
        Ws, bs = list(saved[2:2 + nl]), list(saved[2 + nl:2 + 2 * nl])
        if dy is None and dy0 is None:
            return (None,) * (7 + 2 * nl)
        if dy is None:
            dy = torch.zeros((x.shape[0], dims[-1]), device=x.device, dtype=torch.float32)
        dy = _c(dy)
        dy0 = _c(dy0).reshape(-1) if dy0 is not None else None  # added to column 0 of dy inside the kernel
        N = x.shape[0]
        dx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        zs, ret = _grad_buffers(ctx.sinks, Ws + bs)
        dWs, dbs = zs[:nl], zs[nl:]
        def launch(lo, n):
            off = lambda t, width: None if t is None else t.data_ptr() + 4 * width * lo
            if tc:
                call("nkbs_mlp_tc_bwd", off(x, dims[0]), off(dy, dims[-1]), n, nl, host_i32(dims), host_ptrs(Ws), host_ptrs(bs),
                     relu_mask, out_act, bwd_fmt, off(selector, 1), off(dy0, 1), off(dx, dims[0]), host_ptrs(dWs),
                     host_ptrs(dbs), stream())
            else:
                call("nkbs_mlp_bwd", off(x, dims[0]), off(dy, dims[-1]), n, nl, host_i32(dims), host_ptrs(Ws), host_ptrs(bs),
                     relu_mask, out_act, off(selector, 1), off(dy0, 1), off(dx, dims[0]), host_ptrs(dWs), host_ptrs(dbs), stream())

        frac = _BWD_SPLIT["fraction"]
        if ctx.split and dx is not None and 0.0 < frac < 1.0 and N >= _BWD_SPLIT["min_rows"] and not _TEMPS:
            k = max(128, int(N * frac) // 128 * 128)  # whole 128-sample tiles in part one
            launch(0, k)
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(x.device))
            launch(k, N - k)
            _BWD_SPLIT["pending"].clear()  # one producer / consumer pair at a time
            _BWD_SPLIT["pending"][dx.data_ptr()] = (ev, k)
        else:
            launch(0, N)
        _finish_temps()
        grads = []
        for w, b in zip(ret[:nl], ret[nl:]):
            grads += [w, b]
        return (dx, None, None, None, None, None, None, *grads)


def mlp(x, weights: Sequence[torch.Tensor], biases: Sequence[torch.Tensor], relu_mask: int, out_act: int = ACT_NONE,
        selector: Optional[torch.Tensor] = None, col0: bool = False, split_bwd: bool = False):
    """y = MLP(x) with up to three nn.Linear layers; x [N,K0] -> y [N,Kn].  ``col0=True`` returns (y, y[:, 0] as a
    contiguous [N] tensor) from one autograd node."""
    params = []
    for w, b in zip(weights, biases):
        params += [w, b]
    return _Mlp.apply(x, selector, relu_mask, out_act, bool(col0), bool(split_bwd), _sinks(list(weights) + list(biases)), *params)


# ----------------------------------------------------------------------------------------------
# K4 head-input assembly
# ----------------------------------------------------------------------------------------------
class _Assemble(torch.autograd.Function):
    @staticmethod
    def forward(ctx, geo, emb, emb_const, directions, cam, S, sh_dim, dir_per_ray, cam_per_ray, sink=None):
        geo = _c(geo)
        ctx.sink = sink
        N, geo_dim = geo.shape
        emb_dim = emb.shape[1] if emb is not None else emb_const.shape[0]
        h = torch.empty((N, sh_dim + geo_dim + emb_dim), device=geo.device, dtype=torch.float32)
        use_cam = cam is not None and emb_const is None
        call("nkbs_assemble_fwd", N, S, sh_dim, geo_dim, emb_dim, ptr(directions), int(dir_per_ray), ptr(geo),
             ptr(cam) if use_cam else None, int(cam_per_ray), ptr(emb) if use_cam else None, ptr(emb_const),
             emb.shape[0] if emb is not None else 0, ptr(h), stream())
        ctx.cfg = (N, S, sh_dim, geo_dim, emb_dim, cam_per_ray, use_cam, emb.shape if emb is not None else None)
        ctx.save_for_backward(cam if use_cam else None)
        return h

    @staticmethod
    def backward(ctx, dh):
        N, S, sh_dim, geo_dim, emb_dim, cam_per_ray, use_cam, emb_shape = ctx.cfg
        (cam,) = ctx.saved_tensors
        dh = _c(dh)
        dgeo = torch.empty((N, geo_dim), device=dh.device, dtype=torch.float32) if ctx.needs_input_grad[0] else None
        demb, ret = None, None
        if use_cam and ctx.needs_input_grad[1]:
            if ctx.sink is not None:  # the step driver's gradient arena (accumulated into; autograd sees no gradient)
                demb = ctx.sink
            else:
                demb = ret = torch.zeros(emb_shape, device=dh.device, dtype=torch.float32)
                _temp_shadow(ret)
        call("nkbs_assemble_bwd", N, S, sh_dim, geo_dim, emb_dim, ptr(dh), ptr(dgeo), ptr(cam) if demb is not None else None,
             int(cam_per_ray), ptr(demb), stream())
        _finish_temps()
        return dgeo, ret, None, None, None, None, None, None, None, None


def assemble(geo, emb_table, emb_const, directions, cam, S: int, sh_dim: int, dir_per_ray: bool, cam_per_ray: bool):
    """h = [SH16((d+1)/2) | geo | embedding[cam] or emb_const]; see nkbs_assemble_fwd."""
    return _Assemble.apply(geo, emb_table, emb_const, _c(directions), _c(cam, torch.int64), S, sh_dim, dir_per_ray,
                           cam_per_ray, getattr(emb_table, "_nkbs_grad_sink", None) if emb_table is not None else None)


# ----------------------------------------------------------------------------------------------
# K3+K4 head MLP with in-kernel input assembly (tensor-core path only)
# ----------------------------------------------------------------------------------------------
_FUSED_HEAD = {"enabled": True}


def set_fused_head(enabled: bool) -> None:
    """Head MLPs build their input inside the tensor-core kernel (default) or go through nkbs_assemble + nkbs_mlp_tc."""
    _FUSED_HEAD["enabled"] = bool(enabled)


def fused_head_enabled() -> bool:
    return _FUSED_HEAD["enabled"]


class _HeadMlp(torch.autograd.Function):
    """y = MLP([SH16 | geo | emb[cam]]) with the input row built inside the tensor-core kernel.  ``src`` holds the
    geometry features in columns [col, col+15) of its rows (the base MLP's output is passed whole: no slice copy)."""

    @staticmethod
    def forward(ctx, src, col, emb, emb_const, directions, cam, S, sh_dim, dir_per_ray, cam_per_ray, relu_mask, out_act,
                sinks, *params):
        ctx.sinks, ctx.emb_sink = sinks  # (W.., b.. sinks or None, embedding-table sink or None)
        nl = len(params) // 2
        Ws = [_c(p) for p in params[0::2]]
        bs = [_c(p) for p in params[1::2]]
        src = _c(src)
        require_cuda(src, *Ws)
        N, stride = src.shape
        emb_dim = emb.shape[1] if emb is not None else (emb_const.shape[0] if emb_const is not None else 0)
        dims = [Ws[0].shape[1]] + [w.shape[0] for w in Ws]
        use_cam = cam is not None and emb_const is None
        y = torch.empty((N, dims[-1]), device=src.device, dtype=torch.float32)
        geo = src.reshape(-1)[col:]  # element offset of the first geometry column
        call("nkbs_head_tc_fwd", sh_dim, 15, emb_dim, ptr(geo), stride, ptr(directions), int(dir_per_ray),
             ptr(cam) if use_cam else None, int(cam_per_ray), ptr(emb) if use_cam else None, ptr(emb_const), S, N, nl,
             host_i32(dims), host_ptrs(Ws), host_ptrs(bs), relu_mask, out_act, _MLP_CFG["fwd_fmt"], ptr(y), stream())
        ctx.cfg = (col, S, sh_dim, emb_dim, dir_per_ray, cam_per_ray, relu_mask, out_act, nl, dims, use_cam,
                   _MLP_CFG["bwd_fmt"], emb.shape if emb is not None else None)
        ctx.save_for_backward(src, emb, emb_const, directions, cam, *Ws, *bs)
        return y

    @staticmethod
    def backward(ctx, dy):
        col, S, sh_dim, emb_dim, dpr, cpr, relu_mask, out_act, nl, dims, use_cam, bwd_fmt, emb_shape = ctx.cfg
        saved = ctx.saved_tensors
        src, emb, emb_const, directions, cam = saved[:5]
        Ws, bs = list(saved[5:5 + nl]), list(saved[5 + nl:5 + 2 * nl])
        dy = _c(dy)
        N, stride = src.shape
        # a [N, col + 15] source (the base MLP's [N,16] output, col = 1) is written whole by the kernel: geometry
        # columns from dX_0, the `col` leading ones zeroed (dgeo_lead); any other layout starts from a zero fill
        whole = col + 15 == stride
        dsrc = (torch.empty_like(src) if whole else torch.zeros_like(src)) if ctx.needs_input_grad[0] else None
        demb, demb_ret = None, None
        if use_cam and ctx.needs_input_grad[2]:
            if ctx.emb_sink is not None:
                demb = ctx.emb_sink
            else:
                demb = demb_ret = torch.zeros(emb_shape, device=src.device, dtype=torch.float32)
                _temp_shadow(demb_ret)
        zs, ret = _grad_buffers(ctx.sinks, Ws + bs)
        dWs, dbs = zs[:nl], zs[nl:]
        geo = src.reshape(-1)[col:]
        dgeo = dsrc.reshape(-1)[col:] if dsrc is not None else None
        call("nkbs_head_tc_bwd", sh_dim, 15, emb_dim, ptr(geo), stride, ptr(directions), int(dpr),
             ptr(cam) if use_cam else None, int(cpr), ptr(emb) if use_cam else None, ptr(emb_const), S, ptr(dy), N, nl,
             host_i32(dims), host_ptrs(Ws), host_ptrs(bs), relu_mask, out_act, bwd_fmt, ptr(dgeo), stride,
             col if (whole and dgeo is not None) else 0, ptr(demb), host_ptrs(dWs), host_ptrs(dbs), stream())
        _finish_temps()
        grads = []
        for w, b in zip(ret[:nl], ret[nl:]):
            grads += [w, b]
        return (dsrc, None, demb_ret, None, None, None, None, None, None, None, None, None, None, *grads)


def head_mlp(src, col: int, emb_table, emb_const, directions, cam, S: int, sh_dim: int, dir_per_ray: bool,
             cam_per_ray: bool, weights, biases, relu_mask: int, out_act: int):
    """Head MLP over rows [SH(sh_dim) | src[:, col:col+15] | embedding row]; tensor-core path (nkbs_head_tc_*).
    Without an embedding (``emb_table`` and ``emb_const`` both None) the row is the geometry feature alone."""
    params = []
    for w, b in zip(weights, biases):
        params += [w, b]
    sinks = (_sinks(list(weights) + list(biases)), getattr(emb_table, "_nkbs_grad_sink", None))
    return _HeadMlp.apply(src, col, emb_table, emb_const, _c(directions), _c(cam, torch.int64), S, sh_dim, dir_per_ray,
                          cam_per_ray, relu_mask, out_act, sinks, *params)


# ----------------------------------------------------------------------------------------------
# K5 samplers (no gradient: the reference detaches the bins)
# ----------------------------------------------------------------------------------------------
_LIN_CACHE = {}


def _lin_table(kind: str, n: int, device) -> torch.Tensor:
    """torch.linspace evaluated on the CPU (the oracle's rounding), cached per device."""
    key = (kind, n, str(device))
    t = _LIN_CACHE.get(key)
    if t is None:
        if kind == "bins":  # linspace(0, 1, S+1)
            t = torch.linspace(0.0, 1.0, n + 1)
        else:  # "u": linspace(0, 1 - 1/nb, nb)
            t = torch.linspace(0.0, 1.0 - (1.0 / n), steps=n)
        t = t.to(device)
        _LIN_CACHE[key] = t
    return t


@torch.no_grad()
def initial_sampler(nears, fars, num_samples: int, t_rand=None):
    """UniformLinDispPiecewiseSampler bins.  Returns (spacing bins [R,S+1], euclidean bins [R,S+1])."""
    nears, fars = _c(nears).reshape(-1), _c(fars).reshape(-1)
    require_cuda(nears, fars, t_rand)
    R = nears.shape[0]
    lin = _lin_table("bins", num_samples, nears.device)
    sb = torch.empty((R, num_samples + 1), device=nears.device, dtype=torch.float32)
    eb = torch.empty_like(sb)
    per_bin = 0
    if t_rand is not None:
        t_rand = _c(t_rand)
        per_bin = int(t_rand.numel() != R)
    call("nkbs_initial_sampler", R, num_samples, ptr(lin), ptr(t_rand), per_bin, ptr(nears), ptr(fars), ptr(sb),
         ptr(eb), stream())
    return sb, eb


@torch.no_grad()
def pdf_sample(weights, spacing_bins, nears, fars, num_samples: int, anneal: float = 1.0, rand=None,
               histogram_padding: float = 0.01, eps: float = 1e-5, anneal_dev=None, want_indices=False):
    """PDFSampler: weights [R,S] (pre-anneal), spacing_bins [R,S+1] -> new (spacing, euclid) bins [R,ns+1]."""
    weights = _c(weights.detach())
    R, S = weights.shape
    nb = num_samples + 1
    sbins = _c(spacing_bins)
    nears, fars = _c(nears).reshape(-1), _c(fars).reshape(-1)
    require_cuda(weights, sbins, nears)
    u = _lin_table("u", nb, weights.device)
    ob = torch.empty((R, nb), device=weights.device, dtype=torch.float32)
    oe = torch.empty_like(ob)
    inds = torch.empty((R, nb), device=weights.device, dtype=torch.int64) if want_indices else None
    per_bin = 0
    if rand is not None:
        rand = _c(rand)
        per_bin = int(rand.numel() != R)
    call("nkbs_pdf_sample", R, S, nb, ptr(weights), float(anneal), ptr(anneal_dev), float(histogram_padding),
         float(eps), ptr(sbins), ptr(u), ptr(rand), per_bin, ptr(nears), ptr(fars), ptr(ob), ptr(oe), ptr(inds),
         stream())
    return (ob, oe, inds) if want_indices else (ob, oe)


# ----------------------------------------------------------------------------------------------
# one proposal level in one ray-major pass: [initial sampler] + density field + get_weights + median depth + PDF sampler
# ----------------------------------------------------------------------------------------------
# Measured on the B200 (round 2, same box A/B, 4096-ray captured step): four launches 0.796 ms, one fused launch 0.815 ms
# (4 rays x 4 warps per CTA), 0.85 / 0.89 ms with one field evaluation per thread (2 or 1 rays per 8-warp CTA): the
# per-ray scan + resampling phases keep one warp per ray busy while the CTA's other warps wait, which costs more than
# the three launch boundaries and [R,S] round trips it removes.  Hence off by default; NKBS_FUSED_LEVEL=1 selects it.
_FUSED_LEVEL = {"enabled": __import__("os").environ.get("NKBS_FUSED_LEVEL", "0") == "1"}


def set_fused_level(enabled: bool) -> None:
    """ProposalNetworkSampler runs each proposal level as ONE launch (default) or as sampler + field + compositing
    + PDF launches."""
    _FUSED_LEVEL["enabled"] = bool(enabled)


def fused_level_enabled() -> bool:
    return _FUSED_LEVEL["enabled"]


class _ProposalLevel(torch.autograd.Function):
    """weights of one proposal level and the next level's bins (nkbs_proposal_level_fwd).  Differentiable output:
    the weights (w.r.t. the field's table and MLP; bins and positions are detached in the reference).
    Backward: d weights -> d density (nkbs_composite_bwd) -> table / MLP gradients (nkbs_density_bwd)."""

    @staticmethod
    def forward(ctx, table, W1, b1, W2, b2, cfg, origins, directions, nears, fars, sbins, ebins, t_rand, rand, sinks):
        grid, norm, S, nb, anneal, anneal_dev, hist_pad, eps, dedup = cfg
        table, W1, b1, W2, b2 = _c(table), _c(W1), _c(b1), _c(W2), _c(b2)
        origins, directions = _c(origins), _c(directions)
        nears, fars = _c(nears).reshape(-1), _c(fars).reshape(-1)
        require_cuda(table, W1, origins, nears, sbins, t_rand, rand)
        R, dev = origins.shape[0], origins.device
        gen = sbins is None
        if gen:
            sbins = torch.empty((R, S + 1), device=dev, dtype=torch.float32)
            ebins = torch.empty_like(sbins)
        else:
            sbins, ebins = _c(sbins), _c(ebins)
        f32 = dict(device=dev, dtype=torch.float32)
        density, weights = torch.empty((R, S), **f32), torch.empty((R, S), **f32)
        dmed = torch.empty((R,), **f32)
        nsb, neb = torch.empty((R, nb), **f32), torch.empty((R, nb), **f32)
        t_rand = _c(t_rand).reshape(-1) if t_rand is not None else None
        rand = _c(rand).reshape(-1) if rand is not None else None
        call("nkbs_proposal_level_fwd", R, S, ptr(origins), ptr(directions), ptr(nears), ptr(fars), int(gen),
             ptr(_lin_table("bins", S, dev)) if gen else None, ptr(t_rand), ptr(sbins), ptr(ebins), int(norm.contraction),
             host_floats(norm.aabb), ptr(table), host_floats(grid.scalings), grid.L, grid.log2_T, ptr(W1), ptr(b1), ptr(W2),
             ptr(b2), W1.shape[0], ptr(density), ptr(weights), ptr(dmed), nb, float(anneal), ptr(anneal_dev),
             float(hist_pad), float(eps), ptr(_lin_table("u", nb, dev)), ptr(rand), ptr(nsb), ptr(neb), stream())
        ctx.cfg = (grid, norm, S, R, dedup if dedup >= 0 else grid.L)
        ctx.sinks = sinks
        ctx.set_materialize_grads(False)
        ctx.save_for_backward(table, W1, b1, W2, b2, origins, directions, ebins, density, weights)
        ctx.mark_non_differentiable(dmed, sbins, ebins, nsb, neb)
        return weights, dmed, sbins, ebins, nsb, neb

    @staticmethod
    def backward(ctx, g_w, *_unused):
        grid, norm, S, R, dedup = ctx.cfg
        table, W1, b1, W2, b2, origins, directions, ebins, density, weights = ctx.saved_tensors
        if g_w is None:
            return (None,) * 15
        g_w = _c(g_w)
        ddensity = torch.empty_like(density)
        st, en = ebins.data_ptr(), ebins.data_ptr() + 4
        call("nkbs_composite_bwd", R, S, ptr(density), ptr(weights), st, en, S + 1, None, None, 0, 0, 0, None, None,
             ptr(g_w), None, None, None, None, ptr(ddensity), None, None, None, stream())
        (dtable, dW1, db1, dW2, db2), ret = _grad_buffers(ctx.sinks, [table, W1, b1, W2, b2])
        _density_backward(2, None, origins, directions, st, en, S + 1, R, S, norm, table, grid, W1, b1, W2, b2, ddensity,
                          dedup, dtable, dW1, db1, dW2, db2)
        return (*ret,) + (None,) * 10


def proposal_level(table, grid: GridSpec, norm: NormSpec, W1, b1, W2, b2, origins, directions, nears, fars,
                   num_samples: int, num_next: int, sbins=None, ebins=None, t_rand=None, rand=None, anneal: float = 1.0,
                   anneal_dev=None, histogram_padding: float = 0.01, eps: float = 1e-5, dedup_levels: int = -1):
    """One proposal level: returns (weights [R,S], median depth [R], spacing bins [R,S+1], euclidean bins [R,S+1],
    next spacing bins [R,num_next+1], next euclidean bins [R,num_next+1]).  ``sbins`` / ``ebins`` None: the level's
    bins are the initial sampler's (``t_rand`` [R,1] per-ray jitter or None); ``rand`` [R,1]: per-ray jitter of the
    resampling (None: centred, evaluation)."""
    if (grid.L, W1.shape[0]) not in FUSED_DENSITY_SHAPES or W1.shape[1] != grid.out_dim or W2.shape[0] != 1:
        raise ValueError(f"no fused proposal-level kernel for L={grid.L}, MLP {tuple(W1.shape)} -> {tuple(W2.shape)}")
    cfg = (grid, norm, int(num_samples), int(num_next) + 1, float(anneal), anneal_dev, float(histogram_padding),
           float(eps), int(dedup_levels))
    return _ProposalLevel.apply(table, W1, b1, W2, b2, cfg, origins, directions, nears, fars, sbins, ebins, t_rand, rand,
                                _sinks([table, W1, b1, W2, b2]))


# ----------------------------------------------------------------------------------------------
# K6/K7 compositing
# ----------------------------------------------------------------------------------------------
class _Composite(torch.autograd.Function):
    """One ray-major pass: weights (from density) and every requested render."""

    @staticmethod
    def forward(ctx, density, weights_in, rgb, sem, starts, ends, bg_mode, eval_mode, sem_weight_grad, want):
        ctx.set_materialize_grads(False)
        src = density if density is not None else weights_in
        R, S = src.shape[0], src.shape[1]
        dev = src.device
        density, weights_in, rgb, sem = _c(density), _c(weights_in), _c(rgb), _c(sem)
        require_cuda(src, rgb, sem, starts)
        p_st, p_en, bstride, keep = rows(starts, ends)
        if keep:
            starts, ends = keep
        C = sem.shape[-1] if sem is not None else 0
        f32 = dict(device=dev, dtype=torch.float32)
        weights = torch.empty((R, S), **f32) if density is not None else None
        rgb_out = torch.empty((R, 3), **f32) if (rgb is not None and "rgb" in want) else None
        acc = torch.empty((R,), **f32) if ("acc" in want or rgb_out is not None or "depth_exp" in want) else None
        d_raw = torch.empty((R,), **f32) if "depth_exp" in want else None
        d_exp = torch.empty((R,), **f32) if d_raw is not None else None
        d_mask = torch.empty((R,), **f32) if d_raw is not None else None
        d_med = torch.empty((R,), **f32) if "depth_med" in want else None
        sem_out = torch.empty((R, C), **f32) if (sem is not None and "sem" in want) else None
        mm = torch.empty((2,), device=dev, dtype=torch.int32) if d_raw is not None else None
        call("nkbs_composite_fwd", R, S, ptr(density), ptr(weights_in), p_st, p_en, bstride, ptr(rgb), ptr(sem),
             C, bg_mode, int(eval_mode), ptr(weights), ptr(rgb_out), ptr(acc), ptr(d_raw), ptr(d_exp), ptr(d_mask),
             ptr(d_med), ptr(sem_out), ptr(mm), stream())
        w_saved = weights if density is not None else weights_in
        ctx.cfg = (R, S, C, bg_mode, sem_weight_grad, density is not None)
        ctx.save_for_backward(density, w_saved, starts, ends, rgb, sem, acc, d_raw, d_mask)
        if d_med is not None:
            ctx.mark_non_differentiable(d_med)
        outs = (weights, rgb_out, acc, d_exp, d_med, sem_out)
        return outs

    @staticmethod
    def backward(ctx, g_w, g_rgb, g_acc, g_dexp, _g_dmed, g_sem):
        R, S, C, bg_mode, sem_weight_grad, from_density = ctx.cfg
        density, w, starts, ends, rgb, sem, acc, d_raw, d_mask = ctx.saved_tensors
        dev = w.device
        f32 = dict(device=dev, dtype=torch.float32)
        g_w, g_rgb, g_acc, g_sem = _c(g_w), _c(g_rgb), _c(g_acc), _c(g_sem)
        g_depth = None
        if g_dexp is not None and d_raw is not None:
            g_depth = (_c(g_dexp) * d_mask).contiguous()
        need_src = ctx.needs_input_grad[0] or ctx.needs_input_grad[1]
        d_density = torch.empty((R, S), **f32) if (from_density and need_src) else None
        d_weights = torch.empty((R, S), **f32) if (not from_density and need_src) else None
        d_rgb = torch.empty((R, S, 3), **f32) if (rgb is not None and ctx.needs_input_grad[2] and g_rgb is not None) else None
        d_sem = torch.empty((R, S, C), **f32) if (sem is not None and ctx.needs_input_grad[3] and g_sem is not None) else None
        p_st, p_en, bstride, _keep = rows(starts, ends)
        call("nkbs_composite_bwd", R, S, ptr(density), ptr(w), p_st, p_en, bstride, ptr(rgb), ptr(sem), C, bg_mode,
             int(sem_weight_grad), ptr(acc), ptr(d_raw), ptr(g_w), ptr(g_rgb) if rgb is not None else None,
             ptr(g_acc), ptr(g_depth), ptr(g_sem) if sem is not None else None, ptr(d_density), ptr(d_weights),
             ptr(d_rgb), ptr(d_sem), stream())
        return d_density, d_weights, d_rgb, d_sem, None, None, None, None, None, None


def composite(density=None, weights=None, rgb=None, sem=None, starts=None, ends=None, background="last_sample",
              eval_mode=False, sem_weight_grad=False, want=("rgb", "acc", "depth_exp", "depth_med", "sem")):
    """Fused get_weights + renderers.  [R,S(,C)] inputs (a trailing singleton dim on density / weights /
    starts / ends is accepted).  Returns dict with weights [R,S,1], rgb [R,3], accumulation [R,1],
    expected_depth [R,1], depth [R,1] (median), semantics [R,C] — those requested and possible."""
    def sq(t):
        # squeeze (a view in the backward pass too), not a select (zero fill + copy in its backward)
        return t.squeeze(-1) if (t is not None and t.dim() == 3 and t.shape[-1] == 1) else t
    density, weights, starts, ends = sq(density), sq(weights), sq(starts), sq(ends)
    w, rgb_o, acc, dexp, dmed, sem_o = _Composite.apply(density, weights, rgb, sem, starts, ends, BG_MODES[background],
                                                        eval_mode, sem_weight_grad, frozenset(want))
    out = {}
    if w is not None:
        out["weights"] = w[..., None]
    if rgb_o is not None:
        out["rgb"] = rgb_o
    if acc is not None and "acc" in want:
        out["accumulation"] = acc[..., None]
    if dexp is not None:
        out["expected_depth"] = dexp[..., None]
    if dmed is not None:
        out["depth"] = dmed[..., None]
    if sem_o is not None:
        out["semantics"] = sem_o
    return out


# ----------------------------------------------------------------------------------------------
# get_weights (+ median depth) of a proposal level and the next level's PDF resampling, one ray-major launch
# ----------------------------------------------------------------------------------------------
_FUSED_RESAMPLE = {"enabled": __import__("os").environ.get("NKBS_FUSED_RESAMPLE", "1") != "0"}


def set_fused_resample(enabled: bool) -> None:
    _FUSED_RESAMPLE["enabled"] = bool(enabled)


def fused_resample_enabled() -> bool:
    return _FUSED_RESAMPLE["enabled"]


class _WeightsResample(torch.autograd.Function):
    """(weights [R,S], median depth [R], next spacing bins, next euclidean bins) of one proposal level from its density
    (nkbs_weights_resample_fwd).  Differentiable: weights w.r.t. density (nkbs_composite_bwd); bins are detached in
    the reference."""

    @staticmethod
    def forward(ctx, density, ebins, sbins, nears, fars, cfg, rand):
        nb, anneal, anneal_dev, hist_pad, eps = cfg
        density = _c(density)
        R, S = density.shape
        ebins, sbins = _c(ebins), _c(sbins)
        nears, fars = _c(nears).reshape(-1), _c(fars).reshape(-1)
        require_cuda(density, ebins, sbins, nears, rand)
        dev = density.device
        f32 = dict(device=dev, dtype=torch.float32)
        weights, dmed = torch.empty((R, S), **f32), torch.empty((R,), **f32)
        nsb, neb = torch.empty((R, nb), **f32), torch.empty((R, nb), **f32)
        rand = _c(rand).reshape(-1) if rand is not None else None
        call("nkbs_weights_resample_fwd", R, S, ptr(density), ptr(ebins), ptr(sbins), ptr(nears), ptr(fars), nb,
             float(anneal), ptr(anneal_dev), float(hist_pad), float(eps), ptr(_lin_table("u", nb, dev)), ptr(rand),
             ptr(weights), ptr(dmed), ptr(nsb), ptr(neb), stream())
        ctx.shape = (R, S)
        ctx.set_materialize_grads(False)
        ctx.save_for_backward(density, weights, ebins)
        ctx.mark_non_differentiable(dmed, nsb, neb)
        return weights, dmed, nsb, neb

    @staticmethod
    def backward(ctx, g_w, *_unused):
        R, S = ctx.shape
        density, weights, ebins = ctx.saved_tensors
        if g_w is None or not ctx.needs_input_grad[0]:
            return (None,) * 7
        g_w = _c(g_w)
        ddensity = torch.empty_like(density)
        call("nkbs_composite_bwd", R, S, ptr(density), ptr(weights), ebins.data_ptr(), ebins.data_ptr() + 4, S + 1, None,
             None, 0, 0, 0, None, None, ptr(g_w), None, None, None, None, ptr(ddensity), None, None, None, stream())
        return (ddensity,) + (None,) * 6


def weights_resample(density, euclid_bins, spacing_bins, nears, fars, num_next: int, rand=None, anneal: float = 1.0,
                     anneal_dev=None, histogram_padding: float = 0.01, eps: float = 1e-5):
    """``density`` [R,S(,1)] of a proposal level over its [R,S+1] bin edges -> (weights [R,S], median depth [R], next
    spacing bins [R,num_next+1], next euclidean bins [R,num_next+1]); ``rand`` [R,1] per-ray jitter, None = centred."""
    d = density.squeeze(-1) if density.dim() == 3 else density
    cfg = (int(num_next) + 1, float(anneal), anneal_dev, float(histogram_padding), float(eps))
    return _WeightsResample.apply(d, euclid_bins, spacing_bins, nears, fars, cfg, rand)


# ----------------------------------------------------------------------------------------------
# N1 per-ray losses (interlevel, distortion, colour MSE, semantic cross-entropy) in one launch
# ----------------------------------------------------------------------------------------------
class _RayLosses(torch.autograd.Function):
    """Returns (rgb_mse, interlevel, distortion, cross_entropy) as 0-dim tensors; terms whose inputs are absent
    come back as zeros.  Differentiable inputs: final weights (distortion), proposal weights (interlevel), rgb,
    logits.  The final weights enter the interlevel term detached, as in the reference."""

    @staticmethod
    def forward(ctx, meta, w_final, rgb, logits, *w_props):
        sdist, sdist_props, image, labels, ignore_index, want_dist = meta
        dev = (w_final if w_final is not None else (rgb if rgb is not None else logits)).device
        require_cuda(w_final, rgb, logits, *w_props)
        out = torch.zeros(5, device=dev, dtype=torch.float32)
        hist = w_final is not None
        R = (w_final if hist else (rgb if rgb is not None else logits)).shape[0]
        S = w_final.shape[1] if hist else 1
        w_final = _c(w_final)
        dw = torch.empty_like(w_final) if (hist and want_dist) else None
        w_props = [_c(t) for t in w_props]
        sdist_props = [_c(t) for t in sdist_props]
        dw_props = [torch.empty_like(t) for t in w_props]
        rgb, image = _c(rgb), _c(image)
        drgb = torch.empty_like(rgb) if rgb is not None else None
        logits = _c(logits)
        dlogits = torch.empty_like(logits) if logits is not None else None
        C = logits.shape[1] if logits is not None else 0
        call("nkbs_ray_losses", R, S, ptr(_c(sdist)), ptr(w_final), ptr(dw), len(w_props),
             host_i32([t.shape[1] for t in w_props]), host_ptrs(sdist_props), host_ptrs(w_props), host_ptrs(dw_props),
             ptr(rgb), ptr(image), ptr(drgb), ptr(logits), ptr(_c(labels, torch.int64)), C, int(ignore_index),
             ptr(dlogits), ptr(out), stream())
        ctx.n_prop = len(w_props)
        ctx.has = (dw is not None, drgb is not None, dlogits is not None)
        ctx.save_for_backward(*[t for t in (dw, drgb, dlogits, out) if t is not None], *dw_props)
        ce = out[3] / torch.clamp(out[4], min=1.0) if logits is not None else out[3]
        return out[0], out[1], out[2], ce

    @staticmethod
    def backward(ctx, g_rgb, g_inter, g_dist, g_ce):
        saved = list(ctx.saved_tensors)
        has_dw, has_rgb, has_log = ctx.has
        dw = saved.pop(0) if has_dw else None
        drgb = saved.pop(0) if has_rgb else None
        dlogits = saved.pop(0) if has_log else None
        out = saved.pop(0)
        dw_props = saved
        g_w = dw * g_dist if (dw is not None and g_dist is not None) else None
        g_c = drgb * g_rgb if (drgb is not None and g_rgb is not None) else None
        g_l = dlogits * (g_ce / torch.clamp(out[4], min=1.0)) if (dlogits is not None and g_ce is not None) else None
        g_p = [t * g_inter if g_inter is not None else None for t in dw_props]
        return (None, g_w, g_c, g_l, *g_p)


def ray_losses(w_final=None, sdist_final=None, w_props=(), sdist_props=(), rgb=None, image=None, logits=None,
               labels=None, ignore_index=-100, want_distortion=True):
    """One launch for every per-ray loss term.  ``w_*`` are [R,S] weights, ``sdist_*`` [R,S+1] spacing-space bin
    edges; returns (rgb_mse, interlevel, distortion, cross_entropy)."""
    meta = (sdist_final, list(sdist_props), image, labels, ignore_index, want_distortion)
    return _RayLosses.apply(meta, w_final, rgb, logits, *w_props)


_LOSS_SCRATCH = {}


def loss_scratch(device) -> torch.Tensor:
    """The zero-between-launches accumulator of nkbs_ray_losses_weighted, one per device (loss launches of one device
    must not overlap on the GPU: one training step at a time per device).  Created on first use: a step driver that
    captures CUDA graphs asks for it before the capture (trainer.Trainer.__init__), so that it does not live in a
    graph's private memory pool."""
    key = torch.device(device).index
    t = _LOSS_SCRATCH.get(key)
    if t is None:
        t = _LOSS_SCRATCH[key] = torch.zeros(8, device=device, dtype=torch.float32)
    return t


class _RayLossTotal(torch.autograd.Function):
    """Weighted total of the colour / interlevel / distortion terms as the ROOT of the backward pass (step driver).

    The kernel writes every gradient already multiplied by its term's weight (``nkbs_ray_losses_weighted``), so
    ``total.backward()`` hands them on as they are: no weighting, summing or rescaling launches around the loss.
    ``unit_seed=True`` is the caller's promise that the incoming gradient of ``total`` is 1 (what ``backward()`` on the
    scalar itself seeds); with ``unit_seed=False`` the saved gradients are multiplied by whatever arrives."""

    @staticmethod
    def forward(ctx, meta, w_final, rgb, *w_props):
        sdist, sdist_props, image, term_weights, unit_seed = meta
        require_cuda(w_final, rgb, *w_props)
        R, S = w_final.shape
        # five terms of the kernel, the total, the three weighted terms: written (not accumulated) by the kernel's last
        # CTA, which also re-zeroes the accumulator scratch: no fill launch around the loss
        out = torch.empty(9, device=w_final.device, dtype=torch.float32)
        scratch = loss_scratch(w_final.device)
        w_final, rgb, image = _c(w_final), _c(rgb), _c(image)
        w_props = [_c(t) for t in w_props]
        sdist_props = [_c(t) for t in sdist_props]
        dw, drgb = torch.empty_like(w_final), torch.empty_like(rgb)
        dw_props = [torch.empty_like(t) for t in w_props]
        call("nkbs_ray_losses_weighted", R, S, ptr(_c(sdist)), ptr(w_final), ptr(dw), len(w_props),
             host_i32([t.shape[1] for t in w_props]), host_ptrs(sdist_props), host_ptrs(w_props), host_ptrs(dw_props),
             ptr(rgb), ptr(image), ptr(drgb), None, None, 0, 0, None, ptr(out), host_floats(term_weights),
             out.data_ptr() + 5 * 4, ptr(scratch), stream())
        ctx.unit_seed = bool(unit_seed)
        ctx.set_materialize_grads(False)
        ctx.save_for_backward(dw, drgb, *dw_props)
        terms, weighted = out[:3], out[6:9]
        ctx.mark_non_differentiable(terms, weighted)
        return out[5], terms, weighted

    @staticmethod
    def backward(ctx, g_total, _g_terms, _g_weighted):
        dw, drgb, *dw_props = ctx.saved_tensors
        if g_total is None:
            return (None,) * (3 + len(dw_props))
        if not ctx.unit_seed:
            dw, drgb, dw_props = dw * g_total, drgb * g_total, [t * g_total for t in dw_props]
        return (None, dw, drgb, *dw_props)


def ray_loss_total(w_final, sdist_final, w_props, sdist_props, rgb, image, term_weights, unit_seed=True):
    """(total, terms[3], weighted[3]) with total = w_rgb * mse + w_inter * interlevel + w_dist * distortion; ``terms``
    are the unweighted (mse, interlevel, distortion) and ``weighted`` the same times their weights, neither
    differentiable.  See ``_RayLossTotal`` for ``unit_seed``."""
    meta = (sdist_final, list(sdist_props), image, tuple(float(v) for v in term_weights), unit_seed)
    return _RayLossTotal.apply(meta, w_final, rgb, *w_props)
