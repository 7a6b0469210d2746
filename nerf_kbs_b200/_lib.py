"""ctypes binding of libnkbs.so (include/nkbs.h).  There is no CPU or eager fallback: if the shared
library is missing or an entry point fails, the caller gets an exception."""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import torch

_HERE = Path(__file__).resolve().parent
_LIB_PATH = Path(os.environ["NKBS_LIB"]) if os.environ.get("NKBS_LIB") else _HERE / "libnkbs.so"  # NKBS_LIB: debug builds

_p = C.c_void_p
_i32 = C.c_int32
_i64 = C.c_int64
_f = C.c_float

# name -> argtypes, in the order of include/nkbs.h
SIGNATURES = {
    "nkbs_encode_fwd": [_i32, _p, _p, _p, _p, _p, _i32, _i64, _i32, _i32, _p, _p, _p, _i32, _i32, _i32, _p, _p, _p, _p, _p],
    "nkbs_encode_bwd": [_p, _p, _i64, _p, _i32, _i32, _i32, _i32, _p, _p],
    "nkbs_mlp_fwd": [_p, _i64, _i32, _p, _p, _p, _i32, _i32, _p, _p, _p, _p],
    "nkbs_mlp_bwd": [_p, _p, _i64, _i32, _p, _p, _p, _i32, _i32, _p, _p, _p, _p, _p, _p],
    "nkbs_mlp_tc_fwd": [_p, _i64, _i32, _p, _p, _p, _i32, _i32, _i32, _p, _p, _p, _p],
    "nkbs_mlp_tc_bwd": [_p, _p, _i64, _i32, _p, _p, _p, _i32, _i32, _i32, _p, _p, _p, _p, _p, _p],
    "nkbs_density_fwd": [_i32, _p, _p, _p, _p, _p, _i32, _i64, _i32, _i32, _p, _p, _p, _i32, _i32, _p, _p, _p, _p, _i32, _p, _p],
    "nkbs_density_fwd_save": [_i32, _p, _p, _p, _p, _p, _i32, _i64, _i32, _i32, _p, _p, _p, _i32, _i32, _p, _p, _p, _p, _i32, _p, _p, _p,
                              _p, _p],
    "nkbs_density_bwd_saved": [_i64, _i32, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p],
    "nkbs_density_bwd": [_i32, _p, _p, _p, _p, _p, _i32, _i64, _i32, _i32, _p, _p, _p, _i32, _i32, _p, _p, _p, _p, _i32, _p, _i32,
                         _p, _p, _p, _p, _p, _p, _p],
    "nkbs_encode_bwd_rays": [_i64, _i32, _p, _p, _p, _p, _i32, _i32, _p, _p, _p, _i32, _i32, _p, _p],
    "nkbs_ray_losses": [_i64, _i32, _p, _p, _p, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i32, _i64, _p, _p, _p],
    "nkbs_ray_losses_weighted": [_i64, _i32, _p, _p, _p, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i32, _i64, _p, _p, _p, _p,
                                 _p, _p],
    "nkbs_generate_rays": [_i64, _p, _p, _p, _i64, _p, _p, _p, _p, _p, _p],
    "nkbs_compact_masked": [_i64, _p, _p, _p, _i64, _i64, _p, _p, _p, _p, _p],
    "nkbs_color_to_label": [_i64, _p, _p, _i64, _p, _p],
    "nkbs_semantic_colormap": [_i64, _i32, _p, _p, _p, _p, _p],
    "nkbs_head_tc_fwd": [_i32, _i32, _i32, _p, _i32, _p, _i32, _p, _i32, _p, _p, _i32, _i64, _i32, _p, _p, _p, _i32, _i32, _i32,
                         _p, _p],
    "nkbs_head_tc_bwd": [_i32, _i32, _i32, _p, _i32, _p, _i32, _p, _i32, _p, _p, _i32, _p, _i64, _i32, _p, _p, _p, _i32, _i32,
                         _i32, _p, _i32, _i32, _p, _p, _p, _p],
    "nkbs_assemble_fwd": [_i64, _i32, _i32, _i32, _i32, _p, _i32, _p, _p, _i32, _p, _p, _i64, _p, _p],
    "nkbs_assemble_bwd": [_i64, _i32, _i32, _i32, _i32, _p, _p, _p, _i32, _p, _p],
    "nkbs_initial_sampler": [_i64, _i32, _p, _p, _i32, _p, _p, _p, _p, _p],
    "nkbs_pdf_sample": [_i64, _i32, _i32, _p, _f, _p, _f, _f, _p, _p, _p, _i32, _p, _p, _p, _p, _p, _p],
    "nkbs_composite_fwd": [_i64, _i32, _p, _p, _p, _p, _i32, _p, _p, _i32, _i32, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p],
    "nkbs_adam_step": [_i32, _p, _p, _p, _p, _p, _f, _f, _f, _f, _f, _p, _p, _i32, _p, _p],
    "nkbs_zero_async": [_p, _i64, _p],
    "nkbs_fixed_point_register": [_p, _i64, _p, _i32],
    "nkbs_fixed_point_flush": [_p, _i64, _p],
    "nkbs_weights_resample_fwd": [_i64, _i32, _p, _p, _p, _p, _p, _i32, _f, _p, _f, _f, _p, _p, _p, _p, _p, _p, _p],
    "nkbs_expected_depth_noalloc": [_i64, _i32, _p, _p, _p, _i32, _p, _p],
    "nkbs_proposal_level_fwd": [_i64, _i32, _p, _p, _p, _p, _i32, _p, _p, _p, _p, _i32, _p, _p, _p, _i32, _i32, _p, _p, _p,
                                _p, _i32, _p, _p, _p, _i32, _f, _p, _f, _f, _p, _p, _p, _p, _p],
    "nkbs_apply_colormap": [_i64, _p, _p, _p, _p],
    "nkbs_apply_depth_colormap": [_i64, _p, _p, _p, _i32, _f, _i32, _f, _p, _p, _p],
    "nkbs_dp_adam_step": [_i32, _i32, _p, _p, _i64, _i64, _p, _p, _f, _f, _f, _f, _f, _p, _p, _p, _p, _p],
    "nkbs_composite_bwd": [_i64, _i32, _p, _p, _p, _p, _i32, _p, _p, _i32, _i32, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p],
}

_lib = None


def load() -> C.CDLL:
    """Load libnkbs.so (built in-tree by nerf_kbs_b200.build).  Raises if it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not _LIB_PATH.exists():
        raise RuntimeError(
            f"{_LIB_PATH} is missing: build it with `python -m nerf_kbs_b200.build` "
            "(there is no CPU fallback for the nkbs kernels)")
    lib = C.CDLL(str(_LIB_PATH))
    lib.nkbs_version.restype = C.c_int
    lib.nkbs_last_error.restype = C.c_char_p
    for name, args in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = C.c_int
    _lib = lib
    return lib


def ptr(t):
    """Device pointer of a contiguous tensor (None -> NULL)."""
    if t is None:
        return None
    assert t.is_contiguous(), "nkbs kernels take contiguous tensors"
    return t.data_ptr()


def rows(starts, ends):
    """[R,S] starts / ends with unit inner stride and a common row stride (e.g. the two views bins[:, :-1] and
    bins[:, 1:] of one [R,S+1] bin-edge array) -> (starts pointer, ends pointer, row stride) without a copy;
    anything else is made contiguous first.  (None, None, 0) when absent."""
    if starts is None:
        return None, None, 0, ()
    if not (starts.dim() == 2 and starts.stride(1) == 1 and ends.stride(1) == 1 and starts.stride(0) == ends.stride(0)
            and starts.stride(0) >= starts.shape[1]):
        starts, ends = starts.contiguous(), ends.contiguous()
    return starts.data_ptr(), ends.data_ptr(), starts.stride(0) if starts.shape[0] > 1 else max(starts.stride(0), starts.shape[1]), (starts, ends)


def stream() -> int:
    return torch.cuda.current_stream().cuda_stream


# optional per-launch observer (bench.py's kernel timing pass): callable(name, args) -> context manager
launch_observer = None
# every entry point counts its calls here (bench.py reports the launches of a step from it)
call_counts = {}


def call(name: str, *args) -> None:
    lib = load()
    call_counts[name] = call_counts.get(name, 0) + 1
    if launch_observer is not None:
        with launch_observer(name, args):
            rc = getattr(lib, name)(*args)
    else:
        rc = getattr(lib, name)(*args)
    if rc != 0:
        raise RuntimeError(f"{name} failed ({rc}): {lib.nkbs_last_error().decode()}")


def require_cuda(*tensors) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("nkbs operators run on CUDA tensors only (no CPU fallback)")


def host_floats(vals):
    arr = (C.c_float * len(vals))(*[float(v) for v in vals])
    return arr


def host_i32(vals):
    return (C.c_int32 * len(vals))(*[int(v) for v in vals])


def host_ptrs(tensors):
    return (C.c_void_p * len(tensors))(*[ptr(t) for t in tensors])
