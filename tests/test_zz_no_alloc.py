"""GPU, collected last: the allocation-free expected-depth entry point (nkbs_expected_depth_noalloc: its scratch is a
device array built into the library) against the existing entry point with caller-owned scratch (nkbs_composite_fwd),
both through the C ABI.  Same kernels, same arithmetic: the outputs must be bit-identical, on two calls in a row that
share the built-in scratch (the second smaller, so stale contents of the first would show)."""
import pytest
import torch

from nerf_kbs_b200 import _lib
from nerf_kbs_b200._lib import call, ptr, stream

pytestmark = pytest.mark.gpu
DEV = "cuda"
MAX_RAYS = 1 << 20  # NKBS_NOALLOC_MAX_RAYS (include/nkbs.h)


def _inputs(R, S, seed):
    g = torch.Generator().manual_seed(seed)
    bins = torch.sort(torch.rand(R, S + 1, generator=g) * (6.0 + seed), dim=-1).values.to(DEV)
    w = torch.rand(R, S, generator=g)
    w = (w / w.sum(-1, keepdim=True) * torch.rand(R, 1, generator=g)).to(DEV)
    w[::7] = 0.0  # empty rays: depth 0 / 1e-10, clipped up to the batch minimum
    return w.contiguous(), bins


def _reference(w, bins):
    """nkbs_composite_fwd with the caller's scratch (depth_exp_raw, depth_mask, minmax_ws)."""
    R, S = w.shape
    f32 = dict(device=DEV, dtype=torch.float32)
    acc, raw, out, mask = (torch.empty(R, **f32) for _ in range(4))
    mm = torch.empty(2, device=DEV, dtype=torch.int32)
    st, en = bins[:, :-1], bins[:, 1:]
    call("nkbs_composite_fwd", R, S, None, ptr(w), st.data_ptr(), en.data_ptr(), S + 1, None, None, 0, 0, 0, None, None,
         ptr(acc), ptr(raw), ptr(out), ptr(mask), None, None, ptr(mm), stream())
    return out


def _noalloc(w, bins, out):
    R, S = w.shape
    call("nkbs_expected_depth_noalloc", R, S, ptr(w), bins.data_ptr(), bins.data_ptr() + 4, S + 1, ptr(out), stream())
    return out


def test_noalloc_expected_depth_equals_composite_entry_point_twice():
    cases = [(5000, 48, 1), (777, 96, 2)]
    data = [_inputs(*c) for c in cases]
    want = [_reference(w, b) for w, b in data]
    outs = [torch.full((w.shape[0],), float("nan"), device=DEV) for w, _ in data]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    for (w, b), o in zip(data, outs):   # two calls in a row on one stream, sharing the built-in scratch
        _noalloc(w, b, o)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before  # nothing was allocated through torch either
    for o, ref in zip(outs, want):
        assert torch.equal(o, ref)
    # and once more in the other order
    for (w, b), o, ref in reversed(list(zip(data, outs, want))):
        o.fill_(float("nan"))
        assert torch.equal(_noalloc(w, b, o), ref)


def test_noalloc_rejects_batches_beyond_the_builtin_scratch():
    w, bins = _inputs(8, 4, 3)
    out = torch.empty(8, device=DEV)
    with pytest.raises(RuntimeError, match="exceed the built-in scratch"):
        call("nkbs_expected_depth_noalloc", MAX_RAYS + 1, 4, ptr(w), bins.data_ptr(), bins.data_ptr() + 4, 5, ptr(out), stream())
    assert "nkbs_expected_depth_noalloc" in _lib.SIGNATURES


def test_noalloc_is_capturable_in_a_cuda_graph():
    """The use case: no allocation, so the call can sit inside a stream capture and be replayed."""
    w, bins = _inputs(1024, 48, 4)
    want = _reference(w, bins)
    out = torch.zeros(1024, device=DEV)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        _noalloc(w, bins, out)  # warm-up: the symbol address is looked up outside the capture
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        _noalloc(w, bins, out)
    out.zero_()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, want)
