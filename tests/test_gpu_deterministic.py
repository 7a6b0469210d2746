"""GPU: the deterministic training mode (include/nkbs.h: nkbs_fixed_point_register / nkbs_fixed_point_flush).

Parameter gradients are accumulated in 64-bit fixed point, so they do not depend on the order in which warps, CTAs and
streams arrive.  What that buys, checked here bit for bit (``torch.equal``):
  * the hash-table gradient of a batch does not change when the samples are permuted;
  * two runs of the same training steps end with the same parameters — also when one run is eager and the other a
    CUDA-graph replay, and when one applies the field group's Adam at the end of the step and the other beside the next
    step's proposal levels (the pairs that tests/test_gpu_trainer.py can only hold to 2e-2 / 5e-2 with float atomics:
    Adam with eps = 1e-15 turns rounding noise on a cancelling gradient into a step of +-lr)."""
import ctypes as C
import os

import pytest
import torch

from oracle import nerfstudio_torch as O
import nerf_kbs_b200 as nk
from nerf_kbs_b200 import ops
from nerf_kbs_b200._lib import call, host_floats, ptr, stream
from nerf_kbs_b200.trainer import Trainer
from tests.util import copy_model, make_jitter, to_nk_bundle

pytestmark = pytest.mark.gpu
DEV = "cuda"


def test_fixed_point_table_gradient_does_not_depend_on_the_sample_order():
    """nkbs_encode_bwd into a registered buffer: any permutation of the samples gives the same bits (run merging off:
    it sums neighbouring lanes in float first, and a permutation changes the neighbours), and the result agrees with
    the float-atomic scatter to fp32 rounding."""
    grid = nk.HashEncoding(num_levels=8, min_res=16, max_res=512, log2_hashmap_size=15).grid
    N, L = 20000, grid.L
    g = torch.Generator().manual_seed(3)
    xn = torch.rand(N, 3, generator=g).to(DEV)
    dfeat = (torch.randn(N, 2 * L, generator=g) * 1e-3).to(DEV)
    shape = (L * (1 << grid.log2_T), 2)
    ref = torch.zeros(shape, device=DEV)
    call("nkbs_encode_bwd", ptr(xn), ptr(dfeat), N, host_floats(grid.scalings), L, grid.log2_T, 2, 0, ptr(ref), stream())
    G = torch.zeros(shape, device=DEV)
    F = torch.zeros(G.numel(), dtype=torch.int64, device=DEV)
    call("nkbs_fixed_point_register", G.data_ptr(), G.numel(), F.data_ptr(), 56)
    try:
        outs = []
        for seed in (None, 1, 2):
            if seed is None:
                x, d = xn, dfeat
            else:
                perm = torch.randperm(N, generator=torch.Generator().manual_seed(seed)).to(DEV)
                x, d = xn[perm].contiguous(), dfeat[perm].contiguous()
            call("nkbs_encode_bwd", ptr(x), ptr(d), N, host_floats(grid.scalings), L, grid.log2_T, 2, 0, ptr(G), stream())
            assert not G.any(), "registered buffer: the kernel must not touch the float gradients"
            call("nkbs_fixed_point_flush", G.data_ptr(), G.numel(), stream())
            assert not F.any(), "the flush re-zeroes the shadow"
            outs.append(G.clone())
            G.zero_()
        assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
        assert (outs[0] - ref).abs().max().item() <= 1e-6 * ref.abs().max().item()
        # a sub-range flush outside the registered buffer is refused
        other = torch.zeros(16, device=DEV)
        with pytest.raises(RuntimeError):
            call("nkbs_fixed_point_flush", other.data_ptr(), 16, stream())
    finally:
        call("nkbs_fixed_point_register", G.data_ptr(), G.numel(), None, 0)
    # unregistered again: float reductions straight into the buffer
    call("nkbs_encode_bwd", ptr(xn), ptr(dfeat), N, host_floats(grid.scalings), L, grid.log2_T, 2, 0, ptr(G), stream())
    assert G.any()


def _model(kind):
    torch.manual_seed(0)
    cfg = dict() if kind == "nerfacto" else dict(use_semantic=True, use_transient_embedding=True)
    om = (O.NerfactoModel if kind == "nerfacto" else O.SemanticNerfWModel)(O.ModelConfig(**cfg))
    nm = (nk.NerfactoModel if kind == "nerfacto" else nk.SemanticNerfWModel)(nk.ModelConfig(**cfg))
    with torch.no_grad():
        for p in om.parameters():
            if p.dim() == 2 and p.shape[1] == 2:
                p.mul_(300.0)
    nm = nm.to(DEV)
    copy_model(nm, om)
    return nm.train()


def _run(kind, R, steps, use_graph=True, defer="1", deterministic=True, precision="tc"):
    ops.set_mlp_precision(precision)
    os.environ["NKBS_DEFER_ADAM"] = defer
    try:
        nm = _model(kind)
        tr = Trainer(nm, R, use_graph=use_graph, graph_warmup=1, inject_jitter=True, deterministic=deterministic)
    finally:
        os.environ.pop("NKBS_DEFER_ADAM", None)
    assert tr.deterministic == deterministic and (tr.arena.F is not None) == deterministic
    losses = []
    for step in range(steps):
        rb = O.synthetic_ray_bundle(R, seed=20 + step)
        g = torch.Generator().manual_seed(40 + step)
        batch = {"image": torch.rand(R, 3, generator=g).to(DEV)}
        if kind != "nerfacto":
            batch["semantics_label"] = torch.randint(0, 19, (R,), generator=g).to(DEV)
        losses.append(tr.train_step(to_nk_bundle(rb, DEV), batch, jitter=[j.to(DEV) for j in make_jitter(R, seed=60 + step)]).item())
    tr.flush()
    params = {n: p.detach().clone() for n, p in nm.named_parameters()}
    tr.close()
    return losses, params


def _assert_same(a, b, what):
    for n in a[1]:
        assert torch.equal(a[1][n], b[1][n]), f"{what}: parameter {n} differs (max {(a[1][n] - b[1][n]).abs().max().item():.3e})"
    for x, y in zip(a[0], b[0]):
        assert abs(x - y) <= 1e-6 * abs(x), (what, a[0], b[0])  # the reported loss sums are float atomics


@pytest.mark.parametrize("precision", ["tc", "fp32"])
def test_two_deterministic_runs_end_with_the_same_bits(precision):
    """nerfacto, 2048 rays, 13 steps (crosses step 10, where the proposal gating starts skipping updates)."""
    a = _run("nerfacto", 2048, 13, precision=precision)
    b = _run("nerfacto", 2048, 13, precision=precision)
    _assert_same(a, b, "run 1 vs run 2")


def test_deterministic_eager_steps_equal_graph_replays_and_deferred_adam_bit_for_bit():
    """Schedule changes only (eager launches vs graph replay, Adam at the step's end vs beside the next step's
    proposal levels on a side stream): with order-independent sums the parameters are identical."""
    base = _run("nerfacto", 1024, 8, use_graph=True, defer="1")
    _assert_same(base, _run("nerfacto", 1024, 8, use_graph=False, defer="1"), "graph vs eager")
    _assert_same(base, _run("nerfacto", 1024, 8, use_graph=True, defer="0"), "deferred vs end-of-step Adam")


def test_deterministic_semantic_nerfw_runs_end_with_the_same_bits():
    """semantic-NeRF-W (hidden-64 proposal fields through encode + tensor-core MLP, transient and semantic heads, the
    appearance-embedding gradient)."""
    a = _run("semantic_nerfw", 2048, 4)
    b = _run("semantic_nerfw", 2048, 4)
    _assert_same(a, b, "run 1 vs run 2")


def test_deterministic_mode_tracks_the_default_mode():
    """Same gradients up to fp32 rounding: one step from the same state moves the parameters the same way.  The first
    Adam step is -lr * g / (|g| + 1e-15): entries whose gradient is a cancelling sum (rounding noise of the float
    atomics' order in the default mode — two default runs differ on them as well) may land anywhere in +-lr."""
    a = _run("nerfacto", 1024, 1, deterministic=True)
    b = _run("nerfacto", 1024, 1, deterministic=False)
    assert abs(a[0][0] - b[0][0]) <= 1e-6 * abs(a[0][0])
    for n in a[1]:
        d = (a[1][n] - b[1][n]).abs()
        assert d.max().item() <= 2.1e-3, n
        assert (d > 1e-6).float().mean().item() < 0.1, n
