"""GPU: the tcgen05 tensor-core MLP (nkbs_mlp_tc_fwd/bwd) against a plain torch fp32 evaluation of the same
layers.  Mixed-precision bar of north_star: 1e-2 relative (max-norm); fp16 forward is expected near 1e-3."""
import pytest
import torch

from nerf_kbs_b200 import ops
from tests.util import rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda"

CASES = {
    # name: (dims, relu_mask, out_act, uses selector)
    "proposal": ([10, 16, 1], 0b1, ops.ACT_DENSITY, True),
    "proposal_snw": ([16, 64, 1], 0b1, ops.ACT_DENSITY, True),
    "base": ([32, 64, 16], 0b1, ops.ACT_DENSITY, True),
    "head": ([63, 64, 64, 3], 0b11, ops.ACT_SIGMOID, False),
    "head64": ([64, 64, 64, 3], 0b11, ops.ACT_SIGMOID, False),
    "semantics": ([15, 64, 64, 19], 0b01, ops.ACT_NONE, False),
    "transient": ([31, 64, 64, 5], 0b01, ops.ACT_TRANSIENT, False),
    "linear": ([10, 1], 0, ops.ACT_DENSITY, True),
}


def _q(t, fmt):
    """Round to the kernel's operand format with a straight-through gradient (fmt None: no rounding)."""
    if fmt is None:
        return t
    dt = torch.float16 if fmt == ops.FMT_FP16 else torch.bfloat16
    return t + (t.detach().to(dt).float() - t.detach())


def _torch_mlp(x, Ws, bs, relu_mask, out_act, sel, fmt=None):
    """fp32 evaluation; with ``fmt`` every GEMM operand (inputs, weights, biases, hidden activations) is first
    rounded to that 16-bit format, which is what the tensor-core kernel multiplies (fp32 accumulation)."""
    h = x
    for l, (W, b) in enumerate(zip(Ws, bs)):
        h = torch.nn.functional.linear(_q(h, fmt), _q(W, fmt), _q(b, fmt))
        if (relu_mask >> l) & 1:
            h = torch.relu(h)
    if out_act == ops.ACT_SIGMOID:
        h = torch.sigmoid(h)
    elif out_act == ops.ACT_DENSITY:
        h = torch.cat([torch.exp(h[:, :1]) * sel[:, None], h[:, 1:]], -1)
    elif out_act == ops.ACT_TRANSIENT:
        sp = torch.nn.functional.softplus
        h = torch.cat([sp(h[:, :1]), torch.sigmoid(h[:, 1:4]), sp(h[:, 4:5])], -1)
    return h


def _setup(name, N, seed=0):
    dims, relu_mask, out_act, use_sel = CASES[name]
    g = torch.Generator().manual_seed(seed)
    Ws = [(torch.randn(dims[i + 1], dims[i], generator=g) / dims[i] ** 0.5).to(DEV).requires_grad_() for i in range(len(dims) - 1)]
    bs = [(torch.randn(dims[i + 1], generator=g) * 0.1).to(DEV).requires_grad_() for i in range(len(dims) - 1)]
    x = torch.randn(N, dims[0], generator=g).to(DEV).requires_grad_()
    sel = (torch.rand(N, generator=g) > 0.2).float().to(DEV) if use_sel else None
    return dims, relu_mask, out_act, Ws, bs, x, sel


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("N", [1, 100, 128, 1000, 70001])
@pytest.mark.parametrize("fmt", [ops.FMT_FP16, ops.FMT_BF16])
def test_tc_forward(name, N, fmt):
    ops.set_mlp_precision("tc", fwd_fmt=fmt, bwd_fmt=fmt)
    dims, relu_mask, out_act, Ws, bs, x, sel = _setup(name, N)
    with torch.no_grad():
        got = ops.mlp(x, Ws, bs, relu_mask, out_act, selector=sel)
        want = _torch_mlp(x, Ws, bs, relu_mask, out_act, sel if sel is not None else torch.ones(N, device=DEV))
    assert got.shape == want.shape
    # bf16 operands (8-bit mantissa) under exp(): the error of the raw density is amplified by the output
    # activation, which is why the forward pass defaults to fp16 operands
    tol = 3e-3 if fmt == ops.FMT_FP16 else (4e-2 if out_act == ops.ACT_DENSITY else 1e-2)
    assert rel_err(got, want) <= tol, f"{name} N={N}: {rel_err(got, want):.3e}"


def _l2_err(a, b):
    return ((a - b).norm() / (b.norm() + 1e-20)).item()


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("N", [77, 4096 + 13, 50000])
@pytest.mark.parametrize("fmt", [ops.FMT_FP16, ops.FMT_BF16])
def test_tc_backward(name, N, fmt):
    """Gradients against (a) the same network with operands rounded to the kernel's format (same ReLU masks up
    to fp32 summation order) and (b) the plain fp32 network, where a pre-activation within 16-bit rounding of zero
    takes the other side of the ReLU kink (about 1e-3 of the hidden units in fp16, 1e-2 in bf16).  A flipped unit
    changes one row of dx and one row of dW completely, so the bars are on the relative L2 norm; upstream
    gradients are positive so that the bias sums do not cancel to a badly conditioned value."""
    ops.set_mlp_precision("tc", fwd_fmt=fmt, bwd_fmt=fmt)
    dims, relu_mask, out_act, Ws, bs, x, sel = _setup(name, N, seed=1)
    g = torch.Generator().manual_seed(5)
    dy = (torch.rand(N, dims[-1], generator=g) + 0.25).to(DEV)
    got = ops.mlp(x, Ws, bs, relu_mask, out_act, selector=sel)
    got.backward(dy)
    mine = [x.grad.clone()] + [w.grad.clone() for w in Ws] + [b.grad.clone() for b in bs]
    names = ["dx"] + [f"dW{i}" for i in range(len(Ws))] + [f"db{i}" for i in range(len(bs))]
    ones = torch.ones(N, device=DEV)
    fp16 = fmt == ops.FMT_FP16
    for ref_fmt, metric, tol in ((fmt, _l2_err, 1e-2 if fp16 else 2.5e-2), (None, _l2_err, 3e-2 if fp16 else 1.2e-1)):
        x.grad = None
        for p in Ws + bs:
            p.grad = None
        want = _torch_mlp(x, Ws, bs, relu_mask, out_act, sel if sel is not None else ones, fmt=ref_fmt)
        want.backward(dy)
        ref = [x.grad] + [w.grad for w in Ws] + [b.grad for b in bs]
        for n_, a, b in zip(names, mine, ref):
            assert metric(a, b) <= tol, f"{name} N={N} {n_} vs {'rounded' if ref_fmt is not None else 'fp32'} reference: {metric(a, b):.3e}"


# ---- head MLPs with the input row assembled inside the kernel (nkbs_head_tc_fwd/bwd) ----
@pytest.mark.parametrize("kind,R,S", [("colour", 70, 48), ("colour", 3, 48), ("transient", 70, 48), ("colour", 9, 7)])
def test_head_mlp_fused_input_matches_assemble_route(kind, R, S):
    """Same layers, same samples: input built in the kernel from (directions, geo columns of a [N,16] tensor,
    embedding rows) vs nkbs_assemble + nkbs_mlp_tc.  Both multiply the same 16-bit operands."""
    g = torch.Generator().manual_seed(R * 100 + S)
    N = R * S
    sh_dim, emb_dim, out_dim, act = (16, 32, 3, ops.ACT_SIGMOID) if kind == "colour" else (0, 16, 5, ops.ACT_TRANSIENT)
    relu = 0b11 if kind == "colour" else 0b01
    dims = [sh_dim + 15 + emb_dim, 64, 64, out_dim]
    Ws = [(torch.randn(dims[i + 1], dims[i], generator=g) / dims[i] ** 0.5).to(DEV).requires_grad_() for i in range(3)]
    bs = [(torch.randn(dims[i + 1], generator=g) * 0.1).to(DEV).requires_grad_() for i in range(3)]
    base = torch.randn(N, 16, generator=g).to(DEV).requires_grad_()
    emb = torch.randn(115, emb_dim, generator=g).to(DEV).requires_grad_()
    dirs = torch.nn.functional.normalize(torch.randn(R, 3, generator=g), dim=-1).to(DEV)
    cam = torch.randint(0, 115, (R,), generator=g).to(DEV)
    dy = torch.randn(N, out_dim, generator=g).to(DEV)
    prev = ops.get_mlp_precision()
    ops.set_mlp_precision("tc")
    try:
        res = {}
        for fused in (True, False):
            for t in Ws + bs + [base, emb]:
                t.grad = None
            if fused:
                y = ops.head_mlp(base, 1, emb, None, dirs if sh_dim else None, cam, S, sh_dim, True, True, Ws, bs, relu, act)
            else:
                h = ops.assemble(base[:, 1:], emb, None, dirs if sh_dim else None, cam, S, sh_dim, True, True)
                y = ops.mlp(h, Ws, bs, relu, act)
            y.backward(dy)
            res[fused] = (y.detach().clone(), base.grad.clone(), emb.grad.clone(), [t.grad.clone() for t in Ws + bs])
    finally:
        ops.set_mlp_precision(prev)
    assert_close = lambda a, b, tol, name: (_ for _ in ()).throw(AssertionError(f"{name}: {rel_err(a, b):.3e}")) if rel_err(a, b) > tol else None
    assert_close(res[True][0], res[False][0], 1e-5, "y")
    assert float(res[True][1][:, 0].abs().max()) == 0.0, "column 0 of the base output receives no gradient from the head"
    assert_close(res[True][1], res[False][1], 2e-3, "d base")
    assert_close(res[True][2], res[False][2], 2e-3, "d embedding")
    for i, (a, b) in enumerate(zip(res[True][3], res[False][3])):
        assert_close(a, b, 2e-3, f"param grad {i}")


def test_head_mlp_fused_eval_constant_embedding():
    """Eval: the appearance row is a constant (mean embedding), no embedding gradient path."""
    g = torch.Generator().manual_seed(5)
    R, S = 20, 48
    dims = [63, 64, 64, 3]
    Ws = [(torch.randn(dims[i + 1], dims[i], generator=g) / dims[i] ** 0.5).to(DEV) for i in range(3)]
    bs = [(torch.randn(dims[i + 1], generator=g) * 0.1).to(DEV) for i in range(3)]
    geo = torch.randn(R * S, 15, generator=g).to(DEV)
    const = torch.randn(32, generator=g).to(DEV)
    dirs = torch.nn.functional.normalize(torch.randn(R, 3, generator=g), dim=-1).to(DEV)
    cam = torch.zeros(R, dtype=torch.long, device=DEV)
    prev = ops.get_mlp_precision()
    ops.set_mlp_precision("tc")
    try:
        y = ops.head_mlp(geo, 0, None, const, dirs, cam, S, 16, True, True, Ws, bs, 0b11, ops.ACT_SIGMOID)
        h = ops.assemble(geo, None, const, dirs, cam, S, 16, True, True)
        want = ops.mlp(h, Ws, bs, 0b11, ops.ACT_SIGMOID)
    finally:
        ops.set_mlp_precision(prev)
    assert rel_err(y, want) <= 1e-5


@pytest.mark.parametrize("R,S,with_grad", [(70, 48, True), (70, 48, False), (3, 5, True), (683, 48, False)])
def test_semantics_head_reads_geometry_in_place(R, S, with_grad):
    """Semantics head (15 -> 64 -> 64 -> 19): input rows read from columns 1..15 of the base MLP's [N,16] output inside
    the kernel (no slice copy) vs nkbs_mlp_tc on the contiguous slice; output rows are stored through shared memory."""
    g = torch.Generator().manual_seed(R + S)
    N = R * S
    dims = [15, 64, 64, 19]
    Ws = [(torch.randn(dims[i + 1], dims[i], generator=g) / dims[i] ** 0.5).to(DEV).requires_grad_() for i in range(3)]
    bs = [(torch.randn(dims[i + 1], generator=g) * 0.1).to(DEV).requires_grad_() for i in range(3)]
    base = torch.randn(N, 16, generator=g).to(DEV).requires_grad_(with_grad)
    dy = torch.randn(N, 19, generator=g).to(DEV)
    prev = ops.get_mlp_precision()
    ops.set_mlp_precision("tc")
    try:
        res = {}
        for fused in (True, False):
            for t in Ws + bs + [base]:
                t.grad = None
            if fused:
                y = ops.head_mlp(base, 1, None, None, None, None, S, 0, False, False, Ws, bs, 0b01, ops.ACT_NONE)
            else:
                y = ops.mlp(base[:, 1:], Ws, bs, 0b01, ops.ACT_NONE)
            y.backward(dy)
            res[fused] = (y.detach().clone(), base.grad.clone() if with_grad else None, [t.grad.clone() for t in Ws + bs])
    finally:
        ops.set_mlp_precision(prev)
    assert rel_err(res[True][0], res[False][0]) <= 1e-5, "y"
    if with_grad:
        assert float(res[True][1][:, 0].abs().max()) == 0.0
        assert rel_err(res[True][1], res[False][1]) <= 2e-3, "d base"
    for i, (a, b) in enumerate(zip(res[True][2], res[False][2])):
        assert rel_err(a, b) <= 2e-3, f"param grad {i}"
