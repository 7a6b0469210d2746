"""GPU parity of each nkbs kernel (through the C ABI) against the CPU oracle on the same seeded inputs.
Tolerances follow BASELINE.json north_star: hash indices and sample bins bit-exact, fp32 outputs 1e-4."""
import pytest
import torch

from oracle import nerfstudio_torch as O
import nerf_kbs_b200 as nk
from nerf_kbs_b200 import ops
from tests.util import assert_close, rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda"
FP32_TOL = 1e-4


def _oracle_enc(L=16, max_res=2048, log2=19, seed=0):
    torch.manual_seed(seed)
    return O.HashEncoding(num_levels=L, min_res=16, max_res=max_res, log2_hashmap_size=log2)


@pytest.mark.parametrize("L,max_res,log2", [(16, 2048, 19), (5, 128, 17), (5, 256, 17), (8, 1024, 18)])
def test_hash_indices_bit_exact_and_features(L, max_res, log2):
    enc = _oracle_enc(L, max_res, log2)
    g = torch.Generator().manual_seed(3)
    x = torch.rand(5000, 3, generator=g)
    # exact grid points, zeros (the selector-zeroed case) and the upper edge
    x[:50] = torch.randint(0, 17, (50, 3), generator=g).float() / 16.0
    x[50:60] = 0.0
    x[60:70] = 1.0 - 2**-24
    hs, _ = enc.corner_indices(x)
    want_idx = torch.stack(hs, dim=-1)  # [N, L, 8]
    want = enc(x)
    grid = ops.GridSpec(tuple(enc.scalings.tolist()), log2)
    feats, sel, xn, idx = ops.encode_raw(0, grid, None, enc.hash_table.detach().to(DEV), x=x.to(DEV), want_indices=True)
    assert torch.equal(idx.cpu(), want_idx), "hash indices differ from the oracle"
    assert torch.equal(xn.cpu(), x)
    assert_close(feats, want, 1e-6, "features")


def test_hash_known_answers():
    """Hand-derivable cases: x = 0 hits row l*T for every corner; an exact grid point has zero offset."""
    enc = _oracle_enc(4, 128, 10)
    grid = ops.GridSpec(tuple(enc.scalings.tolist()), 10)
    x = torch.zeros(1, 3)
    _, _, _, idx = ops.encode_raw(0, grid, None, enc.hash_table.detach().to(DEV), x=x.to(DEV), want_indices=True)
    want = (torch.arange(4) * 1024)[None, :, None].expand(1, 4, 8)
    assert torch.equal(idx.cpu(), want)
    # level 0 scale 16: x = (3,5,7)/16 -> ceil == floor == (3,5,7): all 8 corners hash alike
    x = torch.tensor([[3.0, 5.0, 7.0]]) / 16.0
    feats, _, _, idx = ops.encode_raw(0, grid, None, enc.hash_table.detach().to(DEV), x=x.to(DEV), want_indices=True)
    h = (3 ^ (5 * 2654435761) ^ (7 * 805459861)) % 1024
    assert torch.equal(idx[0, 0].cpu(), torch.full((8,), h))
    # with zero offset the blend collapses to corner 6 (f,f,f) == the single row
    assert torch.allclose(feats[0, :2].cpu(), enc.hash_table[h].detach(), atol=0, rtol=0)


def test_encode_backward_matches_autograd():
    enc = _oracle_enc(16, 2048, 19)
    g = torch.Generator().manual_seed(5)
    # points clustered along a few rays so that coarse levels collide inside a warp
    base = torch.rand(64, 1, 3, generator=g)
    x = (base + 0.002 * torch.arange(48)[None, :, None] * torch.rand(64, 1, 3, generator=g)).clamp(0, 0.999).reshape(-1, 3)
    dy = torch.randn(x.shape[0], 32, generator=g)
    enc(x).backward(dy)
    want = enc.hash_table.grad
    table = enc.hash_table.detach().to(DEV).requires_grad_(True)
    grid = ops.GridSpec(tuple(enc.scalings.tolist()), 19)
    for dedup in (-1, 0, 6):
        table.grad = None
        feats, _, _ = ops._HashEncode.apply(table, 0, grid, None, dedup, x.to(DEV), None, None, None, None)
        feats.backward(dy.to(DEV))
        assert_close(table.grad, want, 1e-5, f"dtable dedup={dedup}")
        nz_want = (want.abs().sum(-1) > 0)
        nz_got = (table.grad.abs().sum(-1) > 0).cpu()
        assert (nz_got & ~nz_want).sum() == 0, "gradient written to rows the oracle never touched"


@pytest.mark.parametrize("contraction", [True, False])
def test_front_end_positions_bit_exact(contraction):
    torch.manual_seed(0)
    aabb = torch.tensor([[-1.0, -1.5, -2.0], [1.0, 2.5, 2.0]])
    sc = O.SceneContraction() if contraction else None
    of = O.HashMLPDensityField(aabb, spatial_distortion=sc, hidden_dim=16, num_levels=5, max_res=128, log2_hashmap_size=17)
    rb = O.synthetic_ray_bundle(300, seed=2)
    rb = O.NearFarCollider(0.001, 1000.0)(rb)
    samp = O.UniformLinDispPiecewiseSampler(single_jitter=True)
    samp.train()
    rs = samp(rb, num_samples=64)
    pos = rs.frustums.get_positions()
    want_xn, want_sel = of._normalize(pos)
    hs, _ = of.encoding.corner_indices(want_xn.view(-1, 3))
    want_idx = torch.stack(hs, dim=-1)
    norm = ops.NormSpec(contraction, tuple(aabb.reshape(-1).tolist()))
    grid = ops.GridSpec(tuple(of.encoding.scalings.tolist()), 17)
    table = of.encoding.hash_table.detach().to(DEV)
    # mode 2: per-ray frustum inputs
    f2, sel2, xn2, idx2 = ops.encode_raw(2, grid, norm, table, origins=rb.origins.to(DEV), directions=rb.directions.to(DEV),
                                         starts=rs.frustums.starts[..., 0].contiguous().to(DEV),
                                         ends=rs.frustums.ends[..., 0].contiguous().to(DEV), want_indices=True)
    assert torch.equal(xn2.cpu(), want_xn.reshape(-1, 3)), "normalised positions differ (frustum mode)"
    assert torch.equal(sel2.cpu() > 0, want_sel.reshape(-1))
    assert torch.equal(idx2.cpu(), want_idx)
    # mode 1: materialised positions
    f1, sel1, xn1, idx1 = ops.encode_raw(1, grid, norm, table, x=pos.reshape(-1, 3).contiguous().to(DEV), want_indices=True)
    assert torch.equal(xn1.cpu(), want_xn.reshape(-1, 3))
    assert torch.equal(idx1.cpu(), want_idx)
    assert torch.equal(f1, f2)


MLP_CASES = [
    # dims, relu_mask, act
    ([10, 16, 1], 0b01, ops.ACT_DENSITY),
    ([16, 64, 1], 0b01, ops.ACT_DENSITY),
    ([32, 64, 16], 0b01, ops.ACT_DENSITY),
    ([63, 64, 64, 3], 0b011, ops.ACT_SIGMOID),
    ([15, 64, 64, 19], 0b001, ops.ACT_NONE),
    ([31, 64, 64, 5], 0b001, ops.ACT_TRANSIENT),
    ([10, 1], 0, ops.ACT_DENSITY),
    ([7, 33, 20], 0b01, ops.ACT_NONE),
]


def _torch_mlp(x, Ws, bs, relu_mask, act, sel):
    h = x
    for i, (w, b) in enumerate(zip(Ws, bs)):
        h = torch.nn.functional.linear(h, w, b)
        if (relu_mask >> i) & 1:
            h = torch.relu(h)
    if act == ops.ACT_SIGMOID:
        h = torch.sigmoid(h)
    elif act == ops.ACT_DENSITY:
        d = O.trunc_exp(h[:, :1]) * sel[:, None]
        h = torch.cat([d, h[:, 1:]], dim=-1)
    elif act == ops.ACT_TRANSIENT:
        sp = torch.nn.functional.softplus
        h = torch.cat([sp(h[:, 0:1]), torch.sigmoid(h[:, 1:4]), sp(h[:, 4:5])], dim=-1)
    return h


@pytest.mark.parametrize("dims,relu_mask,act", MLP_CASES)
@pytest.mark.parametrize("N", [1, 127, 1000, 40001])
def test_mlp_forward_backward(dims, relu_mask, act, N):
    g = torch.Generator().manual_seed(N + len(dims))
    Ws = [(torch.randn(dims[i + 1], dims[i], generator=g) / dims[i] ** 0.5).requires_grad_(True) for i in range(len(dims) - 1)]
    bs = [(0.1 * torch.randn(dims[i + 1], generator=g)).requires_grad_(True) for i in range(len(dims) - 1)]
    x = torch.randn(N, dims[0], generator=g).requires_grad_(True)
    sel = (torch.rand(N, generator=g) > 0.2).float()
    dy = torch.randn(N, dims[-1], generator=g)
    want = _torch_mlp(x, Ws, bs, relu_mask, act, sel)
    want.backward(dy)
    Wg = [w.detach().to(DEV).requires_grad_(True) for w in Ws]
    bg = [b.detach().to(DEV).requires_grad_(True) for b in bs]
    xg = x.detach().to(DEV).requires_grad_(True)
    got = ops.mlp(xg, Wg, bg, relu_mask, act, selector=sel.to(DEV) if act == ops.ACT_DENSITY else None)
    assert_close(got, want, FP32_TOL, "mlp forward")
    got.backward(dy.to(DEV))
    assert_close(xg.grad, x.grad, FP32_TOL, "dx")
    for i in range(len(Ws)):
        assert_close(Wg[i].grad, Ws[i].grad, FP32_TOL, f"dW{i}")
        assert_close(bg[i].grad, bs[i].grad, FP32_TOL, f"db{i}")


def test_mlp_rejects_unsupported():
    x = torch.randn(8, 100, device=DEV)
    with pytest.raises(RuntimeError, match="width"):
        ops.mlp(x, [torch.randn(4, 100, device=DEV)], [torch.randn(4, device=DEV)], 0)
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.mlp(torch.randn(8, 4), [torch.randn(4, 4)], [torch.randn(4)], 0)


@pytest.mark.parametrize("S", [48, 96, 256, 33])
@pytest.mark.parametrize("single", [True, False])
def test_initial_sampler_bins_bit_exact(S, single):
    rb = O.NearFarCollider(0.001, 1000.0)(O.synthetic_ray_bundle(777, seed=4))
    rb.nears[:5] = 0.5
    rb.fars[:5] = 0.9     # a bundle entirely inside the linear part
    rb.nears[5:9] = 2.0   # and entirely in the disparity part
    samp = O.UniformLinDispPiecewiseSampler(single_jitter=single)
    for train in (True, False):
        samp.train(train)
        g = torch.Generator().manual_seed(S)
        t_rand = torch.rand(777, 1 if single else S + 1, generator=g)
        rs = samp(rb, num_samples=S, t_rand=t_rand)
        want_s = torch.cat([rs.spacing_starts[..., 0], rs.spacing_ends[..., -1:, 0]], -1)
        want_e = torch.cat([rs.frustums.starts[..., 0], rs.frustums.ends[..., -1:, 0]], -1)
        sb, eb = ops.initial_sampler(rb.nears.to(DEV), rb.fars.to(DEV), S, t_rand.to(DEV) if train else None)
        assert torch.equal(sb.cpu(), want_s), "spacing bins differ"
        assert torch.equal(eb.cpu(), want_e), "euclidean bins differ"
