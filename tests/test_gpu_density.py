"""GPU parity of the fused proposal density field (nkbs_density_fwd/bwd, csrc/density.cu) against the CPU oracle's
HashMLPDensityField and against the unfused nkbs route (encode + fp32 MLP) on the same seeded inputs."""
import pytest
import torch

from oracle import nerfstudio_torch as O
import nerf_kbs_b200 as nk
from nerf_kbs_b200 import ops
from tests.util import assert_close, copy_density_field, density_grad_pairs, rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda"

SHAPES = {
    "nerfacto_prop0": dict(hidden_dim=16, log2_hashmap_size=17, num_levels=5, max_res=128),
    "nerfacto_prop1": dict(hidden_dim=16, log2_hashmap_size=17, num_levels=5, max_res=256),
    "snw_default": dict(hidden_dim=64, log2_hashmap_size=18, num_levels=8, max_res=1024),
}


def _pair(kind, seed=0):
    torch.manual_seed(seed)
    aabb = torch.tensor([[-1.0, -1.0, -1.0], [1.0, 1.0, 1.0]])
    o = O.HashMLPDensityField(aabb, spatial_distortion=O.SceneContraction(order=float("inf")), **SHAPES[kind])
    with torch.no_grad():
        o.encoding.hash_table.mul_(300.0)  # features of order 0.3 instead of 1e-3: the MLP sees real inputs
    n = nk.HashMLPDensityField(aabb, spatial_distortion=nk.SceneContraction(order=float("inf")), **SHAPES[kind]).to(DEV)
    copy_density_field(n, o)
    return o, n


def _samples(R, S, seed=1):
    rb = O.synthetic_ray_bundle(R, seed=seed)
    rb = O.NearFarCollider(0.001, 1000.0)(rb)
    rs = O.UniformLinDispPiecewiseSampler(single_jitter=True)(rb, num_samples=S, t_rand=torch.rand(R, 1, generator=torch.Generator().manual_seed(seed)))
    return rb, rs


@pytest.mark.parametrize("kind", list(SHAPES))
@pytest.mark.parametrize("R,S", [(64, 256), (37, 96), (1, 1), (300, 7)])
def test_fused_density_matches_oracle(kind, R, S):
    o, n = _pair(kind)
    rb, rs_o = _samples(R, S)
    want, _ = o.get_density(rs_o)
    dy = torch.rand(want.shape, generator=torch.Generator().manual_seed(9))
    want.backward(dy)
    fr = rs_o.frustums
    with torch.no_grad():
        assert ops.fused_density_enabled() and n._fused_ok()
    W, b = n.mlp_base[1].weights_biases()
    got = ops.fused_density(n.encoding.hash_table, n.encoding.grid, n._norm, W[0], b[0], W[1], b[1],
                            origins=rb.origins.to(DEV), directions=rb.directions.to(DEV),
                            starts=fr.starts[..., 0].to(DEV), ends=fr.ends[..., 0].to(DEV))
    assert_close(got.reshape(want.shape), want, 1e-4, "density")
    got.backward(dy.reshape(-1).to(DEV))
    # gradients pass through exp(raw): fp32 summation-order noise of the H = 64 field reaches 7e-4 against the CPU
    # order (the unfused nkbs route agrees with the fused one to 1e-4, next test)
    for name, a, bb in density_grad_pairs(n, o):
        assert_close(a, bb, 1e-3, name)


@pytest.mark.parametrize("kind", ["nerfacto_prop0", "snw_default"])
def test_fused_density_equals_unfused_route(kind):
    """Same field, same samples: one kernel vs encode + fp32 MLP kernels; positions mode (density_fn) too."""
    _, n = _pair(kind, seed=3)
    g = torch.Generator().manual_seed(4)
    pos = ((torch.rand(5000, 3, generator=g) - 0.5) * 6.0).to(DEV)  # inside and outside the unit box (contraction)
    dy = torch.rand(5000, 1, generator=g).to(DEV)
    prev = ops.get_mlp_precision()
    prev_train = ops.FUSED_DENSITY_TRAIN_SHAPES
    res = {}
    try:
        ops.set_mlp_precision("fp32")
        ops.FUSED_DENSITY_TRAIN_SHAPES = ops.FUSED_DENSITY_SHAPES  # exercise the fused backward of every shape
        for fused in (True, False):
            ops.set_fused_density(fused)
            for p in n.parameters():
                p.grad = None
            d = n.density_fn(pos)
            d.backward(dy)
            res[fused] = (d.detach().clone(), [p.grad.clone() for p in n.parameters()])
    finally:
        ops.set_fused_density(True)
        ops.set_mlp_precision(prev)
        ops.FUSED_DENSITY_TRAIN_SHAPES = prev_train
    assert_close(res[True][0], res[False][0], 1e-5, "density")
    for a, b in zip(res[True][1], res[False][1]):
        assert_close(a, b, 1e-4, "grad")


def test_fused_density_outside_samples_and_zero_gradient():
    """Samples the selector rejects contribute zero density and no gradient; an all-zero upstream gradient
    leaves every gradient buffer at zero."""
    _, n = _pair("nerfacto_prop0")
    pos = torch.tensor([[0.0, 0.0, 0.0], [2.0, 0.0, 0.0], [1e6, 1e6, 1e6], [-3.0, 5.0, 0.1]]).to(DEV)
    d = n.density_fn(pos)
    assert d.shape == (4, 1)
    assert torch.isfinite(d).all()
    d.backward(torch.zeros_like(d))
    for p in n.parameters():
        assert p.grad is not None and float(p.grad.abs().max()) == 0.0


def test_fused_density_rejects_other_shapes():
    t = torch.zeros(4 * 1024, 2, device=DEV)
    grid = ops.GridSpec((16.0, 32.0, 64.0, 128.0), 10)
    norm = ops.NormSpec(True, (-1.0,) * 3 + (1.0,) * 3)
    with pytest.raises(ValueError):
        ops.fused_density(t, grid, norm, torch.zeros(16, 8, device=DEV), torch.zeros(16, device=DEV),
                          torch.zeros(1, 16, device=DEV), torch.zeros(1, device=DEV), positions=torch.zeros(3, 3, device=DEV))


def test_training_route_of_the_hidden64_field():
    """The semantic-NeRF-W default proposal field trains through the fused forward (nkbs_density_fwd_save: features,
    normalised positions and selectors are left for the backward) and the MLP + encode backward kernels (see
    ops.HYBRID_DENSITY); density and every gradient agree with the encode + MLP route in both directions (ray-sample inputs: test_fused_density_matches_oracle[snw_default])."""
    from nerf_kbs_b200 import _lib
    _, n = _pair("snw_default", seed=5)
    assert n._fused_ok()
    g = torch.Generator().manual_seed(6)
    pos = ((torch.rand(3000, 3, generator=g) - 0.5) * 4.0).to(DEV)
    dy = torch.rand(3000, 1, generator=g).to(DEV)
    prev = ops.get_mlp_precision()
    res = {}
    try:
        for precision, tol in (("fp32", 1e-4), ("tc", 3e-2)):
            ops.set_mlp_precision(precision)
            for hybrid in (True, False):
                ops.HYBRID_DENSITY = hybrid
                assert n._fused_ok() == hybrid
                for p in n.parameters():
                    p.grad = None
                c0 = dict(_lib.call_counts)
                d = n.density_fn(pos)
                d.backward(dy)
                used = {k for k, v in _lib.call_counts.items() if v != c0.get(k, 0)}
                assert ("nkbs_density_fwd_save" in used) == hybrid and ("nkbs_encode_fwd" in used) == (not hybrid)
                assert "nkbs_density_bwd" not in used and "nkbs_encode_bwd" in used
                res[hybrid] = (d.detach().clone(), [p.grad.clone() for p in n.parameters()])
            assert_close(res[True][0], res[False][0], 1e-5 if precision == "fp32" else 1e-2, "density")
            for a, b in zip(res[True][1], res[False][1]):
                assert_close(a, b, tol, f"grad ({precision})")
        with torch.no_grad():
            d_eval = n.density_fn(pos)
        assert_close(res[True][0], d_eval, 1e-5 if precision == "fp32" else 1e-2, "training forward vs evaluation forward")
    finally:
        ops.HYBRID_DENSITY = True
        ops.set_mlp_precision(prev)


@pytest.mark.parametrize("kind", ["nerfacto_prop0", "nerfacto_prop1"])
@pytest.mark.parametrize("R,S", [(200, 256), (333, 96), (50, 48), (7, 33)])
def test_ray_major_scatter_equals_sample_major_scatter(kind, R, S):
    """Backward of the fused field on ray samples: feature gradients + nkbs_encode_bwd_rays (default) against the
    scatter inside nkbs_density_bwd — same sums, different association."""
    if kind not in SHAPES:
        pytest.skip("shape not in this file's table")
    _, n = _pair(kind, seed=7)
    rb, rs_o = _samples(R, S)
    fr = rs_o.frustums
    W, b = n.mlp_base[1].weights_biases()
    dy = torch.rand(R * S, generator=torch.Generator().manual_seed(8)).to(DEV)
    res = {}
    try:
        for ray in (False, True):
            ops.set_ray_scatter(ray)
            for p in n.parameters():
                p.grad = None
            d = ops.fused_density(n.encoding.hash_table, n.encoding.grid, n._norm, W[0], b[0], W[1], b[1],
                                  origins=rb.origins.to(DEV), directions=rb.directions.to(DEV),
                                  starts=fr.starts[..., 0].to(DEV), ends=fr.ends[..., 0].to(DEV))
            d.backward(dy)
            res[ray] = [p.grad.clone() for p in n.parameters()]
    finally:
        ops.set_ray_scatter(True)
    for a, bb in zip(res[True], res[False]):
        assert_close(a, bb, 2e-5, "gradient, ray-major vs sample-major scatter")


@pytest.mark.parametrize("kind", ["nerfacto_prop0", "nerfacto_prop1"])
@pytest.mark.parametrize("R,S", [(200, 256), (333, 96), (7, 33)])
def test_backward_from_saved_features_equals_recomputed_backward(kind, R, S):
    """Default training route of the nerfacto proposal fields on ray samples: the forward stores features and selectors
    (nkbs_density_fwd_save) and the backward reads them (nkbs_density_bwd_saved) — against the backward that recomputes
    positions and gathers (NKBS_SAVED_DENSITY_BWD=0).  Same arithmetic on the same values: identical feature
    gradients, hence identical densities and (up to the order of the float reductions) identical gradients."""
    from nerf_kbs_b200 import _lib
    _, n = _pair(kind, seed=11)
    rb, rs_o = _samples(R, S)
    fr = rs_o.frustums
    W, b = n.mlp_base[1].weights_biases()
    g = torch.Generator().manual_seed(12)
    dy = (torch.rand(R * S, generator=g) * (torch.rand(R * S, generator=g) < 0.7)).to(DEV)  # some exact zeros
    res = {}
    prev = ops.SAVED_DENSITY_BWD
    try:
        for saved in (False, True):
            ops.SAVED_DENSITY_BWD = saved
            for p in n.parameters():
                p.grad = None
            c0 = dict(_lib.call_counts)
            d = ops.fused_density(n.encoding.hash_table, n.encoding.grid, n._norm, W[0], b[0], W[1], b[1],
                                  origins=rb.origins.to(DEV), directions=rb.directions.to(DEV),
                                  starts=fr.starts[..., 0].to(DEV), ends=fr.ends[..., 0].to(DEV))
            d.backward(dy)
            used = {k for k, v in _lib.call_counts.items() if v != c0.get(k, 0)}
            assert ("nkbs_density_bwd_saved" in used) == saved and ("nkbs_density_fwd_save" in used) == saved
            res[saved] = (d.detach().clone(), [p.grad.clone() for p in n.parameters()])
    finally:
        ops.SAVED_DENSITY_BWD = prev
    assert torch.equal(res[True][0], res[False][0])
    for a, bb in zip(res[True][1], res[False][1]):
        assert_close(a, bb, 2e-5, "gradient, saved features vs recomputed")


@pytest.mark.parametrize("S", [96, 256, 64])
def test_unfused_encode_backward_ray_major_scatter(S):
    """contract_encode on proposal-sized ray samples: the backward takes nkbs_encode_bwd_rays; same table gradient as
    the sample-major nkbs_encode_bwd."""
    _, n = _pair("snw_default", seed=9)
    R = 150
    rb, rs_o = _samples(R, S)
    fr = rs_o.frustums
    g = torch.Generator().manual_seed(10)
    dy = torch.randn(R * S, n.encoding.grid.out_dim, generator=g).to(DEV)
    res = {}
    try:
        for ray in (False, True):
            ops.set_ray_scatter(True, encode=ray)
            n.encoding.hash_table.grad = None
            feats, sel = ops.contract_encode(n.encoding.hash_table, n.encoding.grid, n._norm, origins=rb.origins.to(DEV),
                                             directions=rb.directions.to(DEV), starts=fr.starts[..., 0].to(DEV),
                                             ends=fr.ends[..., 0].to(DEV))
            (feats * dy * sel[:, None]).sum().backward()
            res[ray] = n.encoding.hash_table.grad.clone()
    finally:
        ops.set_ray_scatter(True)
    assert_close(res[True], res[False], 2e-5, "table gradient, ray-major vs sample-major scatter")
