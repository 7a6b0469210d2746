"""Known-answer tests of the oracle (CPU only).

The reference ships no tests or fixtures for this path, so besides the minted golden vectors the oracle
is held to answers derived BY HAND from the operator definitions in SURVEY Appendix A (closed forms, integer arithmetic
done with Python ints) — nothing below calls the oracle to produce its own expectation.  The last block holds the host
side of the product (schedules, shard arithmetic) to the same hand-derived answers; no kernel is launched here."""
import math

import numpy as np
import pytest
import torch

from oracle import nerfstudio_torch as O

P1, P2 = 2654435761, 805459861


# ---------------------------------------------------------------------------------- A.2 hash grid
def test_scalings_tables_from_survey_a2():
    assert O.hash_scalings(16, 16, 2048).tolist() == [16, 22, 30, 42, 58, 80, 111, 153, 212, 294, 406, 561, 776, 1072, 1482, 2047]
    assert O.hash_scalings(5, 16, 128).tolist() == [16, 26, 45, 76, 128]
    assert O.hash_scalings(5, 16, 256).tolist() == [16, 32, 64, 128, 256]
    assert O.hash_scalings(8, 16, 1024).tolist() == [16, 28, 52, 95, 172, 312, 565, 1024]


def _py_hash(v, T, level):
    return ((v[0] * 1) ^ (v[1] * P1) ^ (v[2] * P2)) % T + level * T


def test_corner_indices_against_python_integer_arithmetic():
    """x = (0.3, 0.6, 0.9): level 0 (scale 16) → s = (4.8, 9.6, 14.4), ceil (5,10,15), floor (4,9,14); level 1 (scale 26)
    → s = (7.8, 15.6, 23.4).  Corner order h0..h7 per A.2: ccc cfc ffc fcc ccf cff fff fcf (x,y,z picks)."""
    enc = O.HashEncoding(num_levels=5, min_res=16, max_res=128, log2_hashmap_size=17)
    T = 2 ** 17
    x = torch.tensor([[0.3, 0.6, 0.9]])
    hs, _ = enc.corner_indices(x)  # 8 tensors [1, L]
    got = torch.stack(list(hs), 0).reshape(8, -1, 5)[:, 0, :]
    picks = ["ccc", "cfc", "ffc", "fcc", "ccf", "cff", "fff", "fcf"]
    for level, scale in enumerate([16, 26, 45, 76, 128]):
        s = [np.float32(v) * np.float32(scale) for v in (0.3, 0.6, 0.9)]
        c = [int(math.ceil(float(v))) for v in s]
        f = [int(math.floor(float(v))) for v in s]
        for k, pk in enumerate(picks):
            v = [c[d] if pk[d] == "c" else f[d] for d in range(3)]
            assert int(got[k, level]) == _py_hash(v, T, level), (level, pk)
    # the first level by hand: (5 ^ 10·P1 ^ 15·P2) mod 2^17
    assert int(got[0, 0]) == (5 ^ (10 * P1) ^ (15 * P2)) % T


def test_trilinear_weights_sit_on_the_ceil_corner():
    """A table that stores 1.0 only at the row of the all-ceil corner gives o_x·o_y·o_z (A.2: weight o on ceil)."""
    enc = O.HashEncoding(num_levels=1, min_res=16, max_res=16, log2_hashmap_size=19, features_per_level=2)
    T = 2 ** 19
    with torch.no_grad():
        enc.hash_table.zero_()
        enc.hash_table[_py_hash([5, 10, 15], T, 0)] = 1.0
    out = enc(torch.tensor([[0.3, 0.6, 0.9]]))
    ox, oy, oz = (np.float32(0.3) * 16 - 4, np.float32(0.6) * 16 - 9, np.float32(0.9) * 16 - 14)
    assert out.shape == (1, 2)
    assert abs(float(out[0, 0].detach()) - float(ox * oy * oz)) < 1e-6
    # on an exact lattice point ceil == floor and o == 0: every weight collapses onto the "floor" picks
    with torch.no_grad():
        enc.hash_table.zero_()
        enc.hash_table[_py_hash([4, 8, 12], T, 0)] = 3.0
    out = enc(torch.tensor([[0.25, 0.5, 0.75]]))
    assert torch.allclose(out, torch.full((1, 2), 3.0))


# ---------------------------------------------------------------------------------- A.1 contraction, A.5 sampler
def test_scene_contraction_inf_norm():
    sc = O.SceneContraction(order=float("inf"))
    x = torch.tensor([[0.5, -0.25, 0.1], [2.0, 0.0, 0.0], [-4.0, 2.0, 1.0], [1.0, 1.0, 1.0]])
    want = torch.tensor([[0.5, -0.25, 0.1], [1.5, 0.0, 0.0], [-1.75, 0.875, 0.4375], [1.0, 1.0, 1.0]])
    assert torch.allclose(sc(x), want, atol=1e-7)
    far = sc(torch.tensor([[1e6, 0.0, 0.0]]))
    assert float(far[0, 0]) < 2.0 and float(far[0, 0]) > 1.99999


def test_initial_sampler_endpoints_and_midpoint():
    """s→t with near 0.001, far 1000: spacing(near)=0.0005, spacing(far)=1−1/2000; s=0 → t=near, s=1 → t=far, and the
    spacing midpoint 0.5 maps to t = 1 (the joint of the two pieces)."""
    rb = O.RayBundle(origins=torch.zeros(2, 3), directions=torch.tensor([[0.0, 0, 1], [1.0, 0, 0]]),
                     pixel_area=torch.ones(2, 1), camera_indices=torch.zeros(2, 1, dtype=torch.long),
                     nears=torch.full((2, 1), 0.001), fars=torch.full((2, 1), 1000.0))
    smp = O.UniformLinDispPiecewiseSampler(train_stratified=False)
    rs = smp(rb, num_samples=4)
    starts, ends = rs.frustums.starts[..., 0], rs.frustums.ends[..., 0]
    assert starts.shape == (2, 4)
    assert abs(float(starts[0, 0]) - 0.001) < 1e-7
    assert abs(float(ends[0, -1]) - 1000.0) / 1000.0 < 1e-3  # 1/(2-2s) near s≈0.9995 amplifies fp32 rounding
    assert abs(float(rs.spacing_starts[0, 2, 0]) - 0.5) < 1e-7
    assert abs(float(starts[0, 2]) - 1.0) < 1e-5
    assert torch.equal(starts[0], starts[1])  # depends on near/far only
    assert torch.all(ends[:, :-1] == starts[:, 1:])


# ---------------------------------------------------------------------------------- A.4/A.8 weights and renderers
def _uniform_samples(R, S, delta):
    edges = torch.arange(S + 1, dtype=torch.float32) * delta + 1.0
    edges = edges[None, :, None].expand(R, S + 1, 1)
    fr = O.Frustums(origins=torch.zeros(R, S, 3), directions=torch.ones(R, S, 3) / math.sqrt(3.0),
                    starts=edges[:, :-1], ends=edges[:, 1:], pixel_area=torch.ones(R, S, 1))
    return O.RaySamples(frustums=fr, deltas=edges[:, 1:] - edges[:, :-1])


def test_get_weights_geometric_series_for_constant_density():
    S, delta, sigma = 16, 0.25, 0.8
    rs = _uniform_samples(3, S, delta)
    w = rs.get_weights(torch.full((3, S, 1), sigma))
    a = 1.0 - math.exp(-sigma * delta)
    want = torch.tensor([a * math.exp(-i * sigma * delta) for i in range(S)])
    assert torch.allclose(w[0, :, 0], want, rtol=1e-5, atol=1e-7)
    assert abs(float(w[1].sum()) - (1.0 - math.exp(-S * sigma * delta))) < 1e-6
    # inf·0 products are cleaned, not propagated (nan_to_num)
    w = rs.get_weights(torch.tensor([float("inf")]).expand(3, S, 1))
    assert torch.isfinite(w).all() and abs(float(w[0, 0, 0]) - 1.0) < 1e-7 and float(w[0, 1:].abs().sum()) == 0.0


def test_renderers_closed_forms():
    S = 8
    rs = _uniform_samples(2, S, 0.5)
    w = torch.zeros(2, S, 1)
    w[0, 3] = 1.0                      # one-hot: every renderer returns that sample
    w[1, :] = 0.05                     # Σw = 0.4 → 0.6 of the background (last sample)
    rgb = torch.rand(2, S, 3, generator=torch.Generator().manual_seed(3))
    out = O.RGBRenderer(background_color="last_sample")(rgb=rgb, weights=w)
    assert torch.allclose(out[0], rgb[0, 3], atol=1e-7)
    assert torch.allclose(out[1], 0.05 * rgb[1].sum(0) + 0.6 * rgb[1, -1], atol=1e-6)
    assert torch.allclose(O.AccumulationRenderer()(weights=w)[:, 0], torch.tensor([1.0, 0.4]), atol=1e-7)
    steps = (rs.frustums.starts + rs.frustums.ends) / 2
    med = O.DepthRenderer("median")(weights=w, ray_samples=rs)
    assert float(med[0, 0]) == float(steps[0, 3, 0])
    assert float(med[1, 0]) == float(steps[1, S - 1, 0])  # cumsum never reaches 0.5 → clamped to the last sample
    exp = O.DepthRenderer("expected")(weights=w, ray_samples=rs)
    assert abs(float(exp[0, 0]) - float(steps[0, 3, 0])) < 1e-6
    assert abs(float(exp[1, 0]) - float(steps[1, :, 0].mean())) < 1e-5
    logits = torch.randn(2, S, 5, generator=torch.Generator().manual_seed(4))
    sem = O.SemanticRenderer()(logits, weights=w)
    assert torch.allclose(sem[0], logits[0, 3], atol=1e-7) and torch.allclose(sem[1], 0.05 * logits[1].sum(0), atol=1e-6)
    unc = O.UncertaintyRenderer()(torch.full((2, S, 1), 2.0), w)
    assert torch.allclose(unc[:, 0], torch.tensor([2.0, 0.8]), atol=1e-6)


# ---------------------------------------------------------------------------------- A.6 PDF sampler
def test_pdf_sampler_uniform_histogram_gives_uniform_bins():
    """Equal weights → the cdf is the identity on spacing space, so eval-mode bins are u = linspace(0, 1−1/nb, nb) + 1/(2nb)."""
    R, S, nb = 2, 8, 5  # num_samples 4 → nb = 5 bin edges
    rb = O.RayBundle(origins=torch.zeros(R, 3), directions=torch.tensor([[0.0, 0, 1]] * R), pixel_area=torch.ones(R, 1),
                     camera_indices=torch.zeros(R, 1, dtype=torch.long), nears=torch.full((R, 1), 0.001), fars=torch.full((R, 1), 1000.0))
    rs = O.UniformLinDispPiecewiseSampler(train_stratified=False)(rb, num_samples=S)
    pdf = O.PDFSampler(train_stratified=False, include_original=False)
    out = pdf(rb, rs, torch.full((R, S, 1), 0.3), num_samples=nb - 1)
    bins = torch.cat([out.spacing_starts[..., 0], out.spacing_ends[:, -1:, 0]], -1)
    want = torch.linspace(0.0, 1.0 - 1.0 / nb, nb) + 1.0 / (2 * nb)
    assert torch.allclose(bins[0], want, atol=1e-6)
    assert out.frustums.starts.requires_grad is False
    # all mass in one interval → every new edge falls inside it (histogram padding 0.01 spreads a little outside)
    w = torch.zeros(R, S, 1)
    w[:, 5] = 10.0
    out = pdf(rb, rs, w, num_samples=nb - 1)
    mid = torch.cat([out.spacing_starts[..., 0], out.spacing_ends[:, -1:, 0]], -1)[0, 1:-1]
    assert float(mid.min()) >= 5 / 8 - 1e-6 and float(mid.max()) <= 6 / 8 + 1e-6


# ---------------------------------------------------------------------------------- A.3 activations, SH
def test_trunc_exp_clamps_its_gradient_only():
    x = torch.tensor([0.0, 1.0, 20.0], requires_grad=True)
    y = O._TruncExp.apply(x)
    assert torch.allclose(y, torch.exp(x.detach()))
    y.sum().backward()
    assert torch.allclose(x.grad, torch.exp(torch.tensor([0.0, 1.0, 15.0])))


def test_spherical_harmonics_degree4_constants():
    d = torch.tensor([[0.0, 0.0, 1.0], [1.0, 0.0, 0.0]])
    sh = O.components_from_spherical_harmonics(4, d)
    assert sh.shape == (2, 16)
    assert torch.allclose(sh[:, 0], torch.full((2,), 0.28209479177387814))
    # l=1: (y, z, x)·0.4886; l=2, m=0: 0.3154·(3z²−1)
    assert torch.allclose(sh[0, 1:4], torch.tensor([0.0, 0.4886025119029199, 0.0]), atol=1e-7)
    assert torch.allclose(sh[1, 1:4], torch.tensor([0.0, 0.0, 0.4886025119029199]), atol=1e-7)
    assert abs(float(sh[0, 6]) - 0.31539156525252005 * 2.0) < 1e-6
    assert abs(float(sh[1, 6]) + 0.31539156525252005) < 1e-6


# ---------------------------------------------------------------------------------- A.9 losses
def test_distortion_loss_single_interval_is_a_third_of_its_width():
    """w = δ_k: the pairwise term vanishes and the self term is w²·(t1−t0)/3."""
    t = torch.linspace(0, 1, 9)[None]
    w = torch.zeros(1, 8)
    w[0, 2] = 1.0
    assert abs(float(O.lossfun_distortion(t, w)) - (1 / 8) / 3) < 1e-7
    # two unit-mass intervals at distance D between midpoints add 2·(½)(½)·D
    w = torch.zeros(1, 8)
    w[0, 1] = 0.5
    w[0, 6] = 0.5
    want = 2 * 0.25 * (5 / 8) + 2 * 0.25 * (1 / 8) / 3
    assert abs(float(O.lossfun_distortion(t, w)) - want) < 1e-6


def test_interlevel_loss_is_zero_when_the_proposal_bounds_the_histogram():
    t = torch.linspace(0, 1, 9)[None]
    w = torch.rand(1, 8, generator=torch.Generator().manual_seed(0))
    w = w / w.sum()
    assert float(O.lossfun_outer(t, w, t, w).sum()) == 0.0
    # a coarser proposal carrying the summed mass still bounds it
    t_env = t[:, ::2]
    w_env = w.reshape(1, 4, 2).sum(-1)
    assert float(O.lossfun_outer(t, w, t_env, w_env).sum()) < 1e-12
    # envelope [-0.1, 0.3, 1.1] with masses (0.05, 0.2) under a uniform histogram of 1/8: intervals 0 and 1 lie inside
    # the first envelope bin (bound 0.05 < 0.125), interval 2 straddles (bound 0.25), the rest see 0.2 → two terms
    # (0.125 − 0.05)² / 0.125 = 0.045
    w = torch.full((1, 8), 0.125)
    got = O.lossfun_outer(t, w, torch.tensor([[-0.1, 0.3, 1.1]]), torch.tensor([[0.05, 0.2]]))
    assert abs(float(got.sum()) - 0.09) < 1e-6 and float(got[0, 2:].abs().sum()) == 0.0


# ---------------------------------------------------------------------------------- a20 schedules
def test_anneal_and_update_schedule_known_values():
    assert O.anneal_value(0, 1000, 10.0) == 0.0
    assert abs(O.anneal_value(500, 1000, 10.0) - 5.0 / 5.5) < 1e-12
    assert abs(O.anneal_value(100, 1000, 10.0) - 1.0 / 1.9) < 1e-12
    assert O.anneal_value(1000, 1000, 10.0) == 1.0 and O.anneal_value(9999, 1000, 10.0) == 1.0


def _gating_sequence(model, steps):
    """Which steps carry proposal gradients (`updated` of ProposalNetworkSampler) for a model driven like the trainer does."""
    out = []
    s = model.proposal_sampler
    for step in range(steps):
        updated = s._steps_since_update > s.update_sched(s._step) or s._step < 10
        out.append(bool(updated))
        if updated:
            s._steps_since_update = 0
        model.after_step(step)
    return out


def test_proposal_update_gating_matches_the_reference_schedule():
    """update_sched = clip(interp(step, [0, 5000], [0, 5]), 1, 5): every step below 10, then every other step while the
    schedule is 1 (steps_since_update must EXCEED it), every sixth step once it saturates at 5."""
    cfg = O.ModelConfig()
    m = O.NerfactoModel(cfg)
    seq = _gating_sequence(m, 60)
    assert all(seq[:11])
    assert seq[11:21] == [False, True] * 5
    sched = m.proposal_sampler.update_sched
    assert sched(0) == 1 and sched(1000) == 1 and sched(2500) == 2.5 and sched(5000) == 5 and sched(10 ** 6) == 5


def test_host_mirror_follows_the_same_schedules():
    """The product's host side (no kernel launched): same anneal values and the same gating sequence as the oracle."""
    from nerf_kbs_b200 import models as M
    for step in (0, 1, 7, 100, 500, 999, 1000, 5000):
        assert M.anneal_value(step, 1000, 10.0) == O.anneal_value(step, 1000, 10.0)
    want = _gating_sequence(O.NerfactoModel(O.ModelConfig()), 80)
    prod = M.NerfactoModel(M.ModelConfig())
    got = []
    for step in range(80):
        u = prod.proposal_sampler.will_update()
        got.append(bool(u))
        if u:
            prod.proposal_sampler._steps_since_update = 0
        prod.after_step(step)
    assert got == want


def test_host_mirror_has_the_reference_parameter_inventory():
    """Same parameter count as the torch path: 16·2^19·2 main table, two 5·2^17·2 proposal tables, the MLPs and the
    115×32 appearance embedding (≈19.4 M parameters, 77.6 MB)."""
    from nerf_kbs_b200 import models as M
    prod = M.NerfactoModel(M.ModelConfig())
    ref = O.NerfactoModel(O.ModelConfig())
    n_prod = sum(p.numel() for p in prod.parameters())
    n_ref = sum(p.numel() for p in ref.parameters())
    assert n_prod == n_ref
    assert n_ref > 16 * 2 ** 19 * 2 + 2 * 5 * 2 ** 17 * 2 and abs(4 * n_ref - 77.6e6) < 0.5e6
    groups = prod.get_param_groups()
    assert set(groups) == {"proposal_networks", "fields"}


def test_shard_rays_equal_disjoint_slices():
    from nerf_kbs_b200.trainer import shard_rays
    # reference semantics (datamanager_z.py:225-231): train_num_rays_per_batch // world_size rays on every rank,
    # the remainder is dropped
    for total, world in ((4096, 1), (4096, 8), (4097, 8), (7, 8), (465750, 4)):
        parts = [shard_rays(total, r, world) for r in range(world)]
        seen = [i for p in parts for i in p]
        assert all(len(p) == total // world for p in parts)
        assert seen == list(range((total // world) * world))


def test_c_abi_refuses_bad_arguments_without_a_gpu():
    """Argument validation happens before any launch: a null / zero-sized / oversized request comes back as a negative
    code with a message in nkbs_last_error(), never as a crash (SURVEY §8b errors)."""
    from nerf_kbs_b200 import _lib
    lib = _lib.load()
    assert lib.nkbs_version() > 0
    err = lib.nkbs_last_error()
    assert err is None or isinstance(err, (bytes, str))
    fake = 1 << 20  # never dereferenced: every call below is refused (or is host-only) before any launch
    refused = [
        ("nkbs_fixed_point_register", (fake, 16, fake, 61)),       # scale outside [16, 60]
        ("nkbs_fixed_point_register", (None, 16, fake, 52)),       # null buffer
        ("nkbs_fixed_point_flush", (fake + 4096, 16, None)),       # not inside a registered buffer
        ("nkbs_density_bwd_saved", (128, 8, 64) + (fake,) * 12 + (None,)),   # only the (L=5, H=16) field
        ("nkbs_expected_depth_noalloc", ((1 << 20) + 1, 48, fake, fake, fake, 0, fake, None)),  # beyond the scratch
        ("nkbs_adam_step", (-1,) + (None,) * 4 + (None, 1e-3, 0.9, 0.999, 1e-15, 0.0, None, None, 1, None, None)),
    ]
    for name, args in refused:
        with pytest.raises(RuntimeError) as e:
            _lib.call(name, *args)
        assert name.split("(")[0] in str(e.value), (name, str(e.value))
    # the deterministic-mode registry is host state: register, look up through a refused flush, unregister
    _lib.call("nkbs_fixed_point_register", fake, 1024, fake * 2, 56)
    with pytest.raises(RuntimeError):  # odd element offset into the shadow
        _lib.call("nkbs_fixed_point_flush", fake + 4, 8, None)
    with pytest.raises(RuntimeError):  # range runs past the registered buffer
        _lib.call("nkbs_fixed_point_flush", fake + 4 * 1000, 100, None)
    _lib.call("nkbs_fixed_point_register", fake, 1024, None, 0)
    with pytest.raises(RuntimeError):
        _lib.call("nkbs_fixed_point_flush", fake, 8, None)


def test_host_ray_containers_keep_the_reference_attributes():
    """RaySamples.deltas is ends - starts although it is only built on demand; the collider's cached planes carry the
    configured near / far values for every bundle of the same size (ref: NearFarCollider at nerfacto.py:335)."""
    import nerf_kbs_b200 as nk
    R, S = 6, 4
    edges = torch.linspace(1.0, 3.0, S + 1).repeat(R, 1)
    rb = nk.RayBundle(torch.zeros(R, 3), torch.ones(R, 3), torch.ones(R, 1), torch.zeros(R, 1, dtype=torch.long))
    col = nk.NearFarCollider(near_plane=0.001, far_plane=1000.0)
    rb = col(rb)
    assert rb.nears.shape == (R, 1) and float(rb.nears[0, 0]) == pytest.approx(0.001) and float(rb.fars[-1, 0]) == 1000.0
    rb2 = col(nk.RayBundle(torch.ones(R, 3), torch.ones(R, 3), torch.ones(R, 1), torch.zeros(R, 1, dtype=torch.long)))
    assert torch.equal(rb2.nears, rb.nears) and torch.equal(rb2.fars, rb.fars)
    rb3 = col(nk.RayBundle(torch.ones(2 * R, 3), torch.ones(2 * R, 3)))
    assert rb3.nears.shape == (2 * R, 1)
    rs = rb.samples_from_bins(edges / 3.0, edges)
    assert rs.deltas_ is None                                   # nothing built yet
    assert rs.deltas.shape == (R, S, 1) and torch.allclose(rs.deltas, torch.full((R, S, 1), 0.5))
    assert rs.frustums.starts.shape == (R, S, 1) and rs.spacing_bins.shape == (R, S + 1)
    rs.deltas = torch.zeros(R, S, 1)
    assert float(rs.deltas.abs().sum()) == 0.0
