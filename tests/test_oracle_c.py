"""CPU: pins the C restatement (oracle/nkbs_oracle.c) to the torch restatement (oracle/nerfstudio_torch.py)."""
import torch

from oracle import c_oracle as CO
from oracle import nerfstudio_torch as O


def _rays(R, seed):
    return O.NearFarCollider(0.001, 1000.0)(O.synthetic_ray_bundle(R, seed=seed))


def test_positions_and_hash_match_torch_oracle():
    torch.manual_seed(0)
    aabb = torch.tensor([[-1.0, -1.0, -1.0], [1.0, 1.0, 1.0]])
    f = O.HashMLPDensityField(aabb, spatial_distortion=O.SceneContraction(), hidden_dim=16, num_levels=5, max_res=128,
                              log2_hashmap_size=17)
    rb = _rays(200, 1)
    s = O.UniformLinDispPiecewiseSampler(single_jitter=True).train()
    rs = s(rb, num_samples=64)
    want_xn, want_sel = f._normalize(rs.frustums.get_positions())
    xn, sel = CO.positions(2, True, aabb.reshape(-1).tolist(), origins=rb.origins, directions=rb.directions,
                           starts=rs.frustums.starts[..., 0], ends=rs.frustums.ends[..., 0])
    assert torch.equal(xn, want_xn.reshape(-1, 3))
    assert torch.equal(sel > 0, want_sel.reshape(-1))
    hs, _ = f.encoding.corner_indices(want_xn.reshape(-1, 3))
    feats, idx = CO.hash_encode(xn, f.encoding.hash_table, f.encoding.scalings.tolist(), 17)
    assert torch.equal(idx, torch.stack(hs, -1))
    assert torch.equal(feats, f.encoding(want_xn.reshape(-1, 3)).detach())


def test_main_grid_indices_match_torch_oracle():
    torch.manual_seed(1)
    enc = O.HashEncoding(16, 16, 2048, 19)
    x = torch.rand(3000, 3)
    hs, _ = enc.corner_indices(x)
    feats, idx = CO.hash_encode(x, enc.hash_table, enc.scalings.tolist(), 19)
    assert torch.equal(idx, torch.stack(hs, -1))
    assert torch.equal(feats, enc(x).detach())


def test_initial_sampler_matches_torch_oracle():
    rb = _rays(300, 2)
    for single in (True, False):
        s = O.UniformLinDispPiecewiseSampler(single_jitter=single)
        for train in (True, False):
            s.train(train)
            t_rand = torch.rand(300, 1 if single else 257)
            rs = s(rb, num_samples=256, t_rand=t_rand)
            sb, eb = CO.initial_sampler(rb.nears, rb.fars, 256, t_rand if train else None)
            assert torch.equal(sb, torch.cat([rs.spacing_starts[..., 0], rs.spacing_ends[..., -1:, 0]], -1))
            assert torch.equal(eb, torch.cat([rs.frustums.starts[..., 0], rs.frustums.ends[..., -1:, 0]], -1))


def test_pdf_matches_torch_oracle_up_to_the_normaliser():
    """Identical except for torch.sum's summation order: rays whose normaliser agrees are bit-identical."""
    R, S, ns = 2000, 256, 96
    rb = _rays(R, 3)
    s = O.UniformLinDispPiecewiseSampler(single_jitter=True).train()
    rs = s(rb, num_samples=S)
    g = torch.Generator().manual_seed(0)
    w = torch.rand(R, S, 1, generator=g) ** 4
    w = w / w.sum(-2, keepdim=True) * torch.rand(R, 1, 1, generator=g)
    pdf = O.PDFSampler(include_original=False, single_jitter=True).train()
    rand = torch.rand(R, 1, generator=g)
    out = pdf(rb, rs, w, num_samples=ns, rand=rand)
    want = torch.cat([out.spacing_starts[..., 0], out.spacing_ends[..., -1:, 0]], -1)
    sb_in = torch.cat([rs.spacing_starts[..., 0], rs.spacing_ends[..., -1:, 0]], -1)
    sb, eb, inds = CO.pdf_sample(w[..., 0], sb_in, rb.nears, rb.fars, ns, rand=rand)
    wsum_t = (w[..., 0] + 0.01).sum(-1)
    wsum_c = (w[..., 0] + 0.01).double().sum(-1).float()
    same = wsum_t == wsum_c
    assert same.float().mean() > 0.3
    assert torch.equal(sb[same], want[same]), "rays with an identical normaliser must be bit-identical"
    assert torch.equal(inds[same], out._dbg_inds[same])
    assert (sb - want).abs().max() < 2e-6
    assert (inds == out._dbg_inds).float().mean() > 0.999
