"""Shared helpers for the parity tests: parameter transfer oracle -> nkbs, tolerances, ray bundles."""
import torch

from oracle import nerfstudio_torch as O
import nerf_kbs_b200 as nk


def _copy(dst, src):
    with torch.no_grad():
        dst.copy_(src.to(dst.device))


def copy_mlp(nk_mlp, o_mlp):
    for a, b in zip(nk_mlp.layers, o_mlp.layers):
        _copy(a.weight, b.weight)
        _copy(a.bias, b.bias)


def copy_density_field(nk_f, o_f):
    _copy(nk_f.encoding.hash_table, o_f.encoding.hash_table)
    if not o_f.use_linear:
        copy_mlp(nk_f.mlp_base[1], o_f.network)
    else:
        _copy(nk_f.linear.weight, o_f.linear.weight)
        _copy(nk_f.linear.bias, o_f.linear.bias)


def copy_field(nk_f, o_f):
    _copy(nk_f.mlp_base_grid.hash_table, o_f.encoding.hash_table)
    copy_mlp(nk_f.mlp_base_mlp, o_f.mlp_base_mlp)
    copy_mlp(nk_f.mlp_head, o_f.mlp_head)
    _copy(nk_f.embedding_appearance.embedding.weight, o_f.embedding_appearance.weight)
    if o_f.use_semantics:
        copy_mlp(nk_f.mlp_semantics, o_f.mlp_semantics)
        _copy(nk_f.field_head_semantics.net.weight, o_f.field_head_semantics.weight)
        _copy(nk_f.field_head_semantics.net.bias, o_f.field_head_semantics.bias)
    if o_f.use_transient_embedding:
        copy_mlp(nk_f.mlp_transient, o_f.mlp_transient)
        _copy(nk_f.embedding_transient.embedding.weight, o_f.embedding_transient.weight)
        for n in ("uncertainty", "rgb", "density"):
            _copy(getattr(nk_f, f"field_head_transient_{n}").net.weight, getattr(o_f, f"field_head_transient_{n}").weight)
            _copy(getattr(nk_f, f"field_head_transient_{n}").net.bias, getattr(o_f, f"field_head_transient_{n}").bias)


def copy_model(nk_m, o_m):
    copy_field(nk_m.field, o_m.field)
    for a, b in zip(nk_m.proposal_networks, o_m.proposal_networks):
        copy_density_field(a, b)


def field_grad_pairs(nk_f, o_f):
    """(name, nkbs grad, oracle grad) for every parameter of a NerfactoField pair."""
    out = [("hash_table", nk_f.mlp_base_grid.hash_table.grad, o_f.encoding.hash_table.grad),
           ("appearance", nk_f.embedding_appearance.embedding.weight.grad, o_f.embedding_appearance.weight.grad)]
    def mlp_pairs(tag, a, b):
        for i, (x, y) in enumerate(zip(a.layers, b.layers)):
            out.append((f"{tag}.{i}.w", x.weight.grad, y.weight.grad))
            out.append((f"{tag}.{i}.b", x.bias.grad, y.bias.grad))
    mlp_pairs("base", nk_f.mlp_base_mlp, o_f.mlp_base_mlp)
    mlp_pairs("head", nk_f.mlp_head, o_f.mlp_head)
    if o_f.use_semantics:
        mlp_pairs("sem", nk_f.mlp_semantics, o_f.mlp_semantics)
        out.append(("semhead.w", nk_f.field_head_semantics.net.weight.grad, o_f.field_head_semantics.weight.grad))
        out.append(("semhead.b", nk_f.field_head_semantics.net.bias.grad, o_f.field_head_semantics.bias.grad))
    if o_f.use_transient_embedding:
        mlp_pairs("transient", nk_f.mlp_transient, o_f.mlp_transient)
        out.append(("emb_transient", nk_f.embedding_transient.embedding.weight.grad, o_f.embedding_transient.weight.grad))
        for n in ("uncertainty", "rgb", "density"):
            out.append((f"t_{n}.w", getattr(nk_f, f"field_head_transient_{n}").net.weight.grad,
                        getattr(o_f, f"field_head_transient_{n}").weight.grad))
    return out


def density_grad_pairs(nk_f, o_f):
    out = [("hash_table", nk_f.encoding.hash_table.grad, o_f.encoding.hash_table.grad)]
    for i, (x, y) in enumerate(zip(nk_f.mlp_base[1].layers, o_f.network.layers)):
        out.append((f"net.{i}.w", x.weight.grad, y.weight.grad))
        out.append((f"net.{i}.b", x.bias.grad, y.bias.grad))
    return out


def rel_err(a, b):
    """max |a-b| / (max|b| + tiny): the 'relative error' of north_star, scale taken over the tensor."""
    a = a.detach().float().cpu()
    b = b.detach().float().cpu()
    return ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()


def assert_close(a, b, tol, name=""):
    e = rel_err(a, b)
    assert e <= tol, f"{name}: relative error {e:.3e} > {tol:.1e}"


def to_nk_bundle(rb: "O.RayBundle", device="cuda"):
    mv = lambda t: t.to(device) if t is not None else None
    return nk.RayBundle(origins=mv(rb.origins), directions=mv(rb.directions), pixel_area=mv(rb.pixel_area),
                        camera_indices=mv(rb.camera_indices), nears=mv(rb.nears), fars=mv(rb.fars),
                        metadata={k: mv(v) for k, v in rb.metadata.items()})


def make_jitter(R, levels=(256, 96, 48), seed=1, single=True):
    g = torch.Generator().manual_seed(seed)
    return [torch.rand(R, 1 if single else (n + 1), generator=g) for n in levels]


# ------------------------------------------------------------------------------------------------
# analytic scene for the fixed-step PSNR run: a normal-shaded sphere over a checkered ground plane under a
# graded sky, rendered by a closed-form ray caster (no images on disk, nothing learned)
# ------------------------------------------------------------------------------------------------
def analytic_scene_colors(origins, directions):
    """rgb [N,3] seen along each ray.  Sphere: centre 0, radius 0.35, colour 0.5 + 0.5 n.  Plane z = -0.35 with a
    0.25-wide checker (inside |x|,|y| < 1.5).  Sky: vertical gradient."""
    o, d = origins.double(), directions.double()
    b = (o * d).sum(-1)
    c = (o * o).sum(-1) - 0.35 ** 2
    disc = b * b - c
    t_s = -b - torch.sqrt(disc.clamp_min(0))
    hit_s = (disc > 0) & (t_s > 0)
    t_p = (-0.35 - o[:, 2]) / torch.where(d[:, 2].abs() < 1e-9, torch.full_like(d[:, 2], 1e-9), d[:, 2])
    pp = o + t_p[:, None] * d
    hit_p = (t_p > 0) & (pp[:, 0].abs() < 1.5) & (pp[:, 1].abs() < 1.5)
    sky = torch.stack([0.55 + 0.2 * d[:, 2], 0.65 + 0.2 * d[:, 2], 0.85 + 0.1 * d[:, 2]], -1)
    n = (o + t_s[:, None] * d) / 0.35
    col_s = 0.5 + 0.5 * n
    chk = ((torch.floor(pp[:, 0] / 0.25) + torch.floor(pp[:, 1] / 0.25)) % 2)[:, None]
    col_p = chk * torch.tensor([0.85, 0.8, 0.7], dtype=torch.float64) + (1 - chk) * torch.tensor([0.2, 0.25, 0.3], dtype=torch.float64)
    use_p = hit_p & (~hit_s | (t_p < t_s))
    rgb = torch.where(use_p[:, None], col_p, torch.where(hit_s[:, None], col_s, sky))
    return rgb.float().clamp(0, 1)


def analytic_scene_rays(n, num_images=8, seed=0):
    """n rays of ``num_images`` pinhole-like views on a shell around the sphere: (O.RayBundle, rgb [n,3])."""
    g = torch.Generator().manual_seed(seed)
    cam = torch.randint(0, num_images, (n, 1), generator=g)
    gc = torch.Generator().manual_seed(12345)  # the views are the same for every ray set
    az = torch.rand(num_images, generator=gc) * 6.2831853
    el = 0.15 + 0.5 * torch.rand(num_images, generator=gc)
    centers = 0.85 * torch.stack([torch.cos(az) * torch.cos(el), torch.sin(az) * torch.cos(el), torch.sin(el)], -1)
    o = centers[cam[:, 0]]
    target = (torch.rand(n, 3, generator=g) - 0.5) * torch.tensor([1.0, 1.0, 0.8])
    d = target - o
    d = d / d.norm(dim=-1, keepdim=True)
    rb = O.RayBundle(origins=o, directions=d, pixel_area=torch.full((n, 1), 1e-6), camera_indices=cam, metadata={})
    return rb, analytic_scene_colors(o, d)


def psnr(pred, target):
    mse = ((pred.detach().float().cpu() - target.float().cpu()) ** 2).mean().item()
    return -10.0 * __import__("math").log10(max(mse, 1e-12))


PSNR_PROPOSAL_ARGS = ({"hidden_dim": 16, "log2_hashmap_size": 13, "num_levels": 5, "max_res": 128, "use_linear": False},
                      {"hidden_dim": 16, "log2_hashmap_size": 13, "num_levels": 5, "max_res": 256, "use_linear": False})
PSNR_CFG = dict(log2_hashmap_size=15, proposal_net_args_list=PSNR_PROPOSAL_ARGS, num_images=8)
