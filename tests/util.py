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
