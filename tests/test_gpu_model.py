"""End-to-end parity of the two get_outputs paths (BASELINE configs 1-3 at oracle-sized batches): forward
outputs, sample bins, and every parameter gradient against the CPU oracle on identical rays and draws."""
import pytest
import torch

from oracle import nerfstudio_torch as O
import nerf_kbs_b200 as nk
from tests.util import (assert_close, copy_model, density_grad_pairs, field_grad_pairs, make_jitter, rel_err,
                        to_nk_bundle)

pytestmark = pytest.mark.gpu
DEV = "cuda"
FP32_TOL = 1e-4


def _pair(kind, **cfg):
    torch.manual_seed(0)
    if kind == "nerfacto":
        om, nm = O.NerfactoModel(O.ModelConfig(**cfg)), nk.NerfactoModel(nk.ModelConfig(**cfg))
    else:
        om, nm = O.SemanticNerfWModel(O.ModelConfig(**cfg)), nk.SemanticNerfWModel(nk.ModelConfig(**cfg))
    # larger table values than the 1e-3 init so that densities and gradients are not vanishing
    with torch.no_grad():
        for p in om.parameters():
            if p.dim() == 2 and p.shape[1] == 2:
                p.mul_(300.0)
    nm = nm.to(DEV)
    copy_model(nm, om)
    return om, nm


def _batch(R, C, seed=5):
    g = torch.Generator().manual_seed(seed)
    return {"image": torch.rand(R, 3, generator=g), "semantics_label": torch.randint(0, C, (R,), generator=g)}


def _check_outputs(got, want, keys, tol=FP32_TOL):
    for k in keys:
        assert got[k].shape == want[k].shape, k
        assert_close(got[k], want[k], tol, k)


@pytest.mark.parametrize("train", [True, False])
def test_nerfacto_forward(train):
    R = 512
    om, nm = _pair("nerfacto", use_semantic=True)
    om.train(train), nm.train(train)
    rb = O.synthetic_ray_bundle(R, seed=1)
    jit = make_jitter(R)
    want = om(rb, jitter=jit)
    got = nm(to_nk_bundle(rb), jitter=[j.to(DEV) for j in jit])
    _check_outputs(got, want, ["rgb", "accumulation", "expected_depth", "semantics"])
    for k in ("depth", "prop_depth_0", "prop_depth_1"):
        frac = (rel_close := ((got[k].cpu() - want[k]).abs() <= 1e-4 * want[k].abs().clamp_min(1e-3))).float().mean().item()
        assert frac > 0.99, f"{k}: only {frac:.4f} of the median depths agree"
    if train:
        for lvl in range(3):
            sb_want = O.ray_samples_to_sdist(want["ray_samples_list"][lvl])
            sb_got = got["ray_samples_list"][lvl].spacing_bins
            assert_close(sb_got, sb_want, 1e-5, f"spacing bins level {lvl}")
            assert_close(got["weights_list"][lvl], want["weights_list"][lvl], FP32_TOL * 10, f"weights level {lvl}")
        assert torch.equal(got["ray_samples_list"][0].spacing_bins.cpu(), O.ray_samples_to_sdist(want["ray_samples_list"][0]))
    else:
        assert "weights_list" not in got


def _grad_check(pairs, tol):
    worst = []
    for name, g, w in pairs:
        assert g is not None and w is not None, name
        worst.append((rel_err(g, w), name))
        assert rel_err(g, w) <= tol, f"grad {name}: rel err {rel_err(g, w):.3e}"
    return worst


def test_nerfacto_train_step_gradients():
    R = 384
    om, nm = _pair("nerfacto", use_semantic=True, pass_semantic_gradients=False)
    om.train(), nm.train()
    rb = O.synthetic_ray_bundle(R, seed=2)
    jit = make_jitter(R, seed=3)
    batch = _batch(R, 19)
    for step in (0, 500):
        om.zero_grad(), nm.zero_grad()
        om.set_step(step), nm.set_step(step)
        lo = om.get_loss_dict(om(rb, jitter=jit), batch)
        ln = nm.get_loss_dict(nm(to_nk_bundle(rb), jitter=[j.to(DEV) for j in jit]), {k: v.to(DEV) for k, v in batch.items()})
        for k in lo:
            assert_close(ln[k], lo[k], 2e-4, f"loss {k} @step {step}")
        sum(lo.values()).backward()
        sum(ln.values()).backward()
        _grad_check(field_grad_pairs(nm.field, om.field), 2e-3)
        for a, b in zip(nm.proposal_networks, om.proposal_networks):
            _grad_check(density_grad_pairs(a, b), 2e-3)


def test_proposal_update_gating():
    """After step 10 the proposal networks only receive gradients on scheduled steps."""
    R = 64
    om, nm = _pair("nerfacto")
    om.train(), nm.train()
    rb = O.synthetic_ray_bundle(R, seed=2)
    batch = _batch(R, 19)
    pattern = []
    for step in range(0, 16):
        nm.zero_grad(), om.zero_grad()
        nm.set_step(step), om.set_step(step)
        jit = make_jitter(R, seed=step)
        sum(nm.get_loss_dict(nm(to_nk_bundle(rb), jitter=[j.to(DEV) for j in jit]), {k: v.to(DEV) for k, v in batch.items()}).values()).backward()
        sum(om.get_loss_dict(om(rb, jitter=jit), batch).values()).backward()
        g_n = nm.proposal_networks[0].encoding.hash_table.grad
        g_o = om.proposal_networks[0].encoding.hash_table.grad
        pattern.append((g_n is not None and g_n.abs().sum().item() > 0, g_o is not None and g_o.abs().sum().item() > 0))
        nm.after_step(step), om.after_step(step)
    assert all(a == b for a, b in pattern), pattern
    assert not all(a for a, _ in pattern), "gating never skipped an update"


@pytest.mark.parametrize("transient", [False, True])
def test_semantic_nerfw_train_step(transient):
    R = 256
    om, nm = _pair("semantic", use_semantic=True, use_transient_embedding=transient)
    om.train(), nm.train()
    rb = O.synthetic_ray_bundle(R, seed=7)
    jit = make_jitter(R, seed=8)
    batch = _batch(R, 19)
    wo = om(rb, jitter=jit)
    wn = nm(to_nk_bundle(rb), jitter=[j.to(DEV) for j in jit])
    keys = ["rgb", "accumulation", "semantics"] + (["uncertainty", "density_transient"] if transient else [])
    _check_outputs(wn, wo, keys)
    lo = om.get_loss_dict(wo, batch)
    ln = nm.get_loss_dict(wn, {k: v.to(DEV) for k, v in batch.items()})
    for k in lo:
        assert_close(ln[k], lo[k], 2e-4, f"loss {k}")
    sum(lo.values()).backward()
    sum(ln.values()).backward()
    _grad_check(field_grad_pairs(nm.field, om.field), 2e-3)
    for a, b in zip(nm.proposal_networks, om.proposal_networks):
        _grad_check(density_grad_pairs(a, b), 2e-3)


def test_state_dict_round_trip():
    _, nm = _pair("nerfacto", use_semantic=True)
    sd = nm.state_dict()
    assert "field.mlp_base_grid.hash_table" in sd and sd["field.mlp_base_grid.hash_table"].shape == (16 * 2**19, 2)
    assert "proposal_networks.0.mlp_base.1.layers.0.weight" in sd
    nm2 = nk.NerfactoModel(nk.ModelConfig(use_semantic=True)).to(DEV)
    nm2.load_state_dict(sd)
    groups = nm.get_param_groups()
    assert set(groups) == {"proposal_networks", "fields"}
    assert sum(p.numel() for p in groups["fields"]) + sum(p.numel() for p in groups["proposal_networks"]) == \
        sum(p.numel() for p in nm.parameters())
