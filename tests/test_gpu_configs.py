"""GPU: parity at the BENCHMARKED configurations, through the code path bench.py times — tcgen05 mixed-precision MLPs,
``Trainer`` with CUDA-graph replay, the fused loss route, gradient arena + Adam — against the CPU oracle's own
optimisation loop on the same rays, targets, parameters and jitter draws.

  configs[1]  nerfacto training step, 4096 rays, (256, 96) + 48 samples      (ref:train.py:42-101, nerfacto.py:396-489)
  configs[2]  semantic-NeRF-W step, 8192 rays, semantic + transient heads     (ref:semantic_nerfw.py:186-293)
  X1          fixed-step training run on an analytic scene, PSNR of both arms (ref:nerfacto.py:355,458-460)

Bars (north_star): sample bins bit-exact, mixed-precision outputs / losses within 1e-2 relative, matching PSNR."""
import os

import pytest
import torch

from oracle import nerfstudio_torch as O
import nerf_kbs_b200 as nk
from nerf_kbs_b200 import ops
from nerf_kbs_b200.trainer import Trainer
from tests.util import (PSNR_CFG, analytic_scene_rays, copy_model, field_grad_pairs, density_grad_pairs, make_jitter,
                        psnr, rel_err, to_nk_bundle)

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rel_l2(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return ((a - b).norm() / (b.norm() + 1e-30)).item()


def _pair(kind, scale_tables=300.0, **cfg):
    torch.manual_seed(0)
    if kind == "nerfacto":
        om, nm = O.NerfactoModel(O.ModelConfig(**cfg)), nk.NerfactoModel(nk.ModelConfig(**cfg))
    else:
        om, nm = O.SemanticNerfWModel(O.ModelConfig(**cfg)), nk.SemanticNerfWModel(nk.ModelConfig(**cfg))
    if scale_tables != 1.0:
        # the 1e-3 hash-table init gives a nearly empty scene; scaled tables put real structure on every level
        with torch.no_grad():
            for p in om.parameters():
                if p.dim() == 2 and p.shape[1] == 2:
                    p.mul_(scale_tables)
    nm = nm.to(DEV)
    copy_model(nm, om)
    return om.train(), nm.train()


def _oracle_step(om, opt, step, rb, batch, jit):
    om.set_step(step)
    opt.zero_grad()
    out = om(rb, jitter=jit)
    loss = sum(om.get_loss_dict(out, batch).values())
    loss.backward()
    grads = {n: p.grad.detach().clone() for n, p in om.named_parameters() if p.grad is not None}
    opt.step()
    om.after_step(step)
    return out, loss.item(), grads


def _run_config(kind, R, steps, cfg, labels):
    om, nm = _pair(kind, **cfg)
    opt = torch.optim.Adam(om.parameters(), lr=1e-3, eps=1e-15)
    ops.set_mlp_precision("tc")
    tr = Trainer(nm, R, use_graph=True, graph_warmup=1, inject_jitter=True)
    assert tr.fused_loss
    report = []
    for step in range(steps):
        rb = O.synthetic_ray_bundle(R, seed=20 + step)
        g = torch.Generator().manual_seed(40 + step)
        batch = {"image": torch.rand(R, 3, generator=g)}
        if labels:
            batch["semantics_label"] = torch.randint(0, 19, (R,), generator=g)
        jit = make_jitter(R, seed=60 + step)
        o_out, o_loss, _ = _oracle_step(om, opt, step, rb, batch, jit)
        n_loss = tr.train_step(to_nk_bundle(rb, DEV), {k: v.to(DEV) for k, v in batch.items()},
                               jitter=[j.to(DEV) for j in jit]).item()
        n_out = tr.outputs
        if step >= 1:
            assert tr._graphs, "steps after the warm-up must be CUDA-graph replays"
        # level-0 bins: bit-exact (spacing and euclidean domain)
        o_rs0, n_rs0 = o_out["ray_samples_list"][0], n_out["ray_samples_list"][0]
        o_sb = torch.cat([o_rs0.spacing_starts[..., 0], o_rs0.spacing_ends[..., -1:, 0]], -1)
        o_eb = torch.cat([o_rs0.frustums.starts[..., 0], o_rs0.frustums.ends[..., -1:, 0]], -1)
        assert torch.equal(n_rs0.spacing_bins.cpu(), o_sb), f"step {step}: level-0 spacing bins differ"
        assert torch.equal(n_rs0.euclid_bins.cpu(), o_eb), f"step {step}: level-0 euclidean bins differ"
        e_loss = abs(n_loss - o_loss) / abs(o_loss)
        errs = {k: rel_err(n_out[k], o_out[k]) for k in ("rgb", "accumulation", "expected_depth", "depth", "semantics")
                if k in o_out and k in n_out}
        report.append((step, o_loss, n_loss, e_loss, errs))
        assert e_loss <= 1e-2, f"step {step}: loss {n_loss} vs oracle {o_loss}"
        for k in ("rgb", "accumulation", "expected_depth", "semantics"):
            if k in errs:
                assert errs[k] <= 1e-2, f"step {step}: {k} relative error {errs[k]:.3e}"
    for r in report:
        print(f"[{kind} R={R}] step {r[0]}: loss oracle {r[1]:.6f} nkbs {r[2]:.6f} (rel {r[3]:.2e}); "
              + ", ".join(f"{k} {v:.2e}" for k, v in r[4].items()))
    return om, nm, tr


def test_config1_nerfacto_4096_rays_tc_graph_trainer_tracks_oracle():
    """BASELINE configs[1] exactly as bench.py runs it (4096 rays, (256,96)+48, tcgen05 MLPs, graph replay, fused loss
    route), 4 optimisation steps against the oracle's Adam loop."""
    _run_config("nerfacto", 4096, 4, dict(), labels=False)


def test_config2_semantic_nerfw_8192_rays_tc_graph_trainer_tracks_oracle():
    """BASELINE configs[2]: 8192 rays, semantic + transient heads, default (8-level, 2^18, hidden 64) proposal fields."""
    _run_config("semantic_nerfw", 8192, 3, dict(use_semantic=True, use_transient_embedding=True), labels=True)


def _round_field_mlps_to_fp16(om):
    """The oracle with the weights of the main field's MLPs rounded to fp16 (what the tensor-core path feeds its
    GEMMs): the change of each gradient tensor under that perturbation is the tensor's mixed-precision noise floor."""
    import copy
    pm = copy.deepcopy(om)
    with torch.no_grad():
        for name, p in pm.field.named_parameters():
            if p.dim() == 2 and p.shape[1] != 2 and "embedding" not in name:
                p.copy_(p.half().float())
    return pm


def _all_grad_pairs(nm, om):
    pairs = [("field." + n, a, b) for n, a, b in field_grad_pairs(nm.field, om.field)]
    for i, (a, b) in enumerate(zip(nm.proposal_networks, om.proposal_networks)):
        pairs += [(f"prop{i}." + n, x, y) for n, x, y in density_grad_pairs(a, b)]
    return pairs


@pytest.mark.parametrize("kind,R", [("nerfacto", 4096), ("semantic_nerfw", 8192)])
def test_config_gradients_per_tensor(kind, R):
    """Gradients of the configurations above, per parameter tensor, accumulated into the step driver's arena by the
    backward kernels: relative L2 against the oracle's autograd.  fp32 MLP kernels: 2e-2 on every tensor (the bar of the
    fp32 end-to-end tests).  Mixed precision: 9e-2 (2x the 4.75e-2 measured on the main table by the round-1 driver
    smoke) or 4x the tensor's own mixed-precision noise floor, whichever is larger — the floor is what the ORACLE's
    gradient moves by when the field MLP weights are rounded to fp16: tensors whose gradient is a near-cancelling sum
    (proposal networks through the clipped interlevel loss, first-layer weights behind the table) move by tens of
    percent under that rounding in the oracle itself, so no 16-bit path can be held tighter there.  All numbers are
    printed."""
    cfg = dict() if kind == "nerfacto" else dict(use_semantic=True, use_transient_embedding=True)
    om, nm = _pair(kind, **cfg)
    rb = O.synthetic_ray_bundle(R, seed=20)
    g = torch.Generator().manual_seed(40)
    batch = {"image": torch.rand(R, 3, generator=g)}
    if kind != "nerfacto":
        batch["semantics_label"] = torch.randint(0, 19, (R,), generator=g)
    jit = make_jitter(R, seed=60)
    sum(om.get_loss_dict(om(rb, jitter=jit), batch).values()).backward()
    pm = _round_field_mlps_to_fp16(om)
    pm.zero_grad()
    sum(pm.get_loss_dict(pm(rb, jitter=jit), batch).values()).backward()
    floor = {n: _rel_l2(gp, go) for (n, gp), (_, go) in zip(
        [(n, p.grad) for n, p in pm.named_parameters() if p.grad is not None],
        [(n, p.grad) for n, p in om.named_parameters() if p.grad is not None])}
    # noise floor by tensor in the naming of the grad pairs: match through the oracle gradient tensor itself
    floor_by_id = {id(p.grad): floor[n] for n, p in om.named_parameters() if p.grad is not None}
    tr = Trainer(nm, R, use_graph=False, inject_jitter=True, keep_grads=True)  # parameters + gradients in the arena
    nb, nbatch = to_nk_bundle(rb, DEV), {k: v.to(DEV) for k, v in batch.items()}
    njit = [j.to(DEV) for j in jit]
    failures = []
    for prec in ("fp32", "tc"):
        ops.set_mlp_precision(prec)
        tr.zero_grad()
        out = nm(nb, jitter=njit)
        fused = nm.get_loss_total(out, nbatch) if hasattr(nm, "get_loss_total") else None
        loss = fused[0] if fused is not None else sum(nm.get_loss_dict(out, nbatch).values())
        loss.backward()
        torch.cuda.synchronize()
        for name, got, want in _all_grad_pairs(nm, om):
            if want is None:
                continue
            assert got is not None, name
            e = _rel_l2(got, want)
            fl = floor_by_id.get(id(want), 0.0)
            bar = 2e-2 if prec == "fp32" else max(9e-2, 4.0 * fl)
            print(f"[{kind} R={R} {prec}] grad {name}: relative L2 {e:.3e} (|oracle| {want.norm().item():.3e}, "
                  f"fp16-weight noise floor {fl:.3e}, bar {bar:.1e})")
            if e > bar:
                failures.append((prec, name, e, bar))
    assert not failures, failures


def test_config1_resampled_bins_against_oracle_fp32():
    """north_star "sample bins bit-exact" at configs[1], measured end to end with the fp32 kernels: level 0 (initial
    sampler) is bit-identical; the resampled levels inherit the last-ulp differences of the proposal densities (GPU vs
    CPU summation order inside the MLP), so they are compared in absolute spacing units and the bit-identical fraction is
    REPORTED (the PDF kernel itself is bit-exact on identical weights: tests/test_gpu_render.py, test_oracle_c.py)."""
    R = 4096
    om, nm = _pair("nerfacto")
    ops.set_mlp_precision("fp32")
    rb = O.synthetic_ray_bundle(R, seed=20)
    jit = make_jitter(R, seed=60)
    with torch.no_grad():
        want = om.proposal_sampler(om.collider(rb), om.density_fns, jitter=jit)
        got = nm.proposal_sampler(nm.collider(to_nk_bundle(rb, DEV)), density_fns=nm.density_fns,
                                  jitter=[j.to(DEV) for j in jit])
    pairs = [(0, got[2][0], want[2][0]), (1, got[2][1], want[2][1]), (2, got[0], want[0])]
    for lvl, g, w in pairs:
        gb, wb = g.spacing_bins.cpu(), O.ray_samples_to_sdist(w)
        same_rays = (gb == wb).all(dim=-1).float().mean().item()
        same_bins = (gb == wb).float().mean().item()
        err = (gb - wb).abs().max().item()
        print(f"[bins R={R} fp32] level {lvl}: {100 * same_rays:.1f} % of rays / {100 * same_bins:.1f} % of bin edges "
              f"bit-identical to the oracle, max |diff| {err:.2e} (spacing units in [0,1])")
        if lvl == 0:
            assert torch.equal(gb, wb)
        else:
            assert err <= 2e-5, (lvl, err)


# ------------------------------------------------------------------------------------------------
# X1: fixed-step run on an analytic scene, both arms, PSNR on held-out rays
# ------------------------------------------------------------------------------------------------
PSNR_STEPS = int(os.environ.get("NKBS_PSNR_STEPS", "300"))
PSNR_RAYS = 1024


def _slice(rb, a, b):
    return O.RayBundle(origins=rb.origins[a:b], directions=rb.directions[a:b], pixel_area=rb.pixel_area[a:b],
                       camera_indices=rb.camera_indices[a:b], metadata={})


def _heldout_psnr(model, rb, rgb, jit, dev=None):
    """(train-mode forward on held-out rays of the trained views with fixed jitter, eval-mode forward) PSNR."""
    with torch.no_grad():
        if dev is None:
            a = model(rb, jitter=jit)["rgb"]
            model.eval()
            b = model(rb)["rgb"]
        else:
            a = model(to_nk_bundle(rb, dev), jitter=[j.to(dev) for j in jit])["rgb"]
            model.eval()
            b = model(to_nk_bundle(rb, dev))["rgb"]
        model.train()
    return psnr(a, rgb), psnr(b, rgb)


@pytest.fixture(scope="module")
def psnr_oracle_run():
    """The oracle arm: PSNR_STEPS Adam steps of the CPU torch path on the analytic scene (about two minutes)."""
    torch.manual_seed(0)
    om = O.NerfactoModel(O.ModelConfig(**PSNR_CFG)).train()
    init = {k: v.detach().clone() for k, v in om.state_dict().items()}
    opt = torch.optim.Adam(om.parameters(), lr=1e-3, eps=1e-15)
    R, pool = PSNR_RAYS, 64
    rb_all, rgb_all = analytic_scene_rays(R * pool, seed=1)
    rb_te, rgb_te = analytic_scene_rays(4096, seed=2)
    losses = []
    for step in range(PSNR_STEPS):
        k = step % pool
        _, l, _ = _oracle_step(om, opt, step, _slice(rb_all, k * R, (k + 1) * R), {"image": rgb_all[k * R:(k + 1) * R]},
                               make_jitter(R, seed=1000 + step))
        losses.append(l)
    jit_te = make_jitter(4096, seed=5)
    return dict(init=init, rb_all=rb_all, rgb_all=rgb_all, rb_te=rb_te, rgb_te=rgb_te, jit_te=jit_te, losses=losses,
                psnr=_heldout_psnr(om, rb_te, rgb_te, jit_te), pool=pool)


@pytest.mark.parametrize("precision", ["fp32", "tc"])
def test_fixed_step_psnr_matches_oracle(psnr_oracle_run, precision):
    """north_star: "matching PSNR after a fixed-step run".  Same initial parameters, ray batches, targets and jitter
    draws in both arms; nkbs trains through the Trainer (CUDA-graph replay, fused losses, arena Adam)."""
    o = psnr_oracle_run
    torch.manual_seed(0)
    om0 = O.NerfactoModel(O.ModelConfig(**PSNR_CFG))
    om0.load_state_dict(o["init"])
    nm = nk.NerfactoModel(nk.ModelConfig(**PSNR_CFG)).to(DEV)
    copy_model(nm, om0)
    nm.train()
    ops.set_mlp_precision(precision)
    R = PSNR_RAYS
    tr = Trainer(nm, R, use_graph=True, graph_warmup=1, inject_jitter=True)
    rb_all = to_nk_bundle(o["rb_all"], DEV)
    rgb_all = o["rgb_all"].to(DEV)
    losses = []
    for step in range(PSNR_STEPS):
        k = step % o["pool"]
        rb = nk.RayBundle(rb_all.origins[k * R:(k + 1) * R], rb_all.directions[k * R:(k + 1) * R],
                          rb_all.pixel_area[k * R:(k + 1) * R], rb_all.camera_indices[k * R:(k + 1) * R])
        loss = tr.train_step(rb, {"image": rgb_all[k * R:(k + 1) * R]},
                             jitter=[j.to(DEV) for j in make_jitter(R, seed=1000 + step)])
        if step % 25 == 0 or step == PSNR_STEPS - 1:
            losses.append((step, loss.item()))
    got = _heldout_psnr(nm, o["rb_te"], o["rgb_te"], o["jit_te"], DEV)
    print(f"[psnr {precision}] {PSNR_STEPS} steps x {R} rays: held-out PSNR (train-mode, eval-mode) oracle "
          f"{o['psnr'][0]:.3f} / {o['psnr'][1]:.3f} dB, nkbs {got[0]:.3f} / {got[1]:.3f} dB; final loss oracle "
          f"{o['losses'][-1]:.5f}, nkbs {losses[-1][1]:.5f}")
    for step, l in losses:
        assert abs(l - o["losses"][step]) <= 5e-2 * abs(o["losses"][step]), (step, l, o["losses"][step])
    assert abs(got[0] - o["psnr"][0]) <= 0.3, (got, o["psnr"])
    assert abs(got[1] - o["psnr"][1]) <= 0.3, (got, o["psnr"])
    assert got[0] > 14.0, "the run must actually learn the scene"
