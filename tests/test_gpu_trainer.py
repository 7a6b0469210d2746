"""GPU: the Adam kernel against torch.optim.Adam, and the captured-graph training step against the eager
step and against the CPU oracle's training loop."""
import pytest
import torch

from oracle import nerfstudio_torch as O
import nerf_kbs_b200 as nk
from nerf_kbs_b200.optim import Adam
from nerf_kbs_b200.trainer import Trainer
from tests.util import copy_model, make_jitter, rel_err, to_nk_bundle

pytestmark = pytest.mark.gpu
DEV = "cuda"


def test_adam_matches_torch_adam():
    g = torch.Generator().manual_seed(0)
    shapes = [(4096 * 3 + 5, 2), (64, 32), (16,), (1,), (300, 7)]
    ref = [torch.nn.Parameter(torch.randn(s, generator=g)) for s in shapes]
    mine = [torch.nn.Parameter(p.detach().clone().to(DEV)) for p in ref]
    o_ref = torch.optim.Adam(ref, lr=1e-3, eps=1e-15)
    o_mine = Adam(mine, lr=1e-3, eps=1e-15)
    for step in range(5):
        for p, q in zip(ref, mine):
            gr = torch.randn(p.shape, generator=g) * (10.0 ** (step - 2))
            if step == 3:
                gr[::2] = 0.0  # untouched hash-table rows: zero gradient, eps = 1e-15
            p.grad = gr
            q.grad = gr.to(DEV)
        o_ref.step()
        o_mine.step()
    for p, q in zip(ref, mine):
        assert rel_err(q, p) < 2e-6


def test_adam_skips_parameters_without_gradient():
    a = torch.nn.Parameter(torch.ones(10, device=DEV))
    b = torch.nn.Parameter(torch.ones(10, device=DEV))
    o = Adam([a, b])
    a.grad = torch.ones(10, device=DEV)
    o.step()
    assert torch.all(b == 1.0) and torch.all(a < 1.0)


def _train_pair(R, steps, use_graph):
    torch.manual_seed(0)
    om = O.NerfactoModel(O.ModelConfig()).train()
    nm = nk.NerfactoModel(nk.ModelConfig()).to(DEV).train()
    with torch.no_grad():
        for p in om.parameters():
            if p.dim() == 2 and p.shape[1] == 2:
                p.mul_(300.0)
    copy_model(nm, om)
    return om, nm


def test_graph_step_equals_eager_step():
    """Same parameters, same batch, same jitter stream: replayed steps must reproduce eager steps."""
    R = 512
    rb = O.synthetic_ray_bundle(R, seed=1)
    batch = {"image": torch.rand(R, 3, generator=torch.Generator().manual_seed(2)).to(DEV)}
    losses = {}
    for mode in (False, True):
        _, nm = _train_pair(R, 0, mode)
        tr = Trainer(nm, R, use_graph=mode, graph_warmup=2)
        torch.cuda.manual_seed(7)
        ls = []
        for _ in range(14):  # crosses step 10, where proposal gating starts skipping updates
            ls.append(tr.train_step(to_nk_bundle(rb, DEV), batch).item())
        losses[mode] = ls
        if mode:
            assert set(tr._graphs) == {True, False}
    # the jitter draws differ between eager and replayed RNG consumption only if the philox offsets diverge
    for a, b in zip(losses[False], losses[True]):
        assert abs(a - b) <= 2e-4 * abs(a), (losses[False], losses[True])


def test_fused_loss_route_equals_loss_dict_route():
    """The step driver's route (gradients weighted inside the loss kernel, total as the backward root) against the
    loss-dict route (torch weights, sums and rescales): same losses, same terms, same parameters after the steps."""
    R = 512
    rb = O.synthetic_ray_bundle(R, seed=5)
    batch = {"image": torch.rand(R, 3, generator=torch.Generator().manual_seed(6)).to(DEV)}
    runs = {}
    for fused in (False, True):
        _, nm = _train_pair(R, 0, False)
        tr = Trainer(nm, R, use_graph=False, fused_loss=fused)
        torch.cuda.manual_seed(11)
        ls = []
        for _ in range(6):
            ls.append(tr.train_step(to_nk_bundle(rb, DEV), batch).item())
        runs[fused] = (ls, {k: v.item() for k, v in tr.loss_dict.items()}, nm)
    for a, b in zip(runs[False][0], runs[True][0]):
        assert abs(a - b) <= 1e-5 * abs(a), (runs[False][0], runs[True][0])
    for k in ("rgb_loss", "interlevel_loss", "distortion_loss"):
        assert abs(runs[False][1][k] - runs[True][1][k]) <= 1e-5 * abs(runs[False][1][k]) + 1e-9, k
    pa, pb = runs[False][2], runs[True][2]
    # Adam (eps 1e-15) turns every gradient entry into a step of about lr, so rows whose gradient is rounding noise of
    # the atomics order move by +-lr in either run (as in test_training_loop_tracks_oracle): loose bar on parameters
    assert rel_err(pb.field.mlp_base_grid.hash_table, pa.field.mlp_base_grid.hash_table) < 2e-2
    assert rel_err(pb.proposal_networks[0].encoding.hash_table, pa.proposal_networks[0].encoding.hash_table) < 2e-2


def test_weighted_loss_kernel_scales_its_gradients():
    """nkbs_ray_losses_weighted: gradients = weight x the unweighted kernel's, terms unweighted, total = weighted sum."""
    from nerf_kbs_b200 import ops
    g = torch.Generator().manual_seed(0)
    R, S, Sp = 64, 48, 96
    def hist(n):
        e = torch.sort(torch.rand(R, n + 1, generator=g), dim=-1).values
        w = torch.rand(R, n, generator=g)
        return (w / w.sum(-1, keepdim=True)).to(DEV).requires_grad_(True), e.to(DEV)
    w, sd = hist(S)
    wp, sdp = hist(Sp)
    rgb = torch.rand(R, 3, generator=g).to(DEV).requires_grad_(True)
    img = torch.rand(R, 3, generator=g).to(DEV)
    tw = (1.0, 0.7, 0.002)
    mse, inter, dist, _ = ops.ray_losses(w_final=w, sdist_final=sd, w_props=[wp], sdist_props=[sdp], rgb=rgb, image=img)
    (tw[0] * mse + tw[1] * inter + tw[2] * dist).backward()
    want = [t.grad.clone() for t in (w, wp, rgb)]
    for t in (w, wp, rgb):
        t.grad = None
    total, terms = ops.ray_loss_total(w, sd, [wp], [sdp], rgb, img, tw)
    total.backward()
    assert abs(total.item() - (tw[0] * mse + tw[1] * inter + tw[2] * dist).item()) <= 1e-6 * abs(total.item())
    assert torch.allclose(terms, torch.stack([mse, inter, dist]).detach(), rtol=1e-6, atol=0)
    for got, ref in zip((w.grad, wp.grad, rgb.grad), want):
        assert rel_err(got, ref) < 1e-6
    # a non-unit seed is honoured when the caller says so
    for t in (w, wp, rgb):
        t.grad = None
    total, _ = ops.ray_loss_total(w, sd, [wp], [sdp], rgb, img, tw, unit_seed=False)
    (3.0 * total).backward()
    assert rel_err(rgb.grad, 3.0 * want[2]) < 1e-6


def test_training_loop_tracks_oracle():
    """Five optimisation steps with injected jitter: losses follow the CPU oracle's Adam loop."""
    R = 256
    om, nm = _train_pair(R, 0, False)
    opt = torch.optim.Adam(om.parameters(), lr=1e-3, eps=1e-15)
    tr = Trainer(nm, R, use_graph=False)
    rb = O.synthetic_ray_bundle(R, seed=3)
    image = torch.rand(R, 3, generator=torch.Generator().manual_seed(4))
    nrb = to_nk_bundle(rb, DEV)
    for step in range(5):
        jit = make_jitter(R, seed=10 + step)
        om.set_step(step)
        opt.zero_grad()
        lo = sum(om.get_loss_dict(om(rb, jitter=jit), {"image": image}).values())
        lo.backward()
        opt.step()
        om.after_step(step)
        # same draws on the device
        nm.set_step(step)
        for o in tr.optimizers.values():
            o.zero_grad()
        ln = sum(nm.get_loss_dict(nm(nrb, jitter=[j.to(DEV) for j in jit]), {"image": image.to(DEV)}).values())
        ln.backward()
        for o in tr.optimizers.values():
            o.step()
        nm.after_step(step)
        assert abs(ln.item() - lo.item()) <= 1e-3 * abs(lo.item()), (step, ln.item(), lo.item())
    # Adam normalises every gradient entry to a step of about lr: entries whose gradient is rounding noise move
    # by +-lr in either run, so the weights are compared loosely
    assert rel_err(nm.field.mlp_head.layers[0].weight, om.field.mlp_head.layers[0].weight) < 5e-2


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_fused_dp_step_equals_nccl_route():
    """torchrun over 2 GPUs: reduce-Adam-broadcast over NVLink peer memory vs NCCL allreduce + Adam (scripts/dp_check.py)."""
    import os, re, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", "29577", os.path.join(root, "scripts", "dp_check.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = re.findall(r"max\|dp\| = ([0-9.e+-]+) .*rank spread ([0-9.e+-]+); losses ([0-9.e+-]+) vs ([0-9.e+-]+)", r.stdout)
    assert len(lines) == 2
    for d, spread, l_fused, l_nccl in lines:
        # ranks hold bit-identical parameters; the two routes reach the same loss.  Parameters are compared loosely:
        # Adam (eps 1e-15) moves an entry whose gradient is summation-order noise by about lr per step in either
        # direction, so after the 8 steps of the script two correct runs differ by up to 8 lr on such entries
        assert float(spread) == 0.0
        assert abs(float(l_fused) - float(l_nccl)) <= 1e-5 * abs(float(l_nccl))
        assert float(d) < 8e-3
