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


@pytest.mark.parametrize("ieee", ["0", "1"])
def test_adam_matches_torch_adam(ieee, monkeypatch):
    """Default kernel (hardware sqrt / division approximations on the update term) and the IEEE one."""
    monkeypatch.setenv("NKBS_ADAM_IEEE", ieee)
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


def test_adam_without_gradients_does_not_advance_the_step_count():
    """a group whose parameters all have ``grad is None`` (the proposal networks on a gated step) must
    leave the step count alone, as torch.optim.Adam does per parameter — the bias correction depends on it."""
    g = torch.Generator().manual_seed(1)
    ref = [torch.nn.Parameter(torch.randn(33, 5, generator=g))]
    mine = [torch.nn.Parameter(ref[0].detach().clone().to(DEV))]
    o_ref, o_mine = torch.optim.Adam(ref, lr=1e-3, eps=1e-15), Adam(mine, lr=1e-3, eps=1e-15)
    for step in range(9):
        if step % 3 == 1:  # gated step: no gradient at all
            ref[0].grad, mine[0].grad = None, None
        else:
            gr = torch.randn(33, 5, generator=g)
            ref[0].grad, mine[0].grad = gr, gr.to(DEV)
        o_ref.step()
        o_mine.step()
    assert float(o_mine.step_t) == 6.0 == float(o_ref.state[ref[0]]["step"])
    assert rel_err(mine[0], ref[0]) < 2e-6


def test_arena_adam_matches_torch_adam_with_gated_group():
    """The step driver's optimizer (flat arena, one launch per group, gradients re-zeroed by the kernel) against
    torch.optim.Adam with a group that skips every other step."""
    from nerf_kbs_b200.optim import ArenaAdam, ParamArena
    g = torch.Generator().manual_seed(2)
    shapes = {"proposal_networks": [(4099, 2), (16, 10), (16,)], "fields": [(8195, 2), (64, 32), (3,)]}
    ref = {k: [torch.nn.Parameter(torch.randn(s, generator=g)) for s in v] for k, v in shapes.items()}
    mine = {k: [torch.nn.Parameter(p.detach().clone().to(DEV)) for p in v] for k, v in ref.items()}
    o_ref = {k: torch.optim.Adam(v, lr=1e-3, eps=1e-15) for k, v in ref.items()}
    arena = ParamArena(mine)
    opt = ArenaAdam(arena)
    for step in range(6):
        active = ["fields"] + (["proposal_networks"] if step % 2 == 0 else [])
        for k in shapes:
            for p, q in zip(ref[k], mine[k]):
                if k in active:
                    gr = torch.randn(p.shape, generator=g)
                    p.grad = gr
                    q._nkbs_grad_sink.add_(gr.to(DEV))  # what an accumulating backward kernel does
                else:
                    p.grad = None
        for k in active:
            o_ref[k].step()
        opt.step([k for k in arena.layout if k in active])
        assert float(arena.G.abs().max()) == 0.0  # re-zeroed by the kernel that read it
    assert float(opt.step_t["proposal_networks"]) == 3.0 and float(opt.step_t["fields"]) == 6.0
    for k in shapes:
        for p, q in zip(ref[k], mine[k]):
            assert rel_err(q, p) < 2e-6
            assert q.grad is not None and q.grad.data_ptr() == q._nkbs_grad_sink.data_ptr()


def test_gated_steps_leave_proposal_optimizer_state_alone():
    """Trainer steps with the proposal gating off (captured graph variant included): the proposal group's step count,
    moments and parameters do not move; the field group advances."""
    R = 256
    _, nm = _train_pair(R, 0, True)
    tr = Trainer(nm, R, use_graph=True, graph_warmup=1)
    rb = O.synthetic_ray_bundle(R, seed=1)
    batch = {"image": torch.rand(R, 3, generator=torch.Generator().manual_seed(2)).to(DEV)}
    tr.train_step(to_nk_bundle(rb, DEV), batch)  # eager, updated (step 0 < 10)
    tr.flush()  # the field group's Adam is deferred into the next step: apply it now
    before = nm.proposal_networks[0].encoding.hash_table.detach().clone()
    t_prop, t_field = float(tr.opt.step_t["proposal_networks"]), float(tr.opt.step_t["fields"])
    nm.proposal_sampler.force_updated = False
    try:
        for _ in range(3):  # capture + replays of the gated variant
            tr.train_step(to_nk_bundle(rb, DEV), batch)
    finally:
        nm.proposal_sampler.force_updated = None
    tr.flush()
    assert float(tr.opt.step_t["proposal_networks"]) == t_prop
    assert float(tr.opt.step_t["fields"]) == t_field + 3
    assert torch.equal(nm.proposal_networks[0].encoding.hash_table.detach(), before)
    assert float(tr.arena.G.abs().max()) == 0.0


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
            assert set(tr._graphs) == {(True, True), (False, True)}  # (proposal gating, deferred Adam pending)
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
        tr.flush()
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
    total, terms, weighted = ops.ray_loss_total(w, sd, [wp], [sdp], rgb, img, tw)
    total.backward()
    assert torch.allclose(weighted, torch.tensor(tw, device=DEV) * terms, rtol=1e-6, atol=0)
    assert abs(total.item() - (tw[0] * mse + tw[1] * inter + tw[2] * dist).item()) <= 1e-6 * abs(total.item())
    assert torch.allclose(terms, torch.stack([mse, inter, dist]).detach(), rtol=1e-6, atol=0)
    for got, ref in zip((w.grad, wp.grad, rgb.grad), want):
        assert rel_err(got, ref) < 1e-6
    # a non-unit seed is honoured when the caller says so
    for t in (w, wp, rgb):
        t.grad = None
    total, _, _ = ops.ray_loss_total(w, sd, [wp], [sdp], rgb, img, tw, unit_seed=False)
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
        ln = sum(nm.get_loss_dict(nm(nrb, jitter=[j.to(DEV) for j in jit]), {"image": image.to(DEV)}).values())
        ln.backward()          # the kernels accumulate into the trainer's (zeroed) gradient arena
        tr.optimizer_step(True)  # Adam, which leaves the arena zeroed again
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
    import json
    lines = [json.loads(l[len("DP_CHECK "):]) for l in r.stdout.splitlines() if l.startswith("DP_CHECK ")]
    assert len(lines) == 3
    # deterministic accumulation on both routes, two ranks: bit-identical parameters
    assert lines[2]["deterministic"] and lines[2]["max_param_diff"] == 0.0, lines[2]
    for d in lines:
        # ranks hold bit-identical parameters; the two routes reach the same loss.  Parameters are compared loosely:
        # Adam (eps 1e-15) moves an entry whose gradient is summation-order noise by about lr per step in either
        # direction (nerf_kbs_b200/dp_check.py)
        assert d["rank_spread"] == 0.0 and d["ok"], d


def test_deferred_adam_equals_adam_at_step_end():
    """The field group's Adam overlapped with the next step's proposal levels (default) against Adam at the end of the
    step (NKBS_DEFER_ADAM=0): same updates in the same order, so the losses of every step and the parameters after a
    flush agree to rounding of the atomics order; an evaluation between steps sees flushed parameters."""
    import os
    R = 512
    rb = O.synthetic_ray_bundle(R, seed=8)
    batch = {"image": torch.rand(R, 3, generator=torch.Generator().manual_seed(9)).to(DEV)}
    runs = {}
    for defer in ("0", "1"):
        os.environ["NKBS_DEFER_ADAM"] = defer
        try:
            _, nm = _train_pair(R, 0, True)
            tr = Trainer(nm, R, use_graph=True, graph_warmup=1, inject_jitter=True)
        finally:
            os.environ.pop("NKBS_DEFER_ADAM", None)
        assert tr.defer_adam == (defer == "1")
        ls = []
        for i in range(8):
            ls.append(tr.train_step(to_nk_bundle(rb, DEV), batch, jitter=[j.to(DEV) for j in make_jitter(R, seed=30 + i)]).item())
            if i == 4:  # an evaluation between two steps must see the updated parameters
                nm.eval()
                with torch.no_grad():
                    ev = nm(to_nk_bundle(rb, DEV))["rgb"].clone()
                nm.train()
                assert not tr._pending
        tr.flush()
        runs[defer] = (ls, ev, nm.field.mlp_head.layers[0].weight.detach().clone(), nm.field.mlp_base_grid.hash_table.detach().clone())
    for a, b in zip(runs["0"][0], runs["1"][0]):
        assert abs(a - b) <= 2e-4 * abs(a), (runs["0"][0], runs["1"][0])
    assert rel_err(runs["1"][1], runs["0"][1]) < 2e-3
    assert rel_err(runs["1"][2], runs["0"][2]) < 5e-2   # Adam eps = 1e-15 noise, see test_training_loop_tracks_oracle
    assert rel_err(runs["1"][3], runs["0"][3]) < 2e-2


def test_packed_batch_equals_tensor_batch():
    """A batch packed in the layout of the step's input buffer (one copy: pinned host -> device, or device -> device)
    feeds the step the same inputs as the per-tensor route, host tensors included."""
    R = 256
    rb = O.synthetic_ray_bundle(R, seed=11)
    img = torch.rand(R, 3, generator=torch.Generator().manual_seed(12))
    _, nm = _train_pair(R, 0, False)
    tr = Trainer(nm, R, use_graph=False)
    nrb_host = nk.RayBundle(rb.origins, rb.directions, rb.pixel_area, rb.camera_indices)
    host_packed = tr.pack_batch(nrb_host, {"image": img})
    assert host_packed.is_pinned() and host_packed.dtype == torch.uint8
    dev_packed = tr.pack_batch(to_nk_bundle(rb, DEV), {"image": img.to(DEV)}, device=DEV)
    want = {}
    tr.load_batch(to_nk_bundle(rb, DEV), {"image": img.to(DEV)})
    torch.cuda.synchronize()
    want = {k: v.clone() for k, v in tr.buf.items()}
    for route in ("host tensors", "host packed", "device packed"):
        tr._flat.zero_()
        if route == "host tensors":
            tr.load_batch(nrb_host, {"image": img})
        else:
            tr.load_packed(host_packed if route == "host packed" else dev_packed)
        torch.cuda.synchronize()
        for k, v in want.items():
            assert torch.equal(tr.buf[k], v), (route, k)
    # prefetched on the copy stream, consumed by identity
    tr._flat.zero_()
    tr.prefetch_packed(host_packed)
    tr.load_packed(host_packed)
    torch.cuda.synchronize()
    for k, v in want.items():
        assert torch.equal(tr.buf[k], v), ("prefetched", k)
    torch.cuda.manual_seed(3)
    a = tr.train_step_host(packed=host_packed)
    assert a == a and tr.h2d_bytes() == R * (12 + 12 + 8 + 4 + 12)


def test_lagged_loss_readback_returns_every_loss_one_step_late():
    R = 256
    rb = O.synthetic_ray_bundle(R, seed=13)
    batch = {"image": torch.rand(R, 3, generator=torch.Generator().manual_seed(14)).to(DEV)}
    runs = []
    for lag in (0, 1):
        _, nm = _train_pair(R, 0, False)
        tr = Trainer(nm, R, use_graph=True, graph_warmup=1, inject_jitter=True)
        ls = []
        for i in range(6):
            tr.train_step(jitter=[j.to(DEV) for j in make_jitter(R, seed=50 + i)]) if False else None
            for dst, src in zip(tr.jitter, make_jitter(R, seed=50 + i)):
                dst.copy_(src.reshape(dst.shape))
            ls.append(tr.train_step_host(to_nk_bundle(rb, DEV), batch, lag=lag))
        if lag:
            assert ls[0] != ls[0]  # NaN: nothing to report yet
            ls = ls[1:] + [tr.drain_host()]
        runs.append(ls)
    for a, b in zip(*runs):
        assert abs(a - b) <= 2e-4 * abs(a), runs


def test_state_dict_sees_the_deferred_update():
    """A checkpoint taken right after train_step includes the optimizer update the driver defers into the next step."""
    R = 256
    rb = O.synthetic_ray_bundle(R, seed=15)
    batch = {"image": torch.rand(R, 3, generator=torch.Generator().manual_seed(16)).to(DEV)}
    _, nm = _train_pair(R, 0, False)
    tr = Trainer(nm, R, use_graph=False)
    before = nm.field.mlp_head.layers[0].weight.detach().clone()
    tr.train_step(to_nk_bundle(rb, DEV), batch)
    assert tr._pending and torch.equal(nm.field.mlp_head.layers[0].weight.detach(), before)
    sd = nm.state_dict()
    assert not tr._pending
    assert not torch.equal(sd["field.mlp_head.layers.0.weight"], before)
    tr.close()
    assert "field.mlp_base_grid.hash_table" in nm.state_dict()


def test_split_backward_of_base_mlp_and_hash_grid_equals_the_single_launches():
    """ops._BWD_SPLIT: the base MLP's backward in two launches with the first part's gradient scatter on a side stream
    (default) against one launch each (fraction 0).  Same sums, other partial-sum grouping in the MLP kernel: losses
    agree to rounding, parameters at the Adam-noise bar of the neighbouring tests; the split route itself is
    reproducible bit for bit in deterministic mode, eager against captured."""
    from nerf_kbs_b200 import ops, _lib
    R = 2048  # 98 304 samples: above the split threshold
    rb = O.synthetic_ray_bundle(R, seed=3)
    batch = {"image": torch.rand(R, 3, generator=torch.Generator().manual_seed(4)).to(DEV)}
    prev = ops._BWD_SPLIT["fraction"]
    runs = {}
    try:
        for key, frac, graph in (("single", 0.0, False), ("split", 0.5, False), ("split_graph", 0.5, True)):
            ops.set_bwd_split(frac)
            _, nm = _train_pair(R, 0, graph)
            tr = Trainer(nm, R, use_graph=graph, graph_warmup=1, inject_jitter=True, deterministic=True)
            c0 = _lib.call_counts.get("nkbs_encode_bwd", 0)
            ls = [tr.train_step(to_nk_bundle(rb, DEV), batch, jitter=[j.to(DEV) for j in make_jitter(R, seed=50 + i)]).item()
                  for i in range(4)]
            calls = _lib.call_counts.get("nkbs_encode_bwd", 0) - c0
            tr.flush()
            runs[key] = (ls, {n: p.detach().clone() for n, p in nm.named_parameters()}, calls)
            tr.close()
    finally:
        ops.set_bwd_split(prev)
    assert runs["split"][2] == runs["single"][2] + 4, "the split route launches the main grid's scatter twice per step"
    for n, p in runs["split"][1].items():
        assert torch.equal(runs["split_graph"][1][n], p), (n, (runs["split_graph"][1][n] - p).abs().max().item())
    for a, b in zip(runs["split"][0], runs["single"][0]):
        assert abs(a - b) <= 2e-4 * abs(b), (runs["split"][0], runs["single"][0])
    for n in ("field.mlp_base_grid.hash_table", "field.mlp_head.layers.0.weight"):
        assert rel_err(runs["split"][1][n], runs["single"][1][n]) < 5e-2, n
