#!/usr/bin/env python
"""bench.py — train rays/s of the nerfacto ray-batch hot path (BASELINE.json configs[1]).

    python bench.py --gpus N --steps K --warmup W            # this repo's sm_100a path
    python bench.py --impl reference --gpus N --steps K ...   # the CPU torch path (oracle), host cores

A step = one training iteration over one synthetic ray batch: proposal sampling (256/96) + 48 final
samples, field, compositing, losses, backward, Adam (ref:train.py:42-101).  Prints ONE JSON line.
For N > 1 launch under torchrun: one rank per GPU, every rank steps its own batch of the same size
(weak scaling), gradients are all-reduced over NCCL before Adam.
"""
from __future__ import annotations

import argparse
import contextlib
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "train_rays_per_sec"
UNIT = "rays/s"
SAMPLES = (256, 96, 48)
L2_BYTES = 126 * 2**20


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="nkbs", choices=["nkbs", "reference"])
    ap.add_argument("--rays", type=int, default=4096, help="rays per GPU per step (configs[1]: 4096; configs[4]: 65536)")
    ap.add_argument("--model", default="nerfacto", choices=["nerfacto", "semantic_nerfw"])
    ap.add_argument("--no-graph", action="store_true", help="eager launches instead of one CUDA graph per step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-kernel-pass", action="store_true")
    ap.add_argument("--fp32-mlp", action="store_true", help="CUDA-core fp32 MLP kernels instead of the tcgen05 path")
    ap.add_argument("--pool", type=int, default=8, help="distinct ray batches cycled through")
    ap.add_argument("--no-render", action="store_true", help="skip the full-frame render leg (BASELINE configs[3])")
    ap.add_argument("--render-frames", type=int, default=10)
    return ap.parse_args()


def workload_name(args):
    if args.model == "nerfacto":
        return (f"nerfacto training step, {args.rays} rays/GPU, 16-level 2^19 hash grid, proposal 256/96 + 48 samples, "
                "MSE+interlevel+distortion losses, Adam lr 1e-3 eps 1e-15")
    return (f"semantic_nerfw training step (semantic + transient heads), {args.rays} rays/GPU, 8-level 2^18 proposal "
            "grids, proposal 256/96 + 48 samples, Adam lr 1e-3 eps 1e-15")


# ------------------------------------------------------------------------------------------------
# synthetic ray batches: origins U([-0.5,0.5]^3), unit directions, 115 images
# ------------------------------------------------------------------------------------------------
def make_batches(n, R, seed, labels, pin):
    import torch
    out = []
    for i in range(n):
        g = torch.Generator().manual_seed(seed * 1000 + i)
        o = torch.rand(R, 3, generator=g) - 0.5
        d = torch.randn(R, 3, generator=g)
        d = d / d.norm(dim=-1, keepdim=True)
        cam = torch.randint(0, 115, (R, 1), generator=g)
        b = {"origins": o, "directions": d, "camera_indices": cam, "pixel_area": torch.full((R, 1), 1e-6),
             "image": torch.rand(R, 3, generator=g)}
        if labels:
            b["semantics_label"] = torch.randint(0, 19, (R,), generator=g)
        if pin:
            b = {k: v.pin_memory() for k, v in b.items()}
        out.append(b)
    return out


# ------------------------------------------------------------------------------------------------
# clocks during the timed region
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def __enter__(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                       "-i", str(self.idx)], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None
        return self

    def __exit__(self, *a):
        if self.p is not None:
            time.sleep(0.15)
            self.p.terminate()
            try:
                self.p.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.p.kill()

    def summary(self):
        sm, mx, reasons = [], [], set()
        try:
            self.f.flush()
            rows = open(self.f.name).read().strip().splitlines()
        except OSError:
            rows = []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            c = [x.strip() for x in r.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])), mx.append(float(c[2]))
            except ValueError:
                continue
            for nme, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# per-kernel device time + algorithmic work (roofline numerators; formulas: DESIGN.md §5)
# ------------------------------------------------------------------------------------------------
def kernel_work(name, a):
    """(kind, amount) for one launch: kind 'hbm' -> algorithmic bytes, 'tensor' -> algorithmic flops."""
    def i32s(arr, n):
        return list(arr)[:n]
    if name == "nkbs_encode_fwd":
        N, L, F = a[7] * a[8], a[13], a[15]
        return "hbm", N * (12 + 8 * L * F * 4 + 4 * L * F)
    if name == "nkbs_encode_bwd":
        N, L, F = a[2], a[4], a[6]
        return "hbm", N * (12 + 4 * L * F + 8 * L * F * 4)
    if name in ("nkbs_mlp_fwd", "nkbs_mlp_bwd", "nkbs_mlp_tc_fwd", "nkbs_mlp_tc_bwd"):
        fwd = name.endswith("fwd")
        N, nl = (a[1], a[2]) if fwd else (a[2], a[3])
        dims = i32s(a[3] if fwd else a[4], nl + 1)
        mac = sum(dims[i] * dims[i + 1] for i in range(nl))
        return "tensor", N * mac * (2 if fwd else 4)  # bwd: data-gradient + weight-gradient contractions
    if name in ("nkbs_head_tc_fwd", "nkbs_head_tc_bwd"):
        fwd = name.endswith("fwd")
        N, nl = (a[12], a[13]) if fwd else (a[13], a[14])
        dims = i32s(a[14] if fwd else a[15], nl + 1)
        mac = sum(dims[i] * dims[i + 1] for i in range(nl))
        return "tensor", N * mac * (2 if fwd else 4)
    if name == "nkbs_composite_fwd":
        R, S, C_ = a[0], a[1], a[9]
        per = S * (4 * (a[2] is not None or a[3] is not None) + 8 * (a[4] is not None) + 12 * (a[7] is not None and a[13] is not None)
                   + 4 * C_ * (a[8] is not None and a[19] is not None)) + 4 * S * (a[12] is not None) + 4 * (3 + 1 + 1 + 1 + C_)
        return "hbm", R * per
    if name == "nkbs_composite_bwd":
        R, S, C_ = a[0], a[1], a[9]
        per = S * (4 + 4 + 8 + 12 * (a[7] is not None) + 4 * C_ * (a[8] is not None)) + S * (4 + 12 * (a[21] is not None) + 4 * C_ * (a[22] is not None))
        return "hbm", R * per
    if name == "nkbs_pdf_sample":
        R, S, nb = a[0], a[1], a[2]
        return "hbm", R * 4 * (S + (S + 1) + 2 * nb)
    if name == "nkbs_initial_sampler":
        return "hbm", a[0] * 4 * 2 * (a[1] + 1)
    if name in ("nkbs_assemble_fwd", "nkbs_assemble_bwd"):
        N, w = a[0], a[2] + a[3] + a[4]
        return "hbm", N * 4 * (w + a[3])
    if name == "nkbs_adam_step":
        return "hbm", 7 * 4 * sum(list(a[5])[:a[0]])
    if name in ("nkbs_density_fwd", "nkbs_density_bwd"):
        # fused proposal field: bin edges in, 8 corner rows of F=2 floats per level gathered, density out;
        # bwd additionally reads d(density) and scatters the same 8*L rows once each
        N, L = a[7] * a[8], a[13]
        gathers = 8 * L * 2 * 4
        return "hbm", N * ((8 + gathers + 4) if name.endswith("fwd") else (8 + 4 + 2 * gathers))
    if name in ("nkbs_ray_losses", "nkbs_ray_losses_weighted"):
        R, S, n_prop = a[0], a[1], a[5]
        sp = list(a[6])[:n_prop]
        per = 4 * (S + 1) + 8 * S + sum(4 * (x + 1) + 8 * x for x in sp) + 36 * (a[10] is not None)
        return "hbm", R * per
    return "hbm", 0


NCU_NAMES = {"nkbs_density_bwd": "density_bwd_kernel", "nkbs_density_fwd": "density_fwd_kernel", "nkbs_encode_bwd": "encode_bwd_kernel",
             "nkbs_encode_fwd": "encode_fwd_kernel", "nkbs_mlp_tc_bwd": "mlp_tc_bwd_kernel", "nkbs_mlp_tc_fwd": "mlp_tc_fwd_kernel",
             "nkbs_head_tc_bwd": "mlp_tc_bwd_kernel", "nkbs_head_tc_fwd": "mlp_tc_fwd_kernel", "nkbs_adam_step": "adam_kernel",
             "nkbs_composite_fwd": "composite_fwd_kernel", "nkbs_composite_bwd": "composite_bwd_kernel",
             "nkbs_ray_losses": "ray_losses_kernel", "nkbs_ray_losses_weighted": "ray_losses_kernel", "nkbs_pdf_sample": "pdf_sample"}
_NCU_BYTES = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def ncu_traffic(entry, rays):
    """DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) of `entry`'s kernel, from the committed
    `ncu --set full` capture of the 4096-ray step (profiles/*_top_kernels_ncu_full.csv); None for other workloads."""
    import csv
    import glob
    if rays != 4096 or entry not in NCU_NAMES:
        return None
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_top_kernels_ncu_full.csv")))
    if not files:
        return None
    rows = list(csv.reader(open(files[-1])))
    if len(rows) < 3:
        return None
    h, units = rows[0], rows[1]
    try:
        ik, ir, iw = h.index("Kernel Name"), h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum")
    except ValueError:
        return None
    vals = [float(r[ir]) * _NCU_BYTES.get(units[ir], 1.0) + float(r[iw]) * _NCU_BYTES.get(units[iw], 1.0)
            for r in rows[2:] if len(r) == len(h) and NCU_NAMES[entry] in r[ik]]
    return round(sum(vals) / len(vals)) if vals else None


def kernel_pass(trainer, batches, steps):
    """Eager (graph-less) steps with a CUDA-event pair around every nkbs launch, on the launching stream."""
    import torch
    from nerf_kbs_b200 import _lib
    from nerf_kbs_b200.rays import RayBundle
    rec = []

    @contextlib.contextmanager
    def observer(name, a):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        yield
        e.record()
        rec.append((name, kernel_work(name, a), s, e))

    was = trainer.use_graph
    trainer.use_graph = False
    # per-kernel times are taken with the proposal levels on the main stream: on their side streams they overlap the
    # main field's backward and an event pair around one launch would also time its neighbours
    sampler = trainer.model.proposal_sampler
    was_side, sampler.side_streams = sampler.side_streams, False
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    _lib.launch_observer = observer
    try:
        t0.record()
        for i in range(steps):
            b = batches[i % len(batches)]
            trainer.train_step(RayBundle(b["origins"], b["directions"], b["pixel_area"], b["camera_indices"]), b)
        t1.record()
        torch.cuda.synchronize()
    finally:
        _lib.launch_observer = None
        trainer.use_graph = was
        sampler.side_streams = was_side
    agg = {}
    for name, (kind, amount), s, e in rec:
        d = agg.setdefault(name, {"ms": 0.0, "launches": 0, "work": 0.0, "kind": kind})
        d["ms"] += s.elapsed_time(e)
        d["launches"] += 1
        d["work"] += amount
    return agg, t0.elapsed_time(t1) / steps


# ------------------------------------------------------------------------------------------------
# CPU torch path (the oracle): used for the cpu_baseline leg and for --impl reference
# ------------------------------------------------------------------------------------------------
def cpu_reference_run(args, steps, warmup, budget_s):
    """Training steps of the CPU oracle (the reference's pure-torch path restated, oracle/nerfstudio_torch.py) on
    all host cores.  Each step is a bounded sample: rays per step are cut so the whole run fits `budget_s`."""
    import torch
    from oracle import nerfstudio_torch as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    if args.model == "nerfacto":
        m = O.NerfactoModel(O.ModelConfig()).train()
    else:
        m = O.SemanticNerfWModel(O.ModelConfig(use_semantic=True, use_transient_embedding=True)).train()
    opt = torch.optim.Adam(m.parameters(), lr=1e-3, eps=1e-15)
    labels = args.model != "nerfacto"

    def one(R, i):
        b = make_batches(1, R, 77 + i, labels, False)[0]
        rb = O.RayBundle(origins=b["origins"], directions=b["directions"], pixel_area=b["pixel_area"],
                         camera_indices=b["camera_indices"], metadata={})
        t = time.perf_counter()
        m.set_step(i)
        opt.zero_grad()
        loss = sum(m.get_loss_dict(m(rb), b).values())
        loss.backward()
        opt.step()
        m.after_step(i)
        return time.perf_counter() - t

    one(256, 0)                      # page in / thread pools
    t_cal = one(256, 1)
    t_fix = 0.25                     # dense Adam over 19.4 M parameters, roughly constant per step
    per_ray = max(t_cal - t_fix, 0.02) / 256
    R = args.rays
    total = steps + warmup
    while R > 256 and total * (per_ray * R + t_fix) > budget_s:
        R //= 2
    for i in range(warmup):
        one(R, 2 + i)
    ts = [one(R, 2 + warmup + i) for i in range(steps)]
    dt = sum(ts)
    return {"value": R * steps / dt, "ms_per_step": 1e3 * dt / steps, "cores": cores, "rays_per_step": R,
            "sample": f"{steps} timed + {warmup} warm-up steps of {R} rays (of the {args.rays}-ray workload), fp32, "
                      f"torch {torch.__version__} CPU, {cores} threads"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = cpu_reference_run(args, args.steps, args.warmup, budget_s=150.0)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args), "rays_per_step_sampled": r["rays_per_step"]},
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


KITTI_H, KITTI_W = 375, 1242


def render_run(dev, frames, warmup, rank=0, world=1, chunk=1 << 16):
    """BASELINE configs[3]: full-frame 1242x375 inference (rgb, depth, semantics) from synthetic poses on a straight
    KITTI-like trajectory, rays generated on the device, 65 536-ray chunks, eval mode.  Returns rays/s over all ranks
    (chunks are dealt round-robin) and the end-to-end figure that includes the D2H copy of the three images."""
    import torch
    import torch.distributed as dist
    import nerf_kbs_b200 as nk
    torch.manual_seed(0)
    model = nk.NerfactoModel(nk.ModelConfig(use_semantic=True)).to(dev).eval()
    n_cam = 8
    c2w = torch.eye(4)[:3].repeat(n_cam, 1, 1)
    c2w[:, 2, 3] = torch.linspace(-0.4, 0.4, n_cam)  # drive along the optical axis
    f = torch.full((n_cam,), 721.5)
    cams = nk.Cameras(c2w, f, f.clone(), torch.full((n_cam,), 621.0), torch.full((n_cam,), 187.5), KITTI_H, KITTI_W)
    gen = nk.RayGenerator(cams).to(dev)
    palette = nk.SemanticPalette([((37 * i) % 256, (91 * i) % 256, (173 * i) % 256) for i in range(19)], dev)
    n_rays = KITTI_H * KITTI_W
    host = {k: torch.empty(s, dtype=t, pin_memory=True) for k, s, t in
            (("rgb", (KITTI_H, KITTI_W, 3), torch.float32), ("depth", (KITTI_H, KITTI_W, 1), torch.float32),
             ("semantics_colormap", (KITTI_H, KITTI_W, 3), torch.uint8))}

    def frame(i, to_host):
        out = nk.render_camera(model, gen, i % n_cam, num_rays_per_chunk=chunk, palette=palette, rank=rank, world_size=world)
        if to_host:
            for k, h in host.items():
                h.copy_(out[k], non_blocking=True)
            torch.cuda.current_stream().synchronize()
        return out

    def timed(to_host):
        for i in range(warmup):
            frame(i, to_host)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(frames):
            frame(i, to_host)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    ms = timed(False)
    ms_e2e = timed(True)
    d2h = sum(h.numel() * h.element_size() for h in host.values())
    return {"metric": "render_rays_per_sec", "value": n_rays * frames / (ms * 1e-3), "unit": UNIT,
            "ms_per_frame": ms / frames, "frames": frames, "rays_per_frame": n_rays, "chunk": chunk,
            "workload": "nerfacto + semantic head, 1242x375 frame, eval mode, rays generated on device, "
                        "outputs rgb + median depth + semantic logits/argmax colour map",
            "e2e": {"value": n_rays * frames / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_frame": ms_e2e / frames,
                    "h2d_bytes_per_step": 0, "d2h_bytes_per_step": d2h}}


def _finish(dist):
    """Leave a multi-rank run: final barrier, then exit without tearing NCCL down (destroy_process_group was seen
    to hang on a communicator that captured CUDA graphs still reference)."""
    import torch
    dist.barrier()
    torch.cuda.synchronize()
    sys.stdout.flush()
    sys.stderr.flush()
    os._exit(0)


# ------------------------------------------------------------------------------------------------
def run_nkbs(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the nkbs path has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    saved_stdout = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL prints its version banner on stdout: keep stdout clean for the one JSON line
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=dev)
    import nerf_kbs_b200 as nk
    from nerf_kbs_b200 import _lib
    from nerf_kbs_b200.rays import RayBundle
    from nerf_kbs_b200.trainer import Trainer

    from nerf_kbs_b200 import ops as nk_ops
    nk_ops.set_mlp_precision("fp32" if args.fp32_mlp else "tc")
    torch.manual_seed(0)  # identical parameters on every rank
    if args.model == "nerfacto":
        model = nk.NerfactoModel(nk.ModelConfig()).to(dev).train()
    else:
        model = nk.SemanticNerfWModel(nk.ModelConfig(use_semantic=True, use_transient_embedding=True)).to(dev).train()
    labels = args.model != "nerfacto"
    torch.cuda.manual_seed(1234 + rank)  # per-rank jitter
    R = args.rays
    host = make_batches(args.pool, R, 100 + rank, labels, pin=True)
    devb = [{k: v.to(dev) for k, v in b.items()} for b in host]
    tr = Trainer(model, R, use_graph=not args.no_graph)

    def bundle(b):
        return RayBundle(b["origins"], b["directions"], b["pixel_area"], b["camera_indices"])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # warm-up: eager steps, graph capture of both proposal-gating variants, then W replayed steps
    tr.load_batch(bundle(devb[0]), devb[0])
    tr.prepare()
    W = max(args.warmup, 3)
    for i in range(W):
        tr.train_step(bundle(devb[i % len(devb)]), devb[i % len(devb)])
    barrier()
    launches0 = tr.nkbs_calls

    # (1) device-resident throughput: K steps between two events
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        barrier()
        e0.record()
        for i in range(args.steps):
            b = devb[i % len(devb)]
            tr.train_step(bundle(b), b)
        e1.record()
        barrier()
    ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = t.item()
    value = world * R * args.steps / (ms * 1e-3)
    final_loss = float(tr.loss_dict["loss"])
    nk_launches = tr.nkbs_calls - launches0

    # (2) end to end through Trainer.train_step_host: pinned host batch -> H2D -> step -> D2H loss
    for i in range(3):
        tr.train_step_host(bundle(host[i % len(host)]), host[i % len(host)])
    barrier()
    e0.record()
    for i in range(args.steps):
        b = host[i % len(host)]
        tr.train_step_host(bundle(b), b)
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms = t.item()
    e2e = {"value": world * R * args.steps / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms / args.steps,
           "h2d_bytes_per_step": tr.h2d_bytes() + 4, "d2h_bytes_per_step": 4}

    # (2b) render rays/s (BASELINE configs[3]); every rank takes part (chunks are dealt round-robin)
    render = None
    if not args.no_render:
        del devb
        render = render_run(dev, args.render_frames, 2, rank, world)
        devb = [{k: v.to(dev) for k, v in b.items()} for b in host]

    if rank != 0:
        if world > 1:
            _finish(dist)
        return

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        peak_src = "measured (MEASURED_PEAKS.json)"
    except (OSError, ValueError):
        peak_src = "fallback (B200_PROFILING.md)"
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    tc_peak = float(peaks.get("bf16_tflops_sustained", 1400.0))

    # (3) per-kernel device time (eager pass, CUDA events on the launching stream) -> dominant kernel roofline
    roofline, kernels, launches_per_step = None, {}, None
    if not args.no_kernel_pass and world == 1:
        nsteps = 5
        c0 = sum(_lib.call_counts.values())
        agg, eager_ms = kernel_pass(tr, devb, nsteps)
        launches_per_step = (sum(_lib.call_counts.values()) - c0) / nsteps
        tot = sum(d["ms"] for d in agg.values())
        for name, d in sorted(agg.items(), key=lambda kv: -kv[1]["ms"]):
            per = d["ms"] / d["launches"]
            ach = d["work"] / d["launches"] / (per * 1e-3) / (1e9 if d["kind"] == "hbm" else 1e12)
            kernels[name] = {"ms_per_step": round(d["ms"] / nsteps, 4), "launches_per_step": d["launches"] / nsteps,
                             "share": round(d["ms"] / tot, 4), "bound": d["kind"],
                             "achieved": round(ach, 2), "unit": "GB/s" if d["kind"] == "hbm" else "TFLOP/s"}
        top = next(iter(kernels))
        k = kernels[top]
        peak = hbm_peak if k["bound"] == "hbm" else tc_peak
        roofline = {"kernel": top, "bound": k["bound"], "achieved": k["achieved"], "peak": peak, "unit": k["unit"],
                    "frac": round(k["achieved"] / peak, 4), "traffic": ncu_traffic(top, R),
                    "traffic_note": "DRAM bytes per launch from profiles/*_top_kernels_ncu_full.csv (ncu --set full, same step)",
                    "peak_source": peak_src,
                    "launch_ms": round(k["ms_per_step"] / k["launches_per_step"], 4),
                    "share_of_nkbs_kernel_time": k["share"], "eager_ms_per_step": round(eager_ms, 3)}

    # (4) the CPU torch path on this box's host cores, bounded sample
    cpu = None
    if not args.no_cpu_baseline and world == 1:
        r = cpu_reference_run(args, steps=2, warmup=1, budget_s=45.0)
        cpu = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]}

    working_set = sum(p.numel() * 4 for p in model.parameters()) * 4  # params + grads + 2 Adam moments
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": W,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f16/bf16 tensor-core MLPs (fp32 accumulate), f32 elsewhere" if not args.fp32_mlp else "f32",
        "data": "synthetic",
        "config": {"workload": workload_name(args), "rays_per_gpu": R, "global_rays_per_step": world * R,
                   "samples_per_ray": list(SAMPLES), "parallelism": f"dp{world}", "cuda_graph": not args.no_graph,
                   "l2": f"no flush: the step's working set (parameters + gradients + Adam moments = {working_set / 2**20:.0f} MiB, "
                         f"all streamed every step) exceeds the {L2_BYTES // 2**20} MiB L2; {args.pool} distinct ray batches rotate"},
        "e2e": e2e, "gpu_launches": nk_launches, "nkbs_entry_calls_per_step": launches_per_step,
        "roofline": roofline, "cpu_baseline": cpu, "render": render, "clocks": clocks.summary(), "kernels": kernels,
        "final_loss": final_loss,
    }
    if saved_stdout is not None:
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
    print(json.dumps(line))
    if world > 1:
        _finish(dist)


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_nkbs(a)
