#!/usr/bin/env python
"""bench.py — train rays/s of the nerfacto ray-batch hot path (BASELINE.json configs[1]).

    python bench.py --gpus N --steps K --warmup W            # this repo's sm_100a path
    python bench.py --impl reference --gpus N --steps K ...   # the CPU torch path (oracle), host cores

A step = one training iteration over one synthetic ray batch: proposal sampling (256/96) + 48 final
samples, field, compositing, losses, backward, Adam (ref:train.py:42-101).  Prints ONE JSON line.
For N > 1 launch under torchrun: one rank per GPU, every rank steps its own batch of the same size
(weak scaling); the gradient exchange is fused with Adam in one kernel over NVLink peer (symmetric)
memory (reduce-scatter + Adam + all-gather, nkbs_dp_adam_step) and NCCL only carries barriers and the
timing reduce (NKBS_DP_FUSED=0 selects an NCCL allreduce + local Adam instead).

The headline `value` stays on BASELINE.json configs[1] (4096 rays per GPU).  The same JSON line carries
sub-legs for the other configurations: `cfg3` (configs[2]: semantic-NeRF-W, 8192 motion-masked rays with ray
generation and mask compaction inside the timed region), `render` (configs[3]: full 1242x375 frame),
`cfg5` (configs[4]: the nerfacto step at 65 536 rays per GPU at this run's N), and for N > 1 `dp_check`
(fused exchange vs the NCCL route, outside the timed regions).
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
L2_BYTES_PER_CYCLE = 6300  # whole-chip L2 (LTS) throughput cap, /opt/skills/guides/B300_MICROARCH.md
# Kernels whose tables are cache-resident are reported against L2 throughput, not HBM: the proposal grids are 5-16 MiB
# (L1/L2 hits, ncu: DRAM 2 %).  The main-grid gather and scatter (64 MiB table + 64 MiB gradient, evicted by the
# 620 MB optimizer stream of every step) and everything ray-major stay HBM kernels.
L2_BOUND = {"nkbs_density_fwd", "nkbs_density_fwd_save", "nkbs_density_bwd", "nkbs_proposal_level_fwd", "nkbs_encode_bwd_rays"}
BOUND_NOTE = {"l2": "table traffic of this kernel is served by L1/L2 (tables of 5-64 MiB against a 126 MiB L2): achieved = "
                    "algorithmic gather/scatter bytes per second against the L2 throughput cap, not HBM",
              "hbm": "achieved = algorithmic bytes per second against the measured HBM copy bandwidth",
              "tensor": "achieved = 2 (fwd) or 4 (bwd) x MAC flops per second against the sustained cuBLAS bf16 rate"}


def kernel_bound(name, kind):
    return "l2" if name in L2_BOUND else kind


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
    ap.add_argument("--no-legs", action="store_true", help="skip the cfg3 / cfg5 sub-legs and the dp_check")
    ap.add_argument("--leg-steps", type=int, default=0, help="timed steps of the cfg3 / cfg5 legs (default: min(--steps, 30))")
    return ap.parse_args()


def workload_name(args):
    if args.model == "nerfacto":
        return (f"nerfacto training step, {args.rays} rays/GPU, 16-level 2^19 hash grid, proposal 256/96 + 48 samples, "
                "MSE+interlevel+distortion losses, Adam lr 1e-3 eps 1e-15")
    return (f"semantic_nerfw training step (semantic + transient heads), {args.rays} rays/GPU, 8-level 2^18 proposal "
            "grids, proposal 256/96 + 48 samples, Adam lr 1e-3 eps 1e-15")


PARAM_WORDS = 16 * (1 << 19) * 2 + 2 * 5 * (1 << 17) * 2 + 21000  # nerfacto defaults: tables + MLPs + embedding


def bench_config(args, world):
    """The `config` object of the JSON line — the same for this repo's arm and for `--impl reference` (the driver
    compares them): what one step is, not how an arm executes it."""
    ws = PARAM_WORDS * 4 * 4 / 2**20  # parameters + gradients + two Adam moments
    return {"workload": workload_name(args), "rays_per_gpu": args.rays, "global_rays_per_step": world * args.rays,
            "samples_per_ray": list(SAMPLES), "parallelism": f"dp{world}", "collective": _collective(world),
            "l2": f"no flush: the step's working set (parameters + gradients + Adam moments = {ws:.0f} MiB, all streamed "
                  f"every step) exceeds the {L2_BYTES // 2**20} MiB L2; {args.pool} distinct ray batches rotate"}


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
    if name == "nkbs_weights_resample_fwd":
        # density in, both bin arrays of the level in, weights + median depth out, both bin arrays of the next level out
        R, S, nb = a[0], a[1], a[7]
        return "hbm", R * 4 * (S + 2 * (S + 1) + S + 1 + 2 * nb)
    if name == "nkbs_pdf_sample":
        R, S, nb = a[0], a[1], a[2]
        return "hbm", R * 4 * (S + (S + 1) + 2 * nb)
    if name == "nkbs_initial_sampler":
        return "hbm", a[0] * 4 * 2 * (a[1] + 1)
    if name in ("nkbs_assemble_fwd", "nkbs_assemble_bwd"):
        N, w = a[0], a[2] + a[3] + a[4]
        return "hbm", N * 4 * (w + a[3])
    if name == "nkbs_adam_step":
        # p, g, m, v read; p, m, v written; + g re-zeroed when the step driver asks for it
        return "hbm", (7 + (1 if a[13] else 0)) * 4 * sum(list(a[5])[:a[0]])
    if name in ("nkbs_density_fwd", "nkbs_density_bwd", "nkbs_density_fwd_save"):
        # fused proposal field: bin edges in, 8 corner rows of F=2 floats per level gathered, density out;
        # bwd additionally reads d(density) and scatters the same 8*L rows once each
        N, L = a[7] * a[8], a[13]
        gathers = 8 * L * 2 * 4
        if name.endswith("fwd"):
            return "hbm", N * (8 + gathers + 4)
        if name.endswith("fwd_save"):  # + features, normalised positions and selectors stored for the backward
            return "hbm", N * (8 + gathers + 4 + 2 * L * 4 + 16)
        if a[22] is None:  # feature gradients written out, scattered by nkbs_encode_bwd_rays
            return "hbm", N * (8 + 4 + gathers + 2 * L * 4)
        return "hbm", N * (8 + 4 + 2 * gathers)
    if name == "nkbs_proposal_level_fwd":
        # per sample: 8 corner rows of 2 floats on each of L levels gathered, density and weight written; per ray: the
        # level's bin edges (both domains) and the next level's
        R, S, L, nb = a[0], a[1], a[15], a[25]
        return "hbm", R * (S * (8 * L * 2 * 4 + 8) + 8 * (S + 1) + 8 * nb)
    if name == "nkbs_density_bwd_saved":  # features + selector + d(density) in, feature gradients out
        N, L = a[0], a[1]
        return "hbm", N * (2 * L * 4 + 8 + 2 * L * 4)
    if name == "nkbs_encode_bwd_rays":
        # feature gradients in, every corner row of every sample scattered once (upper bound before the in-register merge)
        N, L = a[0] * a[1], a[11]
        return "hbm", N * (8 + 2 * L * 4 + 8 * L * 2 * 4)
    if name == "nkbs_fixed_point_flush":  # deterministic mode: the 8-byte shadow is streamed once (writes are sparse)
        return "hbm", a[1] * 8
    if name in ("nkbs_ray_losses", "nkbs_ray_losses_weighted"):
        R, S, n_prop = a[0], a[1], a[5]
        sp = list(a[6])[:n_prop]
        per = 4 * (S + 1) + 8 * S + sum(4 * (x + 1) + 8 * x for x in sp) + 36 * (a[10] is not None)
        return "hbm", R * per
    return "hbm", 0


NCU_NAMES = {"nkbs_weights_resample_fwd": "composite_fwd_kernel", "nkbs_proposal_level_fwd": "proposal_level_fwd_kernel", "nkbs_density_bwd": "density_bwd_kernel", "nkbs_density_fwd": "density_fwd_kernel", "nkbs_encode_bwd": "encode_bwd_kernel",
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
    from nerf_kbs_b200 import ops as _ops
    was_split = _ops._BWD_SPLIT["fraction"]
    _ops.set_bwd_split(0.0)  # same reason: the split backward overlaps the scatter with the base MLP's backward
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    try:
        # one untimed eager step first: after graph replays the caching allocator has no free blocks for the eager
        # intermediates, and the first eager step pays for their cudaMalloc calls
        b = batches[0]
        for _ in range(2):
            trainer.train_step(RayBundle(b["origins"], b["directions"], b["pixel_area"], b["camera_indices"]), b)
        torch.cuda.synchronize()
        _lib.launch_observer = observer
        calls0 = sum(_lib.call_counts.values())
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
        _ops.set_bwd_split(was_split)
    agg = {}
    for name, (kind, amount), s, e in rec:
        d = agg.setdefault(name, {"ms": 0.0, "launches": 0, "work": 0.0, "kind": kind})
        d["ms"] += s.elapsed_time(e)
        d["launches"] += 1
        d["work"] += amount
    kernel_pass.calls_per_step = (sum(_lib.call_counts.values()) - calls0) / steps
    return agg, t0.elapsed_time(t1) / steps


def load_peaks(sm_mhz=None):
    """({"hbm", "l2", "tensor"} -> peak, source note, SM MHz).  HBM / tensor: MEASURED_PEAKS.json (driver-written) or the
    profiling guide's fallback; L2: ~6300 B/cycle over the chip (B300_MICROARCH.md) at the SM clock seen in this run."""
    pk = {}
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        src = "measured (MEASURED_PEAKS.json)"
    except (OSError, ValueError):
        src = "fallback (B200_PROFILING.md)"
    mhz = float(sm_mhz or pk.get("sm_max_mhz", 1965.0))
    return ({"hbm": float(pk.get("hbm_gbs", 6650.0)), "tensor": float(pk.get("bf16_tflops_sustained", 1400.0)),
             "l2": L2_BYTES_PER_CYCLE * mhz * 1e6 / 1e9}, src, mhz)


def kernel_table(agg, nsteps, PEAK):
    """Per-kernel lines of the JSON: device time per step and launch, algorithmic work per second against its roofline."""
    kernels = {}
    tot = sum(d["ms"] for d in agg.values())
    for name, d in sorted(agg.items(), key=lambda kv: -kv[1]["ms"]):
        per = d["ms"] / d["launches"]
        bound = kernel_bound(name, d["kind"])
        ach = d["work"] / d["launches"] / (per * 1e-3) / (1e12 if bound == "tensor" else 1e9)
        kernels[name] = {"ms_per_step": round(d["ms"] / nsteps, 4), "launches_per_step": d["launches"] / nsteps,
                         "launch_ms": round(per, 4), "share": round(d["ms"] / tot, 4), "bound": bound,
                         "achieved": round(ach, 2), "unit": "TFLOP/s" if bound == "tensor" else "GB/s",
                         "peak": round(PEAK[bound], 1), "frac": round(ach / PEAK[bound], 4)}
    return kernels


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
            "config": bench_config(args, int(os.environ.get("WORLD_SIZE", "1"))),
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"],
                             "rays_per_step_sampled": r["rays_per_step"]},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


KITTI_H, KITTI_W = 375, 1242


def render_run(dev, frames, warmup, rank=0, world=1, chunk=1 << 16):
    """BASELINE configs[3]: full-frame 1242x375 inference (rgb, depth, semantics) from synthetic poses on a straight
    KITTI-like trajectory, rays generated on the device, 65 536-ray chunks, eval mode.  Returns rays/s over all ranks
    (every rank renders a contiguous block of the rays, one all-gather merges them) and the end-to-end figure that includes the D2H copy of the three images."""
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
    """Leave a multi-rank run: every Trainer has dropped its captured graphs (Trainer.close: a graph that captured a
    collective keeps the communicator busy and destroy_process_group then waits forever), final barrier, NCCL torn
    down properly.  A watchdog turns a teardown that still does not return within 30 s into a plain exit, so a hung
    destroy can never cost the run its (already printed) result."""
    import gc
    import threading
    import torch
    gc.collect()
    dist.barrier()
    torch.cuda.synchronize()
    sys.stdout.flush()
    sys.stderr.flush()
    dog = threading.Timer(30.0, lambda: os._exit(0))
    dog.daemon = True
    dog.start()
    dist.destroy_process_group()
    dog.cancel()



COLLECTIVE = ("symmetric-memory reduce-scatter + Adam + all-gather in one kernel over NVLink (nkbs_dp_adam_step: peer "
              "loads / stores at N = 2, in-switch multimem.ld_reduce / multimem.st from N = 4); NCCL: barriers and the "
              "timing reduce only")


def _collective(world):
    if world == 1:
        return "none (1 GPU)"
    if os.environ.get("NKBS_DP_FUSED", "1") == "0":
        return "NCCL sum-allreduce of the flat gradient slice per parameter group, then local Adam (NKBS_DP_FUSED=0)"
    return COLLECTIVE


def _max_over_ranks(ms, dev, world):
    import torch
    import torch.distributed as dist
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def _barrier(world):
    import torch
    import torch.distributed as dist
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def timed_steps(step_fn, steps, dev, world):
    """K calls of step_fn(i) between two CUDA events, barrier + synchronize on both sides, max over ranks (ms)."""
    import torch
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    _barrier(world)
    e0.record()
    for i in range(steps):
        step_fn(i)
    e1.record()
    _barrier(world)
    return _max_over_ranks(e0.elapsed_time(e1), dev, world)


def cfg5_leg(args, dev, rank, world, steps, warmup, n1_value=None):
    """BASELINE configs[4]: the nerfacto training step at 65 536 rays per GPU, data parallel over this run's N GPUs."""
    import torch
    import nerf_kbs_b200 as nk
    from nerf_kbs_b200.rays import RayBundle
    from nerf_kbs_b200.trainer import Trainer
    R = 65536
    torch.manual_seed(0)
    model = nk.NerfactoModel(nk.ModelConfig()).to(dev).train()
    torch.cuda.manual_seed(4321 + rank)
    host = make_batches(4, R, 300 + rank, False, pin=True)
    devb = [{k: v.to(dev) for k, v in b.items()} for b in host]
    tr = Trainer(model, R)
    bundle = lambda b: RayBundle(b["origins"], b["directions"], b["pixel_area"], b["camera_indices"])
    dev_packed = [tr.pack_batch(bundle(b), b, device=dev) for b in devb]
    host_packed = [tr.pack_batch(bundle(b), b) for b in host]
    devb = devb[:1]
    tr.load_packed(dev_packed[0])
    tr.prepare()
    for i in range(max(warmup, 3)):
        tr.train_step(packed=dev_packed[i % 4])
    c0 = tr.nkbs_calls
    ms = timed_steps(lambda i: tr.train_step(packed=dev_packed[i % 4]), steps, dev, world)
    launches = tr.nkbs_calls - c0
    for i in range(3):
        tr.train_step_host(packed=host_packed[i % 4])
    def host_step(i):
        tr.train_step_host(packed=host_packed[i % 4], lag=1, prefetch=host_packed[(i + 1) % 4] if i + 1 < steps else None)
        if i == steps - 1:
            tr.drain_host()
    ms_e2e = timed_steps(host_step, steps, dev, world)
    final_loss = float(tr.loss_dict["loss"])
    kernels = None
    if world == 1 and not args.no_kernel_pass:
        # the rooflines of the large batch: north_star's >= 50 % targets for the encode / compositing kernels are meant
        # for batches that fill the machine
        agg, _ = kernel_pass(tr, devb, 3)
        kernels = kernel_table(agg, 3, load_peaks()[0])
    tr.close()
    return {"metric": METRIC, "unit": UNIT, "workload": "nerfacto training step, 65536 rays/GPU (BASELINE configs[4])",
            "kernels": kernels,
            "rays_per_gpu": R, "n_gpus": world, "steps": steps, "value": world * R * steps / (ms * 1e-3),
            "ms_per_step": ms / steps, "scaling": "weak", "collective": _collective(world), "gpu_launches": launches,
            "e2e": {"value": world * R * steps / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e / steps,
                    "h2d_bytes_per_step": tr.h2d_bytes() + 4, "d2h_bytes_per_step": 4},
            "final_loss": final_loss}


def cfg3_leg(args, dev, rank, world, steps, warmup):
    """BASELINE configs[2]: semantic-NeRF-W training step with semantic + transient heads on 8192 MOTION-MASKED rays.
    Every step draws candidate pixels, keeps those whose motion mask is non-zero (Bernoulli keep 0.8; nonzero = keep,
    ref:dataparser/dataset_z.py:186-191,226-232), compacts them to exactly 8192 (nkbs_compact_masked), generates the
    pinhole rays on the device (nkbs_generate_rays), gathers colour and label targets and runs the step — all inside
    the timed region.  e2e: the candidate list comes from pinned host memory, the loss goes back."""
    import torch
    import nerf_kbs_b200 as nk
    from nerf_kbs_b200 import data as D
    from nerf_kbs_b200.trainer import Trainer
    R, n_img, H, W = 8192, 8, KITTI_H, KITTI_W
    M = int(R * 1.5)
    torch.manual_seed(0)
    model = nk.SemanticNerfWModel(nk.ModelConfig(use_semantic=True, use_transient_embedding=True, num_images=n_img)).to(dev).train()
    torch.cuda.manual_seed(999 + rank)
    g = torch.Generator().manual_seed(500 + rank)
    c2w = torch.eye(4)[:3].repeat(n_img, 1, 1)
    c2w[:, 2, 3] = torch.linspace(-0.4, 0.4, n_img)
    f = torch.full((n_img,), 721.5)
    cams = nk.Cameras(c2w, f, f.clone(), torch.full((n_img,), 621.0), torch.full((n_img,), 187.5), H, W)
    gen = nk.RayGenerator(cams).to(dev)
    images = torch.rand(n_img, H * W, 3, generator=g).to(dev)
    labels = torch.randint(0, 19, (n_img, H * W), generator=g).to(dev)
    masks = (torch.rand(n_img, H * W, generator=g) < 0.8).to(torch.uint8).to(dev)
    pool = 8
    cand_host = [torch.stack([torch.randint(0, n_img, (M,), generator=g), torch.randint(0, H * W, (M,), generator=g)], 0).pin_memory()
                 for _ in range(pool)]
    cand_dev = [c.to(dev) for c in cand_host]
    stage = torch.empty(2, M, dtype=torch.int64, device=dev)
    tr = Trainer(model, R)
    kept_min = torch.full((1,), 1 << 30, device=dev, dtype=torch.int32)

    def step(cand):
        img_i, pix_i, kept = D.compact_masked(cand[0], cand[1], masks, R)
        torch.minimum(kept_min, kept, out=kept_min)
        idx = torch.stack([img_i, pix_i // W, pix_i % W], dim=-1)
        rb = gen(idx)
        batch = {"image": images[img_i, pix_i], "semantics_label": labels[img_i, pix_i]}
        return tr.train_step(rb, batch)

    step(cand_dev[0])
    tr.prepare()
    for i in range(max(warmup, 3)):
        step(cand_dev[i % pool])
    c0 = tr.nkbs_calls
    d0 = sum(nk._lib.call_counts.get(k, 0) for k in ("nkbs_compact_masked", "nkbs_generate_rays"))
    ms = timed_steps(lambda i: step(cand_dev[i % pool]), steps, dev, world)
    launches = tr.nkbs_calls - c0 + sum(nk._lib.call_counts.get(k, 0) for k in ("nkbs_compact_masked", "nkbs_generate_rays")) - d0

    loss_host = torch.zeros((), pin_memory=True)

    def step_host(i):
        stage.copy_(cand_host[i % pool], non_blocking=True)
        loss = step(stage)
        loss_host.copy_(loss, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for i in range(3):
        step_host(i)
    ms_e2e = timed_steps(step_host, steps, dev, world)
    final_loss = float(tr.loss_dict["loss"])
    kept = int(kept_min.item())
    kernels = None
    if world == 1 and not args.no_kernel_pass:
        # per-kernel rooflines of this configuration (hidden-64 proposal fields, semantic + transient heads)
        img_i, pix_i, _ = D.compact_masked(cand_dev[0][0], cand_dev[0][1], masks, R)
        rb = gen(torch.stack([img_i, pix_i // W, pix_i % W], dim=-1))
        kb = {"origins": rb.origins, "directions": rb.directions, "pixel_area": rb.pixel_area,
              "camera_indices": rb.camera_indices, "image": images[img_i, pix_i], "semantics_label": labels[img_i, pix_i]}
        agg, _ = kernel_pass(tr, [kb], 3)
        kernels = kernel_table(agg, 3, load_peaks()[0])
    tr.close()
    return {"metric": METRIC, "unit": UNIT,
            "workload": "semantic_nerfw training step (semantic + transient heads), 8192 motion-masked rays/GPU "
                        "(Bernoulli keep 0.8, candidates compacted and rays generated on the device inside the timed "
                        "region), 8-level 2^18 hidden-64 proposal fields (BASELINE configs[2])",
            "kernels": kernels,
            "rays_per_gpu": R, "candidates_per_step": M, "min_kept_candidates": kept, "n_gpus": world, "steps": steps,
            "value": world * R * steps / (ms * 1e-3), "ms_per_step": ms / steps, "scaling": "weak",
            "collective": _collective(world), "gpu_launches": launches,
            "e2e": {"value": world * R * steps / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e / steps,
                    "h2d_bytes_per_step": stage.numel() * 8, "d2h_bytes_per_step": 4},
            "final_loss": final_loss}


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

    # batches in the layout of the step's input buffer: one copy per step (D2D from the device-resident pool, H2D from
    # the pinned host pool)
    dev_packed = [tr.pack_batch(bundle(b), b, device=dev) for b in devb]
    host_packed = [tr.pack_batch(bundle(b), b) for b in host]
    # warm-up: eager steps, graph capture of both proposal-gating variants, then W replayed steps
    tr.load_batch(bundle(devb[0]), devb[0])
    tr.prepare()
    W = max(args.warmup, 3)
    for i in range(W):
        tr.train_step(packed=dev_packed[i % len(dev_packed)])
    barrier()
    launches0 = tr.nkbs_calls

    # (1) device-resident throughput: K steps between two events
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        barrier()
        e0.record()
        for i in range(args.steps):
            tr.train_step(packed=dev_packed[i % len(dev_packed)])
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
        tr.train_step_host(packed=host_packed[i % len(host_packed)])
    barrier()
    e0.record()
    for i in range(args.steps):
        # every step: one H2D transfer of its pinned batch, the step, and the D2H read of its loss; the read-back is
        # double-buffered (the host waits for the loss of step k-1 while step k runs), the last one is drained here
        # ... and the transfer of step k+1's batch is started while step k runs (copy stream, second device buffer)
        nxt = host_packed[(i + 1) % len(host_packed)] if i + 1 < args.steps else None
        tr.train_step_host(packed=host_packed[i % len(host_packed)], lag=1, prefetch=nxt)
    last_loss_host = tr.drain_host()
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms = t.item()
    e2e = {"value": world * R * args.steps / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms / args.steps,
           "h2d_bytes_per_step": tr.h2d_bytes() + 4, "d2h_bytes_per_step": 4,
           "readback": "loss of every step copied to pinned host memory; the host reads it one step late (double "
                       "buffered) and drains the last one inside the timed region",
           "h2d": "one transfer per step of the step's packed pinned batch; the transfer of step k+1 runs on a copy stream "
                  "while step k computes (the first one is not overlapped)", "last_loss": last_loss_host}

    # (2b) render rays/s (BASELINE configs[3]); every rank renders its block of every frame
    render = None
    if not args.no_render:
        del devb
        render = render_run(dev, args.render_frames, 2, rank, world)
        devb = [{k: v.to(dev) for k, v in b.items()} for b in host]

    clk = clocks.summary()
    PEAK, peak_src, sm_mhz = load_peaks(clk.get("sm_mhz"))
    hbm_peak, l2_peak = PEAK["hbm"], PEAK["l2"]

    # (3) per-kernel device time (eager pass, CUDA events on the launching stream) -> dominant kernel roofline
    roofline, kernels, launches_per_step = None, {}, None
    if not args.no_kernel_pass and world == 1:
        nsteps = 5
        agg, eager_ms = kernel_pass(tr, devb, nsteps)
        launches_per_step = kernel_pass.calls_per_step
        kernels = kernel_table(agg, nsteps, PEAK)
        top = next(iter(kernels))
        k = kernels[top]
        roofline = {"kernel": top, "bound": k["bound"], "achieved": k["achieved"], "peak": k["peak"], "unit": k["unit"],
                    "frac": k["frac"], "traffic": ncu_traffic(top, R),
                    "traffic_note": "DRAM bytes per launch from profiles/*_top_kernels_ncu_full.csv (ncu --set full, same step)",
                    "peak_source": peak_src if k["bound"] != "l2" else
                    f"L2: {L2_BYTES_PER_CYCLE} B/cycle (B300_MICROARCH.md) x {l2_peak / L2_BYTES_PER_CYCLE * 1e3:.0f} MHz SM clock of this run",
                    "bound_note": BOUND_NOTE.get(k["bound"], ""),
                    "launch_ms": k["launch_ms"], "share_of_nkbs_kernel_time": k["share"],
                    "eager_ms_per_step": round(eager_ms, 3)}
    if render is not None:
        # forward-only traffic of a rendered ray: hash-grid gathers 161 792 B + PDF 3408 B + compositing
        # 5092 B; the proposal tables stay in L1/L2, the 64 MiB main table is re-read from HBM by every 65 536-ray chunk
        per_ray = 161792 + 2440 + 968 + 5092
        ach = per_ray * render["value"] / 1e9
        render["roofline"] = {"bound": "hbm", "achieved": round(ach, 1), "peak": hbm_peak, "unit": "GB/s",
                              "frac": round(ach / hbm_peak, 4), "traffic": None,
                              "note": "whole render leg (all kernels of a frame), algorithmic gather + sampler + compositing "
                                      "bytes per ray x rays/s; the MLPs' tensor work (1.22 MFLOP/ray) is not counted"}

    # (2c) the other BASELINE configurations as sub-legs, and the data-parallel correctness check (every rank)
    legs = {}
    if not args.no_legs and args.model == "nerfacto" and R == 4096:
        tr.close()
        del devb
        tr = None
        torch.cuda.empty_cache()
        leg_steps = args.leg_steps or min(args.steps, 30)
        legs["cfg3"] = cfg3_leg(args, dev, rank, world, leg_steps, 5)
        torch.cuda.empty_cache()
        legs["cfg5"] = cfg5_leg(args, dev, rank, world, leg_steps, 5)
        torch.cuda.empty_cache()
        if world > 1:
            from nerf_kbs_b200.dp_check import run_dp_check
            legs["dp_check"] = run_dp_check(dev, rank, world, rays=1024, steps=8, graph=True)

    if tr is not None:
        tr.close()
    if rank != 0:
        if world > 1:
            _finish(dist)
        return

    # (4) the CPU torch path on this box's host cores, bounded sample
    cpu = None
    if not args.no_cpu_baseline and world == 1:
        r = cpu_reference_run(args, steps=2, warmup=1, budget_s=45.0)
        cpu = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": W,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f16/bf16 tensor-core MLPs (fp32 accumulate), f32 elsewhere" if not args.fp32_mlp else "f32",
        "data": "synthetic",
        "config": bench_config(args, world),
        "execution": {"cuda_graph": not args.no_graph, "collective": _collective(world),
                      "mlp": "fp32 CUDA cores" if args.fp32_mlp else "tcgen05 (fp16 forward / bf16 backward operands, fp32 accumulate)"},
        "e2e": e2e, "gpu_launches": nk_launches, "nkbs_entry_calls_per_step": launches_per_step,
        "roofline": roofline, "cpu_baseline": cpu, "render": render, "clocks": clk, "kernels": kernels,
        "final_loss": final_loss,
    }
    line.update(legs)
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
