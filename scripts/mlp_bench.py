"""Micro-benchmark of the MLP kernels (CUDA events): python scripts/mlp_bench.py [case ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_kbs_b200 import ops
from tests.test_gpu_mlp_tc import CASES, _setup
DEV = "cuda"
SIZES = {"proposal": 65536 * 256, "proposal_snw": 65536 * 96, "base": 65536 * 48, "head": 65536 * 48, "head64": 65536 * 48,
         "semantics": 65536 * 48, "transient": 65536 * 48, "linear": 65536 * 96}
names = sys.argv[1:] or ["proposal", "base", "head64", "semantics"]
reps = int(os.environ.get("REPS", "5"))
for prec in ("tc", "fp32") if os.environ.get("FP32") else ("tc",):
    for name in names:
        N = int(os.environ.get("N", SIZES[name]))
        ops.set_mlp_precision(prec)
        dims, relu_mask, out_act, Ws, bs, x, sel = _setup(name, N)
        dy = torch.rand(N, dims[-1], device=DEV)
        def fwd():
            return ops.mlp(x, Ws, bs, relu_mask, out_act, selector=sel)
        y = fwd(); y.backward(dy); torch.cuda.synchronize()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        tf = tb = 0.0
        for _ in range(reps):
            x.grad = None
            for p in Ws + bs: p.grad = None
            e[0].record(); y = fwd(); e[1].record(); y.backward(dy); e[2].record(); torch.cuda.synchronize()
            tf += e[0].elapsed_time(e[1]); tb += e[1].elapsed_time(e[2])
        mac = sum(dims[i] * dims[i + 1] for i in range(len(dims) - 1))
        print(f"{prec} {name} N={N}: fwd {tf/reps:.3f} ms ({2*mac*N/(tf/reps)/1e9:.1f} TFLOP/s, {N/(tf/reps)/1e6:.2f} Gsamp/s)  "
              f"bwd {tb/reps:.3f} ms ({4*mac*N/(tb/reps)/1e9:.1f} TFLOP/s, {N/(tb/reps)/1e6:.2f} Gsamp/s)")
