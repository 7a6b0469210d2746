"""Per-phase clock64 trace of the tensor-core backward kernel (needs libnkbs_trace.so: scripts/build_trace.sh)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
NOTRACE = len(sys.argv) > 3 and sys.argv[3] == "notrace"  # product library, no trace read-back (for ncu)
if not NOTRACE:
    os.environ["NKBS_LIB"] = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "nerf_kbs_b200", "libnkbs_trace.so")
import torch
from nerf_kbs_b200 import ops, _lib
from tests.test_gpu_mlp_tc import CASES, _setup
name = sys.argv[1] if len(sys.argv) > 1 else "head64"
FWD_ONLY = len(sys.argv) > 2 and sys.argv[2] == "fwd"  # trace the forward kernel instead
N = 148 * (4 if FWD_ONLY else 2) * 128 * 6
ops.set_mlp_precision("tc")
dims, relu_mask, out_act, Ws, bs, x, sel = _setup(name, N)
dy = torch.rand(N, dims[-1], device="cuda")
for it in range(2):
    x.grad = None
    if FWD_ONLY:
        with torch.no_grad():
            y = ops.mlp(x, Ws, bs, relu_mask, out_act, selector=sel)
    else:
        y = ops.mlp(x, Ws, bs, relu_mask, out_act, selector=sel)
        y.backward(dy)
    torch.cuda.synchronize()
    if NOTRACE:
        continue
    lib = _lib.load()
    buf = (C.c_longlong * (2 * 4096))()
    n = (C.c_int * 2)()
    lib.nkbs_debug_trace(buf, n)
for role in (() if NOTRACE else (0, 1)):
    ev = [(buf[role * 4096 + 2 * i], buf[role * 4096 + 2 * i + 1]) for i in range(n[role])]
    if not ev:
        continue
    t0 = ev[0][1]
    print("role", role, "events", len(ev))
    prev = t0
    for tag, t in ev[:70]:
        print(f"  tag {tag:4d}  t={t - t0:8d}  dt={t - prev:7d}")
        prev = t
