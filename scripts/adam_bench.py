"""nkbs_adam_step on the main field's parameter group (16.8 M table entries + MLPs): CUDA-event time per launch of the
default kernel (hardware sqrt / division approximations) and of the IEEE one (NKBS_ADAM_IEEE=1), and the largest
parameter difference between them after 5 steps."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_kbs_b200._lib import call, stream

dev = torch.device("cuda")
n = (1 << 19) * 16 * 2 + 40000
g0 = torch.Generator(device="cuda").manual_seed(0)
P0 = torch.randn(n, device=dev, generator=g0) * 1e-2
grads = [torch.randn(n, device=dev, generator=g0) * (10.0 ** -(3 + 2 * i)) * (torch.rand(n, device=dev, generator=g0) < 0.5) for i in range(3)]
one = lambda x: (C.c_void_p * 1)(x)
res = {}
for variant in ("1", "0"):
    os.environ["NKBS_ADAM_IEEE"] = variant
    P, M, V, G = P0.clone(), torch.zeros(n, device=dev), torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    t = torch.zeros((), device=dev)
    cnt = torch.zeros((), device=dev, dtype=torch.int32)
    def step():
        call("nkbs_adam_step", 1, one(P.data_ptr()), one(G.data_ptr()), one(M.data_ptr()), one(V.data_ptr()), (C.c_int64 * 1)(n),
             1e-3, 0.9, 0.999, 1e-15, 0.0, t.data_ptr(), None, 1, cnt.data_ptr(), stream())
    for i in range(5):
        G.copy_(grads[i % 3]); step()
    res[variant] = P.clone()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    times = []
    for i in range(30):
        G.copy_(grads[i % 3])
        s.record(); step(); e.record(); torch.cuda.synchronize()
        times.append(s.elapsed_time(e) * 1e3)
    times.sort()
    med = times[len(times) // 2]
    print(f"NKBS_ADAM_IEEE={variant}: median {med:.1f} us, min {times[0]:.1f} us  ({32 * n / med / 1e6:.2f} TB/s on 32 B/param)")
d = (res["0"] - res["1"]).abs()
print(f"default vs IEEE after 5 steps: max abs diff {d.max().item():.3e}, fraction different {(d > 0).float().mean().item():.4f}, rel L2 {(d.norm() / res['1'].norm()).item():.3e}")
