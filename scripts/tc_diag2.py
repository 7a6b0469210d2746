import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_kbs_b200 import ops
from tests.test_gpu_mlp_tc import CASES, _setup, _torch_mlp
DEV = "cuda"
torch.set_printoptions(linewidth=200, precision=3, sci_mode=False)
for name, N in (("base", 77), ("head", 77), ("proposal_snw", 256)):
    ops.set_mlp_precision("tc", fwd_fmt=0, bwd_fmt=0)
    dims, relu_mask, out_act, Ws, bs, x, sel = _setup(name, N, seed=1)
    g = torch.Generator().manual_seed(5)
    dy = torch.randn(N, dims[-1], generator=g).to(DEV)
    got = ops.mlp(x, Ws, bs, relu_mask, out_act, selector=sel)
    got.backward(dy)
    mine = [x.grad.clone()] + [w.grad.clone() for w in Ws]
    x.grad = None
    for p in Ws + bs:
        p.grad = None
    want = _torch_mlp(x, Ws, bs, relu_mask, out_act, sel if sel is not None else torch.ones(N, device=DEV))
    want.backward(dy)
    ref = [x.grad] + [w.grad for w in Ws]
    e = (mine[0] - ref[0]).abs()
    scale = ref[0].abs().max()
    print(name, N, "dx err per row (max over cols)/scale:")
    print((e.max(1).values / scale)[:min(N, 128)])
    print("dx err per col:", e.max(0).values / scale)
    e = (mine[1] - ref[1]).abs() / ref[1].abs().max()
    print("dW0 err per out-row:", e.max(1).values)
    print("dW0 err per in-col:", e.max(0).values)
