import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_kbs_b200 import ops
from tests.test_gpu_mlp_tc import CASES, _setup, _torch_mlp
from tests.util import rel_err
DEV = "cuda"
for fmt in (0, 1):
    for name in CASES:
        for N in (77, 128, 256, 4109):
            ops.set_mlp_precision("tc", fwd_fmt=fmt, bwd_fmt=fmt)
            dims, relu_mask, out_act, Ws, bs, x, sel = _setup(name, N, seed=1)
            g = torch.Generator().manual_seed(5)
            dy = torch.randn(N, dims[-1], generator=g).to(DEV)
            got = ops.mlp(x, Ws, bs, relu_mask, out_act, selector=sel)
            got.backward(dy)
            mine = [x.grad.clone()] + [w.grad.clone() for w in Ws] + [b.grad.clone() for b in bs]
            x.grad = None
            for p in Ws + bs:
                p.grad = None
            want = _torch_mlp(x, Ws, bs, relu_mask, out_act, sel if sel is not None else torch.ones(N, device=DEV))
            want.backward(dy)
            ref = [x.grad] + [w.grad for w in Ws] + [b.grad for b in bs]
            names = ["dx"] + [f"dW{i}" for i in range(len(Ws))] + [f"db{i}" for i in range(len(bs))]
            print(fmt, name, N, " ".join(f"{n}={rel_err(a, b):.1e}" for n, a, b in zip(names, mine, ref)))
