import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_kbs_b200 import ops
from tests.util import rel_err
from tests.test_gpu_ops import _torch_mlp
for dims, relu, act in [([63,64,64,3],3,ops.ACT_SIGMOID), ([32,64,16],1,ops.ACT_DENSITY), ([15,64,64,19],1,ops.ACT_NONE)]:
    g = torch.Generator().manual_seed(1)
    N = 5000
    Ws = [(torch.randn(dims[i+1], dims[i], generator=g)/dims[i]**0.5).double().requires_grad_(True) for i in range(len(dims)-1)]
    bs = [(0.1*torch.randn(dims[i+1], generator=g)).double().requires_grad_(True) for i in range(len(dims)-1)]
    x = torch.randn(N, dims[0], generator=g).double().requires_grad_(True)
    sel = torch.ones(N).double()
    dy = torch.randn(N, dims[-1], generator=g).double()
    want = _torch_mlp(x, Ws, bs, relu, act, sel); want.backward(dy)
    # fp32 torch cpu
    W32 = [w.detach().float().requires_grad_(True) for w in Ws]; b32=[b.detach().float().requires_grad_(True) for b in bs]
    x32 = x.detach().float().requires_grad_(True)
    y32 = _torch_mlp(x32, W32, b32, relu, act, sel.float()); y32.backward(dy.float())
    Wg = [w.detach().float().cuda().requires_grad_(True) for w in Ws]; bg=[b.detach().float().cuda().requires_grad_(True) for b in bs]
    xg = x.detach().float().cuda().requires_grad_(True)
    got = ops.mlp(xg, Wg, bg, relu, act, selector=sel.float().cuda() if act==ops.ACT_DENSITY else None); got.backward(dy.float().cuda())
    print(dims, "y: cpu32 %.2e nkbs %.2e | dx: cpu32 %.2e nkbs %.2e" % (rel_err(y32,want), rel_err(got,want), rel_err(x32.grad,x.grad), rel_err(xg.grad,x.grad)))
    for i in range(len(Ws)):
        print("   dW%d cpu32 %.2e nkbs %.2e  db cpu32 %.2e nkbs %.2e" % (i, rel_err(W32[i].grad,Ws[i].grad), rel_err(Wg[i].grad,Ws[i].grad), rel_err(b32[i].grad,bs[i].grad), rel_err(bg[i].grad,bs[i].grad)))
