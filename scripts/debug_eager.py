import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import nerf_kbs_b200 as nk
from nerf_kbs_b200 import _lib
from nerf_kbs_b200.trainer import Trainer
from nerf_kbs_b200.rays import RayBundle
from bench import make_batches
dev = torch.device("cuda", 0)
torch.manual_seed(0)
m = nk.NerfactoModel(nk.ModelConfig()).to(dev).train()
R = 4096
b = {k: v.to(dev) for k, v in make_batches(1, R, 1, False, False)[0].items()}
rb = lambda: RayBundle(b["origins"], b["directions"], b["pixel_area"], b["camera_indices"])
tr = Trainer(m, R, use_graph=False)
for mode in ("plain", "observer"):
    if mode == "observer":
        import contextlib
        @contextlib.contextmanager
        def obs(name, a):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); yield; e.record()
        _lib.launch_observer = obs
    for i in range(3): tr.train_step(rb(), b)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(5):
        t1 = time.perf_counter(); tr.train_step(rb(), b); t2 = time.perf_counter()
        torch.cuda.synchronize(); t3 = time.perf_counter()
        print(mode, i, "cpu %.2f ms  +sync %.2f ms" % ((t2 - t1) * 1e3, (t3 - t2) * 1e3))
_lib.launch_observer = None
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for i in range(3): tr.train_step(rb(), b)
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(25)
