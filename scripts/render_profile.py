"""Per-kernel device time of one full-frame render (CUDA events around every nkbs launch)."""
import os, sys, contextlib, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import nerf_kbs_b200 as nk
from nerf_kbs_b200 import _lib
dev = torch.device("cuda")
torch.manual_seed(0)
model = nk.NerfactoModel(nk.ModelConfig(use_semantic=True)).to(dev).eval()
H, W, n_cam = 375, 1242, 4
c2w = torch.eye(4)[:3].repeat(n_cam, 1, 1); c2w[:, 2, 3] = torch.linspace(-0.4, 0.4, n_cam)
f = torch.full((n_cam,), 721.5)
cams = nk.Cameras(c2w, f, f.clone(), torch.full((n_cam,), 621.0), torch.full((n_cam,), 187.5), H, W)
gen = nk.RayGenerator(cams).to(dev)
pal = nk.SemanticPalette([((37 * i) % 256, (91 * i) % 256, (173 * i) % 256) for i in range(19)], dev)
for i in range(2): nk.render_camera(model, gen, i, palette=pal)
torch.cuda.synchronize()
rec = []
@contextlib.contextmanager
def obs(name, a):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); yield; e.record(); rec.append((name, s, e))
_lib.launch_observer = obs
t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0.record(); nk.render_camera(model, gen, 2, palette=pal); t1.record(); torch.cuda.synchronize()
_lib.launch_observer = None
agg = collections.Counter(); cnt = collections.Counter()
for n, s, e in rec: agg[n] += s.elapsed_time(e); cnt[n] += 1
print("frame ms (with events)", t0.elapsed_time(t1), "sum nkbs", sum(agg.values()))
for n, v in agg.most_common(): print(f"  {n:26s} {v:8.3f} ms  x{cnt[n]}")
