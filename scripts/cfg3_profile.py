"""Where the semantic-NeRF-W step (BASELINE configs[2], 8192 rays, semantic + transient heads) spends its device time:
every CUDA kernel of three eager steps from the torch profiler (CUPTI), grouped by name — nkbs kernels and the ATen
launches between them."""
import os, sys, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import nerf_kbs_b200 as nk
from nerf_kbs_b200.rays import RayBundle
from nerf_kbs_b200.trainer import Trainer

dev = torch.device("cuda")
R = int(os.environ.get("R", "8192"))
torch.manual_seed(0)
model = nk.SemanticNerfWModel(nk.ModelConfig(use_semantic=True, use_transient_embedding=True, num_images=8)).to(dev).train()
g = torch.Generator().manual_seed(1)
o = torch.randn(R, 3, generator=g) * 0.3
d = torch.nn.functional.normalize(torch.randn(R, 3, generator=g), dim=-1)
rb = RayBundle(o.to(dev), d.to(dev), torch.full((R, 1), 1e-6, device=dev), torch.randint(0, 8, (R, 1), generator=g).to(dev))
batch = {"image": torch.rand(R, 3, generator=g).to(dev), "semantics_label": torch.randint(0, 19, (R,), generator=g).to(dev)}
tr = Trainer(model, R, use_graph=False)
for _ in range(12):  # past step 10: the steady state (gated proposal updates) is what the bench times
    tr.train_step(rb, batch)
tr.model.proposal_sampler.side_streams = False
for _ in range(2):
    tr.train_step(rb, batch)
torch.cuda.synchronize()
N = 3
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(N):
        tr.train_step(rb, batch)
    torch.cuda.synchronize()
agg, cnt = collections.Counter(), collections.Counter()
for ev in prof.events():
    if ev.device_type == torch.autograd.DeviceType.CUDA:
        agg[ev.name] += ev.device_time if hasattr(ev, "device_time") else ev.cuda_time
        cnt[ev.name] += 1
tot = sum(agg.values())
print(f"device time per step {tot / N / 1e3:.3f} ms over {sum(cnt.values()) / N:.1f} kernels / memops")
ours = sum(v for k, v in agg.items() if "nkbs" in k or "_kernel" in k and "at::" not in k)
for k, v in agg.most_common(60):
    print(f"  {v / N:9.1f} us  x{cnt[k] / N:5.1f}  {k[:150]}")
tr.close()
