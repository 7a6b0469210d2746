#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
python bench.py --steps 50 --warmup 10 --no-cpu-baseline --no-render > gpurun_out/n_bench.json 2>gpurun_out/n_bench.err
python - <<PY
import json
d=json.load(open("gpurun_out/n_bench.json"))
print("ms/step %.4f e2e %.4f" % (d["ms_per_step"], d["e2e"]["ms_per_step"]))
for n in ("nkbs_density_bwd","nkbs_encode_bwd_rays","nkbs_encode_bwd","nkbs_density_fwd"):
    k=d["kernels"][n]; k5=d["cfg5"]["kernels"][n]
    print(n, k["launch_ms"], k["launches_per_step"], k["frac"], "| 65536:", k5["launch_ms"], k5["launches_per_step"], k5["frac"])
PY
