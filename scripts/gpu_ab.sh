#!/bin/bash
# quick pass after a kernel change: encode / model parity tests, then the kernel tables of both bench sizes
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
for R in 4096 65536; do
  timeout 300 python bench.py --rays $R --steps 50 --warmup 5 --no-cpu-baseline --no-render > gpurun_out/ab_$R.json 2> gpurun_out/ab_$R.err; echo "bench $R exit $?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/ab_$R.json").read().strip().splitlines()[-1])
print($R, "ms/step", round(d["ms_per_step"],4), "e2e", round(d["e2e"]["ms_per_step"],4))
for k,v in d["kernels"].items(): print("  ",k, v["ms_per_step"], v["achieved"], v["unit"])
PY
done
