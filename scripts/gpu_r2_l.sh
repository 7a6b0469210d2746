#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ops.py tests/test_gpu_render.py tests/test_gpu_model.py tests/test_gpu_level.py tests/test_golden.py tests/test_gpu_fullsize.py -m gpu -q -x --timeout 600 > gpurun_out/l_tests.log 2>&1
echo "tests rc=$?"; tail -3 gpurun_out/l_tests.log
python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-render > gpurun_out/l_bench.json 2>gpurun_out/l_bench.err
python - <<PY
import json
d=json.load(open("gpurun_out/l_bench.json"))
print("ms/step %.4f e2e %.4f" % (d["ms_per_step"], d["e2e"]["ms_per_step"]))
for n in ("nkbs_initial_sampler","nkbs_weights_resample_fwd"):
    print(n, d["kernels"][n]["launch_ms"], d["kernels"][n]["frac"], "| 65536:", d["cfg5"]["kernels"][n]["launch_ms"], d["cfg5"]["kernels"][n]["frac"])
print("cfg5", d["cfg5"]["ms_per_step"])
PY
