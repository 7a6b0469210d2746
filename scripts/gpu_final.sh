#!/bin/bash
# what the driver runs at round end: GPU tests (-x), smoke, bench at N=1
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
( time timeout 2400 python -m pytest tests -x -q -m gpu ) > gpurun_out/final_pytest.log 2>&1
echo "pytest rc=$?"; tail -6 gpurun_out/final_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1
echo "smoke rc=$?"; tail -3 gpurun_out/final_smoke.log
( time timeout 900 python bench.py ) > gpurun_out/final_bench.log 2>&1
echo "bench rc=$?"; tail -c 600 gpurun_out/final_bench.log
