#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_density.py tests/test_gpu_configs.py tests/test_gpu_ops.py -m gpu -q -x --timeout 600 -k "not psnr" > gpurun_out/o_tests.log 2>&1
echo "tests rc=$?"; tail -3 gpurun_out/o_tests.log
bash scripts/gpu_ab.sh --args "--model semantic_nerfw --rays 8192 --steps 50" "NKBS_RAY_SCATTER=0" "NKBS_RAY_SCATTER=1"
