#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_trainer.py tests/test_gpu_configs.py -m gpu -q -x --timeout 600 -k "not psnr" > gpurun_out/j_tests.log 2>&1
echo "tests rc=$?"; tail -3 gpurun_out/j_tests.log
bash scripts/gpu_ab.sh "NKBS_X=1"
bash scripts/gpu_ab.sh --args "--rays 65536 --steps 30" "NKBS_X=1"
