#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_trainer.py tests/test_gpu_level.py -m gpu -q --timeout 600 -x > gpurun_out/c_tests.log 2>&1
echo "tests rc=$?"; tail -4 gpurun_out/c_tests.log
bash scripts/gpu_ab.sh "NKBS_FUSED_LEVEL=0 NKBS_DEFER_ADAM=0" "NKBS_FUSED_LEVEL=0 NKBS_DEFER_ADAM=1" "NKBS_FUSED_LEVEL=1 NKBS_DEFER_ADAM=1" "NKBS_FUSED_LEVEL=1 NKBS_DEFER_ADAM=1 NKBS_LEVEL_SHAPE=44" "NKBS_FUSED_LEVEL=1 NKBS_DEFER_ADAM=1 NKBS_LEVEL_SHAPE=28"
