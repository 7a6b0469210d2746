#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_density.py tests/test_gpu_level.py tests/test_gpu_model.py tests/test_gpu_configs.py -m gpu -q -x --timeout 600 -k "not psnr" > gpurun_out/m_tests.log 2>&1
echo "tests rc=$?"; tail -3 gpurun_out/m_tests.log
for rs in 0 1; do for R in 4096 65536; do
echo "== NKBS_RAY_SCATTER=$rs R=$R"; NKBS_RAY_SCATTER=$rs R=$R timeout 300 python scripts/density_bench.py 2>&1 | grep "fused=True"
done; done
bash scripts/gpu_ab.sh "NKBS_RAY_SCATTER=0" "NKBS_RAY_SCATTER=1"
bash scripts/gpu_ab.sh --args "--rays 65536 --steps 30" "NKBS_RAY_SCATTER=0" "NKBS_RAY_SCATTER=1"
