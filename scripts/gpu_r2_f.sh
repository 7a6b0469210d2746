#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ops.py tests/test_gpu_density.py tests/test_gpu_model.py tests/test_gpu_fullsize.py -m gpu -q -x --timeout 500 > gpurun_out/f_tests.log 2>&1
echo "tests rc=$?"; tail -3 gpurun_out/f_tests.log
for R in 4096 65536; do
echo "== R=$R v4 paired reductions (default)" | tee -a gpurun_out/f_encode.txt
R=$R timeout 300 python scripts/encode_bench.py 2>&1 | tee -a gpurun_out/f_encode.txt
echo "== R=$R 8-byte reductions only (NKBS_SCATTER_V4=0 build)" | tee -a gpurun_out/f_encode.txt
NKBS_LIB=nerf_kbs_b200/build/libnkbs_nov4.so R=$R timeout 300 python scripts/encode_bench.py 2>&1 | tee -a gpurun_out/f_encode.txt
done
bash scripts/gpu_ab.sh "NKBS_LIB=nerf_kbs_b200/build/libnkbs_nov4.so" "NKBS_X=1"
