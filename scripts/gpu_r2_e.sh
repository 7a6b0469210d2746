#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_density.py tests/test_gpu_configs.py tests/test_gpu_model.py tests/test_gpu_level.py -m gpu -q -x --timeout 500 -k "not psnr" -s > gpurun_out/e_tests.log 2>&1
echo "tests rc=$?"; tail -4 gpurun_out/e_tests.log
echo "== H=64 fused backward, MMA route (default build)" | tee gpurun_out/e_density.txt
NKBS_FUSED_DENSITY_H64_BWD=1 R=8192 MODEL=semantic_nerfw timeout 300 python scripts/density_bench.py 2>&1 | grep "fused=True" | tee -a gpurun_out/e_density.txt
echo "== H=64 fused backward, transposed shared-memory sums (build with NKBS_DENSITY_BWD_MMA_MAXH=16)" | tee -a gpurun_out/e_density.txt
NKBS_LIB=nerf_kbs_b200/build/libnkbs_nomma64.so NKBS_FUSED_DENSITY_H64_BWD=1 R=8192 MODEL=semantic_nerfw timeout 300 python scripts/density_bench.py 2>&1 | grep "fused=True" | tee -a gpurun_out/e_density.txt
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-render --no-kernel-pass > gpurun_out/e_bench.json 2> gpurun_out/e_bench.err
python -c "
import json; d=json.load(open('gpurun_out/e_bench.json'))
print('cfg2', d['ms_per_step'], 'cfg3', d['cfg3']['ms_per_step'], d['cfg3']['e2e']['ms_per_step'], 'cfg5', d['cfg5']['ms_per_step'])"
