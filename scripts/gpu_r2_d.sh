#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_density.py tests/test_gpu_configs.py -m gpu -q -x --timeout 500 -k "not psnr" > gpurun_out/d_tests.log 2>&1
echo "tests rc=$?"; tail -4 gpurun_out/d_tests.log
R=8192 MODEL=semantic_nerfw timeout 300 python scripts/density_bench.py 2>&1 | tee gpurun_out/d_density.txt
R=65536 timeout 300 python scripts/density_bench.py 2>&1 | tee -a gpurun_out/d_density.txt
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-render --no-kernel-pass > gpurun_out/d_bench.json 2> gpurun_out/d_bench.err
python -c "
import json; d=json.load(open('gpurun_out/d_bench.json'))
print('cfg2', d['ms_per_step'], 'cfg3', d['cfg3']['ms_per_step'], d['cfg3']['e2e']['ms_per_step'], 'cfg5', d['cfg5']['ms_per_step'])"
