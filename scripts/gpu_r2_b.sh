#!/bin/bash
# round 2, call B: new tests (fused level, config gradients) + whole suite + bench
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_level.py tests/test_gpu_configs.py -m gpu -q --timeout 600 -s -k "not psnr" > gpurun_out/b_new.log 2>&1
echo "new rc=$?"; tail -3 gpurun_out/b_new.log
timeout 1200 python -m pytest tests -m gpu -q --maxfail 12 --timeout 900 -k "not psnr and not test_gpu_configs and not test_gpu_level" > gpurun_out/b_pytest.log 2>&1
echo "suite rc=$?"; tail -3 gpurun_out/b_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/b_bench.json 2> gpurun_out/b_bench.err
echo "bench rc=$?"; tail -3 gpurun_out/b_bench.err
