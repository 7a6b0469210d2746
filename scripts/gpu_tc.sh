#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_mlp_tc.py -m gpu -q -x 2>&1 > gpurun_out/tc_test_full.log; grep -E "AssertionError:|Error|passed|failed" gpurun_out/tc_test_full.log | sort | uniq | head -30
timeout 250 python scripts/mlp_bench.py 2>&1 | tail
timeout 100 python scripts/tc_trace.py head64 2>&1 | head -32
