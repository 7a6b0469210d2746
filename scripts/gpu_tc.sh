#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_mlp_tc.py -m gpu -q 2>&1 > gpurun_out/tc_test_full.log; grep -E "AssertionError:|passed|failed" gpurun_out/tc_test_full.log | sort | uniq | head -60
