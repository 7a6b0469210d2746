#!/bin/bash
# state-of-the-path pass: GPU parity tests, per-shape MLP timings, encode backward sweep, bench at 4096 rays
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python scripts/mlp_bench.py proposal proposal_snw base head64 semantics transient > gpurun_out/mlp_bench.log 2>&1; cat gpurun_out/mlp_bench.log
N=1048576 timeout 300 python scripts/mlp_bench.py proposal base head64 > gpurun_out/mlp_bench_small.log 2>&1; cat gpurun_out/mlp_bench_small.log
R=65536 timeout 300 python scripts/encode_bench.py > gpurun_out/encode_bench.log 2>&1; cat gpurun_out/encode_bench.log
timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
tail -3 gpurun_out/bench.err; cat gpurun_out/bench.json
