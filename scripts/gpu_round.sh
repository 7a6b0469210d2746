#!/bin/bash
# full GPU pass: all parity tests (no -x), bench at 4096 and 65536 rays
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
grep -E "AssertionError:|Error|passed|failed" gpurun_out/pytest_gpu.log | sort | uniq | head -40
timeout 600 python bench.py --steps 100 --warmup 10 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
tail -3 gpurun_out/bench.err; cat gpurun_out/bench.json
timeout 300 python bench.py --rays 65536 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_64k.json 2> gpurun_out/bench_64k.err; echo "bench 64k exit $?"
tail -3 gpurun_out/bench_64k.err; cat gpurun_out/bench_64k.json
