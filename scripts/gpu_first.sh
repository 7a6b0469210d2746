#!/bin/bash
# first GPU pass: parity tests, bench, launch list
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 100 --warmup 10 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
tail -3 gpurun_out/bench.err; cat gpurun_out/bench.json
timeout 300 python bench.py --no-graph --steps 50 --warmup 5 --no-cpu-baseline --no-kernel-pass > gpurun_out/bench_nograph.json 2> gpurun_out/bench_nograph.err; echo "bench nograph exit $?"
cat gpurun_out/bench_nograph.json
timeout 300 python bench.py --rays 65536 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_64k.json 2> gpurun_out/bench_64k.err; echo "bench 64k exit $?"
cat gpurun_out/bench_64k.json
timeout 300 python bench.py --no-graph --no-cpu-baseline --no-kernel-pass --steps 2 --warmup 3 > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 900 -c 600 --csv --log-file gpurun_out/launches.csv \
  python bench.py --no-graph --no-cpu-baseline --no-kernel-pass --steps 2 --warmup 3 > gpurun_out/ncu.log 2>&1; echo "ncu exit $?"
tail -2 gpurun_out/ncu.log
