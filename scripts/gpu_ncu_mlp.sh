#!/bin/bash
mkdir -p gpurun_out
export N=1048576 REPS=2
timeout 120 python scripts/mlp_bench.py head64 > gpurun_out/mlp_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mlp_tc_bwd -s 1 -c 1 -o gpurun_out/mlp_tc_bwd -f \
  python scripts/mlp_bench.py head64 > gpurun_out/mlp_ncu.log 2>&1
echo "exit $?"; cat gpurun_out/mlp_plain.log; tail -3 gpurun_out/mlp_ncu.log; ls -la gpurun_out/*.ncu-rep
