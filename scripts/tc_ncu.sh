#!/bin/bash
# ncu warp-state / source counters of the tensor-core MLP kernels at 3.1 M samples (scripts/tc_trace.py shapes)
# usage (on the GPU box): bash scripts/tc_ncu.sh base fwd
set -e
kind=${1:-base}; dir=${2:-fwd}
mkdir -p gpurun_out
ncu --set full --import-source on --clock-control none -k regex:mlp_tc_${dir} -s 1 -c 1 -o /tmp/tc_${kind}_${dir} \
  python scripts/tc_trace.py $kind $dir notrace > gpurun_out/tc_ncu_${kind}_${dir}.log 2>&1
ncu -i /tmp/tc_${kind}_${dir}.ncu-rep --page raw --csv > gpurun_out/tc_ncu_${kind}_${dir}_raw.csv
ncu -i /tmp/tc_${kind}_${dir}.ncu-rep --page source --csv > gpurun_out/tc_ncu_${kind}_${dir}_source.csv
ls -la gpurun_out/tc_ncu_${kind}_${dir}*
