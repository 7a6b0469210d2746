#!/bin/bash
# ncu --set full with source counters of the four tensor-core MLP launches of one 65 536-ray training step
# (base forward, colour-head forward, colour-head backward, base backward); raw + source pages exported as CSV.
set -e
mkdir -p gpurun_out
ncu --set full --import-source on --clock-control none -k regex:mlp_tc -s 8 -c 4 -o /tmp/tc_step \
  python bench.py --rays 65536 --steps 1 --warmup 3 --no-graph --no-cpu-baseline --no-kernel-pass --no-render > gpurun_out/tc_step.log 2>&1
ncu -i /tmp/tc_step.ncu-rep --page raw --csv > gpurun_out/tc_step_raw.csv
ncu -i /tmp/tc_step.ncu-rep --page source --csv > gpurun_out/tc_step_source.csv
ls -la gpurun_out/tc_step*
