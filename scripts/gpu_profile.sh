#!/bin/bash
# evidence pass: bench JSON, ncu launch list of the eager step, full ncu capture of the top kernels
mkdir -p gpurun_out
CMD="python bench.py --no-graph --no-cpu-baseline --no-kernel-pass --steps 2 --warmup 3"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 1400 -c 500 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "launch list exit $?"
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'mlp_tc_bwd|encode_bwd|encode_fwd|mlp_tc_fwd|composite_fwd' -s 60 -c 12 -o gpurun_out/top_kernels -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture exit $?"; tail -2 gpurun_out/ncu_full.log
