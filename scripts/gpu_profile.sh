#!/bin/bash
# evidence pass (one GPU): bench JSON lines, ncu launch list of the eager step, full ncu capture of the top kernels
mkdir -p gpurun_out
python bench.py --steps 100 --warmup 10 > gpurun_out/bench_4096.json 2> gpurun_out/bench_4096.err; echo "bench 4096 exit $?"
python bench.py --rays 65536 --steps 20 --warmup 3 --no-cpu-baseline --no-render > gpurun_out/bench_65536.json 2> gpurun_out/bench_65536.err; echo "bench 65536 exit $?"
python bench.py --model semantic_nerfw --rays 8192 --steps 30 --warmup 5 --no-cpu-baseline --no-render > gpurun_out/bench_snw_8192.json 2> gpurun_out/bench_snw.err; echo "bench snw exit $?"
CMD="python bench.py --no-graph --no-cpu-baseline --no-kernel-pass --no-render --steps 2 --warmup 3"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "launch list exit $?"
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none -k regex:'mlp_tc_bwd|mlp_tc_fwd|encode_bwd|encode_fwd|density_bwd|density_fwd|composite_fwd|composite_bwd|ray_losses|pdf_sample|adam' -s 150 -c 30 -o /tmp/top_kernels -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture exit $?"; tail -2 gpurun_out/ncu_full.log
# the report itself is too large to travel back (64 MiB cap): export the raw page as CSV and keep that
ncu -i /tmp/top_kernels.ncu-rep --page raw --csv > gpurun_out/top_kernels_raw.csv 2>/dev/null; ls -la gpurun_out
