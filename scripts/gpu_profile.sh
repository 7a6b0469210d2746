#!/bin/bash
# evidence pass (one GPU), one part per gpurun call (at most one profiler pass per call):
#   bench : the bench JSON line (all legs) and the reference arm
#   list  : ncu launch list of the eager step (after the same command has exited 0 without ncu)
#   full  : full ncu capture of the top kernels (same)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
CMD="python bench.py --no-graph --no-cpu-baseline --no-kernel-pass --no-render --no-legs --steps 2 --warmup 3"
case "${1:-bench}" in
bench)
  python bench.py --steps 200 --warmup 20 > gpurun_out/bench_4096.json 2> gpurun_out/bench_4096.err; echo "bench 4096 exit $?"
  python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "bench reference exit $?"
  tail -c 400 gpurun_out/bench_4096.json
  ;;
list)
  timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
  echo "launch list exit $?"; wc -l gpurun_out/launches.csv
  ;;
full)
  timeout 300 $CMD > gpurun_out/plain2.log 2>&1 && \
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'mlp_tc_bwd|mlp_tc_fwd|encode_bwd|encode_fwd|density_bwd|density_fwd|composite_fwd|composite_bwd|ray_losses|pdf_sample|adam|initial_sampler' -s 110 -c 46 -o /tmp/top_kernels -f $CMD > gpurun_out/ncu_full.log 2>&1
  echo "full capture exit $?"; tail -2 gpurun_out/ncu_full.log
  # the report itself is too large to travel back (64 MiB cap): export the raw page as CSV and keep that
  ncu -i /tmp/top_kernels.ncu-rep --page raw --csv > gpurun_out/top_kernels_raw.csv 2>/dev/null
  ls -la /tmp/top_kernels.ncu-rep gpurun_out | head -20
  ;;
esac
