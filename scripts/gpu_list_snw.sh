#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
CMD="python bench.py --model semantic_nerfw --rays 8192 --no-graph --no-cpu-baseline --no-kernel-pass --no-render --no-legs --steps 2 --warmup 3"
timeout 300 $CMD > gpurun_out/plain_snw.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches_snw.csv $CMD > gpurun_out/ncu_list_snw.log 2>&1
echo "launch list exit $?"
