#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
rm -f gpurun_out/g_v4.txt
for R in 4096 65536; do
for ms in 0 20 40 70 100 200 400 1000 100000; do
echo "== R=$R NKBS_V4_MIN_SCALE=$ms" | tee -a gpurun_out/g_v4.txt
NKBS_V4_MIN_SCALE=$ms DL_ONLY=1 R=$R timeout 300 python scripts/encode_bench.py 2>&1 | tee -a gpurun_out/g_v4.txt
NKBS_V4_MIN_SCALE=$ms R=$R timeout 300 python scripts/density_bench.py 2>&1 | grep "fused=True" | tee -a gpurun_out/g_v4.txt
done
done
