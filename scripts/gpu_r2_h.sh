#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/h_all.log 2>&1
echo "all rc=$?"; tail -4 gpurun_out/h_all.log
bash scripts/gpu_list.sh
bash scripts/gpu_ab.sh "NKBS_X=1"
