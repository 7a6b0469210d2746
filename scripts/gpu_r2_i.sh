#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 -k "not psnr" > gpurun_out/i_all.log 2>&1
echo "all rc=$?"; tail -4 gpurun_out/i_all.log
bash scripts/gpu_ab.sh "NKBS_FUSED_RESAMPLE=0" "NKBS_FUSED_RESAMPLE=1"
bash scripts/gpu_ab.sh --args "--rays 65536 --steps 30" "NKBS_FUSED_RESAMPLE=0" "NKBS_FUSED_RESAMPLE=1"
