#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x --timeout 200 > gpurun_out/na_alone.log 2>&1
echo "alone rc=$?"; tail -4 gpurun_out/na_alone.log
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/na_all.log 2>&1
echo "all rc=$?"; tail -4 gpurun_out/na_all.log
