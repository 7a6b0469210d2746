#!/bin/bash
# 2-GPU check: fused DP step vs NCCL route, then the bench under torchrun
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
N=${1:-2}
timeout 600 python -m pytest tests/test_gpu_trainer.py -m gpu -q -x --timeout 500 -k "fused_dp" -s > gpurun_out/n2_dp.log 2>&1
echo "dp test rc=$?"; tail -5 gpurun_out/n2_dp.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 50 --warmup 10 --render-frames 4 > gpurun_out/n2_bench.json 2> gpurun_out/n2_bench.err
echo "bench rc=$?"; tail -c 1800 gpurun_out/n2_bench.json; tail -5 gpurun_out/n2_bench.err
