#!/bin/bash
# N-GPU bench line (all legs, dp_check) under torchrun
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
N=${1:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 50 --warmup 10 --render-frames 4 > gpurun_out/n${N}_bench.json 2> gpurun_out/n${N}_bench.err
echo "bench rc=$?"; tail -c 700 gpurun_out/n${N}_bench.json; tail -3 gpurun_out/n${N}_bench.err
