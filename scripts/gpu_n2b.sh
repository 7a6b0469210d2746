#!/bin/bash
# N-GPU: dp correctness + A/B of the multimem exchange (headline leg only)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
N=${1:-2}
timeout 600 python -m pytest tests/test_gpu_trainer.py -m gpu -q -x --timeout 500 -k "fused_dp" -s > gpurun_out/n2_dp.log 2>&1
echo "dp test rc=$?"; tail -3 gpurun_out/n2_dp.log
for mm in 0 1 0 1; do
NKBS_DP_MULTIMEM=$mm timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 100 --warmup 10 --no-render --no-legs > gpurun_out/nb_$mm.json 2> gpurun_out/nb_$mm.err
python -c "
import json; d=json.load(open('gpurun_out/nb_$mm.json')); print('N=$N multimem=$mm: ms/step %.4f e2e %.4f loss %.5f' % (d['ms_per_step'], d['e2e']['ms_per_step'], d['final_loss']))" || tail -5 gpurun_out/nb_$mm.err
done
