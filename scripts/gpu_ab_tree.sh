#!/bin/bash
# same-box A/B of two source trees: the repo (new) and ab_old/ (a built export of an earlier commit)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
ARGS="--steps 200 --warmup 20 --no-cpu-baseline --no-render --no-legs --no-kernel-pass $@"
for rep in 1 2 3; do
  for tree in ab_old .; do
    (cd $tree && timeout 600 python bench.py $ARGS 2>/dev/null | python -c "
import json,sys; d=json.load(sys.stdin); print('$tree rep $rep: ms/step %.4f e2e %.4f launches/step %.1f' % (d['ms_per_step'], d['e2e']['ms_per_step'], d['gpu_launches']/d['steps']))")
  done
done
