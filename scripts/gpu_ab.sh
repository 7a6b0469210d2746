#!/bin/bash
# A/B of env toggles on ONE box: usage gpu_ab.sh [--args "<bench args>"] "<VAR=a VAR2=b>" "<VAR=c>" ...
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
EXTRA=""
if [ "$1" == "--args" ]; then EXTRA="$2"; shift 2; fi
for rep in 1 2; do
  for cfg in "$@"; do
    tag=$(echo "$cfg" | tr -c 'A-Za-z0-9=' '_')
    env $cfg timeout 600 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-render --no-legs --no-kernel-pass $EXTRA > gpurun_out/ab_${tag}_${rep}.json 2> gpurun_out/ab_${tag}_${rep}.err
    python - <<PY
import json
try:
    d=json.load(open("gpurun_out/ab_${tag}_${rep}.json"))
    print("$cfg rep $rep: ms/step %.4f  e2e %.4f  loss %.5f" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["final_loss"]))
except Exception as e:
    print("$cfg rep $rep: FAILED", e)
    print(open("gpurun_out/ab_${tag}_${rep}.err").read()[-600:])
PY
  done
done
