#!/bin/bash
# split backward of the main field (base MLP in two launches, first part's scatter on a side stream): off / 0.5 / 0.375
mkdir -p gpurun_out
python -m pytest tests/test_gpu_trainer.py tests/test_gpu_deterministic.py tests/test_gpu_configs.py -x -q 2>&1 | tail -12 > gpurun_out/split_tests.log
for f in 0 0.5 0.375 0.625; do
  NKBS_BWD_SPLIT=$f python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-kernel-pass > gpurun_out/split_$f.json 2> gpurun_out/split_$f.err
done
tail -5 gpurun_out/split_tests.log
