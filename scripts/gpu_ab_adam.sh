#!/bin/bash
# Adam kernel A/B inside the step: IEEE kernel vs default (approximate sqrt / division), same build, same box
mkdir -p gpurun_out
python scripts/adam_bench.py > gpurun_out/adam_bench.txt 2>&1
python -m pytest tests/test_gpu_trainer.py tests/test_gpu_deterministic.py tests/test_gpu_configs.py -x -q 2>&1 | tail -5 > gpurun_out/adam_tests.log
NKBS_ADAM_IEEE=1 python bench.py --steps 200 --warmup 10 > gpurun_out/adam_ieee.json 2> gpurun_out/adam_ieee.err
python bench.py --steps 200 --warmup 10 > gpurun_out/adam_fast.json 2> gpurun_out/adam_fast.err
cat gpurun_out/adam_bench.txt; tail -3 gpurun_out/adam_tests.log
