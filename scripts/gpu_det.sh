#!/bin/bash
# deterministic mode: its tests, then the bench with the mode off / on (same build, same box)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_deterministic.py -x -q 2>&1 | tail -25 > gpurun_out/det_tests.log
python -m pytest tests/test_gpu_trainer.py tests/test_gpu_density.py tests/test_gpu_ops.py tests/test_gpu_mlp_tc.py -x -q 2>&1 | tail -8 >> gpurun_out/det_tests.log
python bench.py --steps 200 --warmup 10 > gpurun_out/det_off.json 2> gpurun_out/det_off.err
NKBS_DETERMINISTIC=1 python bench.py --steps 200 --warmup 10 > gpurun_out/det_on.json 2> gpurun_out/det_on.err
tail -5 gpurun_out/det_tests.log
