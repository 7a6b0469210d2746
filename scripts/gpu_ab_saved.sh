#!/bin/bash
# nerfacto proposal fields: backward from saved features (default) vs recomputed gathers (NKBS_SAVED_DENSITY_BWD=0)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_density.py tests/test_gpu_deterministic.py tests/test_gpu_configs.py tests/test_gpu_level.py -x -q 2>&1 | tail -8 > gpurun_out/saved_tests.log
NKBS_SAVED_DENSITY_BWD=0 python bench.py --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/saved_0.json 2> gpurun_out/saved_0.err
python bench.py --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/saved_1.json 2> gpurun_out/saved_1.err
tail -4 gpurun_out/saved_tests.log
