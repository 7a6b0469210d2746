#!/bin/bash
# hidden-64 proposal fields: fused forward + unfused backward (default) vs encode + MLP both ways (NKBS_HYBRID_DENSITY=0)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_density.py tests/test_gpu_configs.py tests/test_gpu_deterministic.py tests/test_gpu_model.py -x -q 2>&1 | tail -12 > gpurun_out/hyb_tests.log
NKBS_HYBRID_DENSITY=0 python bench.py --steps 100 --warmup 10 > gpurun_out/hyb_off.json 2> gpurun_out/hyb_off.err
python bench.py --steps 100 --warmup 10 > gpurun_out/hyb_on.json 2> gpurun_out/hyb_on.err
tail -5 gpurun_out/hyb_tests.log
