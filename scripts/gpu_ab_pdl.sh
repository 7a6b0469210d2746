#!/bin/bash
# programmatic dependent launch A/B: NKBS_PDL=0 vs default, same build, same box; then the tests on the default
mkdir -p gpurun_out
NKBS_PDL=0 python bench.py --steps 200 --warmup 10 > gpurun_out/pdl_off.json 2> gpurun_out/pdl_off.err
python bench.py --steps 200 --warmup 10 > gpurun_out/pdl_on.json 2> gpurun_out/pdl_on.err
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -8 > gpurun_out/pdl_tests.log
tail -4 gpurun_out/pdl_tests.log; tail -3 gpurun_out/pdl_on.err
