#!/bin/bash
# debug build of the library with in-kernel clock64 trace points (NKBS_TRACE) -> nerf_kbs_b200/libnkbs_trace.so
set -e
cd "$(dirname "$0")/.."
mkdir -p /tmp/nkbs_trace
for f in nerf_kbs_b200/csrc/*.cu; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr --extended-lambda -DNKBS_TRACE \
    -I nerf_kbs_b200/csrc -I include -c $f -o /tmp/nkbs_trace/$(basename $f .cu).o &
done
wait
nvcc -shared -o nerf_kbs_b200/libnkbs_trace.so /tmp/nkbs_trace/*.o
