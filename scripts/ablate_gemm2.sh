#!/bin/bash
# Builds timing-only variants of the library (G2_ABL bits, csrc/gemm2.cu) into scripts/_bin/ (run here, before gpurun).
set -e
cd "$(dirname "$0")/.."
mkdir -p scripts/_bin
for v in "$@"; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared -DG2_ABL=$v \
    -o scripts/_bin/libvipt_g2abl$v.so multi-modal-trakcing-bechmark_b200/csrc/*.cu &
done
wait
ls -la scripts/_bin/
