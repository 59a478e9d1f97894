#!/bin/bash
# Build libk7det.so in-tree for sm_100a.  Usage: csrc/build.sh [extra nvcc flags]
set -euo pipefail
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
OUT=../libk7det.so
FLAGS=(-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false
       -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -shared "$@")
"$NVCC" "${FLAGS[@]}" -o "$OUT" k7_abi.cu k7_anchor_match.cu k7_loss.cu k7_detect.cu k7_nms_v1.cu k7_eval.cu
echo "built $(realpath "$OUT")"
