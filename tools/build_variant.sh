#!/bin/bash
# tools/build_variant.sh NAME "EXTRA NVCC FLAGS": particles.cu rebuilt with the flags, linked with the current objects into
# build/variants/libminipic_NAME.so (select at run time with MPIC_LIB=...).  A/B tool for kernel tuning knobs.
set -e
NAME=$1; shift
ROOT=$(cd "$(dirname "$0")/.." && pwd)
SRC="$ROOT/mini-pic--2d-pic-code-with-tiling-load-balance_b200/csrc"
OUT="$ROOT/build/variants"; mkdir -p "$OUT"
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC -Xptxas -v "$@" \
    -c "$SRC/particles.cu" -o "$OUT/particles_$NAME.o" 2> "$OUT/particles_$NAME.ptxas.log"
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o "$OUT/libminipic_$NAME.so" "$SRC/api.o" "$SRC/yee.o" \
    "$OUT/particles_$NAME.o" "$SRC/ring.o" "$SRC/peer.o" "$SRC/reslab.o" "$SRC/h5dump.o"
grep -A2 "push_packed_kernelILi16ELi16E" "$OUT/particles_$NAME.ptxas.log" | grep "Used\|spill" | head -4
