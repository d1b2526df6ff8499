#!/usr/bin/env bash
# Development aid: altlib/lib<name>.so with the variant-7 BN254 kernels compiled with extra flags, the host code and the
# BN254 t = 3 / t = 5 integer kernels -- a 30 MB library for same-box A/B runs (POSEIDON_B200_LIB=...), not a product.
# usage: tools/build_thin.sh <name> [extra nvcc flags...]
set -euo pipefail
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p altlib build/thin
O=build/obj
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -DPSD_TU_FIELD=0 -DPSD_POOL_SHARED=1 "$@" \
  -c -o build/thin/tu_tc5_$name.o sponge_b200/csrc/tu_tc5.cu
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o altlib/lib$name.so build/thin/tu_tc5_$name.o $O/api.o $O/misc_kernels.o \
  $O/params.o $O/opt_constants.o $O/fp52_tables.o $O/tc_tables.o $O/tu_tc_0.o $O/tu_0_3.o $O/tu_0_3_smem.o $O/tu_0_5.o $O/tu_0_5_smem.o -Xcompiler -fPIC
ls -la altlib/lib$name.so
