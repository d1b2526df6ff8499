#!/usr/bin/env bash
# Development aid: altlib/lib<name>.so with the variant-7 kernels of ONE field compiled with extra flags, the host code and
# that field's variant-6 unit and t = 3 / t = 5 integer kernels -- a 40 MB library for same-box A/B runs
# (POSEIDON_B200_LIB=..., tools/gpu_ab.sh), not a product.
# usage: [FIELD=0|1] tools/build_thin.sh <name> [extra nvcc flags...]      (0 = BN254, 1 = BLS12-381)
set -euo pipefail
cd "$(dirname "$0")/.."
name=$1; shift
f=${FIELD:-0}
mkdir -p altlib build/thin
O=build/obj
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -DPSD_TU_FIELD=$f -DPSD_POOL_SHARED=1 "$@" \
  -c -o build/thin/tu_tc5_$name.o sponge_b200/csrc/tu_tc5.cu
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o altlib/lib$name.so build/thin/tu_tc5_$name.o $O/api.o $O/misc_kernels.o \
  $O/params.o $O/opt_constants.o $O/fp52_tables.o $O/tc_tables.o $O/tu_tc_$f.o $O/tu_${f}_3.o $O/tu_${f}_3_smem.o $O/tu_${f}_5.o \
  $O/tu_${f}_5_smem.o -Xcompiler -fPIC
ls -la altlib/lib$name.so
