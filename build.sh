#!/usr/bin/env bash
# Builds sponge_b200/lib/libposeidon_b200.so (sm_100a) and the CPU oracle oracle/liboracle.so.
# Usage: ./build.sh [--fast]    (--fast: only the BN254/BLS t=3 and t=5 kernels, for development)
set -euo pipefail
cd "$(dirname "$0")"
SRC=sponge_b200/csrc
OBJ=build/obj
LIB=sponge_b200/lib
mkdir -p "$OBJ" "$LIB"
NVCC=${NVCC:-nvcc}
NVFLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC"
TS="2 3 4 5 6 7 8 9"
if [[ "${1:-}" == "--fast" ]]; then TS="3 5"; fi

pids=()
run() { "$@" & pids+=($!); }
newer() { # newer <target> <deps...>: true if target missing or older than any dep
  local t=$1; shift
  [[ ! -f $t ]] && return 0
  for d in "$@"; do [[ $d -nt $t ]] && return 0; done
  return 1
}
HDRS=$(ls $SRC/*.cuh $SRC/*.hpp include/*.h)
for f in 0 1; do
  for t in $TS; do
    o=$OBJ/tu_${f}_${t}.o
    if newer $o $SRC/tu_sponge.cu $HDRS; then
      run $NVCC $NVFLAGS -DPSD_TU_FIELD=$f -DPSD_TU_T=$t -c -o $o $SRC/tu_sponge.cu
    fi
    # the same unit with constants staged through shared memory (no constant-bank limit)
    o=$OBJ/tu_${f}_${t}_smem.o
    if newer $o $SRC/tu_sponge.cu $HDRS; then
      run $NVCC $NVFLAGS -DPSD_TU_FIELD=$f -DPSD_TU_T=$t -DPSD_POOL_SHARED=1 -c -o $o $SRC/tu_sponge.cu
    fi
  done
done
# variant-5 unit: BN254 only (for BLS12-381 it measured slower than variant 3: 140 M vs 168 M perm/s on c5;
# -DPSD_TU_FIELD=1 still builds it for experiments)
for f in 0; do
  o=$OBJ/tu_fp52_${f}.o
  if newer $o $SRC/tu_fp52.cu $HDRS; then run $NVCC $NVFLAGS -DPSD_TU_FIELD=$f -c -o $o $SRC/tu_fp52.cu; fi
  o=$OBJ/tu_fp52_${f}_smem.o
  if newer $o $SRC/tu_fp52.cu $HDRS; then run $NVCC $NVFLAGS -DPSD_TU_FIELD=$f -DPSD_POOL_SHARED=1 -c -o $o $SRC/tu_fp52.cu; fi
done
# variant-6 unit (tensor-core linear layer): BN254 and BLS12-381, shared-memory form only
for f in 0 1; do
  o=$OBJ/tu_tc_${f}.o
  if newer $o $SRC/tu_tc.cu $HDRS; then run $NVCC $NVFLAGS -DPSD_TU_FIELD=$f -DPSD_POOL_SHARED=1 -c -o $o $SRC/tu_tc.cu; fi
done
# variant-7 unit (tcgen05 / TMEM form of the tensor-core linear layer)
for f in 0 1; do
  o=$OBJ/tu_tc5_${f}.o
  if newer $o $SRC/tu_tc5.cu $HDRS; then run $NVCC $NVFLAGS -DPSD_TU_FIELD=$f -DPSD_POOL_SHARED=1 -c -o $o $SRC/tu_tc5.cu; fi
done
for u in misc_kernels api; do
  o=$OBJ/$u.o
  if newer $o $SRC/$u.cu $HDRS; then run $NVCC $NVFLAGS -c -o $o $SRC/$u.cu; fi
done
for u in params opt_constants fp52_tables tc_tables; do
  o=$OBJ/$u.o
  if newer $o $SRC/$u.cpp $HDRS; then run g++ -O2 -std=c++17 -fPIC -c -o $o $SRC/$u.cpp; fi
done
fail=0
for p in "${pids[@]:-}"; do [[ -n "$p" ]] && { wait "$p" || fail=1; }; done
[[ $fail -eq 0 ]] || { echo "build failed" >&2; exit 1; }
OBJS="$OBJ/api.o $OBJ/misc_kernels.o $OBJ/tu_fp52_0.o $OBJ/tu_fp52_0_smem.o $OBJ/tu_tc_0.o $OBJ/tu_tc_1.o $OBJ/tu_tc5_0.o $OBJ/tu_tc5_1.o $OBJ/params.o $OBJ/opt_constants.o $OBJ/fp52_tables.o $OBJ/tc_tables.o"
for f in 0 1; do for t in $TS; do OBJS="$OBJS $OBJ/tu_${f}_${t}.o $OBJ/tu_${f}_${t}_smem.o"; done; done
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o $LIB/libposeidon_b200.so $OBJS -Xcompiler -fPIC
# static facts about the hot kernels (SASS fingerprints + instruction mix) for bench.py's roofline block
python3 tools/kernel_sass.py $LIB/kernel_sass.json $OBJ/tu_tc5_0.o $OBJ/tu_tc5_1.o $OBJ/tu_tc_0.o $OBJ/tu_tc_1.o $OBJ/tu_fp52_0.o $OBJ/tu_fp52_0_smem.o $OBJ/tu_0_3.o $OBJ/tu_0_3_smem.o \
  $OBJ/tu_0_5.o $OBJ/tu_0_5_smem.o $OBJ/tu_1_3.o $OBJ/tu_1_3_smem.o
# CPU oracle (test infrastructure)
if newer oracle/liboracle.so oracle/poseidon_oracle.c; then
  gcc -O3 -fopenmp -shared -fPIC -o oracle/liboracle.so oracle/poseidon_oracle.c
fi
# host emulation of the device arithmetic (test infrastructure)
if newer tests/emu/libfp_emu.so tests/emu/fp_emu.cpp $SRC/opt_constants.cpp $SRC/fp52_tables.cpp $SRC/tc_tables.cpp $SRC/params.cpp $HDRS; then
  g++ -O2 -std=c++17 -frounding-math -ffp-contract=off -shared -fPIC -o tests/emu/libfp_emu.so tests/emu/fp_emu.cpp \
    $SRC/opt_constants.cpp $SRC/fp52_tables.cpp $SRC/tc_tables.cpp $SRC/params.cpp
fi
echo "built $LIB/libposeidon_b200.so"
