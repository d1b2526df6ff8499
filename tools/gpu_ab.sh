# same-box A/B of thin alternative libraries (tools/build_thin.sh): usage gpu_ab.sh name1 name2 ...
# the LAST name is also parity-tested (variant-7 permutation tests of the BN254 t = 3 / t = 5 shapes the thin library holds)
O=gpurun_out/ab
rm -rf $O; mkdir -p $O
for last in "$@"; do :; done
POSEIDON_B200_LIB=$PWD/altlib/lib$last.so timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -x -q -k "permute and (bn254_t3 or bn254_t5) and 7" > $O/pytest_$last.log 2>&1; echo "rc=$?" >> $O/pytest_$last.log
for round in 1 2; do
for n in "$@"; do
  POSEIDON_B200_LIB=$PWD/altlib/lib$n.so timeout 300 python bench.py --workload ${WL:-c2} --steps 5 --warmup 3 --no-cpu --no-e2e > $O/bench_${n}_$round.json 2> $O/bench_${n}_$round.err
done
done
true
