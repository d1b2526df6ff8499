# same-box A/B of thin alternative libraries (tools/build_thin.sh): usage gpu_ab.sh name1 name2 ...
O=gpurun_out/ab
rm -rf $O; mkdir -p $O
for round in 1 2; do
for n in "$@"; do
  POSEIDON_B200_LIB=$PWD/altlib/lib$n.so timeout 300 python bench.py --workload c2 --steps 5 --warmup 3 --no-cpu --no-e2e > $O/bench_${n}_$round.json 2> $O/bench_${n}_$round.err
done
done
true
