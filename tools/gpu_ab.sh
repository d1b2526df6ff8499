set -x
O=gpurun_out/ab
mkdir -p $O
for i in 1 2 3; do
  timeout 300 python bench.py --workload c2 --steps 5 --warmup 3 --no-cpu --no-e2e > $O/bench_A$i.json 2> $O/bench_A$i.err
  POSEIDON_B200_LIB=$PWD/altlib/libB.so timeout 300 python bench.py --workload c2 --steps 5 --warmup 3 --no-cpu --no-e2e > $O/bench_B$i.json 2> $O/bench_B$i.err
done
true
