set -x
O=gpurun_out/g5
mkdir -p $O
for w in c2 c3 c4 c5; do
  timeout 300 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu --no-e2e > $O/bench_${w}_v6.json 2> $O/bench_${w}_v6.err
  tail -c 300 $O/bench_${w}_v6.err
done
for w in c2 c3; do
  timeout 300 python bench.py --workload $w --variant 5 --steps 5 --warmup 3 --no-cpu --no-e2e > $O/bench_${w}_v5.json 2> $O/bench_${w}_v5.err
done
timeout 300 python bench.py --workload c5 --variant 3 --steps 3 --warmup 3 --no-cpu --no-e2e > $O/bench_c5_v3.json 2> $O/bench_c5_v3.err
true
