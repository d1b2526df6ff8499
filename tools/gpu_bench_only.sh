set -x
O=gpurun_out/benchonly
mkdir -p $O
timeout 900 python bench.py --steps 20 --warmup 5 > $O/bench_c2.json 2> $O/bench_c2.err
for w in c3 c4 c5; do
  timeout 900 python bench.py --workload $w --steps 5 --warmup 3 --cpu-seconds 8 > $O/bench_$w.json 2> $O/bench_$w.err
done
true
