set -x
O=gpurun_out/g13
mkdir -p $O
POSEIDON_B200_LIB=$PWD/sponge_b200/lib/alt/lib_640.so timeout 300 python bench.py --workload c2 --steps 5 --warmup 3 --no-cpu --no-e2e > $O/bench_c2_640.json 2> $O/bench_c2_640.err
CMD2="python bench.py --log-n 22 --steps 2 --warmup 3 --no-cpu --no-e2e"
$CMD2 > $O/plain_c2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sponge_kernel -c 1 -o $O/c2_pair_full $CMD2 > $O/ncu_c2_full.log 2>&1
true
