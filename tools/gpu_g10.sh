set -x
O=gpurun_out/g10
mkdir -p $O
export POSEIDON_B200_LIB=$PWD/sponge_b200/lib/alt/lib_bls768.so
timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -x -q -k "bls" > $O/pytest.log 2>&1; tail -1 $O/pytest.log
timeout 300 python bench.py --workload c5 --log-n 18 --variant 6 --steps 3 --warmup 3 --no-cpu --no-e2e > $O/bench_c5_v6_768.json 2> $O/bench_c5_v6_768.err
unset POSEIDON_B200_LIB
timeout 300 python bench.py --workload c5 --log-n 18 --variant 6 --steps 3 --warmup 3 --no-cpu --no-e2e > $O/bench_c5_v6_512.json 2> $O/bench_c5_v6_512.err
timeout 300 python bench.py --workload c5 --log-n 18 --steps 3 --warmup 3 --no-cpu --no-e2e > $O/bench_c5_v3.json 2> $O/bench_c5_v3.err
true
