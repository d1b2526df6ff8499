set -x
O=gpurun_out/g11
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_sponge.py tests/test_gpu_parity.py tests/test_gpu_ragged.py -m gpu -x -q > $O/pytest.log 2>&1; tail -1 $O/pytest.log
timeout 300 python bench.py --workload c5 --log-n 18 --variant 6 --steps 3 --warmup 3 --no-cpu --no-e2e > $O/bench_c5_v6.json 2> $O/bench_c5_v6.err
timeout 300 python bench.py --workload c5 --log-n 18 --steps 3 --warmup 3 --no-cpu --no-e2e > $O/bench_c5_v3.json 2> $O/bench_c5_v3.err
timeout 300 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu --no-e2e > $O/bench_c5_full.json 2> $O/bench_c5_full.err
timeout 300 python bench.py --workload c2 --steps 3 --warmup 3 --no-cpu --no-e2e > $O/bench_c2.json 2> $O/bench_c2.err
true
