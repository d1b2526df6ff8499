set -x
O=gpurun_out/c5cap
mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_tc.py tests/test_gpu_worstcase.py -m gpu -x -q -k "bls" > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
CMD5="python bench.py --workload c5 --log-n 16 --steps 2 --warmup 3 --no-cpu --no-e2e"
$CMD5 > $O/plain_c5.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sponge_kernel -c 1 -o $O/c5_full $CMD5 > $O/ncu_c5_full.log 2>&1
timeout 900 python bench.py --workload c5 --steps 5 --warmup 3 --cpu-seconds 8 > $O/bench_c5.json 2> $O/bench_c5.err
true
