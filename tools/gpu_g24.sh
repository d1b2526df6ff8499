set -x
O=gpurun_out/g24
mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -x -q -k "bn254_t3 and 7" > $O/pytest_first.log 2>&1; rc=$?; echo "rc=$rc" >> $O/pytest_first.log; tail -5 $O/pytest_first.log
timeout 400 python bench.py --workload c2 --steps 5 --warmup 3 --no-cpu > $O/bench_c2_waves.json 2> $O/bench_c2_waves.err
PSD_STAGE_WAVES=0 timeout 400 python bench.py --workload c2 --steps 5 --warmup 3 --no-cpu > $O/bench_c2_pow2.json 2> $O/bench_c2_pow2.err
timeout 400 python bench.py --workload c3 --steps 5 --warmup 3 --no-cpu > $O/bench_c3_waves.json 2> $O/bench_c3_waves.err
PSD_STAGE_WAVES=0 timeout 400 python bench.py --workload c3 --steps 5 --warmup 3 --no-cpu > $O/bench_c3_pow2.json 2> $O/bench_c3_pow2.err
true
