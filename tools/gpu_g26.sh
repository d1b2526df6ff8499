set -x
O=gpurun_out/g26
mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -x -q -k "permute and bn254_t3 and 7" > $O/pytest_first.log 2>&1; rc=$?; echo "rc=$rc" >> $O/pytest_first.log; tail -15 $O/pytest_first.log
if [ $rc -ne 0 ]; then exit 0; fi
timeout 300 python bench.py --workload c2 --steps 5 --warmup 3 --no-cpu --no-e2e > $O/bench_c2.json 2> $O/bench_c2.err
timeout 300 python bench.py --workload c3 --steps 5 --warmup 3 --no-cpu --no-e2e > $O/bench_c3.json 2> $O/bench_c3.err
echo "pytest rc=skipped" > $O/pytest.log
true
