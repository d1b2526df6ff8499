set -x
O=gpurun_out/g27
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_tc.py tests/test_gpu_worstcase.py tests/test_gpu_concurrent.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
tail -5 $O/pytest.log
for w in c2 c3 c4 c5; do
timeout 300 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu --no-e2e > $O/bench_$w.json 2> $O/bench_$w.err
done
true
