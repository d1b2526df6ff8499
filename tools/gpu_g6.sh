set -x
O=gpurun_out/g6
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_tc.py tests/test_gpu_worstcase.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
tail -3 $O/pytest.log
for st in 0 500 1000 1500 2500; do
  PSD_TC_STAGGER=$st timeout 300 python bench.py --workload c2 --steps 5 --warmup 3 --no-cpu --no-e2e > $O/bench_c2_st$st.json 2> $O/bench_c2_st$st.err
done
for st in 0 1500 3000; do
  PSD_TC_STAGGER=$st timeout 300 python bench.py --workload c3 --steps 5 --warmup 3 --no-cpu --no-e2e > $O/bench_c3_st$st.json 2> $O/bench_c3_st$st.err
done
true
