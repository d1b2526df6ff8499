set -x
O=gpurun_out/g8
mkdir -p $O
for cfg in 768_0 768_1; do
  export POSEIDON_B200_LIB=$PWD/sponge_b200/lib/alt/lib_$cfg.so
  timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -x -q -k "permute and (bn254_t3 or bn254_t5)" > $O/pytest_$cfg.log 2>&1; tail -1 $O/pytest_$cfg.log
  for st in 0 1500; do
    PSD_TC_STAGGER=$st timeout 300 python bench.py --workload c2 --steps 5 --warmup 3 --no-cpu --no-e2e > $O/bench_c2_${cfg}_st$st.json 2> $O/bench_c2_${cfg}_st$st.err
  done
  PSD_TC_STAGGER=1500 timeout 300 python bench.py --workload c3 --steps 5 --warmup 3 --no-cpu --no-e2e > $O/bench_c3_${cfg}.json 2> $O/bench_c3_${cfg}.err
done
true
