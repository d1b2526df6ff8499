set -x
mkdir -p gpurun_out/g1
nvidia-smi -L > gpurun_out/g1/smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/g1/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/g1/pytest.log
for w in c2 c3; do
  timeout 300 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/g1/bench_$w.json 2> gpurun_out/g1/bench_$w.err
done
PSD_FORCE_SMEM=1 timeout 300 python bench.py --workload c2 --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/g1/bench_c2_smem.json 2> gpurun_out/g1/bench_c2_smem.err
timeout 300 python bench.py --workload c3 --variant 4 --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/g1/bench_c3_v4.json 2> gpurun_out/g1/bench_c3_v4.err
PSD_FORCE_SMEM=1 timeout 300 python bench.py --workload c3 --variant 4 --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/g1/bench_c3_v4_smem.json 2> gpurun_out/g1/bench_c3_v4_smem.err
PSD_FORCE_SMEM=1 timeout 300 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/g1/bench_c5_smem.json 2> gpurun_out/g1/bench_c5_smem.err
timeout 300 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/g1/bench_c5.json 2> gpurun_out/g1/bench_c5.err
tail -3 gpurun_out/g1/pytest.log
