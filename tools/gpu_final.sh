# Round-end measurement pass on one GPU: GPU tests, smoke, the four bench lines, ncu launch list + full captures of the
# kernels behind C2 / C3 / C5 (each only after the same command exited 0 without ncu).  Output: gpurun_out/final/
set -x
O=gpurun_out/final
mkdir -p $O
nvidia-smi -L > $O/smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
tail -4 $O/pytest.log
timeout 300 python __graft_entry__.py --smoke > $O/smoke.log 2>&1; tail -2 $O/smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > $O/bench_c2.json 2> $O/bench_c2.err
for w in c3 c4 c5; do
  timeout 900 python bench.py --workload $w --steps 5 --warmup 3 --cpu-seconds 8 > $O/bench_$w.json 2> $O/bench_$w.err
done
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err
CMD2="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e"
$CMD2 > $O/plain_c2.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/c2_launches.csv $CMD2 > $O/ncu_c2_list.log 2>&1
$CMD2 > $O/plain_c2b.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sponge_kernel -c 1 -o $O/c2_full $CMD2 > $O/ncu_c2_full.log 2>&1
CMD3="python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu --no-e2e"
$CMD3 > $O/plain_c3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sponge_kernel -c 1 -o $O/c3_full $CMD3 > $O/ncu_c3_full.log 2>&1
CMD5="python bench.py --workload c5 --log-n 16 --steps 2 --warmup 3 --no-cpu --no-e2e"
$CMD5 > $O/plain_c5.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sponge_kernel -c 1 -o $O/c5_full $CMD5 > $O/ncu_c5_full.log 2>&1
ls -la $O
