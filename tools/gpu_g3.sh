set -x
O=gpurun_out/g3
mkdir -p $O
nvidia-smi -L > $O/smi.txt
timeout 1200 python -m pytest tests -m gpu -x -q --durations=12 > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
tail -5 $O/pytest.log
timeout 120 tools/ubench4 > $O/ubench4.txt 2>&1; echo "rc=$?" >> $O/ubench4.txt
cat $O/ubench4.txt
timeout 600 python bench.py --steps 10 --warmup 3 > $O/bench_c2.json 2> $O/bench_c2.err
for w in c3 c4 c5; do
  timeout 600 python bench.py --workload $w --steps 5 --warmup 3 --cpu-seconds 8 > $O/bench_$w.json 2> $O/bench_$w.err
done
# ncu: launch list + full captures (each only after the same command exited 0 without ncu)
CMD2="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e"
$CMD2 > $O/plain_c2.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/c2_launches.csv $CMD2 > $O/ncu_c2_list.log 2>&1
$CMD2 > $O/plain_c2b.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sponge_kernel -c 1 -o $O/c2_full $CMD2 > $O/ncu_c2_full.log 2>&1
CMD3="python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu --no-e2e"
$CMD3 > $O/plain_c3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sponge_kernel -c 1 -o $O/c3_full $CMD3 > $O/ncu_c3_full.log 2>&1
CMD5="python bench.py --workload c5 --log-n 16 --steps 2 --warmup 3 --no-cpu --no-e2e"
$CMD5 > $O/plain_c5.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sponge_kernel -c 1 -o $O/c5_full $CMD5 > $O/ncu_c5_full.log 2>&1
ls -la $O
