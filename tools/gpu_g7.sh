set -x
O=gpurun_out/g7
mkdir -p $O
CMD2="python bench.py --log-n 22 --steps 2 --warmup 3 --no-cpu --no-e2e"
$CMD2 > $O/plain_c2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sponge_kernel -c 1 -o $O/c2_v6_full $CMD2 > $O/ncu_c2_full.log 2>&1
tail -3 $O/ncu_c2_full.log
true
