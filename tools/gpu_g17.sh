set -x
O=gpurun_out/g17
mkdir -p $O
CMD2="python bench.py --workload c2 --steps 2 --warmup 1 --no-cpu --no-e2e"
timeout 300 $CMD2 > $O/plain_c2.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:sponge_kernel -c 1 -o $O/c2_full $CMD2 > $O/ncu_c2_full.log 2>&1
ls -la $O
true
