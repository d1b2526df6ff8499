O=gpurun_out/ab5; rm -rf $O; mkdir -p $O
POSEIDON_B200_LIB=$PWD/altlib/libblsnew.so timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -x -q -k "permute and bls and 7" > $O/pytest.log 2>&1; echo "rc=$?" >> $O/pytest.log
for r in 1 2; do for n in blsold blsnew; do
POSEIDON_B200_LIB=$PWD/altlib/lib$n.so timeout 300 python bench.py --workload c5 --log-n 17 --steps 3 --warmup 2 --no-cpu --no-e2e > $O/bench_${n}_$r.json 2> $O/bench_${n}_$r.err
done; done
true
