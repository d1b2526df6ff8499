set -x
O=gpurun_out/n2
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > $O/pytest.log 2>&1; tail -1 $O/pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 2 --steps 5 --warmup 3 > $O/bench_c2_n2.json 2> $O/bench_c2_n2.err
tail -c 400 $O/bench_c2_n2.err
true
