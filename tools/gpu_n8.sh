set -x
O=gpurun_out/n8
mkdir -p $O
nvidia-smi -L > $O/smi.txt
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > $O/pytest.log 2>&1; tail -1 $O/pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 5 --warmup 3 > $O/bench_c2_n8.json 2> $O/bench_c2_n8.err
tail -c 400 $O/bench_c2_n8.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus 8 --steps 5 --warmup 3 --workload c4 --no-merkle > $O/bench_c4_n8.json 2> $O/bench_c4_n8.err
true
