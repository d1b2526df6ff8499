set -x
mkdir -p gpurun_out/g2
nvidia-smi -L > gpurun_out/g2/smi.txt
timeout 1200 python -m pytest tests -m gpu -x -q --durations=15 > gpurun_out/g2/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/g2/pytest.log
tail -5 gpurun_out/g2/pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/g2/bench_c2_n2.json 2> gpurun_out/g2/bench_c2_n2.err
tail -c 600 gpurun_out/g2/bench_c2_n2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3 --workload c4 --no-merkle > gpurun_out/g2/bench_c4_n2.json 2> gpurun_out/g2/bench_c4_n2.err
tail -c 600 gpurun_out/g2/bench_c4_n2.err
