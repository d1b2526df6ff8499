set -x
O=gpurun_out/g23
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_concurrent.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
tail -15 $O/pytest.log
true
