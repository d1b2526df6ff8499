O=gpurun_out/tests
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
timeout 300 python __graft_entry__.py --smoke > $O/smoke.log 2>&1
tail -3 $O/pytest.log; tail -1 $O/smoke.log
