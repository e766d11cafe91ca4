mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1b_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r1b_pytest.log
tail -5 gpurun_out/r1b_pytest.log
timeout 900 python scripts/sweep.py --out gpurun_out/r1b_sweep.jsonl > gpurun_out/r1b_sweep.log 2>&1; echo "sweep exit $?" >> gpurun_out/r1b_sweep.log
tail -60 gpurun_out/r1b_sweep.log
