mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r3c_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r3c_pytest.log
tail -n 3 gpurun_out/r3c_pytest.log
timeout 900 python scripts/sweep_spmv.py --out gpurun_out/r3c_sweep_spmv.jsonl 2>&1 | tail -n 20
