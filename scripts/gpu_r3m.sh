mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r3m_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r3m_pytest.log
tail -n 3 gpurun_out/r3m_pytest.log
timeout 600 python scripts/sweep.py --tiles 512 --cfgs 0 --blocks 1024 --no-spmv --out gpurun_out/r3m_sweep.jsonl 2>&1 | tail -n 4
