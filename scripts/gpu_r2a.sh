mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu.py -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2a_pytest.log
tail -n 12 gpurun_out/r2a_pytest.log
timeout 1500 python scripts/sweep.py --tiles 512,384,768,1024 --cfgs 2,3,6,7 --depths 2,3,4 --blocks 1024 --no-spmv --out gpurun_out/r2a_sweep.jsonl > gpurun_out/r2a_sweep.log 2>&1
cut -c1-150 gpurun_out/r2a_sweep.jsonl
