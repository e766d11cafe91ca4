mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu.py -m gpu -x -q > gpurun_out/r1r_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r1r_pytest.log
tail -n 15 gpurun_out/r1r_pytest.log
timeout 1200 python scripts/sweep.py --tiles 512,384,768,1024 --cfgs 1,2,3,4,5 --blocks 1024 --no-spmv --out gpurun_out/r1r_sweep.jsonl > gpurun_out/r1r_sweep.log 2>&1
cut -c1-230 gpurun_out/r1r_sweep.jsonl
