mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1f_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r1f_pytest.log
tail -n 3 gpurun_out/r1f_pytest.log
timeout 900 python scripts/sweep.py --tiles 256,384,512 --cfgs 0,1,2,3,4 --blocks 2048 --stages 2 --out gpurun_out/r1f_sweep.jsonl 2>&1 | grep -v spmv
