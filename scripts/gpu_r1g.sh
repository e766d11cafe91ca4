mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1g_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r1g_pytest.log
tail -n 3 gpurun_out/r1g_pytest.log
timeout 900 python scripts/sweep.py --tiles 512 --cfgs 0 --blocks 1,256,1024 --stages 0,4,6 --out gpurun_out/r1g_sweep.jsonl 2>&1 | tail -n 20
