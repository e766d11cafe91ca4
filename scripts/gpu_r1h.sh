mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1i_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r1i_pytest.log
tail -n 3 gpurun_out/r1i_pytest.log
timeout 900 python scripts/sweep.py --tiles 512 --cfgs 0 --blocks 1024 --stages 0,-1,0,-1 --out gpurun_out/r1i_sweep.jsonl 2>&1 | tail -n 20
