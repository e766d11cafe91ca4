mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu.py -m gpu -x -q > gpurun_out/r1z_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r1z_pytest.log
tail -n 12 gpurun_out/r1z_pytest.log
FK_DBG=0,1,2,3 timeout 600 python scripts/asm_phases.py 119 > gpurun_out/r1z_phases.log 2>&1; tail -4 gpurun_out/r1z_phases.log
timeout 1200 python scripts/sweep.py --tiles 512,448,384 --cfgs 1,2,3 --blocks 1024 --no-spmv --out gpurun_out/r1z_sweep.jsonl > gpurun_out/r1z_sweep.log 2>&1
cut -c1-130 gpurun_out/r1z_sweep.jsonl
