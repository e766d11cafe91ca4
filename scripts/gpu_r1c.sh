mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1c_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r1c_pytest.log
tail -3 gpurun_out/r1c_pytest.log
timeout 300 python scripts/prof_asm.py 119 tile_max_elems=512 > gpurun_out/r1c_prof_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:assemble_kernel -s 3 -c 3 -f -o gpurun_out/r1c_asm python scripts/prof_asm.py 119 tile_max_elems=512 > gpurun_out/r1c_ncu.log 2>&1
tail -3 gpurun_out/r1c_prof_plain.log gpurun_out/r1c_ncu.log
timeout 600 python scripts/sweep.py --tiles 256,384,512 --cfgs 0,1 --blocks 2048 --stages 2 --out gpurun_out/r1c_sweep.jsonl 2>&1 | grep asm
