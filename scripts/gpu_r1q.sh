mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1q_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r1q_pytest.log
tail -n 3 gpurun_out/r1q_pytest.log
timeout 300 python scripts/prof_asm.py 119 > gpurun_out/r1q_prof_plain.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:'assemble_kernel' -s 3 -c 3 -f -o gpurun_out/r1q_asm python scripts/prof_asm.py 119 > gpurun_out/r1q_ncu.log 2>&1
tail -n 2 gpurun_out/r1q_prof_plain.log; tail -n 3 gpurun_out/r1q_ncu.log
ls -la gpurun_out | tail -n 6
