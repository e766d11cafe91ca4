mkdir -p gpurun_out
FK_DBG=0,3 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'assemble_kernel' -s 3 -c 3 -f -o gpurun_out/r1v_asm python scripts/asm_phases.py 119 > gpurun_out/r1v_ncu.log 2>&1
FK_DBG=3 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'assemble_kernel' -s 0 -c 1 -f -o gpurun_out/r1v_floor python scripts/asm_phases.py 119 > gpurun_out/r1v_ncu2.log 2>&1
tail -n 3 gpurun_out/r1v_ncu.log gpurun_out/r1v_ncu2.log
