#!/bin/bash
# round 2, call f (1 GPU): where does the rescale sweep spend its time?  event times per launch shape, then ncu on the same script
mkdir -p gpurun_out
python scripts/asm_bench.py 119 20 > gpurun_out/r2f_asm_bench.jsonl 2> gpurun_out/r2f_asm_bench.err; echo "asm_bench rc=$?"
cat gpurun_out/r2f_asm_bench.jsonl
python scripts/asm_bench.py 119 2 short > gpurun_out/r2f_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"geo_rescale|picard_element|material_update|geo_scale" -c 8 -o gpurun_out/r2f_prof python scripts/asm_bench.py 119 2 short > gpurun_out/r2f_ncu.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/ | tail -5
