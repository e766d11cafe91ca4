#!/bin/bash
# final launch list of the bench command (ncu per-launch durations; serialised and cold: shares, not absolutes)
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --full-solve 0 --trace 0"
$CMD > gpurun_out/r2aq_plain.json 2> gpurun_out/r2aq_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --graph-profiling node --csv --log-file gpurun_out/r2aq_launches.csv $CMD > gpurun_out/r2aq_ncu.log 2>&1; echo "ncu rc=$?"
wc -l gpurun_out/r2aq_launches.csv; tail -2 gpurun_out/r2aq_ncu.log | cut -c1-200
