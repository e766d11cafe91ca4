#!/bin/bash
# the sweep behind a PCG solve: persisting-L2 set-aside left alone / lines reset / set-aside removed for the sweep
mkdir -p gpurun_out
run() { local name=$1; shift
  timeout 600 python bench.py --steps 10 --warmup 3 --full-solve 0 --no-cpu-baseline --trace 0 "$@" > gpurun_out/r2x_$name.json 2> gpurun_out/r2x_$name.err; echo "$name rc=$?"
  python - <<P
import json
d=json.loads(open("gpurun_out/r2x_$name.json").read().strip().splitlines()[-1])
print("$name", d["value"], "ms/step", d["ms_per_step"], "asm_ms", d["components"]["picard_assembly_ms"], d["roofline_assembly"]["frac"], "pcg_us", round(d["roofline"]["avg_launch_ms"]*1e3,2), "e2e", d["e2e"]["value"])
P
}
run mode0
run mode2 --tune 15=2
run mode3 --tune 15=3
run nopersist --tune 2=0
run n150_mode0 --n 150
run n150_mode2 --n 150 --tune 15=2
