#!/bin/bash
# round 2, call j (1 GPU): the sweep inside the bench step: L2 window off the stream (new default) vs on it, persisting lines reset
mkdir -p gpurun_out
run() { local name=$1; shift
  timeout 600 python bench.py --steps 10 --warmup 3 --full-solve 0 --no-cpu-baseline "$@" > gpurun_out/r2j_$name.json 2> gpurun_out/r2j_$name.err; echo "$name rc=$?"
  python - <<P
import json
d=json.loads(open("gpurun_out/r2j_$name.json").read().strip().splitlines()[-1])
print("$name", d["value"], "asm_ms", d["components"]["picard_assembly_ms"], d["roofline_assembly"]["frac"], "pcg_us", d["roofline"]["avg_launch_ms"]*1e3, "e2e", d["e2e"]["value"])
P
}
run default
run on_stream --tune 14=1
run reset --tune 15=1
run tile --tune 10=0
run tile_on_stream --tune 10=0,14=1
run n150 --n 150
run n150_reset --n 150 --tune 15=1
