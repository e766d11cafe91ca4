#!/bin/bash
# set-aside released for assembly work by default: GPU suite, default bench line, N=150
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r2y_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2y_pytest.log
tail -4 gpurun_out/r2y_pytest.log
run() { local name=$1; shift
  timeout 900 python bench.py --steps 10 --warmup 3 "$@" > gpurun_out/r2y_$name.json 2> gpurun_out/r2y_$name.err; echo "$name rc=$?"
  python - <<P
import json
d=json.loads(open("gpurun_out/r2y_$name.json").read().strip().splitlines()[-1])
print("$name", d["value"], "ms/step", d["ms_per_step"], "asm_ms", d["components"]["picard_assembly_ms"], d["roofline_assembly"]["frac"], "pcg_us", round(d["roofline"]["avg_launch_ms"]*1e3,2), "e2e", d["e2e"]["value"], d["e2e"]["pageable"]["value"], "nonlin", (d.get("nonlinear_solve") or {}).get("seconds"))
P
}
run default
run n150 --n 150 --no-cpu-baseline --full-solve 0
run mode0 --tune 15=0 --no-cpu-baseline --full-solve 0
