#!/bin/bash
# demag workload bench line at N=119
mkdir -p gpurun_out
timeout 900 python bench.py --workload demag > gpurun_out/r2aj_bench_demag.json 2> gpurun_out/r2aj_bench_demag.err; echo "bench rc=$?"
tail -3 gpurun_out/r2aj_bench_demag.err
python - <<P
import json
for ln in open("gpurun_out/r2aj_bench_demag.json"):
    if ln.startswith("{"):
        d=json.loads(ln)
        print(d["value"], d["ms_per_step"], d["components"], d["e2e"], d["roofline"]["frac"], d.get("cpu_baseline"), d.get("parity"))
P
