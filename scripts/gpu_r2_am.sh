#!/bin/bash
# final state: whole GPU suite (1 GPU), smoke(), short bench line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2am_pytest_gpu.log 2>&1; echo "pytest -m gpu rc=$?" | tee -a gpurun_out/r2am_pytest_gpu.log
tail -4 gpurun_out/r2am_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 600 python bench.py --steps 10 --warmup 3 --full-solve 0 --no-cpu-baseline > gpurun_out/r2am_bench.json 2> gpurun_out/r2am_bench.err; echo "bench rc=$?"
python - <<P
import json
d=json.loads([l for l in open("gpurun_out/r2am_bench.json") if l.startswith("{")][-1])
print(d["value"], "ms/step", d["ms_per_step"], "asm_ms", d["components"]["picard_assembly_ms"], d["roofline_assembly"]["frac"], "pcg_us", round(d["roofline"]["avg_launch_ms"]*1e3,2), "e2e", d["e2e"]["value"])
P
