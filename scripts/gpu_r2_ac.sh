#!/bin/bash
# diagonal from row sums (6 contributions per tet): parity, CTA-slot splits, bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_rescale.py tests/test_gpu_round2.py tests/test_gpu.py -m gpu -q -x > gpurun_out/r2ac_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2ac_pytest.log
tail -6 gpurun_out/r2ac_pytest.log
FK_WINDOWS=65536 FK_SPLITS=0,41,31,32,21,22,42,33 python scripts/asm_bench.py 119 10 > gpurun_out/r2ac_asm_bench.jsonl 2> gpurun_out/r2ac.err; echo rc=$?
grep "split\|v0_win\|geo_stats" gpurun_out/r2ac_asm_bench.jsonl | cut -c1-150; tail -2 gpurun_out/r2ac.err
timeout 600 python bench.py --steps 10 --warmup 3 --full-solve 0 --no-cpu-baseline --trace 0 > gpurun_out/r2ac_bench.json 2> gpurun_out/r2ac_bench.err; echo "bench rc=$?"
python - <<P
import json
d=json.loads(open("gpurun_out/r2ac_bench.json").read().strip().splitlines()[-1])
print(d["value"], "ms/step", d["ms_per_step"], "asm_ms", d["components"]["picard_assembly_ms"], d["roofline_assembly"]["frac"], "pcg_us", round(d["roofline"]["avg_launch_ms"]*1e3,2), "e2e", d["e2e"]["value"])
P
