#!/bin/bash
# round 2, call o (1 GPU): field recovery beside the streamed sums (CTA-slot splits), parity of the new sweep, bench line
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_rescale.py tests/test_gpu_round2.py -m gpu -q -x > gpurun_out/r2o_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2o_pytest.log
tail -5 gpurun_out/r2o_pytest.log
FK_WINDOWS=65536 python scripts/asm_bench.py 119 10 > gpurun_out/r2o_asm_bench.jsonl 2> gpurun_out/r2o_asm_bench.err; echo "asm_bench rc=$?"
cut -c1-200 gpurun_out/r2o_asm_bench.jsonl; tail -3 gpurun_out/r2o_asm_bench.err
timeout 600 python bench.py --steps 10 --warmup 3 --full-solve 0 --no-cpu-baseline > gpurun_out/r2o_bench.json 2> gpurun_out/r2o_bench.err; echo "bench rc=$?"
python - <<P
import json
d=json.loads(open("gpurun_out/r2o_bench.json").read().strip().splitlines()[-1])
print(d["value"], "asm_ms", d["components"]["picard_assembly_ms"], d["roofline_assembly"]["frac"], "pcg_us", d["roofline"]["avg_launch_ms"]*1e3, "e2e", d["e2e"]["value"])
P
