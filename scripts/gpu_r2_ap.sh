#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_edge_cases.py -m gpu -q > gpurun_out/r2ap_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2ap_pytest.log
grep -E "^E  |passed|failed|Error" gpurun_out/r2ap_pytest.log | head -20
timeout 600 python bench.py --steps 10 --warmup 3 --full-solve 0 --no-cpu-baseline --trace 0 > gpurun_out/r2ap_bench.json 2> gpurun_out/r2ap_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r2ap_bench.err
python - <<P
import json
d=json.loads([l for l in open("gpurun_out/r2ap_bench.json") if l.startswith("{")][-1])
print(d["value"], d["e2e"]["value"], d["e2e"]["pageable"]["value"], d["e2e"]["registered"])
P
