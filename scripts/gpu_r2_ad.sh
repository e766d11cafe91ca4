#!/bin/bash
# final state: the allocation-free entry point alone in a fresh process, the whole GPU suite, the default 1-GPU bench line, N=150
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r2ad_pytest_no_alloc_alone.log 2>&1; echo "no_alloc alone rc=$?" | tee -a gpurun_out/r2ad_pytest_no_alloc_alone.log
tail -4 gpurun_out/r2ad_pytest_no_alloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2ad_pytest_gpu.log 2>&1; echo "pytest -m gpu rc=$?" | tee -a gpurun_out/r2ad_pytest_gpu.log
tail -5 gpurun_out/r2ad_pytest_gpu.log
timeout 900 python bench.py > gpurun_out/r2ad_bench_1gpu.json 2> gpurun_out/r2ad_bench_1gpu.err; echo "bench rc=$?"
timeout 600 python bench.py --n 150 --steps 10 --warmup 3 --full-solve 0 --no-cpu-baseline > gpurun_out/r2ad_bench_n150.json 2> gpurun_out/r2ad_bench_n150.err; echo "bench150 rc=$?"
python - <<P
import json
for f in ("gpurun_out/r2ad_bench_1gpu.json", "gpurun_out/r2ad_bench_n150.json"):
    for ln in open(f):
        if ln.startswith("{"):
            d=json.loads(ln)
            print(d["value"], "ms/step", d["ms_per_step"], "asm_ms", d["components"]["picard_assembly_ms"], d["roofline_assembly"]["frac"], "pcg_us", round(d["roofline"]["avg_launch_ms"]*1e3,2), "e2e", d["e2e"]["value"], "parity", (d.get("parity") or {}).get("ok"), "cpu", (d.get("cpu_baseline") or {}).get("value"))
P
