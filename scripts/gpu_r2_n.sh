#!/bin/bash
# round 2, call n (1 GPU): the default bench line (CPU baseline, parity, solve blocks) + the 1-GPU point of the strong curve (100 M tets)
mkdir -p gpurun_out
timeout 1700 python bench.py --steps 10 --warmup 3 --strong-single > gpurun_out/r2n_bench_1gpu.json 2> gpurun_out/r2n_bench_1gpu.err; echo "bench rc=$?"
tail -3 gpurun_out/r2n_bench_1gpu.err
python - <<P
import json
for ln in open("gpurun_out/r2n_bench_1gpu.json"):
    if ln.startswith("{"):
        d=json.loads(ln)
        print(d["n_gpus"], d["value"], d["ms_per_step"], "asm", d["components"]["picard_assembly_ms"], d["roofline_assembly"]["frac"], "pcg_us", d["roofline"]["avg_launch_ms"]*1e3, d["roofline"]["frac"], d["roofline"]["frac_of_floor"], "e2e", d["e2e"]["value"], d["e2e"]["pageable"]["value"])
        print(" split", d["roofline"].get("kernel_split_us"))
        s=d.get("strong") or {}
        print(" strong", s.get("mtets_per_s"), s.get("ms_per_step"), s.get("pcg_iteration_us"), s.get("picard_assembly_ms"), s.get("setup_s_untimed"), s.get("mesh_generation_s"))
        print(" nonlin", d.get("nonlinear_solve"))
        print(" cpu", d.get("cpu_baseline"))
        print(" parity", d.get("parity"))
        print(" solve", d.get("solve"))
P
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2n_bench_reference.json 2> gpurun_out/r2n_bench_reference.err; echo "ref rc=$?"
cut -c1-400 gpurun_out/r2n_bench_reference.json
