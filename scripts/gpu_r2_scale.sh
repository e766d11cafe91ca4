#!/bin/bash
# round 2: one bench line at N GPUs (weak headline + multi_parity + strong block on the N=256 cube), as the driver launches it
N=$1
mkdir -p gpurun_out
timeout 1700 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2s_bench_${N}gpu.json 2> gpurun_out/r2s_bench_${N}gpu.err; echo "bench $N rc=$?"
tail -3 gpurun_out/r2s_bench_${N}gpu.err
python - <<P
import json
for ln in open("gpurun_out/r2s_bench_${N}gpu.json"):
    if ln.startswith("{"):
        d=json.loads(ln)
        print(d["n_gpus"], d["value"], d["ms_per_step"], "asm", d["components"]["picard_assembly_ms"], "pcg_us", d["roofline"]["avg_launch_ms"]*1e3, "e2e", d["e2e"]["value"])
        print(" split", d["roofline"].get("kernel_split_us"))
        print(" parity", d.get("multi_parity", {}).get("ok"), d.get("multi_parity", {}).get("relerr_x"))
        s=d.get("strong") or {}
        print(" strong", s.get("mtets_per_s"), s.get("ms_per_step"), s.get("pcg_iteration_us"), s.get("picard_assembly_ms"), s.get("setup_s_untimed"), s.get("mesh_generation_s"))
        print(" nonlin", d.get("nonlinear_solve"))
P
