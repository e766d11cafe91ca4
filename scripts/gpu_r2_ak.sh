#!/bin/bash
# e2e leg at N GPUs with and without the ranks bound to the cores next to their GPU (bench.py: Ctx.host_affinity)
N=$1
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2ak_topo_${N}gpu.txt 2>&1
for a in 0 1; do
  FK_BENCH_AFFINITY=$a timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$a bench.py --gpus $N --steps 10 --warmup 3 --strong-n 0 --parity-n 0 --full-solve 0 --trace 0 --no-cpu-baseline > gpurun_out/r2ak_bench_${N}gpu_affinity$a.json 2> gpurun_out/r2ak_${N}gpu_affinity$a.err; echo "affinity $a rc=$?"
done
python - <<P
import json
for a in (0, 1):
    for ln in open("gpurun_out/r2ak_bench_${N}gpu_affinity%d.json" % a):
        if ln.startswith("{"):
            d=json.loads(ln)
            print(a, d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], "pageable", d["e2e"]["pageable"]["value"], d["config"].get("host_affinity"), d["components"]["setup_s_untimed"])
P
