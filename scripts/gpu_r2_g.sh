#!/bin/bash
# round 2, call g (2 GPUs): fused element kernel + pipelined rescale kernel (parity, event times), halo timeline at 2 GPUs
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_rescale.py tests/test_gpu_round2.py -m gpu -q -x > gpurun_out/r2g_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2g_pytest.log
tail -5 gpurun_out/r2g_pytest.log
python scripts/asm_bench.py 119 20 > gpurun_out/r2g_asm_bench.jsonl 2> gpurun_out/r2g_asm_bench.err; echo "asm_bench rc=$?"
cat gpurun_out/r2g_asm_bench.jsonl
python scripts/asm_bench.py 150 10 > gpurun_out/r2g_asm_bench_n150.jsonl 2> gpurun_out/r2g_asm_bench_n150.err; echo "asm_bench150 rc=$?"
cat gpurun_out/r2g_asm_bench_n150.jsonl
run() { local name=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --full-solve 0 "$@" > gpurun_out/r2g_$name.json 2> gpurun_out/r2g_$name.err; echo "$name rc=$?"
}
run rec0_fused --recurrence 0 --strong-n 0 --parity-n 0
run rec0_strong --recurrence 0 --strong-n 128
tail -3 gpurun_out/r2g_rec0_strong.err
