#!/bin/bash
# round 2, call h (2 GPUs): rescale plan windows (gather locality), 256-bit node loads, acquire load instead of fence at 2 GPUs,
# single-process multi-GPU entry points
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_rescale.py tests/test_gpu.py -m gpu -q -x > gpurun_out/r2h_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2h_pytest.log
tail -4 gpurun_out/r2h_pytest.log
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q -x -k single_process > gpurun_out/r2h_pytest_multi.log 2>&1; echo "pytest multi rc=$?" >> gpurun_out/r2h_pytest_multi.log
tail -15 gpurun_out/r2h_pytest_multi.log
python scripts/asm_bench.py 119 15 > gpurun_out/r2h_asm_bench.jsonl 2> gpurun_out/r2h_asm_bench.err; echo "asm_bench rc=$?"
cat gpurun_out/r2h_asm_bench.jsonl
run() { local name=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --full-solve 0 "$@" > gpurun_out/r2h_$name.json 2> gpurun_out/r2h_$name.err; echo "$name rc=$?"
}
run rec0_fused --recurrence 0 --strong-n 0 --parity-n 8
run rec1 --recurrence 1 --strong-n 0 --parity-n 0
