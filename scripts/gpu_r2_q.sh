#!/bin/bash
# round 2, call q (2 GPUs): whole GPU suite on a 2-GPU box (multi-GPU tests not skipped), heat workload on 2 GPUs (configs[2]),
# single-process fk_multi bench on 1 and 2 GPUs
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --durations=8 > gpurun_out/r2q_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2q_pytest.log
tail -14 gpurun_out/r2q_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --workload heat --n 150 --steps 20 --warmup 3 > gpurun_out/r2q_heat_2gpu.json 2> gpurun_out/r2q_heat_2gpu.err; echo "heat2 rc=$?"
grep "^{" gpurun_out/r2q_heat_2gpu.json | cut -c1-600
timeout 900 python bench.py --single-process --gpus 2 --n 150 --steps 5 --warmup 3 > gpurun_out/r2q_single_process_2gpu.json 2> gpurun_out/r2q_single_process_2gpu.err; echo "sp2 rc=$?"
grep "^{" gpurun_out/r2q_single_process_2gpu.json | cut -c1-900; tail -3 gpurun_out/r2q_single_process_2gpu.err
timeout 900 python bench.py --single-process --gpus 1 --n 150 --steps 5 --warmup 3 --sp-check 0 > gpurun_out/r2q_single_process_1gpu.json 2> gpurun_out/r2q_single_process_1gpu.err; echo "sp1 rc=$?"
grep "^{" gpurun_out/r2q_single_process_1gpu.json | cut -c1-500
