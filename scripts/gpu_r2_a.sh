#!/bin/bash
# round 2, call a: the whole GPU suite (new parity-at-size tests included) + one bench line per recurrence
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/r2a_smi.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r2a_smoke.log
timeout 1500 python -m pytest tests -m gpu -q -s --durations=15 > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2a_bench_rec1.json 2> gpurun_out/r2a_bench_rec1.err; echo "bench rc=$?"
timeout 600 python bench.py --steps 10 --warmup 3 --recurrence 0 --no-cpu-baseline --full-solve 0 > gpurun_out/r2a_bench_rec0.json 2> gpurun_out/r2a_bench_rec0.err; echo "bench0 rc=$?"
tail -5 gpurun_out/r2a_smoke.log; tail -40 gpurun_out/r2a_pytest.log
