mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r3n_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r3n_pytest.log
tail -n 3 gpurun_out/r3n_pytest.log
timeout 600 python bench.py --no-cpu-baseline --full-solve 0 > gpurun_out/r3n_bench.json 2> gpurun_out/r3n_bench.err; echo "bench exit $?" >> gpurun_out/r3n_bench.err
cut -c1-1500 gpurun_out/r3n_bench.json; tail -n 3 gpurun_out/r3n_bench.err
