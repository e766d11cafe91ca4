mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r3l_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r3l_pytest.log
tail -n 3 gpurun_out/r3l_pytest.log
for T in "7=0" "7=1"; do
  timeout 600 python bench.py --no-cpu-baseline --full-solve 0 --tune "$T" > gpurun_out/r3l_bench_$T.json 2> gpurun_out/r3l_bench_$T.err; echo "bench $T exit $?"
  python -c "
import json
d=json.loads(open('gpurun_out/r3l_bench_$T.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value'], d['components'])"
done
