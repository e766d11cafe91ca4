mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r3s_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r3s_pytest.log
tail -n 3 gpurun_out/r3s_pytest.log
for N in 119 150; do
  timeout 600 python bench.py --n $N --no-cpu-baseline --full-solve 0 > gpurun_out/r3s_bench_n$N.json 2> gpurun_out/r3s_bench_n$N.err; echo "n$N exit $?"
  python -c "
import json
d=json.loads(open('gpurun_out/r3s_bench_n$N.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value'], d['components'])"
done
