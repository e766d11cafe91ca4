#!/bin/bash
# side-by-side sweep: launch shapes of the streamed sums x CTA-slot splits
mkdir -p gpurun_out
for v in 0 1 3; do
  FK_GEO_VARIANT=$v FK_WINDOWS=65536 FK_SPLITS=41,51,61,52,42 python scripts/asm_bench.py 119 10 > gpurun_out/r2ai_asm_v$v.jsonl 2> gpurun_out/r2ai_v$v.err; echo "v$v rc=$?"
  grep "split" gpurun_out/r2ai_asm_v$v.jsonl | cut -c1-140
done
