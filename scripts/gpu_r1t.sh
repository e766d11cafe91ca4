mkdir -p gpurun_out
FK_DBG=0,4,8,1,3 timeout 600 python scripts/asm_phases.py 119 > gpurun_out/r1t_phases.log 2>&1; tail -6 gpurun_out/r1t_phases.log
FK_DBG=0,3 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'assemble_kernel|max_array|fold' --csv --log-file gpurun_out/r1t_launches.csv python scripts/asm_phases.py 119 > gpurun_out/r1t_ncu.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/r1t_launches.csv')) if len(r)>5]
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value')
import collections
d=collections.defaultdict(list)
for r in rows[1:]:
    d[r[ki][:60]].append(float(r[vi].replace(',','')))
for k,v in d.items(): print(k, len(v), [round(x/1000,1) for x in v[:40:3]])
PY
