#!/bin/bash
mkdir -p gpurun_out
T=${1:-r2r}
ncu --metrics gpu__time_duration.sum,launch__grid_size,launch__block_size --clock-control none --kernel-name-base demangled -c 200 --csv --log-file gpurun_out/${T}_launches.csv \
    python tools/wide_quick.py contract 2368 SA_PRECISION=fp32 > gpurun_out/${T}_ncu.log 2>&1
python - <<'PY'
import csv,collections
rows=[r for r in csv.reader(open('gpurun_out/'+'%s'%'r2r'+'_launches.csv')) if len(r)>10]
h=rows[0]; ik=h.index('Kernel Name'); im=h.index('Metric Name'); iv=h.index('Metric Value'); iid=h.index('ID')
d=collections.OrderedDict()
for r in rows[1:]:
    d.setdefault(r[iid],{'k':r[ik][:70]})[r[im]]=r[iv]
for k,v in list(d.items())[-40:]:
    print(k, v['k'], v.get('gpu__time_duration.sum'), v.get('launch__grid_size'), v.get('launch__block_size'))
PY
