#!/bin/bash
# Per-launch durations of the MLP kernels alone (ncu launch list of tools/prof_mlp.py).  Usage: bash tools/gpu_launches.sh <tag> [fwd|train]
TAG=$1; MODE=${2:-train}
OUT=gpurun_out; mkdir -p $OUT
timeout 300 python tools/prof_mlp.py $MODE 2 > $OUT/prof_plain_$TAG.log 2>&1 || { tail -5 $OUT/prof_plain_$TAG.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed --clock-control none -c 40 --csv --log-file $OUT/ll_$TAG.csv python tools/prof_mlp.py $MODE 2 > $OUT/ll_$TAG.log 2>&1
python - <<PY
import csv
rows=list(csv.reader(open("$OUT/ll_$TAG.csv")))
hi=[i for i,r in enumerate(rows) if 'Kernel Name' in r][0]
h=rows[hi]; kn=h.index('Kernel Name'); mn=h.index('Metric Name'); mv=h.index('Metric Value'); idc=h.index('ID')
d={}
for r in rows[hi+1:]:
    if len(r)!=len(h): continue
    d.setdefault((int(r[idc]), r[kn].split('(')[0].replace('<unnamed>::','')), {})[r[mn]]=r[mv]
for (i,k),m in sorted(d.items()):
    print(f"{i:3d} {k[:40]:40s} " + "  ".join(f"{a.split('.')[0][-22:]}={b}" for a,b in m.items()))
PY
