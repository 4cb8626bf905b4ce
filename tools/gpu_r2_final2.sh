#!/bin/bash
# End-of-round-2 refresh after the EG3D tensor-core decoder and the register-resident compositing kernels (one GPU): smoke,
# headline bench (+ reference arm), every other configuration under bench.py's own clock, ncu launch list of the bench command.
# The tcgen05 kernels did not change since tools/gpu_r2_final.sh: their ncu captures stand.  Usage: bash tools/gpu_r2_final2.sh <tag>
TAG=${1:-r2}
OUT=gpurun_out; mkdir -p $OUT
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke_$TAG.log 2>&1; echo "smoke exit $?"; tail -2 $OUT/smoke_$TAG.log
timeout 900 python bench.py --steps 20 --warmup 3 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench exit $?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref_$TAG.json 2> $OUT/bench_ref_$TAG.err; echo "bench ref exit $?"
for C in render siren_render cls_train cls_train_strong eg3d eg3d_train; do
  timeout 600 python bench.py --config $C --steps 5 --warmup 3 > $OUT/bench_${C}_$TAG.json 2> $OUT/bench_${C}_$TAG.err; echo "bench $C exit $?"
done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches_$TAG.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra --sustain 0 > $OUT/ncu_launch_$TAG.log 2>&1; echo "ncu launches exit $?"
python - <<PY
import json, glob
for f in sorted(glob.glob("$OUT/bench_*_$TAG.json")) + ["$OUT/bench_$TAG.json"]:
    try:
        txt = open(f).read().strip().splitlines()
        d = json.loads(txt[-1])
        print(f.split("/")[-1], "| lines", len(txt), "|", d.get("metric"), "| value", round(d.get("value")), d.get("unit"), "ms", round(d.get("ms_per_step"), 3), "e2e", round((d.get("e2e") or {}).get("value") or 0),
              "frac", (d.get("roofline") or {}).get("frac"), "cpu", (d.get("cpu_baseline") or {}).get("value"))
    except Exception as e:
        print(f, "failed", e)
PY
