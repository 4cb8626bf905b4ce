#!/bin/bash
# Round-2 measurement pass (one GPU): whole GPU suite, smoke, headline bench (with extra.render / sustained / CPU baselines),
# reference arm, the other configurations under bench.py's own clock, streaming kernels, launch list.
# Usage (via gpurun, repo root): bash tools/gpu_r2_b.sh <tag>
TAG=${1:-r2b}
OUT=gpurun_out; mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q --tb=line > $OUT/pytest_$TAG.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest_$TAG.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke_$TAG.log 2>&1; echo "smoke exit $?"; tail -2 $OUT/smoke_$TAG.log
timeout 900 python bench.py --steps 20 --warmup 3 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench exit $?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref_$TAG.json 2> $OUT/bench_ref_$TAG.err; echo "bench ref exit $?"
for C in render siren_render cls_train cls_train_strong eg3d; do
  timeout 600 python bench.py --config $C --steps 5 --warmup 3 > $OUT/bench_${C}_$TAG.json 2> $OUT/bench_${C}_$TAG.err; echo "bench $C exit $?"
done
timeout 600 python tools/bench_stream_kernels.py > $OUT/stream_$TAG.json 2> $OUT/stream_$TAG.err; echo "stream exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches_$TAG.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra --sustain 0 > $OUT/ncu_launch_$TAG.log 2>&1; echo "ncu launches exit $?"
python - <<PY
import json, glob
for f in sorted(glob.glob("$OUT/bench_*_$TAG.json")) + ["$OUT/bench_$TAG.json"]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split("/")[-1], d.get("metric"), "value", d.get("value"), d.get("unit"), "ms", d.get("ms_per_step"), "e2e", (d.get("e2e") or {}).get("value"),
              "frac", (d.get("roofline") or {}).get("frac"), "cpu", (d.get("cpu_baseline") or {}).get("value"))
    except Exception as e:
        print(f, "failed", e)
PY
