#!/bin/bash
# Round-2 evidence pass (one GPU): whole GPU suite, smoke, headline bench (+ reference arm), every other configuration under
# bench.py's own clock, streaming kernels, ncu launch list of the bench command, ncu full captures + metric tables of the
# tcgen05 kernels.  Usage (via gpurun, repo root): bash tools/gpu_r2_final.sh <tag>
TAG=${1:-r2}
OUT=gpurun_out; mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q --tb=line > $OUT/pytest_$TAG.log 2>&1; echo "pytest exit $?"; tail -2 $OUT/pytest_$TAG.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke_$TAG.log 2>&1; echo "smoke exit $?"; tail -2 $OUT/smoke_$TAG.log
timeout 900 python bench.py --steps 20 --warmup 3 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench exit $?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref_$TAG.json 2> $OUT/bench_ref_$TAG.err; echo "bench ref exit $?"
for C in render siren_render cls_train cls_train_strong eg3d; do
  timeout 600 python bench.py --config $C --steps 5 --warmup 3 > $OUT/bench_${C}_$TAG.json 2> $OUT/bench_${C}_$TAG.err; echo "bench $C exit $?"
done
NERFB_FWD_PAIR=0 timeout 600 python bench.py --config render --steps 5 --warmup 3 --no-cpu-baseline --sustain 0 > $OUT/bench_render_single_$TAG.json 2> /dev/null; echo "bench render (single-tile kernel) exit $?"
timeout 600 python tools/bench_stream_kernels.py > $OUT/stream_$TAG.json 2> $OUT/stream_$TAG.err; echo "stream exit $?"
timeout 300 python tools/check_pair.py 4096 20 1,2 > $OUT/check_pair_$TAG.json 2>&1; echo "check_pair exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches_$TAG.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra --sustain 0 > $OUT/ncu_launch_$TAG.log 2>&1; echo "ncu launches exit $?"
for M in fwd train siren; do
  timeout 300 python tools/prof_mlp.py $M 2 > $OUT/prof_plain_${M}_$TAG.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:mlp_tc -c 4 -o $OUT/mlp_tc_${M}_$TAG -f \
    python tools/prof_mlp.py $M 2 > $OUT/ncu_full_${M}_$TAG.log 2>&1; echo "ncu full $M exit $?"
  ncu -i $OUT/mlp_tc_${M}_$TAG.ncu-rep --page raw --csv > $OUT/raw_${M}_$TAG.csv 2>/dev/null && python tools/ncu_metrics.py $OUT/raw_${M}_$TAG.csv "$M ($TAG)" > $OUT/metrics_${M}_$TAG.txt
done
python - <<PY
import json, glob
for f in sorted(glob.glob("$OUT/bench_*_$TAG.json")) + ["$OUT/bench_$TAG.json"]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split("/")[-1], "|", d.get("metric"), "| value", round(d.get("value")), d.get("unit"), "ms", round(d.get("ms_per_step"), 3), "e2e", round((d.get("e2e") or {}).get("value") or 0),
              "frac", (d.get("roofline") or {}).get("frac"), "cpu", (d.get("cpu_baseline") or {}).get("value"))
    except Exception as e:
        print(f, "failed", e)
PY
ls -la $OUT | tail -40
