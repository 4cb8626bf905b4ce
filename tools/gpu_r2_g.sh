#!/bin/bash
# Evidence for the fused-backward default: suite, smoke, headline bench (sustained leg), launch list, ncu full of the training kernels.
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-r2g}
timeout 1500 python -m pytest tests -m gpu -q --tb=line > $OUT/pytest_$TAG.log 2>&1; echo "pytest exit $?"; tail -2 $OUT/pytest_$TAG.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke_$TAG.log 2>&1; echo "smoke exit $?"; tail -1 $OUT/smoke_$TAG.log
timeout 900 python bench.py --steps 20 --warmup 3 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches_$TAG.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra --sustain 0 > $OUT/ncu_launch_$TAG.log 2>&1; echo "ncu launches exit $?"
timeout 300 python tools/prof_mlp.py train 2 > $OUT/prof_plain_train_$TAG.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mlp_tc -c 3 -o $OUT/mlp_tc_train_$TAG -f \
  python tools/prof_mlp.py train 2 > $OUT/ncu_full_train_$TAG.log 2>&1; echo "ncu full train exit $?"
ncu -i $OUT/mlp_tc_train_$TAG.ncu-rep --page raw --csv > $OUT/raw_train_$TAG.csv 2>/dev/null && python tools/ncu_metrics.py $OUT/raw_train_$TAG.csv "train ($TAG)" > $OUT/metrics_train_$TAG.txt
python - <<PY
import json
d = json.loads(open("$OUT/bench_$TAG.json").read().strip().splitlines()[-1])
print("ms/step", round(d["ms_per_step"], 4), "rays/s", round(d["value"]), "e2e", round(d["e2e"]["value"]), "sustained", d["sustained"]["ms_per_step"], d["sustained"]["clocks"]["sm_mhz"],
      "render", round(d["extra"]["render"]["value"]), "roofline", d["roofline"]["kernel"], d["roofline"]["frac"], d["roofline"]["traffic"])
PY
cat $OUT/metrics_train_$TAG.txt | cut -c1-260
