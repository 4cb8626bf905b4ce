#!/bin/bash
# Quick GPU pass: parity tests + bench line (no ncu).  Usage: bash tools/gpu_quick.sh [tag] [pytest -k expr]
TAG=${1:-q}
OUT=gpurun_out
mkdir -p $OUT
if [ -n "$2" ]; then K=(-k "$2"); else K=(); fi
timeout 1200 python -m pytest tests -m gpu -x -q "${K[@]}" > $OUT/pytest_$TAG.log 2>&1; echo "pytest exit $?"
tail -15 $OUT/pytest_$TAG.log
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench exit $?"
python - <<PY
import json
try:
    d=json.loads(open("$OUT/bench_$TAG.json").read().strip().splitlines()[-1])
    print("rays/s", round(d["value"]), "ms/step", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"]))
    for k in d["kernels"]: print("  %-45s %8.3f ms  %8.1f %s  frac %.3f" % (k["kernel"], k["ms"], k["achieved"], k["unit"], k["frac"]))
except Exception as e:
    print("bench parse failed", e); print(open("$OUT/bench_$TAG.err").read()[-2000:])
PY
