#!/bin/bash
# One GPU-box pass: parity tests, bench line, ncu launch list of the bench, full ncu capture of the tcgen05 kernels.
# Usage (from the repo root, via gpurun): bash tools/gpu_check.sh [tag]
TAG=${1:-cur}
OUT=gpurun_out
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_$TAG.log 2>&1; echo "pytest exit $?" | tee -a $OUT/pytest_$TAG.log
tail -3 $OUT/pytest_$TAG.log
timeout 600 python bench.py --steps 20 --warmup 3 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench exit $?"
tail -c 3000 $OUT/bench_$TAG.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches_$TAG.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $OUT/ncu_launch_$TAG.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mlp_tc -c 6 -o $OUT/mlp_tc_$TAG -f \
  python tools/prof_mlp.py train 2 > $OUT/ncu_full_$TAG.log 2>&1; echo "ncu full exit $?"
ls -la $OUT | tail -20
