#!/bin/bash
# End-of-round evidence pass (one GPU): parity tests, smoke, headline bench (+ reference arm), the other configurations,
# EG3D and streaming-kernel benches, ncu launch list of the bench command, ncu full captures of the tcgen05 kernels.
# Usage (via gpurun, repo root): bash tools/gpu_final.sh <tag>
TAG=${1:-final}
OUT=gpurun_out; mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q > $OUT/pytest_$TAG.log 2>&1; echo "pytest exit $?" | tee -a $OUT/pytest_$TAG.log; tail -2 $OUT/pytest_$TAG.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke_$TAG.log 2>&1; echo "smoke exit $?"; tail -2 $OUT/smoke_$TAG.log
timeout 900 python bench.py --steps 20 --warmup 3 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench exit $?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref_$TAG.json 2> $OUT/bench_ref_$TAG.err; echo "bench ref exit $?"
timeout 600 python tools/bench_configs.py render siren cls siren_train render_e2e > $OUT/bench_configs_$TAG.json 2> $OUT/bench_configs_$TAG.err; echo "configs exit $?"
timeout 600 python tools/bench_eg3d.py > $OUT/bench_eg3d_$TAG.json 2> $OUT/bench_eg3d_$TAG.err; echo "eg3d exit $?"
timeout 600 python tools/bench_stream_kernels.py > $OUT/stream_$TAG.json 2> $OUT/stream_$TAG.err; echo "stream exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches_$TAG.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $OUT/ncu_launch_$TAG.log 2>&1; echo "ncu launches exit $?"
for M in train fwd siren siren_train; do
  timeout 300 python tools/prof_mlp.py $M 2 > $OUT/prof_plain_${M}_$TAG.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:mlp_tc -c 4 -o $OUT/mlp_tc_${M}_$TAG -f \
    python tools/prof_mlp.py $M 2 > $OUT/ncu_full_${M}_$TAG.log 2>&1; echo "ncu full $M exit $?"
done
ls -la $OUT | tail -30
