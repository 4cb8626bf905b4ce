#!/bin/bash
# Multi-GPU headline lines only (train weak scaling + NeRF render strong scaling).  Usage (via gpurun --gpus N): bash tools/gpu_r2_scale_min.sh N
N=${1:-8}; OUT=gpurun_out; mkdir -p $OUT
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:2}"; }
run 29521 --steps 20 --warmup 3 --no-extra > $OUT/r2_bench_n$N.json 2> $OUT/r2_bench_n$N.err; echo "train exit $?"
run 29523 --config render --steps 5 --warmup 3 --no-cpu-baseline --sustain 0 > $OUT/r2_bench_render_n$N.json 2> $OUT/r2_bench_render_n$N.err; echo "render exit $?"
python - <<PY
import json, glob
for f in sorted(glob.glob("$OUT/r2_bench_*n$N.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split("/")[-1], "| value", round(d["value"]), d["unit"], "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"]), "n_gpus", d["n_gpus"], d.get("scaling"))
    except Exception as e:
        print(f, "failed", e)
PY
