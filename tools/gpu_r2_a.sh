#!/bin/bash
# Round-2 first GPU pass: new parity tests, fused-backward check (values + timing sweep), headline bench with the fused
# backward off / on, then the whole GPU suite.  Usage (via gpurun, repo root): bash tools/gpu_r2_a.sh
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > $OUT/r2a_gpu.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_round2.py -q -s 2>&1 | tail -60 > $OUT/r2a_tests_new.log; echo "new tests: $(tail -1 $OUT/r2a_tests_new.log)"
for K in nerf cls siren; do
  timeout 300 python tools/check_fused_bwd.py $K 4096 128 "6,8;7,8;5,8;6,12;6,4" > $OUT/r2a_fused_$K.log 2>&1; echo "fused $K exit $?"; tail -7 $OUT/r2a_fused_$K.log
done
timeout 120 python tools/check_fused_bwd.py nerf 4096 64 "6,8" > $OUT/r2a_fused_nerf64.log 2>&1; tail -2 $OUT/r2a_fused_nerf64.log
NERFB_BWD_FUSED=0 timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-extra --sustain 0 > $OUT/r2a_bench_unfused.json 2> $OUT/r2a_bench_unfused.err; echo "bench unfused exit $?"
NERFB_BWD_FUSED=1 timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $OUT/r2a_bench_fused.json 2> $OUT/r2a_bench_fused.err; echo "bench fused exit $?"
python - <<'PY'
import json
for t in ("unfused", "fused"):
    try:
        d = json.loads(open(f"gpurun_out/r2a_bench_{t}.json").read().strip().splitlines()[-1])
        print(t, "ms/step", round(d["ms_per_step"], 4), "rays/s", round(d["value"]), "e2e", round(d["e2e"]["value"]),
              "sustained", d.get("sustained") and round(d["sustained"]["ms_per_step"], 4),
              [(k["kernel"][:28], round(k["ms"], 4)) for k in d["kernels"] if k["bound"] == "tensor"])
    except Exception as e:
        print(t, "failed", e)
PY
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -25 > $OUT/r2a_pytest_all.log; echo "all gpu tests (defaults): $(tail -1 $OUT/r2a_pytest_all.log)"
NERFB_BWD_FUSED=1 NERFB_RENDER_FUSED=1 timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -25 > $OUT/r2a_pytest_all_fused.log; echo "all gpu tests (fused backward + fused render ON): $(tail -1 $OUT/r2a_pytest_all_fused.log)"
NERFB_RENDER_FUSED=0 timeout 300 python bench.py --config render --steps 3 --warmup 3 --no-cpu-baseline --sustain 0 > $OUT/r2a_render_sep.json 2> $OUT/r2a_render_sep.err
NERFB_RENDER_FUSED=1 timeout 300 python bench.py --config render --steps 3 --warmup 3 --no-cpu-baseline --sustain 0 > $OUT/r2a_render_fused.json 2> $OUT/r2a_render_fused.err
python - <<'PY'
import json
for t in ("sep", "fused"):
    try:
        d = json.loads(open(f"gpurun_out/r2a_render_{t}.json").read().strip().splitlines()[-1])
        print("render", t, "ms", round(d["ms_per_step"], 3), "rays/s", round(d["value"]), "e2e", round(d["e2e"]["value"]), "launches", d["gpu_launches"])
    except Exception as e:
        print("render", t, "failed", e)
PY
