#!/bin/bash
# Fused backward (NERFB_BWD_FUSED=1) against the 3-launch backward: whole GPU suite with it on, headline bench off / on (with the
# 2 s sustained leg), NeRF_3D bench off / on.
OUT=gpurun_out; mkdir -p $OUT
NERFB_BWD_FUSED=1 timeout 1500 python -m pytest tests -m gpu -q --tb=line -x 2>&1 | tail -6 > $OUT/r2f_pytest_fused.log; echo "gpu tests (fused bwd): $(tail -1 $OUT/r2f_pytest_fused.log)"
for F in 0 1; do
  NERFB_BWD_FUSED=$F timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-extra > $OUT/r2f_bench_fused$F.json 2> $OUT/r2f_bench_fused$F.err; echo "bench fused=$F exit $?"
  NERFB_BWD_FUSED=$F timeout 600 python bench.py --config cls_train --steps 10 --warmup 3 --no-cpu-baseline --sustain 0 > $OUT/r2f_bench_cls_fused$F.json 2> $OUT/r2f_bench_cls_fused$F.err; echo "cls bench fused=$F exit $?"
done
python - <<'PY'
import json
for t in ("fused0", "fused1", "cls_fused0", "cls_fused1"):
    try:
        d = json.loads(open(f"gpurun_out/r2f_bench_{t}.json").read().strip().splitlines()[-1])
        print(t, "ms/step", round(d["ms_per_step"], 4), "rays/s", round(d["value"]), "e2e", round(d["e2e"]["value"]), "loss", d.get("final_loss"),
              "sustained", d.get("sustained") and (round(d["sustained"]["ms_per_step"], 4), d["sustained"]["clocks"]["sm_mhz"]),
              [(k["kernel"][:24], round(k["ms"], 4)) for k in d["kernels"] if k["bound"] == "tensor"])
    except Exception as e:
        print(t, "failed", e)
PY
