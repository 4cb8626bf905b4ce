#!/bin/bash
# NeRF_3D on the fused backward (default) against the 3-launch backward: GPU suite, cls_train bench both ways, headline bench.
OUT=gpurun_out; mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q --tb=line -x 2>&1 | tail -3 > $OUT/r2h_pytest.log; echo "gpu tests: $(tail -1 $OUT/r2h_pytest.log)"
for F in 0 2; do
  NERFB_BWD_FUSED=$F timeout 600 python bench.py --config cls_train --steps 10 --warmup 3 --no-cpu-baseline > $OUT/r2h_bench_cls_fused$F.json 2> $OUT/r2h_bench_cls_fused$F.err; echo "cls bench fused=$F exit $?"
done
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-extra > $OUT/r2h_bench.json 2> $OUT/r2h_bench.err; echo "bench exit $?"
python - <<'PY'
import json
for t in ("cls_fused0", "cls_fused2", ""):
    try:
        d = json.loads(open(f"gpurun_out/r2h_bench{'_' + t if t else ''}.json").read().strip().splitlines()[-1])
        print(t or "train", "ms/step", round(d["ms_per_step"], 4), "rays/s", round(d["value"]), "e2e", round(d["e2e"]["value"]), "loss", d.get("final_loss"),
              "sustained", d.get("sustained") and (round(d["sustained"]["ms_per_step"], 4), d["sustained"]["clocks"]["sm_mhz"]),
              [(k["kernel"][:24], round(k["ms"], 4)) for k in d["kernels"] if k["bound"] == "tensor"])
    except Exception as e:
        print(t, "failed", e)
PY
