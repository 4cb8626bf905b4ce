#!/bin/bash
# Pair-forward validation: whole GPU suite with NERFB_FWD_PAIR=1, then the headline bench with the pair forward off / on.
OUT=gpurun_out; mkdir -p $OUT
NERFB_FWD_PAIR=1 timeout 1500 python -m pytest tests -m gpu -q --tb=line -x 2>&1 | tail -15 > $OUT/r2c_pytest_pair.log; echo "gpu tests (pair): $(tail -1 $OUT/r2c_pytest_pair.log)"
for P in 0 1; do
  NERFB_FWD_PAIR=$P timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --sustain 0 > $OUT/r2c_bench_pair$P.json 2> $OUT/r2c_bench_pair$P.err; echo "bench pair=$P exit $?"
done
python - <<'PY'
import json
for t in ("0", "1"):
    try:
        d = json.loads(open(f"gpurun_out/r2c_bench_pair{t}.json").read().strip().splitlines()[-1])
        print("pair", t, "ms/step", round(d["ms_per_step"], 4), "rays/s", round(d["value"]), "e2e", round(d["e2e"]["value"]), "loss", d.get("final_loss"),
              "render", round(d["extra"]["render"]["value"]) if d.get("extra") else None,
              [(k["kernel"][:28], round(k["ms"], 4)) for k in d["kernels"] if k["bound"] == "tensor"])
    except Exception as e:
        print(t, "failed", e)
PY
