#!/bin/bash
# Training forward on the tile-pair kernel (NERFB_FWD_PAIR=3): whole GPU suite, then the headline bench with 1 / 3.
OUT=gpurun_out; mkdir -p $OUT
NERFB_FWD_PAIR=3 timeout 1500 python -m pytest tests -m gpu -q --tb=line -x 2>&1 | tail -15 > $OUT/r2e_pytest_pair3.log; echo "gpu tests (pair=3): $(tail -1 $OUT/r2e_pytest_pair3.log)"
for P in 1 3; do
  NERFB_FWD_PAIR=$P timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --sustain 0 --no-extra > $OUT/r2e_bench_pair$P.json 2> $OUT/r2e_bench_pair$P.err; echo "bench pair=$P exit $?"
done
python - <<'PY'
import json
for t in ("1", "3"):
    try:
        d = json.loads(open(f"gpurun_out/r2e_bench_pair{t}.json").read().strip().splitlines()[-1])
        print("pair", t, "ms/step", round(d["ms_per_step"], 4), "rays/s", round(d["value"]), "e2e", round(d["e2e"]["value"]), "loss", d.get("final_loss"),
              [(k["kernel"][:28], round(k["ms"], 4)) for k in d["kernels"] if k["bound"] == "tensor"])
    except Exception as e:
        print(t, "failed", e)
PY
