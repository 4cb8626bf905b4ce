#!/bin/bash
# Quick validation of a kernel change: whole GPU suite (defaults), then the headline bench without CPU legs.
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-r2d}
timeout 1500 python -m pytest tests -m gpu -q --tb=line -x 2>&1 | tail -15 > $OUT/${TAG}_pytest.log; echo "gpu tests: $(tail -1 $OUT/${TAG}_pytest.log)"
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --sustain 0 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench exit $?"
python - <<PY
import json
try:
    d = json.loads(open("$OUT/${TAG}_bench.json").read().strip().splitlines()[-1])
    print("ms/step", round(d["ms_per_step"], 4), "rays/s", round(d["value"]), "e2e", round(d["e2e"]["value"]), "loss", d.get("final_loss"),
          "render", round(d["extra"]["render"]["value"]) if d.get("extra") else None,
          [(k["kernel"][:28], round(k["ms"], 4)) for k in d["kernels"] if k["bound"] == "tensor"])
except Exception as e:
    print("failed", e)
PY
