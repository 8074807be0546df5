#!/bin/bash
# A/B of bench.py argument sets: one line per run in $RUNS ("name|args"), summaries printed, full lines in gpurun_out/${TAG}_<name>.json
TAG=${TAG:-ab}; O=gpurun_out; mkdir -p $O
while IFS= read -r run; do
  [ -z "$run" ] && continue
  name=${run%%|*}; args=${run#*|}
  timeout 900 python bench.py --no-cpu-baseline $args > $O/${TAG}_$name.json 2> $O/${TAG}_$name.err; rc=$?
  python - <<PY
import json
try:
    d = json.loads(open("$O/${TAG}_$name.json").read().strip().splitlines()[-1])
    print("[$name] value", round(d["value"]), "ms/step", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"]), "frac", round(d["step_roofline"]["frac"], 3), "prefill_ms", round(d["prefill_ms"], 1))
    if d.get("kernels"): print("   ", {k: v["ms_per_launch"] for k, v in d["kernels"].items()})
    if d.get("sweep"): print("   ", {k: {a: (round(b, 4) if isinstance(b, float) else b) for a, b in v.items() if a in ("tokens_per_s", "ms_per_step", "hbm_frac", "error")} for k, v in d["sweep"].items()})
except Exception as e:
    print("[$name] rc $rc no bench line:", e); print(open("$O/${TAG}_$name.err").read()[-1200:])
PY
done <<< "$RUNS"
