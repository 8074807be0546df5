#!/bin/bash
# A/B of decode-path switches on the bench workload: prints value / ms per step for each environment setting
O=gpurun_out; mkdir -p $O
TAG=${TAG:-ab}
if [ -n "$TESTS" ]; then timeout 1500 python -m pytest $TESTS -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -5 $O/${TAG}_pytest.log; fi
i=0
while IFS= read -r cfg; do
  [ -z "$cfg" ] && continue
  i=$((i+1))
  env $cfg timeout 600 python bench.py --no-cpu-baseline ${BENCH_ARGS} > $O/${TAG}_bench_$i.json 2> $O/${TAG}_bench_$i.err
  echo "[$cfg] exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("$O/${TAG}_bench_$i.json").read().strip().splitlines()[-1])
    print("  value", round(d["value"]), "ms/step", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"]), "frac", round(d["step_roofline"]["frac"], 3))
    if d.get("kernels"): print("  ", {k: v["ms_per_launch"] for k, v in d["kernels"].items()})
    if d.get("sweep"): print("  ", d["sweep"])
except Exception as e:
    print("  no bench line:", e); print(open("$O/${TAG}_bench_$i.err").read()[-1500:])
PY
done <<< "$CONFIGS"
