#!/bin/bash
# generic visit: TESTS (pytest paths, each its own process), BENCH_ARGS, NCU_K (kernel regex for one --set full capture), TAG
TAG=${TAG:-v}
O=gpurun_out; mkdir -p $O
for f in ${TESTS}; do
  b=$(basename $f .py)
  timeout 1500 python -m pytest $f -m gpu -q --timeout 900 ${PYTEST_ARGS} > $O/${TAG}_$b.log 2>&1; echo "$f exit $?"
  tail -4 $O/${TAG}_$b.log
done
if [ -n "${BENCH_ARGS}" ]; then
  timeout 900 python bench.py ${BENCH_ARGS} > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench exit $?"
  tail -5 $O/${TAG}_bench.err
  python - <<PY
import json
try:
    d = json.loads(open("$O/${TAG}_bench.json").read().strip().splitlines()[-1])
    print("value", round(d["value"]), "ms/step", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"]), "frac", round(d["step_roofline"]["frac"], 3), "prefill_ms", round(d["prefill_ms"], 2))
    if d.get("kernels"): print({k: (v["ms_per_launch"], v["GBps"]) for k, v in d["kernels"].items()})
    if d.get("sweep"): print({k: {a: (round(b, 4) if isinstance(b, float) else b) for a, b in v.items() if a in ("tokens_per_s", "ms_per_step", "hbm_frac", "prefill_ms", "error")} for k, v in d["sweep"].items()})
except Exception as e:
    print("no bench line:", e)
PY
fi
if [ -n "${EXTRA}" ]; then bash -c "${EXTRA}"; fi
if [ -n "${NCU_K}" ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"${NCU_K}" -c ${NCU_C:-4} --launch-skip ${NCU_SKIP:-40} -o $O/${TAG}_full -f \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-breakdown --sweep '' > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full exit $?"
  ncu -i $O/${TAG}_full.ncu-rep --page raw --csv > $O/${TAG}_full_raw.csv 2>/dev/null
  ncu -i $O/${TAG}_full.ncu-rep --page source --csv > $O/${TAG}_full_source.csv 2>/dev/null
fi
ls -la $O | tail -8
