#!/bin/bash
# quick GPU visit: selected tests + a short bench (TAG, TESTS, BENCH_ARGS from the environment)
TAG=${TAG:-q}
O=gpurun_out
mkdir -p $O
timeout 1200 python -m pytest ${TESTS:-tests} -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?"
tail -15 $O/${TAG}_pytest.log
timeout 600 python bench.py ${BENCH_ARGS:---no-cpu-baseline} > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench exit $?"
tail -5 $O/${TAG}_bench.err
python - <<PY
import json
try:
    d = json.loads(open("$O/${TAG}_bench.json").read().strip().splitlines()[-1])
    print("value", round(d["value"]), "ms/step", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"]), "frac", round(d["step_roofline"]["frac"], 3))
    print(json.dumps(d.get("kernels")))
    print(json.dumps(d.get("sweep")))
except Exception as e:
    print("no bench line:", e)
PY
