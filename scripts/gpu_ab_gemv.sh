#!/bin/bash
# A/B of t3b_options_t.gemv_mma (tensor-core GEMV for <= 8 rows) at 1..4 requests:
# the bit-exactness tests first, then bench.py device-resident ms/step per setting.  Everything lands in gpurun_out/$TAG_*.
TAG=${TAG:-ab}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_round2.py tests/test_gpu_kernels.py -m gpu -x -q -k "gemv or attn or attention or decode or rope or row_order" --timeout 300 > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?"
tail -4 $O/${TAG}_pytest.log
run() {  # name, batch, opts
  timeout 300 python bench.py --batch $2 --steps 100 --warmup 5 --no-cpu-baseline --no-breakdown --sweep '' --opt "$3" > $O/${TAG}_$1.json 2> $O/${TAG}_$1.err
  echo "$1 exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("$O/${TAG}_$1.json").read().strip().splitlines()[-1])
    print("$1", "ms/step", round(d["ms_per_step"], 4), "tokens/s", round(d["value"]), "e2e", round(d["e2e"]["value"]), "step frac", round(d["step_roofline"]["frac"], 3))
except Exception as e:
    print("$1 no bench line:", e)
PY
}
run b5 5 ""
run b6 6 ""
run b8 8 ""
run b2 2 ""
