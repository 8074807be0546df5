#!/bin/bash
# GPU: decode-lane sweep of the bench workload (device-resident value only)
for L in 1 2 4 8; do
  echo "lanes=$L"
  T3B_DECODE_LANES=$L python bench.py --steps 150 --no-cpu-baseline --no-breakdown 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(' value', round(d['value']), 'ms/step', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']), 'frac', round(d['step_roofline']['frac'],3))"
done
