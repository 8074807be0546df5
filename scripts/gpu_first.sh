#!/bin/bash
# first visit of a new tree: every GPU test file in its own process (a trapped context does not poison the rest), smoke, short bench
TAG=${TAG:-r2a}
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks_throttle_reasons.active --format=csv > $O/${TAG}_env.log 2>&1
for f in tests/test_gpu_kernels.py tests/test_gpu_decode.py tests/test_gpu_awq.py tests/test_gpu_round2.py; do
  b=$(basename $f .py)
  timeout 1200 python -m pytest $f -m gpu -q --timeout 600 ${PYTEST_ARGS} > $O/${TAG}_$b.log 2>&1; echo "$f exit $?"
  tail -4 $O/${TAG}_$b.log
done
timeout 600 python __graft_entry__.py --smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -3 $O/${TAG}_smoke.log
timeout 900 python bench.py ${BENCH_ARGS:---no-cpu-baseline} > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench exit $?"
tail -5 $O/${TAG}_bench.err; tail -c 3000 $O/${TAG}_bench.json
