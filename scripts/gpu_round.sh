#!/bin/bash
# One GPU-box visit: parity tests, smoke, the bench (both arms), the ncu launch list and one full-set capture of the
# dominant kernels.  Everything lands in gpurun_out/$TAG_*; nothing printed under ncu is a bench value.
TAG=${1:-r01}
O=gpurun_out
mkdir -p $O
python -c "import torch; print(torch.cuda.get_device_name(0))" > $O/${TAG}_env.log 2>&1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks_throttle_reasons.active --format=csv >> $O/${TAG}_env.log 2>&1
if [ "${SKIP_TESTS:-0}" != "1" ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest.log
  tail -3 $O/${TAG}_pytest.log
fi
timeout 600 python __graft_entry__.py --smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke exit $?"
timeout 900 python bench.py --sweep ${SWEEP:-1,4,128} > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench exit $?"
tail -c 2500 $O/${TAG}_bench.json
if [ "${SKIP_REF:-0}" != "1" ]; then
  timeout 600 python bench.py --impl reference --steps 8 --warmup 1 > $O/${TAG}_bench_reference.json 2>&1; echo "ref exit $?"
fi
if [ "${SKIP_NCU:-0}" != "1" ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file $O/${TAG}_launches.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-breakdown > $O/${TAG}_ncu_launch.log 2>&1; echo "ncu launches exit $?"
  python scripts/summarise_ncu.py launches $O/${TAG}_launches.csv $O/${TAG}_launches_summary.csv
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"paged_attn_decode|linear_umma|add_rmsnorm" -c 8 --launch-skip 400 -o $O/${TAG}_full -f \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-breakdown > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full exit $?"
fi
ls -la $O | tail -12
