#!/bin/bash
# One full GPU-box visit: every parity test file (own process), smoke, the bench (both arms), the ncu launch list and full-set
# captures of the dominant kernels at 64 rows (configs[1]) and of the projections at 256..6336 rows (tensor-pipe evidence).
# Everything lands in gpurun_out/$TAG_*; nothing printed under ncu is a bench value.
TAG=${1:-r2}
O=gpurun_out
mkdir -p $O
python -c "import torch; print(torch.cuda.get_device_name(0))" > $O/${TAG}_env.log 2>&1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks_throttle_reasons.active --format=csv >> $O/${TAG}_env.log 2>&1
if [ "${SKIP_TESTS:-0}" != "1" ]; then
  : > $O/${TAG}_pytest.log
  for f in tests/test_gpu_kernels.py tests/test_gpu_round2.py tests/test_gpu_decode.py tests/test_gpu_awq.py tests/test_gpu_parity_full.py; do
    timeout 1500 python -m pytest $f -m gpu -q --timeout 900 >> $O/${TAG}_pytest.log 2>&1; echo "$f exit $?" | tee -a $O/${TAG}_pytest.log
  done
  # the no-allocation entry point alone in a fresh process, then the whole suite in ONE process the way the round's end runs it
  timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 300 >> $O/${TAG}_pytest.log 2>&1; echo "tests/test_zz_no_alloc.py exit $?" | tee -a $O/${TAG}_pytest.log
  timeout 1500 python -m pytest tests -x -q -m gpu --timeout 900 > $O/${TAG}_pytest_all.log 2>&1; echo "pytest tests -x -q -m gpu exit $?" | tee -a $O/${TAG}_pytest.log
  tail -3 $O/${TAG}_pytest_all.log
  grep -E "passed|failed|error" $O/${TAG}_pytest.log | tail -8
fi
timeout 600 python __graft_entry__.py --smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke exit $?"
timeout 1200 python bench.py ${BENCH_ARGS} > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench exit $?"
tail -c 1500 $O/${TAG}_bench.json
if [ "${SKIP_REF:-0}" != "1" ]; then
  timeout 600 python bench.py --impl reference --steps 8 --warmup 1 > $O/${TAG}_bench_reference.json 2>&1; echo "ref exit $?"
fi
if [ "${SKIP_NCU:-0}" != "1" ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file $O/${TAG}_launches.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-breakdown --sweep '' > $O/${TAG}_ncu_launch.log 2>&1; echo "ncu launches exit $?"
  python scripts/summarise_ncu.py launches $O/${TAG}_launches.csv $O/${TAG}_launches_summary.csv
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"attn_ring|linear_umma|add_rmsnorm|cfg_sample" -c 9 --launch-skip 1500 -o $O/${TAG}_full -f \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-breakdown --sweep '' > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full exit $?"
  python scripts/summarise_ncu.py full $O/${TAG}_full.ncu-rep $O/${TAG}_full_summary.csv
  if [ "${SKIP_GEMM_NCU:-0}" != "1" ]; then
    timeout 600 ncu --set full --clock-control none -k regex:"linear_umma|linear_big" -o $O/${TAG}_gemm -f \
      python scripts/probe/gemm_ncu.py > $O/${TAG}_ncu_gemm.log 2>&1; echo "ncu gemm exit $?"
    python scripts/summarise_ncu.py full $O/${TAG}_gemm.ncu-rep $O/${TAG}_gemm_summary.csv
    rm -f $O/${TAG}_gemm.ncu-rep
  fi
fi
ls -la $O | tail -14
