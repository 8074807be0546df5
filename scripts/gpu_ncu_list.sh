#!/bin/bash
# ncu launch list (gpu__time_duration.sum) of a short bench run per environment setting in $CONFIGS
O=gpurun_out; mkdir -p $O
TAG=${TAG:-nl}
i=0
while IFS= read -r cfg; do
  [ -z "$cfg" ] && continue
  i=$((i+1))
  env $cfg timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file $O/${TAG}_launches_$i.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-breakdown > $O/${TAG}_ncu_$i.log 2>&1; echo "[$cfg] ncu exit $?"
  python scripts/summarise_ncu.py launches $O/${TAG}_launches_$i.csv $O/${TAG}_launches_${i}_summary.csv
  head -14 $O/${TAG}_launches_${i}_summary.csv
done <<< "$CONFIGS"
