O=gpurun_out; mkdir -p $O
for b in 1 4; do
  timeout 300 python bench.py --batch $b --steps 100 --warmup 5 --no-cpu-baseline --sweep '' > $O/r5k_b$b.json 2> $O/r5k_b$b.err
  python - <<PY
import json
d = json.loads(open("$O/r5k_b$b.json").read().strip().splitlines()[-1])
print("batch $b", round(d["ms_per_step"], 4), {k: round(v["ms_per_launch"]*1e3, 2) for k, v in d["kernels"].items()})
PY
done
