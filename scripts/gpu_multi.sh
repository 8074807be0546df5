#!/bin/bash
# multi-GPU visit (gpurun --gpus N): weak-scaling replicas of configs[1] and the strong-scaling configs[4] workload through the dispatcher
N=${N:-2}; TAG=${TAG:-mg}; REQ=${REQ:-1024}; O=gpurun_out; mkdir -p $O
P=$((29500 + RANDOM % 200))
launch() { if [ "$N" = "1" ]; then python bench.py --gpus 1 "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N "$@"; fi; }
launch --steps 100 --warmup 5 --no-cpu-baseline --no-breakdown --sweep= > $O/${TAG}_cfg2_${N}gpu.json 2> $O/${TAG}_cfg2_${N}gpu.err; echo "cfg2 x$N exit $?"
tail -c 600 $O/${TAG}_cfg2_${N}gpu.json | head -c 600; echo
P=$((P + 1))
launch --workload cfg5 --requests $REQ > $O/${TAG}_cfg5_${N}gpu.json 2> $O/${TAG}_cfg5_${N}gpu.err; echo "cfg5 x$N exit $?"
python - <<PY
import json
for name in ("cfg2", "cfg5"):
    try:
        d = json.loads(open("$O/${TAG}_%s_${N}gpu.json" % name).read().strip().splitlines()[-1])
        print(name, "n_gpus", d["n_gpus"], "value", round(d["value"]), "ms/step", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"]), d["config"].get("requests_rank0"), d["config"].get("preemptions_rank0"))
    except Exception as e:
        print(name, "no line:", e); print(open("$O/${TAG}_%s_${N}gpu.err" % name).read()[-1500:])
PY
