#!/usr/bin/env python
"""gpurun_out/parity.jsonl (written by the -m gpu parity tests on the B200 box) -> profiles/parity_latest.json, the `parity` key of
the bench line.  Later records of a test replace earlier ones."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "parity.jsonl")
dst = os.path.join(ROOT, "profiles", "parity_latest.json")
out = json.load(open(dst)) if os.path.exists(dst) else {}
for line in open(src):
    line = line.strip()
    if line:
        r = json.loads(line)
        name = r.pop("test")
        if "req" in r:
            name = f"{name}[{r['req']}]"
        out[name] = r
out["_about"] = ("CUDA path vs the bf16-mode CPU oracle / its committed fixtures, measured on a B200 by tests/test_gpu_*.py "
                 "(teacher-forced; max-abs / cosine of logits, token agreement, mismatch margins)")
json.dump(out, open(dst, "w"), indent=1, sort_keys=True)
print(f"{len(out) - 1} parity records -> {dst}")
