"""Summarise ncu outputs into small tracked files under profiles/.

  python scripts/summarise_ncu.py launches gpurun_out/X_launches.csv profiles/X_launches_summary.csv
  python scripts/summarise_ncu.py full     gpurun_out/X_prof.ncu-rep profiles/X_full_summary.csv
"""
import collections
import csv
import re
import subprocess
import sys


def short(name: str) -> str:
    m = re.search(r"(\w+)(<[^(]*>)?\(", name)
    return (m.group(1) + (m.group(2) or "")) if m else name[:60]


def launches(src, dst):
    rows = [r for r in csv.reader(l for l in open(src) if l.startswith('"'))]
    h = rows[0]
    kn, mv, mn = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Name")
    grid, block = h.index("Grid Size"), h.index("Block Size")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        if r[mn] != "gpu__time_duration.sum":
            continue
        k = (short(r[kn]), r[grid], r[block])
        a = agg.setdefault(k, [0, 0.0, 1e30, 0.0])
        t = float(r[mv].replace(",", ""))
        a[0] += 1; a[1] += t; a[2] = min(a[2], t); a[3] = max(a[3], t)
    total = sum(a[1] for a in agg.values())
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel", "grid", "block", "launches", "total_ns", "avg_ns", "min_ns", "max_ns", "share"])
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            w.writerow([*k, a[0], round(a[1]), round(a[1] / a[0]), round(a[2]), round(a[3]), round(a[1] / total, 4)])
    print(f"{dst}: {len(agg)} kernels, {sum(a[0] for a in agg.values())} launches, {total / 1e6:.3f} ms")


FULL = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "sm__inst_executed_pipe_tensor.sum", "smsp__cycles_active.avg"]


def full(src, dst):
    out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, units = rows[0], rows[1]
    # every tensor-pipe metric the report holds (names differ between ncu releases), whatever the kernel
    tensor = [m for m in h if "tensor" in m and m not in FULL and ("pipe" in m or "inst_executed" in m) and "dmma" not in m and "imma" not in m]
    cols = [("Kernel Name", h.index("Kernel Name"))] + [(m, h.index(m)) for m in FULL + tensor if m in h]
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow([c if c == "Kernel Name" else f"{c} [{units[i]}]" for c, i in cols])
        for r in rows[2:]:
            w.writerow([short(r[i]) if c == "Kernel Name" else r[i] for c, i in cols])
    print(f"{dst}: {len(rows) - 2} captured launches")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
