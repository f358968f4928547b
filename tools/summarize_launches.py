"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time and share per kernel name.
usage: summarize_launches.py launches.csv [--last-step REGEX]
--last-step keeps only the launches of the last training step: those after the second-to-last run of kernels
matching REGEX (the optimiser kernel that ends a step) up to and including the last such run."""
import csv
import re
import sys
from collections import defaultdict

path = sys.argv[1]
step_re = re.compile(sys.argv[sys.argv.index("--last-step") + 1]) if "--last-step" in sys.argv else None
rows = []
with open(path, newline="") as fh:
    lines = [ln for ln in fh if not ln.startswith("==")]
rd = csv.DictReader(lines)
tot = defaultdict(lambda: [0.0, 0])
recs = [r for r in rd if r.get("Metric Name") == "gpu__time_duration.sum"]
if step_re is not None:
    ends = [i for i, r in enumerate(recs) if step_re.search(r["Kernel Name"])]
    runs = []  # maximal runs of consecutive step-ending kernels
    for i in ends:
        if runs and i == runs[-1][1] + 1:
            runs[-1][1] = i
        else:
            runs.append([i, i])
    if len(runs) >= 2:
        recs = recs[runs[-2][1] + 1: runs[-1][1] + 1]
for r in recs:
    name = r["Kernel Name"]
    name = re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", name)
    name = re.sub(r"\((int|bool)\)", "", name)
    v = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}.get(unit, 1e-3)
    tot[name][0] += v * scale
    tot[name][1] += 1
total = sum(v[0] for v in tot.values())
print(f"# {path}: {sum(v[1] for v in tot.values())} launches, {total/1e3:.3f} ms of kernel time (cold-cache, serialised)")
print(f"{'share':>7} {'ms':>9} {'launches':>8}  kernel")
for name, (us, n) in sorted(tot.items(), key=lambda kv: -kv[1][0]):
    print(f"{100*us/total:6.2f}% {us/1e3:9.3f} {n:8d}  {name[:150]}")
