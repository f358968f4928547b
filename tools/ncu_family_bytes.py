"""Sum gpu time and DRAM bytes per kernel name from an `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,
dram__bytes_write.sum --csv` log.   usage: python tools/ncu_family_bytes.py capture.csv [out.txt]"""
import collections
import csv
import re
import sys

lines = [ln for ln in open(sys.argv[1]) if not ln.startswith("==")]
byid = collections.OrderedDict()
for r in csv.DictReader(lines):
    d = byid.setdefault(r["ID"], {"name": r["Kernel Name"], "us": 0.0, "rd": 0.0, "wr": 0.0})
    v = float(r["Metric Value"].replace(",", ""))
    u = r["Metric Unit"]
    if r["Metric Name"].startswith("dram__bytes"):
        v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
        d["rd" if "read" in r["Metric Name"] else "wr"] = v
    elif r["Metric Name"] == "gpu__time_duration.sum":
        d["us"] = v * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}.get(u, 1e-3)
tot = collections.OrderedDict()
for d in byid.values():
    n = re.sub(r"\(anonymous namespace\)::|<unnamed>::|^void ", "", d["name"])
    n = re.sub(r"\(.*", "", n)
    a = tot.setdefault(n, [0, 0.0, 0.0, 0.0])
    a[0] += 1; a[1] += d["us"]; a[2] += d["rd"]; a[3] += d["wr"]
text = "launches\tus\tdram_read_MB\tdram_write_MB\tkernel\n" + "".join(
    f"{a[0]}\t{a[1]:.1f}\t{a[2] / 1e6:.2f}\t{a[3] / 1e6:.2f}\t{n[:140]}\n" for n, a in tot.items())
if len(sys.argv) > 2:
    open(sys.argv[2], "w").write(text)
print(text)
