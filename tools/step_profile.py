"""Turn one `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv` capture of
`bench.py --steps 1 --warmup 3 --no-graph` into the committed per-step evidence:
  <out>_launches_step.csv   every launch of the LAST training step (name, grid, block, us, DRAM read/write bytes)
  <out>_launches_step.txt   time and share per kernel name
  <out>_dram_step.tsv       DRAM traffic per kernel family
  <out>_traffic.json        per-launch DRAM bytes of the tc_kernel family (bench.py's roofline.traffic)
usage: python tools/step_profile.py capture.csv profiles/r01"""
import collections
import csv
import json
import re
import sys

src, out = sys.argv[1], sys.argv[2]
lines = [ln for ln in open(src) if not ln.startswith("==")]
byid = collections.OrderedDict()
for r in csv.DictReader(lines):
    d = byid.setdefault(r["ID"], {"name": r["Kernel Name"], "grid": r["Grid Size"], "block": r["Block Size"],
                                  "us": 0.0, "rd": 0.0, "wr": 0.0})
    v = float(r["Metric Value"].replace(",", ""))
    u = r["Metric Unit"]
    if r["Metric Name"].startswith("dram__bytes"):
        v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
        d["rd" if "read" in r["Metric Name"] else "wr"] = v
    elif r["Metric Name"] == "gpu__time_duration.sum":
        d["us"] = v * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}.get(u, 1e-3)
recs = list(byid.values())
ends = [i for i, r in enumerate(recs) if "sgd_kernel" in r["name"] or "adamw_kernel" in r["name"]]
runs = []
for i in ends:
    if runs and i == runs[-1][1] + 1:
        runs[-1][1] = i
    else:
        runs.append([i, i])
step = recs[runs[-2][1] + 1: runs[-1][1] + 1]


def short(n):
    n = re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", n)
    n = re.sub(r"\((int|bool)\)", "", n)
    return re.sub(r"^void ", "", n)


with open(out + "_launches_step.csv", "w", newline="") as f:
    w = csv.writer(f)
    w.writerow(["kernel", "grid", "block", "gpu_time_us", "dram_read_bytes", "dram_write_bytes"])
    for r in step:
        w.writerow([short(r["name"]), r["grid"], r["block"], f"{r['us']:.3f}", int(r["rd"]), int(r["wr"])])

tot = collections.defaultdict(lambda: [0.0, 0])
for r in step:
    a = tot[short(r["name"])]
    a[0] += r["us"]; a[1] += 1
total = sum(v[0] for v in tot.values())
with open(out + "_launches_step.txt", "w") as f:
    f.write(f"# {len(step)} launches of one ResNet-50 b256 training step, {total/1e3:.3f} ms of kernel time "
            f"(ncu: cold-cache, serialised -- compare shares, not absolutes)\n")
    f.write(f"{'share':>7} {'ms':>9} {'launches':>8}  kernel\n")
    for name, (us, n) in sorted(tot.items(), key=lambda kv: -kv[1][0]):
        f.write(f"{100*us/total:6.2f}% {us/1e3:9.3f} {n:8d}  {name[:150]}\n")

fam = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
for r in step:
    n = short(r["name"])
    k = "tc_kernel" if n.startswith("tc_kernel") else re.sub(r"\(.*", "", n)
    a = fam[k]
    a[0] += 1; a[1] += r["rd"]; a[2] += r["wr"]; a[3] += r["us"]
dram = sum(a[1] + a[2] for a in fam.values())
with open(out + "_dram_step.tsv", "w") as f:
    f.write(f"# DRAM traffic per kernel family over one ResNet-50 b256 training step: {len(step)} launches, "
            f"{dram/1e9:.2f} GB, {sum(a[3] for a in fam.values())/1e3:.2f} ms kernel time under ncu\n")
    f.write("family\tlaunches\tread_GB\twrite_GB\tms\tGB/s\tbytes_per_launch\n")
    for k, a in sorted(fam.items(), key=lambda kv: -(kv[1][1] + kv[1][2])):
        f.write(f"{k}\t{a[0]}\t{a[1]/1e9:.3f}\t{a[2]/1e9:.3f}\t{a[3]/1e3:.3f}\t{(a[1]+a[2])/max(a[3],1e-9)/1e3:.0f}\t"
                f"{(a[1]+a[2])/a[0]:.0f}\n")
tc = fam["tc_kernel"]
bn = sum(a[1] + a[2] for k, a in fam.items() if k.startswith("bn_"))
json.dump({"source": out + "_dram_step.tsv", "tc_kernel_dram_bytes_per_launch": (tc[1] + tc[2]) / max(tc[0], 1),
           "tc_kernel_launches_per_step": tc[0], "step_dram_bytes": dram, "bn_family_dram_bytes_per_step": bn},
          open(out + "_traffic.json", "w"), indent=1)
print(open(out + "_launches_step.txt").read()[:3000])
print(open(out + "_dram_step.tsv").read()[:1500])
