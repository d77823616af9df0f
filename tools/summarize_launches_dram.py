"""Summarise an ncu CSV holding gpu__time_duration.sum + dram__bytes_{read,write}.sum per launch: per kernel name the
count, total / average time, DRAM bytes per launch and the DRAM GB/s over the kernel's own duration."""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]
ii, ik, im, iv, iu = (hdr.index(c) for c in ("ID", "Kernel Name", "Metric Name", "Metric Value", "Metric Unit"))
per = collections.OrderedDict()
for r in rows[1:]:
    d = per.setdefault(r[ii], {"name": re.sub(r"\(.*", "", r[ik]).replace("void f5b::", "").replace("(anonymous namespace)::", "")})
    v = float(r[iv].replace(",", ""))
    u = r[iu]
    if r[im].startswith("gpu__time"):
        d["us"] = v / 1e3 if u in ("ns", "nsecond") else v * (1e3 if u in ("ms", "msecond") else 1.0)
    else:
        mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
        d["bytes"] = d.get("bytes", 0.0) + v * mult
agg = collections.OrderedDict()
for d in per.values():
    a = agg.setdefault(d["name"], [0, 0.0, 0.0])
    a[0] += 1
    a[1] += d.get("us", 0.0)
    a[2] += d.get("bytes", 0.0)
tot = sum(a[1] for a in agg.values())
print(f"{'kernel':58s} {'n':>5s} {'total_us':>10s} {'avg_us':>8s} {'share':>6s} {'dram_MB/launch':>14s} {'dram_GB/s':>9s}")
for k, (n, t, b) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{k[:58]:58s} {n:5d} {t:10.1f} {t / n:8.2f} {100 * t / tot:5.1f}% {b / n / 1e6:14.2f} {b / max(t, 1e-9) / 1e3:9.0f}")
print(f"{'TOTAL':58s} {sum(a[0] for a in agg.values()):5d} {tot:10.1f}")
