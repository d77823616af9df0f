"""Summarise an ncu launch list (`--metrics gpu__time_duration.sum --csv`) per kernel: count, total us, share."""
import csv, sys, collections, re
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]; ik = hdr.index("Kernel Name"); iv = hdr.index("Metric Value"); iu = hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    name = re.sub(r"\(.*", "", r[ik]); v = float(r[iv].replace(",", ""))
    if r[iu] in ("nsecond", "ns"): v /= 1e3
    elif r[iu] in ("msecond", "ms"): v *= 1e3
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
print(f"{'kernel':60s} {'n':>6s} {'total_us':>10s} {'avg_us':>8s} {'share':>6s}")
for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{k[:60]:60s} {n:6d} {t:10.1f} {t/n:8.2f} {100*t/tot:5.1f}%")
print(f"{'TOTAL':60s} {sum(a[0] for a in agg.values()):6d} {tot:10.1f}")
