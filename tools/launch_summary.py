"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel name launches, total and mean microseconds, share."""
import csv, sys, re, collections
rows = list(csv.reader(open(sys.argv[1], errors="ignore")))
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
H = rows[hdr]; kn, mv, mn = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Name")
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
agg = collections.OrderedDict(); total = 0.0; n = 0
for r in rows[hdr + 1:]:
    if len(r) <= mv or r[mn] != "gpu__time_duration.sum": continue
    n += 1
    if n <= skip: continue
    name = re.sub(r"\(.*", "", r[kn]); name = re.sub(r"^(void )?rb::", "", name)
    us = float(r[mv].replace(",", "")) / 1000.0
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += us; total += us
print(f"{'kernel':90s} {'n':>6s} {'total us':>10s} {'mean us':>8s} {'share':>6s}")
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:90]:90s} {c:6d} {t:10.1f} {t / c:8.2f} {100 * t / total:5.1f}%")
print(f"{'total':90s} {n - skip:6d} {total:10.1f}")
