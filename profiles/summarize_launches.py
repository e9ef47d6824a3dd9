"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total
and share of device time.  Usage: python profiles/summarize_launches.py launches.csv [out.md]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
h = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
hdr = rows[h]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.OrderedDict()
tot = 0.0
for r in rows[h + 1:]:
    if len(r) <= vi:
        continue
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    if r[ui] == "ns":
        v /= 1e3
    elif r[ui] == "ms":
        v *= 1e3
    name = r[ki].split("(")[0].replace("void ", "")[:80]
    c = agg.setdefault(name, [0, 0.0])
    c[0] += 1
    c[1] += v
    tot += v
lines = ["| kernel | launches | total us | avg us | share |", "|---|---:|---:|---:|---:|"]
for k, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    lines.append(f"| `{k}` | {c} | {t:.1f} | {t / c:.1f} | {100 * t / tot:.1f}% |")
out = "\n".join(lines) + f"\n\ntotal device time {tot / 1e3:.2f} ms over {sum(c for c, _ in agg.values())} launches\n"
if len(sys.argv) > 2:
    open(sys.argv[2], "w").write(out)
print(out)
