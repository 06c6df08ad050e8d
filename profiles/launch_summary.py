"""Per-kernel means of an `ncu --csv --metrics gpu__time_duration.sum,...` launch list.
usage: python profiles/launch_summary.py profiles/r1_launches_final.csv"""
import csv, re, sys
from collections import defaultdict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
head = rows[0]
iname, imet, ival = head.index("Kernel Name"), head.index("Metric Name"), head.index("Metric Value")
iid = head.index("ID")
per = defaultdict(lambda: defaultdict(list))
for r in rows[1:]:
    m = re.search(r"(\w+)(<[^>]*>)?\(", r[iname])
    name = (m.group(1) + (m.group(2) or "")) if m else r[iname]
    try:
        per[name][r[imet]].append(float(r[ival].replace(",", "")))
    except ValueError:
        pass
tot = 0.0
lines = []
for name, mets in per.items():
    t = mets.get("gpu__time_duration.sum", [0.0])
    mean = sum(t) / len(t) / 1e3
    tot += mean
    extra = " ".join(f"{k.split('__')[-1]}={sum(v) / len(v):.3g}" for k, v in mets.items() if k != "gpu__time_duration.sum")
    lines.append((name, len(t), mean, min(t) / 1e3, extra))
for name, n, mean, mn, extra in lines:
    print(f"{name:44s} n={n:3d} mean={mean:8.1f} us min={mn:8.1f} share={mean / tot * 100:5.1f}% {extra}")
print(f"sum of means {tot:.1f}")
