"""Per-kernel summary of an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv --log-file X`):
    python tools/launch_summary.py profiles/r2_launches.csv"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
head = rows[hi]
col = {n: i for i, n in enumerate(head)}
agg = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) < len(head) or r[col["Metric Name"]] != "gpu__time_duration.sum":
        continue
    full = r[col["Kernel Name"]]
    m = re.search(r"(k_\w+(<[^>]*>)?)", full)
    name = m.group(1) if m else full[:40]
    v = float(r[col["Metric Value"]].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[col["Metric Unit"]], 1.0)
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
print("%d launches, %.1f us in total (cold-cache, serialised under ncu)\n" % (sum(a[0] for a in agg.values()), tot))
print("| kernel | launches | avg us | share |\n|---|---|---|---|")
for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("| `%s` | %d | %.1f | %.1f %% |" % (n, c, t / c, 100 * t / tot))
