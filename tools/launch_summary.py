"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel launches, total, mean, share."""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
iK, iV, iU = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    name = r[iK].replace("<unnamed>", "").replace("(anonymous namespace)", "")
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(pcp::.*|\(const.*|\(float.*|\(long.*|\(unsigned.*", "", name)
    name = re.sub(r"<\(.*", "", name) if "knn_tile" not in name and "knn_warp" not in name else name
    name = re.sub(r"^.*::", "", name.split("<")[0]) + ("<" + name.split("<", 1)[1] if "<" in name else "")
    v = float(r[iV].replace(",", ""))
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[iU], 1.0)
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
print("kernel,launches,total_us,mean_us,share")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k},{a[0]},{a[1]:.1f},{a[1]/a[0]:.2f},{a[1]/tot:.3f}")
