"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel:
python tools/launch_breakdown.py <csv> [first-kernel-substring: take the launches from its LAST occurrence block on]"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr, data = rows[0], rows[1:]
iK, iV = hdr.index("Kernel Name"), hdr.index("Metric Value")
start = 0
if len(sys.argv) > 2:
    idx = [i for i, r in enumerate(data) if sys.argv[2] in r[iK]]
    block = [i for i in idx if i >= idx[-1] - 3]
    start = block[0]
seg = data[start:]
agg = collections.OrderedDict()
for r in seg:
    n = re.sub(r"^void ", "", r[iK]).replace("b200lab::", "")
    n = re.sub(r"\((?!.*<).*$", "", n)
    a = agg.setdefault(n, [0, 0.0])
    a[0] += 1
    a[1] += float(r[iV].replace(",", ""))
tot = sum(v[1] for v in agg.values())
print(f"{len(seg)} launches, {tot / 1e6:.3f} ms of kernel time")
for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:int(sys.argv[3]) if len(sys.argv) > 3 else 20]:
    print(f"{t / tot * 100:5.1f}%  {t / 1e3:9.1f} us  x{c:4d}  {t / c / 1e3:7.1f} us/launch  {n[:100]}")
