"""Share of the summed kernel time per kernel from an ncu launch list
(`ncu --metrics gpu__time_duration.sum --csv --log-file X.csv ...`):  python tools/launch_shares.py X.csv"""
import csv
import sys
from collections import defaultdict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
h = rows[0]
ik, iv, iu = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
tot, cnt = defaultdict(float), defaultdict(int)
for r in rows[1:]:
    v = float(r[iv].replace(",", ""))
    v = v / 1e3 if r[iu] in ("ns", "nsecond") else v * 1e3 if r[iu] in ("ms", "msecond") else v
    tot[r[ik]] += v
    cnt[r[ik]] += 1
s = sum(tot.values())
print("# share of summed kernel time, launches, mean duration (ncu gpu__time_duration, cold cache, serialised)")
for k in sorted(tot, key=lambda k: -tot[k]):
    print("{:5.1f}%  {:4d}  {:8.2f}us  {}".format(100 * tot[k] / s, cnt[k], tot[k] / cnt[k], k[:64]))
