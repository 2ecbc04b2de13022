"""Instruction mix + stall samples per opcode from an `ncu --page source --csv` export.
    ncu -i X.ncu-rep --page source --csv --kernel-name regex:K > src.csv ; python tools/sass_mix.py src.csv [rows]
"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
div = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
hdr = rows[1]
ia, ie, isamp = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('# Samples')
blk = []
for r in rows[2:]:
    if r[0] == 'Kernel Name':
        break
    if len(r) > ie:
        blk.append(r)
tot = sum(int(r[ie]) for r in blk)
ts = sum(int(r[isamp]) for r in blk) or 1
print(len(blk), 'SASS instrs, executed', tot, '=', tot / div, 'per unit; samples', ts)
op, ops = collections.Counter(), collections.Counter()
for r in blk:
    s = r[ia].strip()
    m = re.match(r'(@!?U?P\d+\s+)?([A-Z0-9_.]+)', s)
    o = m.group(2).split('.')[0] if m else s
    op[o] += int(r[ie])
    ops[o] += int(r[isamp])
for o, c in op.most_common(40):
    print("{:10s} {:9.2f} per unit   samples {:5.1f}%".format(o, c / div, ops[o] / ts * 100))
