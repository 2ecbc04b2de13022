"""Static per-kernel SASS instruction mix of the shipped library (no GPU needed):
    python tools/sass_static_mix.py [collision_b200/libcollision_b200.so] > profiles/rN_sass_mix.txt
Counts SASS mnemonics per kernel from `cuobjdump -sass` and flags the sm_100 features that
matter on this path (bulk-async copies UBLKCP / UTMALDG, cp.async LDGSTS, mbarrier SYNCS,
256-bit loads/stores, packed 16-bit min/max VIMNMX, warp votes / matches, shared atomics)."""
import collections
import re
import subprocess
import sys
from pathlib import Path

lib = sys.argv[1] if len(sys.argv) > 1 else str(Path(__file__).resolve().parent.parent / "collision_b200" / "libcollision_b200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
FLAG = ["UBLKCP", "UTMALDG", "UTMASTG", "LDGSTS", "SYNCS", "LDG.E.ENL2.256", "STG.E.ENL2.256", "LDG.E.128", "STG.E.128",
        "VIMNMX", "VOTE", "MATCH", "ATOMS", "REDS", "ATOMG", "RED.", "BAR.SYNC", "POPC", "SHFL", "REDUX", "PRMT", "R2P", "LDS", "STS"]
kern, mix, flags = None, {}, {}
demangle = {}
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        kern = m.group(1)
        mix[kern] = collections.Counter()
        flags[kern] = collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
    if m and kern:
        full = m.group(2)
        mix[kern][full.split(".")[0]] += 1
        for f in FLAG:
            if full.startswith(f) or (f.endswith(".") and full.startswith(f[:-1] + ".")):
                flags[kern][f] += 1
names = sorted(mix)
try:
    dem = subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.splitlines()
    demangle = dict(zip(names, dem))
except OSError:
    pass
print("# static SASS instruction mix per kernel: {} ({} kernels)".format(Path(lib).name, len(names)))
tot_flags = collections.Counter()
for k in names:
    n = sum(mix[k].values())
    d = demangle.get(k, k)
    d = re.sub(r"\(anonymous namespace\)::", "", d)[:110]
    print("\n== {}  [{} SASS instructions]".format(d, n))
    print("   top: " + ", ".join("{} {}".format(o, c) for o, c in mix[k].most_common(14)))
    if flags[k]:
        print("   features: " + ", ".join("{} x{}".format(f, c) for f, c in sorted(flags[k].items())))
    tot_flags.update(flags[k])
print("\n# library totals of the flagged mnemonics: " + ", ".join("{} {}".format(f, c) for f, c in sorted(tot_flags.items())))
