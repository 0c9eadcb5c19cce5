"""Opcode evidence per kernel from the built library: cuobjdump -sass s2t_b200/libs2t_b200.so | python tools/sass_table.py"""
import collections
import re
import subprocess
import sys

OPS = ("UTCHMMA", "UTCQMMA", "UTCIMMA", "UTCMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "UTCBAR")
cur, counts = None, collections.OrderedDict()
for line in sys.stdin:
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    for o in OPS:
        if re.search(r"\b" + o + r"\b", line):
            counts[cur][o] += 1
names = list(counts)
dem = subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.split("\n")
print("# cuobjdump -sass s2t_b200/libs2t_b200.so (sm_100a): tensor-core / TMA / TMEM / bulk-copy opcode counts per kernel")
print("# UTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG/UTMASTG = TMA tensor load/store, UBLKCP = cp.async.bulk, "
      "SYNCS = mbarrier ops, UTCBAR = tcgen05.commit")
rows = set()
for n, d in zip(names, dem):
    c = counts[n]
    if not any(c[o] for o in OPS if o != "SYNCS"):
        continue
    d = re.sub(r"\(.*", "", d.replace("(anonymous namespace)::", "")).replace("void ", "")
    rows.add((d, "  ".join(f"{o}:{c[o]}" for o in OPS if c[o])))
for d, r in sorted(rows):
    print(f"{d[:64]:64s} {r}")
