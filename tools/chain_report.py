"""Critical-chain report from a tools/timeline.py CSV: for the kernels of the main (chain) stream, the increment of the
completion time each one adds (end_i - end_{i-1}: with programmatic dependent launch the start times overlap, the ends do
not), aggregated per kernel name.  usage: python tools/chain_report.py gpurun_out/timeline.csv"""
import collections
import csv
import sys

rows = [(float(r["start_us"]), float(r["dur_us"]), r["name"].replace("s2t::", ""), r["stream"]) for r in csv.DictReader(open(sys.argv[1]))]
# chain stream = the stream with the largest summed duration per phase; simply take, at every time, the kernel that ends last
rows.sort(key=lambda r: r[0] + r[1])
agg = collections.defaultdict(lambda: [0, 0.0])
prev_end = rows[0][0]
crit = []
for s, d, n, st in rows:
    e = s + d
    inc = e - prev_end
    if inc > 0:
        agg[n][0] += 1
        agg[n][1] += inc
        crit.append((e, inc, n, st))
        prev_end = e
tot = sum(v[1] for v in agg.values())
print(f"span {rows[-1][0] + rows[-1][1] - rows[0][0]:.1f} us; attribution of completion-time increments (kernel that finishes last):")
for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:28]:
    print(f"  {n[:58]:58s} {c:4d} x  {t:8.1f} us  ({100 * t / tot:4.1f} %)  avg {t / c:5.1f}")
if len(sys.argv) > 2:
    lo, hi = float(sys.argv[2]), float(sys.argv[3])
    for e, inc, n, st in crit:
        if lo <= e <= hi:
            print(f"{e:9.1f} +{inc:5.1f} s{st} {n[:60]}")
