"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list into per-kernel shares of one step.
usage: python tools/summarize_launches.py gpurun_out/launches.csv > profiles/<name>.md"""
import collections
import csv
import re
import sys


def main(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    rows = []
    for row in csv.DictReader(lines):
        v = float(row["Metric Value"].replace(",", ""))
        unit = row["Metric Unit"]
        v = v / 1000 if unit in ("ns", "nsecond") else v * 1000 if unit in ("ms", "msecond") else v
        rows.append((row["Kernel Name"], v, row.get("Grid Size", ""), row.get("Block Size", "")))
    marks = [i for i, r in enumerate(rows) if "counter_add" in r[0]]
    step = rows[marks[-2]:marks[-1]] if len(marks) >= 2 else rows
    tot = sum(r[1] for r in step)
    agg = collections.defaultdict(lambda: [0, 0.0])
    for n, v, g, b in step:
        k = re.sub(r"\(.*", "", n).replace("void ", "")[:80]
        agg[k][0] += 1
        agg[k][1] += v
    print(f"# ncu launch list: one training step ({len(step)} launches, {tot:.1f} us summed, cold-cache serialised)\n")
    print(f"source: `{path}` (ncu --metrics gpu__time_duration.sum --clock-control none); shares, not absolutes, are meaningful\n")
    print("| kernel | launches | total us | share | avg us |\n|---|---|---|---|---|")
    for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k}` | {c} | {t:.1f} | {100 * t / tot:.1f}% | {t / c:.1f} |")
    fam = collections.defaultdict(float)
    for k, (c, t) in agg.items():
        fam["gemm_bf16_kernel (all tile shapes)" if "gemm_bf16" in k else "at:: (torch int plumbing)" if k.startswith("at::") else "other s2t kernels"] += t
    print("\n| family | share |\n|---|---|")
    for k, t in sorted(fam.items(), key=lambda kv: -kv[1]):
        print(f"| {k} | {100 * t / tot:.1f}% |")


if __name__ == "__main__":
    main(sys.argv[1])
