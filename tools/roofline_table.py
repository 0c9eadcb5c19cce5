"""Per-kernel roofline table of one training step from an ncu launch list with DRAM byte counters:
  ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum[,sm__pipe_tensor_cycles_active...] --csv ...
usage: python tools/roofline_table.py gpurun_out/launches_mem.csv [MEASURED_PEAKS.json] > profiles/<name>.md
Durations are cold-cache and serialised (ncu flushes caches between replays): DRAM bytes are therefore an upper bound of
what the kernel moves inside the real step, where its inputs are usually L2-resident."""
import collections
import csv
import json
import re
import sys


def main(path, peaks_path=None):
    peak_gbs = 6547.5
    if peaks_path:
        try:
            peak_gbs = float(json.load(open(peaks_path))["hbm_gbs"])
        except Exception:
            pass
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    per = collections.OrderedDict()          # launch id -> dict
    for row in csv.DictReader(lines):
        k = row["ID"]
        d = per.setdefault(k, {"name": row["Kernel Name"]})
        v = float(row["Metric Value"].replace(",", "")) if row["Metric Value"] not in ("", "n/a") else 0.0
        unit = row["Metric Unit"]
        m = row["Metric Name"]
        if m.startswith("gpu__time_duration"):
            v = v / 1000 if unit in ("ns", "nsecond") else v * 1000 if unit in ("ms", "msecond") else v
            d["us"] = v
        elif m.startswith("dram__bytes"):
            scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
            d[m.split(".")[0]] = v * scale
        elif "tensor" in m:
            d["tensor_pct"] = v
    rows = list(per.values())
    marks = [i for i, r in enumerate(rows) if "counter_add" in r["name"]]
    step = rows[marks[-2]:marks[-1]] if len(marks) >= 2 else rows
    tot = sum(r.get("us", 0.0) for r in step)
    agg = collections.OrderedDict()
    for r in step:
        k = re.sub(r"\(.*", "", r["name"]).replace("void ", "").replace("s2t::", "")[:64]
        a = agg.setdefault(k, {"n": 0, "us": 0.0, "rd": 0.0, "wr": 0.0, "tp": 0.0})
        a["n"] += 1
        a["us"] += r.get("us", 0.0)
        a["rd"] += r.get("dram__bytes_read", 0.0)
        a["wr"] += r.get("dram__bytes_write", 0.0)
        a["tp"] += r.get("tensor_pct", 0.0) * r.get("us", 0.0)
    print(f"# per-kernel roofline table: one training step ({len(step)} launches, {tot:.0f} us summed, ncu cold-cache serialised)\n")
    print(f"source: `{path}`; HBM peak {peak_gbs:.0f} GB/s (MEASURED_PEAKS.json).  `GB/s` = DRAM bytes / duration of the same"
          " launch; `% HBM` = that over the measured copy bandwidth; `tensor %` = time-weighted sm__pipe_tensor_cycles_active.\n")
    print("| kernel | launches | total us | share | avg us | DRAM MB / launch (rd + wr) | GB/s | % HBM | tensor % |")
    print("|---|---|---|---|---|---|---|---|---|")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["us"]):
        mb = (a["rd"] + a["wr"]) / a["n"] / 1e6
        gbs = (a["rd"] + a["wr"]) / (a["us"] * 1e-6) / 1e9 if a["us"] else 0.0
        tp = a["tp"] / a["us"] if a["us"] else 0.0
        print(f"| `{k}` | {a['n']} | {a['us']:.1f} | {100 * a['us'] / tot:.1f}% | {a['us'] / a['n']:.1f} | "
              f"{a['rd'] / a['n'] / 1e6:.2f} + {a['wr'] / a['n'] / 1e6:.2f} | {gbs:.0f} | {100 * gbs / peak_gbs:.1f}% | {tp:.1f} |")


if __name__ == "__main__":
    main(*sys.argv[1:3])
