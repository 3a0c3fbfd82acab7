#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time and launch count per kernel.

    python tools/launch_summary.py gpurun_out/launches.csv [--skip N] [--top 40] [--grid]

--skip drops the first N launches (warm-up); --grid keys by (kernel, grid size) as well."""
import argparse
import collections
import csv
import re


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--skip", type=int, default=0)
    ap.add_argument("--top", type=int, default=40)
    ap.add_argument("--grid", action="store_true")
    ap.add_argument("--width", type=int, default=100)
    a = ap.parse_args()
    rows = list(csv.reader(open(a.csv, errors="replace")))
    hdr = None
    agg = collections.defaultdict(lambda: [0, 0.0])
    n = 0
    for r in rows:
        if hdr is None:
            if "Kernel Name" in r:
                hdr = r
            continue
        if len(r) < len(hdr):
            continue
        d = dict(zip(hdr, r))
        if d.get("Metric Name") != "gpu__time_duration.sum":
            continue
        n += 1
        if n <= a.skip:
            continue
        v = float(d["Metric Value"].replace(",", ""))
        u = d["Metric Unit"]
        v = v / 1000 if u in ("ns", "nsecond") else (v * 1000 if u in ("ms", "msecond") else v)
        name = re.sub(r"\(.*", "", d["Kernel Name"])
        if a.grid:
            name += " grid=" + d.get("Grid Size", "?")
        agg[name][0] += 1
        agg[name][1] += v
    tot = sum(v[1] for v in agg.values())
    own = sum(v[1] for k, v in agg.items() if "tony::" in k or k.startswith("void tony") or "tony" in k)
    print(f"launches {n - a.skip}  total {tot:.0f} us  own-kernel share {100 * own / max(tot, 1e-9):.1f} %")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[: a.top]:
        print(f"{v[1]:10.0f} us {100 * v[1] / tot:5.1f}% {v[0]:6d}  {k[:a.width]}")


if __name__ == "__main__":
    main()
