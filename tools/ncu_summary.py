#!/usr/bin/env python
"""Turn an ``ncu --set full`` report into the markdown summary committed under profiles/ (no GPU needed).

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep [--title "..."] [--out profiles/x.md] [--max 12]

One column per captured launch (first ``--max`` launches, de-duplicated by kernel name + grid), rows = the
metrics B200_PROFILING.md asks for: duration, DRAM bytes and throughput, tensor-pipe activity, L2 -> SM
traffic, occupancy, registers, shared memory, plus the achieved DRAM bandwidth as a fraction of the
measured copy peak in MEASURED_PEAKS.json."""
import argparse
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
METRICS = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__registers_per_thread", "registers / thread"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem / block"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM written"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput % of peak"),
    ("l1tex__m_xbar2l1tex_read_bytes.sum", "L2 -> SM bytes"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe active %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
    ("smsp__inst_executed.sum", "instructions"),
    ("sm__cycles_elapsed.max", "SM cycles"),
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("--title", default="")
    ap.add_argument("--out", default="")
    ap.add_argument("--max", type=int, default=12)
    ap.add_argument("--note", default="")
    ap.add_argument("--pick", default="", help="comma-separated launch indices to show instead of de-duplicating by name + grid")
    a = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", a.rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
        src = "of measured"
    except Exception:  # noqa: BLE001
        peak, src = 6650.0, "of fallback"
    iname = hdr.index("Kernel Name")
    igrid = hdr.index("launch__grid_size") if "launch__grid_size" in hdr else None
    seen, picked = set(), []
    for n, r in enumerate(data):
        key = (r[iname], r[igrid] if igrid is not None else "")
        if a.pick:
            if str(n) not in a.pick.split(","):
                continue
        elif key in seen:
            continue
        seen.add(key)
        picked.append(r)
        if len(picked) >= a.max:
            break
    lines = [f"# {a.title or os.path.basename(a.rep)}", "",
             f"`ncu --set full --clock-control none --import-source on` on one B200; summary made by "
             f"`tools/ncu_summary.py` from `{os.path.basename(a.rep)}` ({len(data)} launches captured, "
             f"{len(picked)} distinct shown).  Durations under ncu are cold-cache, serialised launches.", ""]
    if a.note:
        lines += [a.note, ""]
    names = [re.sub(r"\(.*", "", r[iname]).replace("tony::", "").replace("void ", "") for r in picked]
    lines.append("| metric | unit | " + " | ".join(f"`{n[:48]}`" for n in names) + " |")
    lines.append("|---|---|" + "---|" * len(picked))
    for key, label in METRICS:
        if key not in hdr:
            continue
        i = hdr.index(key)
        vals = []
        for r in picked:
            v = r[i]
            try:
                f = float(v.replace(",", ""))
                v = f"{f:.4g}" if abs(f) < 1e6 else f"{f:.6g}"
            except ValueError:
                pass
            vals.append(v)
        lines.append(f"| {label} | {units[i]} | " + " | ".join(vals) + " |")
    # achieved DRAM bandwidth
    try:
        ir, iw, it = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("gpu__time_duration.sum")

        def to_bytes(v, u):
            f = float(v.replace(",", ""))
            return f * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)

        def to_us(v, u):
            f = float(v.replace(",", ""))
            return f * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(u.replace("second", "s").replace("usecond", "us"), 1)

        vals = []
        for r in picked:
            b = to_bytes(r[ir], units[ir]) + to_bytes(r[iw], units[iw])
            t = to_us(r[it], units[it])
            gbs = b / t / 1e3
            vals.append(f"{gbs:.0f} ({gbs / peak:.2f} {src})")
        lines.append("| DRAM traffic / duration | GB/s | " + " | ".join(vals) + " |")
    except (ValueError, ZeroDivisionError):
        pass
    text = "\n".join(lines) + "\n"
    if a.out:
        open(a.out, "w").write(text)
    else:
        sys.stdout.write(text)


if __name__ == "__main__":
    main()
