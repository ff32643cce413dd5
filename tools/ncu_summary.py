"""Summarise .ncu-rep captures (ncu --set full) into the JSON committed under profiles/.

    python tools/ncu_summary.py WORKLOAD=path.ncu-rep [WORKLOAD=path.ncu-rep ...] > profiles/rN_roofline.json

Per captured launch: the demangled kernel name with its template arguments, duration, DRAM
bytes read / written (the `traffic` figure bench.py quotes), registers, grid, achieved
occupancy, issue activity, DRAM throughput in % of ncu's own peak, L2 hit rate, NVLink bytes
where the counters exist, and the three largest warp-stall reasons summed over the kernel's
instructions (source page; needs -lineinfo).  Runs here, without a GPU: it only reads reports.
"""
from __future__ import annotations

import csv
import io
import json
import re
import subprocess
import sys

RAW = {
    "gpu__time_duration.sum": "duration",
    "dram__bytes_read.sum": "dram_bytes_read",
    "dram__bytes_write.sum": "dram_bytes_write",
    "launch__registers_per_thread": "registers",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
    "launch__shared_mem_per_block_dynamic": "dynamic_smem",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_throughput_pct_of_ncu_peak",
    "lts__t_sector_hit_rate.pct": "l2_hit_rate_pct",
    "nvlrx__bytes.sum": "nvlink_rx_bytes",
    "nvltx__bytes.sum": "nvlink_tx_bytes",
}
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1e-6, "ms": 1e-3, "ns": 1e-9,
         "s": 1.0, "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9, "second": 1.0}


def ncu(*args) -> str:
    return subprocess.run(["ncu", *args], capture_output=True, text=True, check=True).stdout


def short_name(full: str) -> str:
    m = re.match(r"(?:void )?([A-Za-z0-9_]+)(<.*>)?\(", full)
    if not m:
        return full
    return m.group(1) + (m.group(2) or "").replace(" ", "")


def stalls(rep: str, index: int):
    try:
        out = ncu("-i", rep, "--page", "source", "--csv", "--launch-skip", str(index),
                  "--launch-count", "1")
    except subprocess.CalledProcessError:
        return None
    rows = list(csv.reader(io.StringIO(out)))
    hdr = next((r for r in rows if "Source" in r and "# Samples" in r), None)
    if hdr is None:
        return None
    cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    agg = {hdr[i]: 0 for i in cols}
    i_s = hdr.index("# Samples")
    seen = set()
    for r in rows:
        if len(r) != len(hdr) or not r[i_s].isdigit() or r[0] in seen:
            continue
        seen.add(r[0])                      # the page lists every instruction twice
        for i in cols:
            if r[i].isdigit():
                agg[hdr[i]] += int(r[i])
    total = sum(agg.values()) or 1
    top = sorted(agg.items(), key=lambda kv: -kv[1])[:3]
    return {k: round(100.0 * v / total, 1) for k, v in top}


def summarise(workload: str, rep: str):
    rows = list(csv.reader(io.StringIO(ncu("-i", rep, "--page", "raw", "--csv"))))
    hdr, units = rows[0], rows[1]
    out = []
    for n, r in enumerate(rows[2:]):
        rec = {"workload": workload, "kernel": short_name(r[hdr.index("Kernel Name")]),
               "report": rep.split("/")[-1], "launch_in_report": n}
        for metric, key in RAW.items():
            if metric not in hdr:
                continue
            i = hdr.index(metric)
            try:
                v = float(r[i].replace(",", ""))
            except ValueError:
                continue
            rec[key] = v * SCALE.get(units[i], 1)
        if "dram_bytes_read" in rec:
            rec["dram_bytes_per_launch"] = int(rec["dram_bytes_read"] + rec["dram_bytes_write"])
            rec["dram_gbs_under_ncu"] = round(rec["dram_bytes_per_launch"] / rec["duration"] / 1e9, 1)
        rec["duration_us_under_ncu"] = round(rec.pop("duration") * 1e6, 2)
        rec["top_stalls_pct_of_samples"] = stalls(rep, n)
        out.append(rec)
    return out


def main() -> None:
    captures = []
    for arg in sys.argv[1:]:
        workload, rep = arg.split("=", 1)
        captures += summarise(workload, rep)
    json.dump({"source": "ncu --set full --clock-control none --import-source on (B200, CUDA 12.9); "
                         "durations are cold-cache and serialised: compare shares and traffic, "
                         "not absolute times", "captures": captures}, sys.stdout, indent=1)


if __name__ == "__main__":
    main()
