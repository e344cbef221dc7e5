#!/usr/bin/env python
"""ncu raw page (CSV) of a `--set full` capture -> profiles/ncu_traffic.json: per kernel, per launch, the numbers
bench.py's `roofline` object quotes (DRAM bytes, duration under ncu, issue-slot utilisation, warp execution
efficiency, L1 / L2 hit rates), stamped with the source hash of the build that was profiled.

usage: ncu_traffic.py <raw.csv | prof.ncu-rep> [more ...] --hash <source hash> [--out profiles/ncu_traffic.json]
Several launches of one kernel are averaged.  Kernel names are reduced to their function name (k_trace, k_fit, ...)."""
import argparse
import csv
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
METRICS = {
    "dram_read_bytes": "dram__bytes_read.sum",
    "dram_write_bytes": "dram__bytes_write.sum",
    "duration_ns": "gpu__time_duration.sum",
    "issue_slots_busy_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "warp_exec_eff_threads": "smsp__thread_inst_executed_per_inst_executed.ratio",
    "warp_inst": "smsp__inst_executed.sum",
    "l1_hit_pct": "l1tex__t_sector_hit_rate.pct",
    "l2_hit_pct": "lts__t_sector_hit_rate.pct",
    "dram_pct_of_peak": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "l1_data_pipe_pct": "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "warps_active_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
    "registers": "launch__registers_per_thread",
    "grid": "launch__grid_size",
    "block": "launch__block_size",
}
UNIT_SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1.0, "us": 1e3, "usecond": 1e3, "ms": 1e6,
              "msecond": 1e6, "nsecond": 1.0, "second": 1e9, "s": 1e9}


def rows_of(path):
    if path.endswith(".ncu-rep"):
        out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        lines = out.splitlines()
    else:
        lines = open(path).read().splitlines()
    lines = [l for l in lines if l.startswith('"')]
    return list(csv.reader(lines))


def short_name(full):
    m = re.search(r"(k_[A-Za-z0-9_]+)", full)
    return m.group(1) if m else full.split("(")[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("inputs", nargs="+")
    ap.add_argument("--hash", required=True)
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "ncu_traffic.json"))
    ap.add_argument("--note", default="")
    args = ap.parse_args()
    acc = {}
    for path in args.inputs:
        rows = rows_of(path)
        hdr, units = rows[0], rows[1]
        col = {h: i for i, h in enumerate(hdr)}
        for vals in rows[2:]:
            name = short_name(vals[col["Kernel Name"]])
            k = acc.setdefault(name, {"launches": 0, "full_name": vals[col["Kernel Name"]]})
            k["launches"] += 1
            for key, metric in METRICS.items():
                if metric not in col:
                    continue
                raw = vals[col[metric]].replace(",", "")
                try:
                    x = float(raw)
                except ValueError:
                    continue
                x *= UNIT_SCALE.get(units[col[metric]], 1.0)
                k[key] = k.get(key, 0.0) + x
    out = {"source_hash": args.hash, "how": "ncu --set full --clock-control none, one capture of the benched build; "
           "per-launch averages; durations under ncu are cold-cache and serialised (shares, not absolutes)",
           "inputs": [os.path.basename(p) for p in args.inputs], "note": args.note, "kernels": {}}
    for name, k in sorted(acc.items()):
        n = k.pop("launches")
        full = k.pop("full_name")
        d = {key: (v / n) for key, v in k.items()}
        d["launches_averaged"] = n
        d["full_name"] = full[:160]
        if "dram_read_bytes" in d and "dram_write_bytes" in d:
            d["dram_bytes"] = d["dram_read_bytes"] + d["dram_write_bytes"]
        out["kernels"][name] = d
    json.dump(out, open(args.out, "w"), indent=1, sort_keys=True)
    sys.stdout.write(f"wrote {args.out}: {', '.join(out['kernels'])}\n")


if __name__ == "__main__":
    main()
