#!/usr/bin/env python
"""One line per kernel launch of an `ncu --page raw --csv` export: duration, DRAM bytes, L2 hit rate, pipe and issue
utilisation, occupancy, top stall reasons.  usage: python tools/ncu_summary.py profiles/r02_ncu_full_*_raw.csv"""
import csv
import json
import sys

KEYS = {
    "ms": "gpu__time_duration.sum", "dram_read_GB": "dram__bytes_read.sum", "dram_write_GB": "dram__bytes_write.sum",
    "dram_pct": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l2_hit_pct": "lts__t_sector_hit_rate.pct",
    "issue_active_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "fmaheavy_pct": "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "alu_pct": "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "tensor_pct": "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "warps_active_pct": "sm__warps_active.avg.pct_of_peak_sustained_active", "regs": "launch__registers_per_thread",
    "warp_inst": "smsp__inst_executed.sum", "grid": "launch__grid_size",
}


def num(x):
    try:
        return float(x)
    except (TypeError, ValueError):
        return None


for path in sys.argv[1:]:
    rows = list(csv.reader(open(path)))
    if len(rows) < 3:
        continue
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        u = dict(zip(hdr, units))
        out = {"file": path.split("/")[-1], "kernel": d.get("Kernel Name", "")[:70]}
        for k, m in KEYS.items():
            v = num(d.get(m))
            if v is None:
                continue
            unit = u.get(m, "")
            if k == "ms":
                v = v / 1e6 if unit in ("ns", "nsecond") else (v / 1e3 if unit in ("us", "usecond") else v)
            if k.endswith("_GB"):
                scale = {"byte": 1e-9, "Kbyte": 1e-6, "Mbyte": 1e-3, "Gbyte": 1.0}.get(unit, 1e-9)
                v = v * scale
            out[k] = round(v, 3) if abs(v) < 1e6 else int(v)
        stalls = {h.split("issue_stalled_")[1].split("_per_issue")[0]: num(d[h]) for h in hdr
                  if "average_warps_issue_stalled_" in h and h.endswith("per_issue_active.ratio") and num(d[h])}
        out["top_stalls"] = [k for k, _ in sorted(stalls.items(), key=lambda kv: -kv[1]) if k != "selected"][:3]
        print(json.dumps(out))
