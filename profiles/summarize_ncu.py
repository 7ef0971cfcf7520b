#!/usr/bin/env python3
"""Per-kernel summary of one `ncu --set full` capture: profiles/summarize_ncu.py <raw.csv> <particles> <out.json> [note]

<raw.csv> = `ncu -i <capture>.ncu-rep --page raw --csv` (wide format: one row per profiled launch, a units row
under the header). Launches are averaged per kernel (template arguments kept). bench.py reads
roofline.traffic from the newest profiles/*_ncu_full_summary.json.
"""
import csv
import json
import re
import sys

WANT = {
    "time_ns": ["gpu__time_duration.sum"],
    "dram_read_bytes": ["dram__bytes_read.sum"],
    "dram_write_bytes": ["dram__bytes_write.sum"],
    "regs": ["launch__registers_per_thread"],
    "issue_active_pct": ["sm__inst_issued.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
                         "sm__issue_active.avg.pct_of_peak_sustained_active"],
    "warps_active_pct": ["sm__warps_active.avg.pct_of_peak_sustained_active"],
    "dram_throughput_pct": ["gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"],
    "inst_executed": ["smsp__inst_executed.sum", "sm__inst_executed.sum"],
    "l1_throughput_pct": ["l1tex__throughput.avg.pct_of_peak_sustained_active"],
    "l1_hit_pct": ["l1tex__t_sector_hit_rate.pct"],
    "l2_hit_pct": ["lts__t_sector_hit_rate.pct"],
}
UNIT_SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9,
              "nsecond": 1.0, "usecond": 1e3, "msecond": 1e6, "second": 1e9}


def short_name(full):
    m = re.match(r"(?:void )?([A-Za-z_0-9]+)(<[^>]*>)?", full)
    return (m.group(1) + (m.group(2) or "")).replace("(bool)", "").replace("(int)", "").replace(" ", "") if m else full


def main():
    raw, particles, out = sys.argv[1], int(sys.argv[2]), sys.argv[3]
    note = sys.argv[4] if len(sys.argv) > 4 else ""
    rows = [r for r in csv.reader(l for l in open(raw, errors="replace") if l.startswith('"'))]
    header, units, data = rows[0], rows[1], rows[2:]
    col = {n: i for i, n in enumerate(header)}
    kn = col["Kernel Name"]
    acc = {}
    for r in data:
        k = acc.setdefault(short_name(r[kn]), {"launches": 0})
        k["launches"] += 1
        for key, names in WANT.items():
            for nme in names:
                if nme in col and r[col[nme]] not in ("", "n/a"):
                    v = float(r[col[nme]].replace(",", "")) * UNIT_SCALE.get(units[col[nme]], 1.0)
                    k[key] = k.get(key, 0.0) + v
                    break
    for k in acc.values():
        n = k["launches"]
        for key in list(k):
            if key != "launches":
                k[key] /= n
        k["time_ms"] = k.pop("time_ns", 0.0) / 1e6
        k["dram_bytes_per_launch"] = k.get("dram_read_bytes", 0.0) + k.get("dram_write_bytes", 0.0)
        if k["time_ms"] > 0:
            k["dram_GBps"] = k["dram_bytes_per_launch"] / k["time_ms"] / 1e6
        if "inst_executed" in k:
            k["warp_inst_per_particle"] = k["inst_executed"] / particles
    json.dump({"source": note, "particles": particles, "kernels": acc}, open(out, "w"), indent=1)
    for name, k in sorted(acc.items(), key=lambda kv: -kv[1]["time_ms"] * kv[1]["launches"]):
        print("%-34s x%d %8.3f ms  %7.1f MB  issue %5.1f%%  L1 %5.1f%%  %6.1f winst/particle" % (
            name, k["launches"], k["time_ms"], k["dram_bytes_per_launch"] / 1e6, k.get("issue_active_pct", -1),
            k.get("l1_throughput_pct", -1), k.get("warp_inst_per_particle", -1)))


if __name__ == "__main__":
    main()
