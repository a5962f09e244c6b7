#!/usr/bin/env python
"""Digest of `ncu --page raw --csv` exports: one line per profiled launch with the metrics the roofline needs.
    python tools/ncu_digest.py gpurun_out/full_*.csv > profiles/ncu_full_rNN.md"""
import csv
import sys

KEYS = [("gpu__time_duration.sum", "us", 1e-3), ("dram__bytes_read.sum", "rd MB", 1e-6),
        ("dram__bytes_write.sum", "wr MB", 1e-6), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %", 1),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor %", 1),
        ("sm__inst_executed_pipe_tensor.sum", "tensor inst", 1),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %", 1),
        ("lts__t_sector_hit_rate.pct", "L2 hit %", 1),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ %", 1),
        ("launch__registers_per_thread", "regs", 1), ("smsp__inst_executed.sum", "Minst", 1e-6),
        ("sm__inst_executed_pipe_xu.sum", "xu Minst", 1e-6),
        ("smsp__cycles_active.avg", "cyc", 1)]


def main():
    print("| file | kernel | grid | block | " + " | ".join(k[1] for k in KEYS) + " |")
    print("|---|---|---|---|" + "---|" * len(KEYS))
    for path in sys.argv[1:]:
        rows = list(csv.reader(open(path, errors="replace")))
        if len(rows) < 3:
            continue
        hdr, units = rows[0], rows[1]
        col = {n: i for i, n in enumerate(hdr)}
        for r in rows[2:]:
            if len(r) < len(hdr):
                continue
            name = r[col["Kernel Name"]]
            name = name.split("(")[0].replace("b200ad::", "")[:44]
            vals = []
            for key, _, scale in KEYS:
                if key in col:
                    try:
                        v = float(r[col[key]].replace(",", ""))
                        u = units[col[key]].lower()
                        if "byte" in u:
                            v *= {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
                        if key.startswith("gpu__time"):
                            v *= {"ns": 1, "nsecond": 1, "us": 1e3, "usecond": 1e3, "ms": 1e6, "msecond": 1e6}.get(u, 1)
                        vals.append(f"{v * scale:.1f}")
                    except ValueError:
                        vals.append(r[col[key]])
                else:
                    vals.append("-")
            print(f"| {path.split('/')[-1].replace('full_', '').replace('.csv', '')} | {name} | {r[col['Grid Size']]} | "
                  f"{r[col['Block Size']]} | " + " | ".join(vals) + " |")


if __name__ == "__main__":
    main()
