#!/usr/bin/env python
"""Summarise an ncu launch list of tools/profile_step.py
    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv ...
    python tools/summarize_launches.py gpurun_out/launches.csv [launches_per_step] [traffic.json] > profiles/launches_rNN.md
Keeps the library's own kernels (namespace b200ad), takes the LAST `launches_per_step` of them (= one warm
denoise step) and prints time (and DRAM traffic when captured) per kernel and per (kernel, grid) -- cold-cache,
serialised times: compare SHARES.  With a third argument also writes {kernel: {launches, ms, dram_bytes}} as JSON
(bench.py reads profiles/traffic.json for roofline.traffic)."""
import collections
import csv
import json
import re
import sys


def short(name: str) -> str:
    m = re.search(r"b200ad::(\w+)(<[^(]*>)?", name)
    return (m.group(1) + (m.group(2) or "")) if m else name[:60]


def to_us(v, unit):
    return v / 1e3 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1e3)


def to_bytes(v, unit):
    u = unit.lower()
    return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)


def main():
    path = sys.argv[1]
    per_step = int(sys.argv[2]) if len(sys.argv) > 2 else 804
    rows = list(csv.reader(open(path, errors="replace")))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    col = {n: i for i, n in enumerate(rows[h])}
    launches = collections.OrderedDict()          # ID -> record
    for r in rows[h + 1:]:
        if len(r) <= col["Metric Value"] or "b200ad" not in r[col["Kernel Name"]]:
            continue
        rec = launches.setdefault(r[col["ID"]], {"k": short(r[col["Kernel Name"]]), "g": r[col["Grid Size"]],
                                                 "b": r[col["Block Size"]], "us": 0.0, "dram": None})
        v = float(r[col["Metric Value"]].replace(",", ""))
        name, unit = r[col["Metric Name"]], r[col["Metric Unit"]]
        if name.startswith("gpu__time_duration"):
            rec["us"] = to_us(v, unit)
        elif name.startswith("dram__bytes"):
            rec["dram"] = (rec["dram"] or 0.0) + to_bytes(v, unit)
    step = list(launches.values())[-per_step:]
    total = sum(x["us"] for x in step)
    have_dram = any(x["dram"] is not None for x in step)
    print(f"# one warm denoise step: {len(step)} launches of libb200ad kernels, {total / 1e3:.2f} ms summed "
          f"(ncu-serialised, cold cache)" + (f", {sum(x['dram'] or 0 for x in step) / 1e9:.2f} GB of DRAM traffic"
                                             if have_dram else "") + "\n")
    by = collections.OrderedDict()
    for x in step:
        d = by.setdefault(x["k"], [0, 0.0, 0.0])
        d[0] += 1
        d[1] += x["us"]
        d[2] += x["dram"] or 0.0
    print("| kernel | launches | ms | share | DRAM GB | GB/s |\n|---|---|---|---|---|---|")
    for k, (n, us, db) in sorted(by.items(), key=lambda kv: -kv[1][1]):
        print(f"| {k} | {n} | {us / 1e3:.3f} | {100 * us / total:.1f}% | {db / 1e9:.3f} | {db / us / 1e3 if us else 0:.0f} |")
    byg = collections.OrderedDict()
    for x in step:
        d = byg.setdefault((x["k"], x["g"], x["b"]), [0, 0.0, 0.0])
        d[0] += 1
        d[1] += x["us"]
        d[2] += x["dram"] or 0.0
    print("\n| kernel | grid | block | launches | ms | avg us | DRAM MB / launch |\n|---|---|---|---|---|---|---|")
    for (k, g, b), (n, us, db) in sorted(byg.items(), key=lambda kv: -kv[1][1])[:40]:
        print(f"| {k} | {g} | {b} | {n} | {us / 1e3:.3f} | {us / n:.1f} | {db / n / 1e6:.1f} |")
    if len(sys.argv) > 3:
        out = {k: {"launches": n, "ms_ncu": us / 1e3, "dram_bytes": db if have_dram else None}
               for k, (n, us, db) in by.items()}
        json.dump(out, open(sys.argv[3], "w"), indent=1)


if __name__ == "__main__":
    main()
