#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list of tools/profile_step.py:
    python tools/summarize_launches.py gpurun_out/launches.csv [launches_per_step] > profiles/launches_rNN.md
Keeps the library's own kernels (namespace b200ad), takes the LAST `launches_per_step` of them (= one warm
denoise step) and prints time per kernel and per (kernel, grid) -- cold-cache, serialised times: compare SHARES."""
import collections
import csv
import re
import sys


def short(name: str) -> str:
    m = re.search(r"b200ad::(\w+)(<[^(]*>)?", name)
    return (m.group(1) + (m.group(2) or "")) if m else name[:60]


def main():
    path = sys.argv[1]
    per_step = int(sys.argv[2]) if len(sys.argv) > 2 else 804
    rows = list(csv.reader(open(path, errors="replace")))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    col = {n: i for i, n in enumerate(rows[h])}
    ours = []
    for r in rows[h + 1:]:
        if len(r) <= col["Metric Value"] or "b200ad" not in r[col["Kernel Name"]]:
            continue
        v = float(r[col["Metric Value"]].replace(",", ""))
        unit = r[col["Metric Unit"]]
        us = v / 1e3 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1e3)
        ours.append((short(r[col["Kernel Name"]]), r[col["Grid Size"]], r[col["Block Size"]], us))
    step = ours[-per_step:]
    total = sum(x[3] for x in step)
    print(f"# one warm denoise step: {len(step)} launches of libb200ad kernels, {total / 1e3:.2f} ms summed "
          f"(ncu-serialised, cold cache)\n")
    by = collections.OrderedDict()
    for k, g, b, us in step:
        d = by.setdefault(k, [0, 0.0])
        d[0] += 1
        d[1] += us
    print("| kernel | launches | ms | share |\n|---|---|---|---|")
    for k, (n, us) in sorted(by.items(), key=lambda kv: -kv[1][1]):
        print(f"| {k} | {n} | {us / 1e3:.3f} | {100 * us / total:.1f}% |")
    byg = collections.OrderedDict()
    for k, g, b, us in step:
        d = byg.setdefault((k, g, b), [0, 0.0])
        d[0] += 1
        d[1] += us
    print("\n| kernel | grid | block | launches | ms | avg us |\n|---|---|---|---|---|---|")
    for (k, g, b), (n, us) in sorted(byg.items(), key=lambda kv: -kv[1][1])[:40]:
        print(f"| {k} | {g} | {b} | {n} | {us / 1e3:.3f} | {us / n:.1f} |")


if __name__ == "__main__":
    main()
