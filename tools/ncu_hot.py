#!/usr/bin/env python
"""Hot spots of an `ncu --page source --csv` export (SASS view): instructions with the most stall samples and the
stall-reason totals of the kernel.   python tools/ncu_hot.py src.csv [top]"""
import csv
import sys
from collections import Counter

rows = list(csv.reader(open(sys.argv[1], errors="replace")))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
col = {n: i for i, n in enumerate(rows[h])}
stall_cols = [n for n in rows[h] if n.startswith("stall_") and "Not Issued" not in n]
tot = Counter()
ins = []
for k, r in enumerate(rows[h + 1:]):
    if len(r) < len(rows[h]):
        continue
    s = int(r[col["# Samples"]] or 0)
    ex = int(r[col["Instructions Executed"]] or 0)
    st = {n: int(r[col[n]] or 0) for n in stall_cols}
    for n, v in st.items():
        tot[n] += v
    ins.append((s, k, r[col["Source"]].strip(), ex, st))
allS = sum(x[0] for x in ins)
print(f"total samples {allS}, instructions executed {sum(x[3] for x in ins)}")
print("stall totals:", ", ".join(f"{n[6:]} {v} ({100 * v / max(allS, 1):.0f}%)" for n, v in tot.most_common(10)))
for s, k, src, ex, st in sorted(ins, reverse=True)[:top]:
    why = ", ".join(f"{n[6:]} {v}" for n, v in sorted(st.items(), key=lambda kv: -kv[1])[:2] if v)
    print(f"{s:7d} {100 * s / max(allS, 1):5.1f}%  #{k:5d} x{ex:9d}  {src[:70]:70s} {why}")
