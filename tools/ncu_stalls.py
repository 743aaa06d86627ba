#!/usr/bin/env python
"""Summarise an ncu report's SASS source page: top instructions by stall samples and the
stall-reason totals.  Usage: tools/ncu_stalls.py <report.ncu-rep> [top_n] [kernel-index]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
# the page holds one table per profiled launch, each introduced by a "Kernel Name" line
blocks, cur = [], None
for line in txt.splitlines():
    if line.startswith('"Kernel Name"'):
        cur = [line]
        blocks.append(cur)
    elif cur is not None:
        cur.append(line)
which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
blk = blocks[which]
print(blk[0][:160])
rows = list(csv.reader(io.StringIO("\n".join(blk[1:]))))
hdr = rows[0]
ix = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = {h: 0 for h in stall_cols}
data = []
for r in rows[1:]:
    if len(r) != len(hdr):
        continue
    s = int(r[ix["# Samples"]] or 0)
    for h in stall_cols:
        tot[h] += int(r[ix[h]] or 0)
    data.append((s, r[ix["Source"]].strip(), int(r[ix["Instructions Executed"]] or 0),
                 {h: int(r[ix[h]] or 0) for h in stall_cols}))
total = sum(d[0] for d in data)
print(f"total samples {total}; instructions executed {sum(d[2] for d in data)}")
print("stall totals:", ", ".join(f"{k[6:]}={v}" for k, v in sorted(tot.items(), key=lambda kv: -kv[1]) if v))
for s, src, n, st in sorted(data, key=lambda d: -d[0])[:top]:
    main = ", ".join(f"{k[6:]}={v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:2] if v)
    print(f"{s:7d} {100.0 * s / max(total, 1):5.1f}%  n={n:9d}  {src[:70]:70s} {main}")
