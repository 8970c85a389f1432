#!/usr/bin/env python3
"""Top stall sites of one kernel of an `ncu --set full --import-source on` report:
    python tools/ncu_hot_sass.py <report.ncu-rep> <kernel name substring> [top N]
Prints the SASS instructions with the most warp-stall samples, their dominant stall reasons and the total per reason."""
import csv, io, subprocess, sys
rep, name = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
blocks, cur = [], None
for line in raw.splitlines():
    if line.startswith('"Kernel Name"'):
        cur = {"name": next(csv.reader([line]))[1], "lines": []}
        blocks.append(cur)
    elif cur is not None:
        cur["lines"].append(line)
for b in blocks:
    if name not in b["name"]:
        continue
    rows = list(csv.reader(io.StringIO("\n".join(b["lines"]))))
    hdr = rows[0]
    si = hdr.index("# Samples"); src = hdr.index("Source")
    reasons = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    tot = sum(int(r[si] or 0) for r in rows[1:] if len(r) > si)
    per = {hdr[i]: 0 for i in reasons}
    for r in rows[1:]:
        for i in reasons:
            per[hdr[i]] += int(r[i] or 0)
    print(f"== {b['name'][:80]}: {tot} samples")
    print("   by reason: " + ", ".join(f"{k[6:]} {100*v/max(tot,1):.0f}%" for k, v in sorted(per.items(), key=lambda x: -x[1])[:7]))
    ranked = sorted(((int(r[si] or 0), n, r) for n, r in enumerate(rows[1:]) if len(r) > si), reverse=True)[:top]
    for s, n, r in sorted(ranked, key=lambda x: x[1]):
        why = sorted(((int(r[i] or 0), hdr[i][6:]) for i in reasons), reverse=True)[:2]
        print(f"   {100*s/max(tot,1):5.1f}%  #{n:4d} {r[src].strip()[:70]:70s} {why[0][1]}:{why[0][0]} {why[1][1]}:{why[1][0]}")
    break
