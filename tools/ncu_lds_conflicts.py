#!/usr/bin/env python
"""Which shared-memory instructions of a kernel pay bank conflicts: per SASS line of an ncu report (source page), the
shared-memory wavefronts it issued and how many of them were excess (more than one per instruction = conflict), with the
CUDA source line (-lineinfo + --import-source on).  usage: ncu_lds_conflicts.py rep [kernel-index] [top]"""
import csv, subprocess, sys
rep = sys.argv[1]; which = int(sys.argv[2]) if len(sys.argv) > 2 else 0; top = int(sys.argv[3]) if len(sys.argv) > 3 else 16
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
starts = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
s = starts[which]; e = starts[which + 1] - 2 if which + 1 < len(starts) else len(rows)
hdr = rows[s]
print("columns with 'Shared':", [h for h in hdr if "hared" in h])
def col(*needles):
    for i, h in enumerate(hdr):
        if all(n.lower() in h.lower() for n in needles): return i
    return -1
iW = col("L1", "Wavefronts", "Shared"); iX = col("Wavefronts", "Shared", "Excessive"); iI = hdr.index("Instructions Executed")
iIdeal = col("Wavefronts", "Shared", "Ideal")
if iW == iX or iW == iIdeal:      # the plain total is the column without a qualifier
    cand = [i for i, h in enumerate(hdr) if "wavefronts shared" in h.lower() and "excessive" not in h.lower() and "ideal" not in h.lower()]
    iW = cand[0] if cand else iW
iSrc = col("Source") if col("Source") >= 0 else 1
num = lambda v: int(float(v.replace(",", ""))) if v and v.replace(",", "").replace(".", "").isdigit() else 0
out = []
tot_w = tot_x = 0
for r in rows[s + 1:e]:
    if len(r) <= max(iW, iX, iI): continue
    w, x = num(r[iW]), num(r[iX])
    if w == 0 and x == 0: continue
    tot_w += w; tot_x += x
    out.append((x, w, num(r[iI]), r[0][-5:], r[1][:70]))
print(f"kernel {which}: shared wavefronts {tot_w}, excess (conflicts) {tot_x} = {100.0 * tot_x / max(tot_w, 1):.1f} %")
for x, w, n, addr, ins in sorted(out, reverse=True)[:top]:
    print(f"{addr}  excess {x:>11d} ({100.0 * x / max(tot_x, 1):5.1f} % of all)  wavefronts {w:>11d}  executed {n:>10d}  {w / max(n, 1):4.2f}/inst  {ins}")
