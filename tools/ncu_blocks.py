#!/usr/bin/env python
"""basic-block level instruction shares from an ncu report (source page, sass): usage ncu_blocks.py rep [kernel-index]"""
import csv, collections, subprocess, sys
rep = sys.argv[1]; which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
starts = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
s = starts[which]; e = starts[which + 1] - 2 if which + 1 < len(starts) else len(rows)
hdr = rows[s]
iI = hdr.index("Instructions Executed"); iT = hdr.index("Thread Instructions Executed"); iSm = hdr.index("# Samples")
blocks = []; cur = None; byop = collections.Counter(); tot = 0
for r in rows[s + 1:e]:
    if len(r) <= iT or not r[iI].isdigit(): continue
    n = int(r[iI]); t = int(r[iT]); sm = int(r[iSm] or 0)
    toks = r[1].split(); op = (toks[1] if toks[0].startswith('@') else toks[0]).split('.')[0]
    byop[op] += n; tot += n
    if cur and cur['n'] == n: cur['cnt'] += 1; cur['thr'] += t; cur['samp'] += sm; cur['ops'][op] += 1
    else:
        cur = {'addr': r[0], 'n': n, 'cnt': 1, 'thr': t, 'samp': sm, 'ops': collections.Counter({op: 1})}; blocks.append(cur)
print("total warp instr", tot)
print(" ".join(f"{k}:{v/tot*100:.1f}%" for k, v in byop.most_common(24)))
for b in blocks:
    share = b['n'] * b['cnt'] / tot * 100
    if share > 0.4:
        print(f"{b['addr'][-5:]} n={b['n']:>9d} x{b['cnt']:3d} share {share:5.1f}% thr/inst {b['thr']/(b['n']*b['cnt']+1e-9):5.1f} samp {b['samp']:6d} {dict(b['ops'].most_common(7))}")
