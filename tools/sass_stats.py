#!/usr/bin/env python
"""Static SASS statistics of one kernel in an object file: opcode histogram split by issue pipe
(sm_100a: FMA pipe = FFMA/FMUL/FADD/IMAD/..., ALU pipe = PRMT/LOP3/SHF/SEL/FMNMX/FSETP/ISETP/IADD3...).
usage: sass_stats.py obj kernel-substring [--dump]"""
import collections, re, subprocess, sys
obj, key = sys.argv[1], sys.argv[2]
txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
blocks = re.split(r"\n\s*Function : ", txt)
FMA = {"FFMA", "FMUL", "FADD", "IMAD", "HFMA2", "FFMA2", "FMUL2", "FADD2", "HADD2", "HMUL2"}
ALU = {"PRMT", "LOP3", "SHF", "SEL", "FSEL", "FMNMX", "FMNMX3", "FSETP", "FSET", "ISETP", "IADD3", "VIADD", "LEA", "MOV",
       "IABS", "VIMNMX", "VIMNMX3", "IMNMX", "PLOP3", "BMSK", "SGXT", "LOP", "FLO", "POPC", "BREV", "P2R", "R2P", "VABSDIFF", "I2FP", "F2FP"}
for b in blocks[1:]:
    name = b.split("\n", 1)[0]
    if key not in name:
        continue
    ops = collections.Counter()
    lines = []
    for ln in b.split("\n"):
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if not m:
            continue
        ins = m.group(2).strip()
        toks = ins.split()
        op = toks[1] if toks[0].startswith("@") else toks[0]
        ops[op.split(".")[0]] += 1
        lines.append((m.group(1), ins))
    tot = sum(ops.values())
    fma = sum(v for k, v in ops.items() if k in FMA)
    alu = sum(v for k, v in ops.items() if k in ALU)
    print(f"{name[:100]}\n  total {tot}  fma-pipe {fma}  alu-pipe {alu}  other {tot - fma - alu}")
    print("  " + " ".join(f"{k}:{v}" for k, v in ops.most_common(30)))
    if "--dump" in sys.argv:
        for a, i in lines:
            print(a, i)
