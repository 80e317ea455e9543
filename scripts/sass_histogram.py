#!/usr/bin/env python
"""SASS opcode histogram of the tent kernels in libngst_b200.so (static counts from `cuobjdump -sass`; the dynamic counts
are in the ncu summaries). usage: sass_histogram.py [substring of the demangled kernel name ...] > profiles/<tag>_sass.txt"""
import collections
import re
import subprocess
import sys

LIB = "ngstrefftz_b200/libngst_b200.so"
want = sys.argv[1:] or ["tent_kernel_fast<3, 4, 0, 1, 0>", "tent_kernel_fast<3, 4, 0, 1, 1>", "tent_kernel_fast<3, 4, 0, 1, 2>"]
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
funcs, cur = {}, None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        funcs[cur] = []
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(.*?);", line)
    if m and cur:
        funcs[cur].append(m.group(1))
names = list(funcs)
dem = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
KEYS = ["DMMA", "DFMA", "DMUL", "DADD", "DSETP", "MUFU", "LDG", "STG", "LDS", "STS", "LDC", "UBLKCP", "SYNCS", "BAR", "SHFL", "ATOM", "RED",
        "NOP", "BRA", "IMAD", "LEA", "ISETP"]
for mangled, name in zip(names, dem):
    if not any(w in name for w in want):
        continue
    ops = collections.Counter()
    for ins in funcs[mangled]:
        toks = ins.split()
        op = toks[1] if toks[0].startswith("@") else toks[0]
        ops[op.split(".")[0]] += 1
    tot = sum(ops.values())
    print(f"{name[:name.index('>(') + 1] if '>(' in name else name}: {tot} SASS instructions")
    print("  " + "  ".join(f"{k}={ops.get(k, 0)}" for k in KEYS))
    rest = [(k, v) for k, v in ops.most_common() if k not in KEYS][:10]
    print("  other: " + "  ".join(f"{k}={v}" for k, v in rest))
