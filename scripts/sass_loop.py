#!/usr/bin/env python
"""Static size of the marching loops: scripts/sass_loop.py [lib.so]
For k1_kernel<false,false> and k2_kernel<false>, finds the innermost backward branch whose body holds a UBLKCP (the
row loop) and prints the number of SASS instructions in it, by opcode class. A tuning aid: the kernels are bound by
instruction issue and dependent-issue latency, so the loop body's size is the first thing to look at."""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "poisson-ellipse-openmp-mpi-cuda-new_b200/libpoisson_b200.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", out)
for f in funcs[1:]:
    name = f.split("\n", 1)[0].strip()
    if not (("k1_kernelILb0ELb0" in name) or ("k2_kernelILb0" in name)):
        continue
    ins = []
    for line in f.split("\n"):
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
    addr_index = {a: i for i, (a, _) in enumerate(ins)}
    best = None
    for i, (a, s) in enumerate(ins):
        m = re.search(r"\bBRA(?:\.\w+)*\s+(?:\w+,\s*)?`?\(?0x([0-9a-f]+)", s)
        if not m:
            continue
        t = int(m.group(1), 16)
        if t < a and t in addr_index:
            body = ins[addr_index[t]:i + 1]
            if sum("UBLKCP" in b[1] for b in body) >= 2:
                if best is None or len(body) < len(best):
                    best = body
    print(name[:60], "total", len(ins), "loop", len(best) if best else None)
    if best:
        c = collections.Counter()
        for _, s in best:
            s = re.sub(r"^@!?U?P\d+\s+", "", s)
            c[s.split()[0].split(".")[0]] += 1
        print("   ", ", ".join(f"{k}:{v}" for k, v in c.most_common(24)))
