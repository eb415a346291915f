#!/usr/bin/env python
"""Cycle stamps of tile_loop_kernel's phases (variant builds libpoisson_b200_tt{1,2}.so: -DPB200_TILE_TIMING=1 records the
last CTA's thread 0, =2 CTA 0's): PB200_LIB=.../libpoisson_b200_tt1.so python scripts/tile_timing.py M N [M N ...]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pb = load_package()
a = [int(x) for x in sys.argv[1:]]
for m, n in zip(a[::2], a[1::2]):
    it = 60
    with pb.Solver(pb.default_params(m, n, fixed_iters=it, trace_len=2 * it)) as s:
        s.solve()
        res = s.solve()
        t = s.trace(2 * it).reshape(-1, 6)[10:res.iterations - 1]
    if os.environ.get("AR"):
        print(f"{m}x{n}: all-reduce 2 cycles: to-sync {np.median(t[:,0]):.0f} post {np.median(t[:,1]):.0f} gather {np.median(t[:,2]):.0f} gathered-sync {np.median(t[:,3]):.0f} sum+mail {np.median(t[:,4]):.0f} poll+exit {np.median(t[:,5]):.0f}", flush=True)
        continue
    med = np.median(t[:, :5], axis=0)
    per_it = np.median(np.diff(t[:, 5]))
    print(f"{m}x{n}: cycles A {med[0]:.0f}  B {med[1]:.0f}  allreduce1 {med[2]:.0f}  C {med[3]:.0f}  allreduce2 {med[4]:.0f}  | iteration {per_it:.0f} cycles"
          f" ({res.t_solve_s / res.iterations * 1e6:.2f} us)", flush=True)
