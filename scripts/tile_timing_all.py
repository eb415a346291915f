#!/usr/bin/env python
"""Per-CTA phase cycles of tile_loop_kernel at iteration 30 (variant -DPB200_TILE_TIMING=5)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pb = load_package()
m, n = int(sys.argv[1]), int(sys.argv[2])
with pb.Solver(pb.default_params(m, n, fixed_iters=300, trace_len=400)) as s:
    s.solve()
    res = s.solve()
    t = s.trace(400).reshape(-1, 6)
t = t[t[:, 5] > 0]
print("ctas", len(t))
comp = t[:, 0] + t[:, 1]
print("A+B: min %.0f med %.0f max %.0f | C: min %.0f med %.0f max %.0f" % (comp.min(), np.median(comp), comp.max(), t[:, 3].min(), np.median(t[:, 3]), t[:, 3].max()))
order = np.argsort(-comp)
for i in order[:12]:
    print("cta %3d tile %6d  A %6.0f B %6.0f ar1 %6.0f C %6.0f ar2 %6.0f" % (i, t[i, 5], t[i, 0], t[i, 1], t[i, 2], t[i, 3], t[i, 4]))
print("...")
for i in order[-5:]:
    print("cta %3d tile %6d  A %6.0f B %6.0f ar1 %6.0f C %6.0f ar2 %6.0f" % (i, t[i, 5], t[i, 0], t[i, 1], t[i, 2], t[i, 3], t[i, 4]))
