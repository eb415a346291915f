#!/usr/bin/env python
"""Per-iteration time of the three loop forms on one GPU: scripts/time_paths.py [M N]...
tile = tile_loop_kernel (shared-memory resident), cell = cell_loop_kernel / cell_loop_stream_kernel, march = K1 + K2."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package

pb = load_package()
args = [int(x) for x in sys.argv[1:]]
grids = list(zip(args[::2], args[1::2])) or [(40, 40), (400, 600), (800, 1200), (1000, 1300), (1200, 800)]
iters = int(os.environ.get("ITERS", "1500"))
for m, n in grids:
    row = []
    for name, loop, tile in (("tile", "1", "1"), ("cell", "1", "0"), ("march", "0", "0")):
        os.environ["PB200_LOOP"], os.environ["PB200_TILE"] = loop, tile
        with pb.Solver(pb.default_params(m, n, fixed_iters=iters)) as s:
            s.solve()
            runs = [s.solve() for _ in range(3)]
            best = min(r.t_solve_s / r.iterations for r in runs)          # a long fixed-count run may end at the |denom| guard
        row.append(f"{name} {best * 1e6:7.2f} us ({runs[0].iterations} it)")
    print(f"{m:5d} x {n:5d}: " + "   ".join(row), flush=True)
