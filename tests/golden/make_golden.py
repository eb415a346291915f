#!/usr/bin/env python
"""Generate tests/golden/* from the REFERENCE ITSELF (oracle/_ref, compiled by oracle/Makefile
from the sources under /root/reference -- nothing else). Run here (the reference tree is
not on the GPU box); the outputs are committed:

    python tests/golden/make_golden.py

golden.json    iteration counts of the unmodified solvers (stage0 `solve`, stage3 `solve_mpi`
               on 1/2/4/8 thread-ranks), decomposition tables, whole-solve norms and
               alpha/beta/diff traces from the reference's operators (ref3_solve_traced).
operators.npz  a, b, B of the reference's fic_reg_local on a full grid and on one block of a
               2x4 decomposition; mat_A_local / mat_D / dot_local on seeded inputs; the 40x40
               solution.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_lib import Reference, seeded_field  # noqa: E402


def main():
    os.environ.setdefault("OMP_NUM_THREADS", "8")
    ref = Reference()
    g = {"generated_by": "tests/golden/make_golden.py from oracle/_ref (reference sources, unmodified)",
         "omp_threads": int(os.environ["OMP_NUM_THREADS"]), "delta": 1e-6}

    # -- iteration counts, unmodified solvers -------------------------------------------
    g["iters_stage0"] = {f"{m}x{n}": ref.solve_stage0(m, n) for m, n in [(10, 10), (20, 20), (40, 40)]}
    g["iters_stage3"] = {f"{m}x{n}": ref.solve(m, n) for m, n in
                         [(10, 10), (20, 20), (40, 40), (400, 600), (800, 1200)]}
    g["iters_stage3_ranks"] = {f"{m}x{n}@{p}": ref.solve(m, n, ranks=p)
                               for m, n in [(40, 40), (400, 600)] for p in (2, 4, 8)}
    g["iters_stage3_ranks"]["37x53@6"] = ref.solve(37, 53, ranks=6)
    g["iters_stage3"]["37x53"] = ref.solve(37, 53)

    # -- decomposition ------------------------------------------------------------------
    g["process_grid"] = {str(p): list(ref.process_grid(p)) for p in (1, 2, 3, 4, 6, 8, 12, 16)}
    dec = {}
    for (m, n, p) in [(40, 40, 8), (400, 600, 4), (2400, 3200, 8), (37, 53, 6), (32768, 32768, 8), (8192, 8192, 2)]:
        px, py = ref.process_grid(p)
        dec[f"{m}x{n}@{p}"] = [list(ref.decompose(m, n, px, py, r)) for r in range(p)]
    g["decompose"] = dec

    # -- whole-solve norms and traces through the reference operators ----------------------
    solves = {}
    arrays = {}
    # the two largest published grids (R4 p.11: 1858 and 2449 iterations) take 7 s and 19 s on 8 cores
    for (m, n, tl) in [(10, 10, 15), (20, 20, 26), (40, 40, 50), (400, 600, 32), (800, 1200, 8), (1600, 2400, 8),
                       (2400, 3200, 8)]:
        out, w, trace = ref.solve_traced(m, n, want_w=(m == 40), trace_len=tl)
        g["iters_stage3"].setdefault(f"{m}x{n}", int(out["iterations"]))
        assert int(out["iterations"]) == g["iters_stage3"][f"{m}x{n}"], (m, n, out)
        out["trace"] = trace.tolist()
        solves[f"{m}x{n}"] = out
        if w is not None:
            arrays["w_40x40"] = w
    g["solves_stage3"] = solves

    # -- operators ---------------------------------------------------------------------
    for tag, (m, n, block) in {"full40": (40, 40, None), "blk": (37, 53, None)}.items():
        a, b, B = ref.setup(m, n, block)
        arrays[f"a_{tag}"], arrays[f"b_{tag}"], arrays[f"B_{tag}"] = a, b, B
    px, py = ref.process_grid(8)
    blk = ref.decompose(400, 600, px, py, 5)
    a, b, B = ref.setup(400, 600, blk)
    g["block_400x600_rank5_of_8"] = list(blk)
    # keep the fixture small: checksums of the big block + a corner window
    g["block_checks"] = {"a_sum": float(a.sum()), "b_sum": float(b.sum()), "B_sum": float(B.sum()),
                         "a_sumsq": float((a * a).sum()), "b_sumsq": float((b * b).sum())}
    arrays["a_blk5_win"] = a[:24, :24].copy()
    arrays["b_blk5_win"] = b[:24, :24].copy()

    m, n = 40, 40
    h1, h2 = 2.0 / m, 1.2 / n
    a, b, B = arrays["a_full40"], arrays["b_full40"], arrays["B_full40"]
    p = seeded_field(m - 1, n - 1, seed=1234)
    r = seeded_field(m - 1, n - 1, seed=4321)
    arrays["Ap_full40"] = ref.apply_A(p, a, b, h1, h2)
    arrays["z_full40"] = ref.apply_Dinv(r, a, b, h1, h2)
    g["dot_full40"] = ref.dot(p, r, h1, h2)

    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(g, f, indent=1, sort_keys=True)
    np.savez_compressed(os.path.join(HERE, "operators.npz"), **arrays)
    print("wrote golden.json, operators.npz")
    print(json.dumps({k: g[k] for k in ("iters_stage0", "iters_stage3", "iters_stage3_ranks")}, indent=1))


if __name__ == "__main__":
    main()
