"""Debug aid: per-iteration time of repeated solves on one handle, for several grid sizes."""
import sys
sys.path.insert(0, '.')
from __graft_entry__ import load_package
pb = load_package()
for m in (2048, 4096, 8192):
    for it in (50, 200):
        with pb.Solver(pb.default_params(m, m, fixed_iters=it)) as s:
            out = []
            for rep in range(3):
                r = s.solve()
                out.append(round(1e6 * r.t_solve_s / r.iterations, 1))
            k1, k2 = s.profile_iteration(20)
            print(m, 'iters', it, 'safe', r.safe_mode, 'us/it per solve', out, 'profile k1,k2 us', round(k1 * 1e3, 1), round(k2 * 1e3, 1))
