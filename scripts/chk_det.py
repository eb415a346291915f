"""Debug aid: repeatability of a converging solve (iteration count, norms) for one grid."""
import sys
sys.path.insert(0, '.')
from __graft_entry__ import load_package
pb = load_package()
M, N = int(sys.argv[1]), int(sys.argv[2])
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
for rep in range(reps):
    r, _ = pb.run(pb.default_params(M, N))
    print(M, N, 'iters', r.iterations, 'w_norm', repr(r.w_norm_E), 'diff', repr(r.diff_norm), 'safe', r.safe_mode)
