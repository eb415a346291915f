"""ctypes bindings for the CPU oracle (oracle/liboracle.so) and, when it was built, the
reference itself (oracle/_ref/libref_stage{0,3}.so, compiled from /root/reference sources).

TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs. The product path never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

c_double_p = C.POINTER(C.c_double)
c_int_p = C.POINTER(C.c_int)


class Problem(C.Structure):
    """po_problem (oracle/poisson_oracle.h)."""

    _fields_ = [("A1", C.c_double), ("B1", C.c_double), ("A2", C.c_double), ("B2", C.c_double),
                ("f_val", C.c_double), ("ell_k", C.c_double), ("eps", C.c_double), ("ell_ax", C.c_double),
                ("rhs_mode", C.c_int)]


class Result(C.Structure):
    """po_result (oracle/poisson_oracle.h)."""

    _fields_ = [("iterations", C.c_int), ("converged", C.c_int), ("denom_break", C.c_int),
                ("diff_norm", C.c_double), ("zr", C.c_double), ("alpha", C.c_double),
                ("w_norm_E", C.c_double), ("r_norm_E", C.c_double), ("w_max", C.c_double),
                ("err_max_in_D", C.c_double), ("err_l2_in_D", C.c_double)]


def _ptr(arr):
    return None if arr is None else arr.ctypes.data_as(c_double_p)


def build_oracle():
    """(Re)build oracle/liboracle.so (and oracle/_ref when /root/reference is mounted)."""
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "all"], check=True, stdout=subprocess.DEVNULL)


class Oracle:
    def __init__(self):
        path = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(path):
            build_oracle()
        self.lib = L = C.CDLL(path)
        L.po_default_problem.argtypes = [C.POINTER(Problem)]
        L.po_choose_process_grid.argtypes = [C.c_int, c_int_p, c_int_p]
        L.po_decompose_2d.argtypes = [C.c_int] * 5 + [c_int_p] * 4
        L.po_setup_local.argtypes = [C.POINTER(Problem)] + [C.c_int] * 6 + [c_double_p] * 3
        L.po_diag.argtypes = [c_double_p] * 3 + [C.c_int, C.c_int, C.c_double, C.c_double]
        L.po_apply_A.argtypes = [c_double_p] * 4 + [C.c_int, C.c_int, C.c_double, C.c_double]
        L.po_apply_Dinv.argtypes = [c_double_p] * 4 + [C.c_int, C.c_int, C.c_double, C.c_double]
        L.po_dot.argtypes = [c_double_p] * 2 + [C.c_int, C.c_int, C.c_double, C.c_double]
        L.po_dot.restype = C.c_double
        L.po_update_w_r.argtypes = [c_double_p] * 4 + [C.c_double, C.c_int, C.c_int]
        L.po_update_w_r.restype = C.c_double
        L.po_update_p.argtypes = [c_double_p] * 2 + [C.c_double, C.c_int, C.c_int]
        L.po_solve.argtypes = [C.POINTER(Problem), C.c_int, C.c_int, C.c_double, C.c_long, C.c_int, C.c_int,
                               C.POINTER(Result), c_double_p, c_double_p, C.c_int]
        L.po_solve.restype = C.c_int
        L.po_solve_blocks.argtypes = [C.POINTER(Problem), C.c_int, C.c_int, C.c_double, C.c_long, C.c_int, C.c_int,
                                      C.c_int, C.c_int, C.POINTER(Result), c_double_p]
        L.po_solve_blocks.restype = C.c_int

    def problem(self, **kw):
        pb = Problem()
        self.lib.po_default_problem(C.byref(pb))
        for k, v in kw.items():
            setattr(pb, k, v)
        return pb

    @staticmethod
    def steps(M, N, pb):
        h1 = (pb.B1 - pb.A1) / M
        h2 = (pb.B2 - pb.A2) / N
        eps = pb.eps if pb.eps > 0 else max(h1, h2) * max(h1, h2)
        return h1, h2, eps

    def process_grid(self, size):
        px, py = C.c_int(), C.c_int()
        self.lib.po_choose_process_grid(size, C.byref(px), C.byref(py))
        return px.value, py.value

    def decompose(self, M, N, Px, Py, rank):
        v = [C.c_int() for _ in range(4)]
        self.lib.po_decompose_2d(M, N, Px, Py, rank, *[C.byref(x) for x in v])
        return tuple(x.value for x in v)

    def setup(self, M, N, block=None, pb=None):
        pb = pb or self.problem()
        i0, i1, j0, j1 = block or (1, M - 1, 1, N - 1)
        shape = (i1 - i0 + 3, j1 - j0 + 3)
        a, b, B = (np.zeros(shape) for _ in range(3))
        self.lib.po_setup_local(C.byref(pb), M, N, i0, i1, j0, j1, _ptr(a), _ptr(b), _ptr(B))
        return a, b, B

    def diag(self, a, b, h1, h2):
        D = np.zeros_like(a)
        self.lib.po_diag(_ptr(a), _ptr(b), _ptr(D), a.shape[0] - 2, a.shape[1] - 2, h1, h2)
        return D

    def apply_A(self, p, a, b, h1, h2):
        Ap = np.zeros_like(p)
        self.lib.po_apply_A(_ptr(p), _ptr(a), _ptr(b), _ptr(Ap), p.shape[0] - 2, p.shape[1] - 2, h1, h2)
        return Ap

    def apply_Dinv(self, r, a, b, h1, h2):
        z = np.zeros_like(r)
        self.lib.po_apply_Dinv(_ptr(r), _ptr(a), _ptr(b), _ptr(z), r.shape[0] - 2, r.shape[1] - 2, h1, h2)
        return z

    def dot(self, u, v, h1, h2):
        return self.lib.po_dot(_ptr(u), _ptr(v), u.shape[0] - 2, u.shape[1] - 2, h1, h2)

    def update_w_r(self, w, r, p, Ap, alpha):
        return self.lib.po_update_w_r(_ptr(w), _ptr(r), _ptr(p), _ptr(Ap), alpha, w.shape[0] - 2, w.shape[1] - 2)

    def update_p(self, p, z, beta):
        self.lib.po_update_p(_ptr(p), _ptr(z), beta, p.shape[0] - 2, p.shape[1] - 2)

    def solve(self, M, N, delta=1e-6, max_iter=0, norm_mode=0, fixed_iters=0, pb=None, want_w=False, trace_len=0):
        pb = pb or self.problem()
        res = Result()
        w = np.zeros((M + 1, N + 1)) if want_w else None
        trace = np.zeros((trace_len, 3)) if trace_len else None
        self.lib.po_solve(C.byref(pb), M, N, delta, max_iter, norm_mode, fixed_iters, C.byref(res), _ptr(w),
                          _ptr(trace), trace_len)
        return res, w, trace

    def solve_blocks(self, M, N, P, Px=0, Py=0, delta=1e-6, max_iter=0, norm_mode=0, pb=None, want_w=False):
        pb = pb or self.problem()
        res = Result()
        w = np.zeros((M + 1, N + 1)) if want_w else None
        self.lib.po_solve_blocks(C.byref(pb), M, N, delta, max_iter, norm_mode, P, Px, Py, C.byref(res), _ptr(w))
        return res, w


class Reference:
    """The reference's own code (oracle/_ref), present when it was compiled from /root/reference."""

    def __init__(self):
        p3 = os.path.join(ORACLE_DIR, "_ref", "libref_stage3.so")
        p0 = os.path.join(ORACLE_DIR, "_ref", "libref_stage0.so")
        if not (os.path.exists(p3) and os.path.exists(p0)):
            raise FileNotFoundError("oracle/_ref not built (needs /root/reference; run `make -C oracle`)")
        self.s3 = S = C.CDLL(p3)
        self.s0 = C.CDLL(p0)
        S.ref3_grid.argtypes = [C.c_int, c_int_p, c_int_p]
        S.ref3_decomp.argtypes = [C.c_int] * 5 + [c_int_p] * 4
        S.ref3_setup.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double, C.c_double] + [C.c_int] * 4 + [c_double_p] * 3
        S.ref3_apply_A.argtypes = [c_double_p] * 4 + [C.c_int, C.c_int, C.c_double, C.c_double]
        S.ref3_apply_Dinv.argtypes = [c_double_p] * 4 + [C.c_int, C.c_int, C.c_double, C.c_double]
        S.ref3_dot.argtypes = [c_double_p] * 2 + [C.c_int, C.c_int, C.c_double, C.c_double]
        S.ref3_dot.restype = C.c_double
        S.ref3_solve.argtypes = [C.c_int, C.c_int, C.c_double, C.c_int, C.c_int]
        S.ref3_solve.restype = C.c_int
        S.ref3_solve_traced.argtypes = [C.c_int, C.c_int, C.c_double, C.c_int, c_double_p, c_double_p, c_double_p, C.c_int]
        S.ref3_solve_traced.restype = C.c_int
        S.ref3_main.argtypes = [C.c_int, C.c_int, C.c_int]
        self.s0.ref0_solve.argtypes = [C.c_int, C.c_int, C.c_double, C.c_int]
        self.s0.ref0_solve.restype = C.c_int

    def process_grid(self, size):
        px, py = C.c_int(), C.c_int()
        self.s3.ref3_grid(size, C.byref(px), C.byref(py))
        return px.value, py.value

    def decompose(self, M, N, Px, Py, rank):
        v = [C.c_int() for _ in range(4)]
        self.s3.ref3_decomp(M, N, Px, Py, rank, *[C.byref(x) for x in v])
        return tuple(x.value for x in v)

    def setup(self, M, N, block=None):
        h1, h2 = 2.0 / M, 1.2 / N
        h1, h2 = (1.0 - -1.0) / M, (0.6 - -0.6) / N
        eps = max(h1, h2) * max(h1, h2)
        i0, i1, j0, j1 = block or (1, M - 1, 1, N - 1)
        shape = (i1 - i0 + 3, j1 - j0 + 3)
        a, b, B = (np.zeros(shape) for _ in range(3))
        self.s3.ref3_setup(M, N, h1, h2, eps, i0, i1, j0, j1, _ptr(a), _ptr(b), _ptr(B))
        return a, b, B

    def apply_A(self, p, a, b, h1, h2):
        Ap = np.zeros_like(p)
        self.s3.ref3_apply_A(_ptr(p), _ptr(a), _ptr(b), _ptr(Ap), p.shape[0] - 2, p.shape[1] - 2, h1, h2)
        return Ap

    def apply_Dinv(self, r, a, b, h1, h2):
        z = np.zeros_like(r)
        self.s3.ref3_apply_Dinv(_ptr(r), _ptr(a), _ptr(b), _ptr(z), r.shape[0] - 2, r.shape[1] - 2, h1, h2)
        return z

    def dot(self, u, v, h1, h2):
        return self.s3.ref3_dot(_ptr(u), _ptr(v), u.shape[0] - 2, u.shape[1] - 2, h1, h2)

    def solve(self, M, N, delta=1e-6, max_iter=0, ranks=1):
        return self.s3.ref3_solve(M, N, delta, max_iter or (M - 1) * (N - 1), ranks)

    def solve_stage0(self, M, N, delta=1e-6, max_iter=0):
        return self.s0.ref0_solve(M, N, delta, max_iter or (M - 1) * (N - 1))

    def solve_traced(self, M, N, delta=1e-6, max_iter=0, want_w=False, trace_len=0):
        out = np.zeros(8)
        w = np.zeros((M + 1, N + 1)) if want_w else None
        trace = np.zeros((trace_len, 3)) if trace_len else None
        self.s3.ref3_solve_traced(M, N, delta, max_iter or (M - 1) * (N - 1), _ptr(out), _ptr(w), _ptr(trace), trace_len)
        keys = ["iterations", "w_norm_E", "w_max", "diff_norm", "zr", "r_norm_E", "alpha", "converged"]
        return dict(zip(keys, out.tolist())), w, trace


def reference_available():
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libref_stage3.so"))


def splitmix64_uniform(n, seed=1234):
    """n doubles uniform in [-1, 1) from splitmix64(seed) -- the seeded inputs SURVEY.md 8(d) names."""
    mask = (1 << 64) - 1
    out = np.empty(n, dtype=np.float64)
    x = seed & mask
    for i in range(n):
        x = (x + 0x9E3779B97F4A7C15) & mask
        z = x
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & mask
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & mask
        z ^= z >> 31
        out[i] = (z >> 11) * (2.0 / (1 << 53)) - 1.0
    return out


def seeded_field(nx, ny, seed=1234, halo=True):
    """(nx+2, ny+2) array; interior (and halo ring if halo) from splitmix64_uniform, vectorised."""
    n = (nx + 2) * (ny + 2)
    mask = np.uint64((1 << 64) - 1)
    idx = np.arange(1, n + 1, dtype=np.uint64)
    with np.errstate(over="ignore"):
        x = (np.uint64(seed) + idx * np.uint64(0x9E3779B97F4A7C15)) & mask
        z = x
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    f = (z >> np.uint64(11)).astype(np.float64) * (2.0 / (1 << 53)) - 1.0
    f = f.reshape(nx + 2, ny + 2)
    if not halo:
        f[0, :] = f[-1, :] = 0.0
        f[:, 0] = f[:, -1] = 0.0
    return np.ascontiguousarray(f)
