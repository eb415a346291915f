"""ctypes binding of libpoisson_b200.so (include/poisson_b200.h) -- the test/bench harness side.

The product is the C ABI library and the `poisson_b200` driver; this module only marshals numpy
arrays and structs across that ABI. It has NO fallback: if the shared library is missing or no
CUDA device is present, calls raise.

The directory name carries hyphens (it mirrors the reference repository's name), so import it
through `__graft_entry__.load_package()` or `importlib` rather than a plain `import`.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PB200_LIB") or os.path.join(HERE, "libpoisson_b200.so")  # PB200_LIB: tuning variants
CLI_PATH = os.path.join(HERE, "poisson_b200")

NORM_WEIGHTED, NORM_STAGE0 = 0, 1
RHS_INDICATOR, RHS_AREA = 0, 1
STOP_MAX_ITER, STOP_CONVERGED, STOP_DENOM, STOP_COMM_TIMEOUT = 0, 1, 2, 3

c_double_p = C.POINTER(C.c_double)


class Params(C.Structure):
    """pb200_params."""

    _fields_ = [("M", C.c_int32), ("N", C.c_int32), ("delta", C.c_double), ("max_iter", C.c_int64),
                ("A1", C.c_double), ("B1", C.c_double), ("A2", C.c_double), ("B2", C.c_double),
                ("f_val", C.c_double), ("ell_k", C.c_double), ("eps", C.c_double),
                ("norm_mode", C.c_int32), ("fixed_iters", C.c_int32), ("device", C.c_int32),
                ("check_every", C.c_int32), ("trace_len", C.c_int32),
                ("nranks", C.c_int32), ("rank", C.c_int32), ("Px", C.c_int32), ("Py", C.c_int32),
                ("use_graph", C.c_int32), ("rhs_mode", C.c_int32), ("reserved0", C.c_int32), ("ell_ax", C.c_double),
                ("reserved", C.c_int32 * 4)]


class Result(C.Structure):
    """pb200_result."""

    _fields_ = [("iterations", C.c_int64), ("stop_reason", C.c_int32), ("first_converged_iter", C.c_int32),
                ("diff_norm", C.c_double), ("zr", C.c_double), ("alpha", C.c_double), ("beta", C.c_double),
                ("w_norm_E", C.c_double), ("r_norm_E", C.c_double), ("w_max", C.c_double),
                ("t_setup_s", C.c_double), ("t_solve_s", C.c_double), ("t_total_s", C.c_double),
                ("pt_it_per_s", C.c_double), ("algo_gbps", C.c_double), ("kernel_launches", C.c_int64),
                ("t_device_s", C.c_double), ("safe_mode", C.c_double), ("err_max_in_D", C.c_double),
                ("err_l2_in_D", C.c_double), ("reserved", C.c_double * 2)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if not k.startswith("reserved")}


class PB200Error(RuntimeError):
    pass


# every symbol include/poisson_b200.h declares
EXPORTS = ["pb200_api_version", "pb200_status_string", "pb200_last_error", "pb200_device_count",
           "pb200_default_params", "pb200_choose_process_grid", "pb200_decompose_2d", "pb200_launch_plan", "pb200_tile_plan", "pb200_create",
           "pb200_comm_unique_id", "pb200_attach_comm", "pb200_peer_export", "pb200_peer_attach", "pb200_solve", "pb200_get_solution", "pb200_get_trace",
           "pb200_get_field", "pb200_profile_iteration", "pb200_destroy", "pb200_run", "pb200_release_cache", "pb200_op_setup", "pb200_op_apply_A",
           "pb200_op_apply_Dinv", "pb200_op_iteration", "pb200_op_dot", "pb200_dot_noalloc", "pb200_selftest_div"]

_lib = None


def lib():
    """Load the shared library (once). Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise PB200Error(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                             "(there is no fallback path)")
        L = C.CDLL(LIB_PATH)
        L.pb200_status_string.restype = C.c_char_p
        L.pb200_last_error.restype = C.c_char_p
        L.pb200_default_params.argtypes = [C.POINTER(Params)]
        L.pb200_default_params.restype = None
        L.pb200_choose_process_grid.argtypes = [C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        L.pb200_decompose_2d.argtypes = [C.c_int32] * 5 + [C.POINTER(C.c_int32)] * 4
        L.pb200_launch_plan.argtypes = [C.c_int32] * 3 + [C.POINTER(C.c_int32)]
        L.pb200_tile_plan.argtypes = [C.c_int32] * 3 + [C.c_int64, C.POINTER(C.c_int32)]
        L.pb200_create.argtypes = [C.POINTER(Params), C.POINTER(C.c_void_p)]
        L.pb200_comm_unique_id.argtypes = [C.c_char_p]
        L.pb200_attach_comm.argtypes = [C.c_void_p, C.c_char_p]
        L.pb200_peer_export.argtypes = [C.c_void_p, C.c_char_p]
        L.pb200_peer_attach.argtypes = [C.c_void_p, C.c_char_p, C.c_int32]
        L.pb200_solve.argtypes = [C.c_void_p, C.POINTER(Result)]
        L.pb200_get_solution.argtypes = [C.c_void_p, c_double_p, C.c_int64]
        L.pb200_get_trace.argtypes = [C.c_void_p, c_double_p, C.c_int32, C.POINTER(C.c_int32)]
        L.pb200_get_field.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_void_p), C.POINTER(C.c_int64),
                                      C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        L.pb200_profile_iteration.argtypes = [C.c_void_p, C.c_int32, c_double_p]
        L.pb200_destroy.argtypes = [C.c_void_p]
        L.pb200_destroy.restype = None
        L.pb200_run.argtypes = [C.POINTER(Params), C.POINTER(Result), c_double_p, C.c_int64]
        L.pb200_release_cache.restype = None
        L.pb200_op_setup.argtypes = [C.POINTER(Params)] + [C.c_int32] * 4 + [c_double_p] * 4
        L.pb200_op_apply_A.argtypes = [C.POINTER(Params)] + [c_double_p] * 4 + [C.c_int32] * 2
        L.pb200_op_apply_Dinv.argtypes = [C.POINTER(Params)] + [c_double_p] * 4 + [C.c_int32] * 2
        L.pb200_op_iteration.argtypes = ([C.POINTER(Params)] + [c_double_p] * 6 + [C.c_int32] * 2 +
                                         [C.c_double, C.c_double, C.c_int32, c_double_p])
        L.pb200_op_dot.argtypes = [C.POINTER(Params)] + [c_double_p] * 2 + [C.c_int32] * 2 + [c_double_p]
        L.pb200_dot_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int64, C.c_double, C.c_double,
                                        C.c_void_p, C.c_void_p]
        L.pb200_selftest_div.argtypes = [C.c_double, C.c_int64, C.c_uint64, C.POINTER(C.c_int64)]
        _lib = L
    return _lib


def _check(st):
    if st != 0:
        L = lib()
        raise PB200Error(f"{L.pb200_status_string(st).decode()}: {L.pb200_last_error().decode()}")


def _ptr(a):
    return None if a is None else a.ctypes.data_as(c_double_p)


def _c(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a


def default_params(M=40, N=40, **kw):
    p = Params()
    lib().pb200_default_params(C.byref(p))
    p.M, p.N = M, N
    for k, v in kw.items():
        if not hasattr(p, k):
            raise AttributeError(k)
        setattr(p, k, v)
    return p


def device_count():
    return lib().pb200_device_count()


def choose_process_grid(size):
    px, py = C.c_int32(), C.c_int32()
    _check(lib().pb200_choose_process_grid(size, C.byref(px), C.byref(py)))
    return px.value, py.value


def decompose_2d(M, N, Px, Py, rank):
    v = [C.c_int32() for _ in range(4)]
    _check(lib().pb200_decompose_2d(M, N, Px, Py, rank, *[C.byref(x) for x in v]))
    return tuple(x.value for x in v)


def launch_plan(nx, ny, resident_ctas=592):
    """(nbA, nxA, nbB, ctas, windows, warps_per_cta): how the iteration kernels cut an nx x ny block into bands."""
    v = (C.c_int32 * 6)()
    _check(lib().pb200_launch_plan(nx, ny, resident_ctas, v))
    return tuple(v)


def tile_plan(nx, ny, sms=148, smem_bytes=231000):
    """(R, C, tr, tc): tiles of the resident loop kernel for an nx x ny block, or None if the block does not fit."""
    v = (C.c_int32 * 4)()
    st = lib().pb200_tile_plan(nx, ny, sms, smem_bytes, v)
    if st == 5:
        return None
    _check(st)
    return tuple(v)


def comm_unique_id():
    buf = C.create_string_buffer(128)
    _check(lib().pb200_comm_unique_id(buf))
    return buf.raw


class Solver:
    """pb200_create .. pb200_destroy."""

    def __init__(self, params):
        self.params = params
        self.h = C.c_void_p()
        _check(lib().pb200_create(C.byref(params), C.byref(self.h)))

    def attach_comm(self, unique_id):
        _check(lib().pb200_attach_comm(self.h, unique_id))

    def peer_export(self):
        """256-byte description of this rank's mailbox / residual array for peer-memory mode."""
        buf = C.create_string_buffer(256)
        _check(lib().pb200_peer_export(self.h, buf))
        return buf.raw

    def peer_attach(self, blobs):
        """blobs: the exports of all ranks, in rank order."""
        _check(lib().pb200_peer_attach(self.h, b"".join(blobs), len(blobs)))

    def solve(self):
        res = Result()
        _check(lib().pb200_solve(self.h, C.byref(res)))
        return res

    def solution(self, out=None):
        """Whole-grid (M+1, N+1) array with this rank's block filled in."""
        p = self.params
        if out is None:
            out = np.zeros((p.M + 1, p.N + 1))
        _check(lib().pb200_get_solution(self.h, _ptr(out), p.N + 1))
        return out

    def trace(self, n):
        buf = np.zeros((max(n, 1), 3))
        rows = C.c_int32()
        _check(lib().pb200_get_trace(self.h, _ptr(buf), n, C.byref(rows)))
        return buf[:rows.value]

    def field_ptr(self, which):
        ptr, pitch, nx, ny = C.c_void_p(), C.c_int64(), C.c_int32(), C.c_int32()
        _check(lib().pb200_get_field(self.h, which, C.byref(ptr), C.byref(pitch), C.byref(nx), C.byref(ny)))
        return ptr.value, pitch.value, nx.value, ny.value

    def profile_iteration(self, iters):
        """Mean milliseconds of K1 and K2 (CUDA events around each launch)."""
        out = np.zeros(2)
        _check(lib().pb200_profile_iteration(self.h, iters, _ptr(out)))
        return float(out[0]), float(out[1])

    def close(self):
        if self.h:
            lib().pb200_destroy(self.h)
            self.h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def run(params, want_w=False):
    """pb200_run: the reference binary's solver phase in one call (host buffers in and out)."""
    res = Result()
    w = np.zeros((params.M + 1, params.N + 1)) if want_w else None
    _check(lib().pb200_run(C.byref(params), C.byref(res), _ptr(w), params.N + 1))
    return res, w


def release_cache():
    """Free the device workspace pb200_run keeps for an identical next call."""
    lib().pb200_release_cache()


def op_setup(params, block=None):
    i0, i1, j0, j1 = block or (1, params.M - 1, 1, params.N - 1)
    shape = (i1 - i0 + 3, j1 - j0 + 3)
    a, b, B, D = (np.zeros(shape) for _ in range(4))
    _check(lib().pb200_op_setup(C.byref(params), i0, i1, j0, j1, _ptr(a), _ptr(b), _ptr(B), _ptr(D)))
    return a, b, B, D


def op_apply_A(params, p, a, b):
    p, a, b = _c(p), _c(a), _c(b)
    Ap = np.zeros_like(p)
    _check(lib().pb200_op_apply_A(C.byref(params), _ptr(p), _ptr(a), _ptr(b), _ptr(Ap), p.shape[0] - 2, p.shape[1] - 2))
    return Ap


def op_apply_Dinv(params, r, a, b):
    r, a, b = _c(r), _c(a), _c(b)
    z = np.zeros_like(r)
    _check(lib().pb200_op_apply_Dinv(C.byref(params), _ptr(r), _ptr(a), _ptr(b), _ptr(z), r.shape[0] - 2, r.shape[1] - 2))
    return z


def op_iteration(params, w, r, p, a, b, beta, zr_old, first=False):
    """Returns (w, r, p, Ap, (denom, alpha, diff2, zr)); inputs are not modified."""
    w, r, p = (np.array(x, dtype=np.float64, order="C", copy=True) for x in (w, r, p))
    a, b = _c(a), _c(b)
    Ap = np.zeros_like(w)
    out = np.zeros(4)
    _check(lib().pb200_op_iteration(C.byref(params), _ptr(w), _ptr(r), _ptr(p), _ptr(Ap), _ptr(a), _ptr(b),
                                    w.shape[0] - 2, w.shape[1] - 2, beta, zr_old, int(first), _ptr(out)))
    return w, r, p, Ap, tuple(out.tolist())


def op_dot(params, u, v):
    u, v = _c(u), _c(v)
    out = C.c_double()
    _check(lib().pb200_op_dot(C.byref(params), _ptr(u), _ptr(v), u.shape[0] - 2, u.shape[1] - 2,
                              C.cast(C.byref(out), c_double_p)))
    return out.value


def dot_noalloc(dev_u, dev_v, nx, ny, pitch, h1, h2, dev_out, stream=None):
    """pb200_dot_noalloc: DEVICE pointers (ints) in and out, nothing allocated, asynchronous on `stream`.
    For callers that cannot call cudaMalloc (memory owned by their own pool, or no-sync loops); capped size,
    shared built-in scratch -- see include/poisson_b200.h."""
    _check(lib().pb200_dot_noalloc(dev_u, dev_v, nx, ny, pitch, h1, h2, dev_out, stream))


def selftest_div(h, n=1 << 20, seed=1234):
    bad = C.c_int64()
    _check(lib().pb200_selftest_div(h, n, seed, C.byref(bad)))
    return bad.value
