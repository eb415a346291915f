// pb200_capi.cu -- the extern "C" surface declared in include/poisson_b200.h.
// Solver entry points forward to pb200::Solver; the operator entry points stage caller arrays
// (reference layout, S4:723-736) into a scratch block and run the SAME kernels the solver runs.
#include "pb200_solver.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <new>

using namespace pb200;

struct pb200_solver {
    Solver impl;
};

namespace {

// The entry points make the handle's device current for their own work and hand the caller's device back on return.
struct DeviceGuard {
    int saved = -1;
    DeviceGuard() {
        if (cudaGetDevice(&saved) != cudaSuccess) {
            cudaGetLastError();
            saved = -1;
        }
    }
    ~DeviceGuard() {
        if (saved >= 0) cudaSetDevice(saved);
    }
};

// A scratch block for the operator entry points: fields + control words, host arrays in/out.
struct OpBlock {
    Fields f;
    Consts c{};
    Ctl *ctl = nullptr;
    double *part = nullptr;
    double *out = nullptr;
    double *Ap = nullptr; // A p' as K1 computes it (the solver never stores it)
    cudaStream_t stream = nullptr;
    dim3 k1_grid, k2_grid, init_grid;
    int wn = 1;
    Bands bands{1, 1, 0};
    bool safe_div = false;

    ~OpBlock() {
        if (ctl) cudaFree(ctl);
        if (part) cudaFree(part);
        if (out) cudaFree(out);
        if (Ap) cudaFree(Ap);
    }

    pb200_status init(const pb200_params &p, int nx, int ny, int gi0 = 0, int gj0 = 0) {
        pb200_status st = validate(p);
        if (st != PB200_OK) return st;
        if (nx < 1 || ny < 1) { set_error("nx, ny must be >= 1"); return PB200_ERR_INVALID; }
        int ndev = 0;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
            cudaGetLastError();
            set_error("no CUDA device: libpoisson_b200 has no CPU path");
            return PB200_ERR_CUDA;
        }
        if (p.device >= 0) CK(cudaSetDevice(p.device));
        Geom g{};
        g.nx = nx;
        g.ny = ny;
        g.pitch = pitch_for(ny);
        g.cpitch = cpitch_for(ny);
        g.gi0 = gi0;
        g.gj0 = gj0;
        g.M = p.M;
        g.N = p.N;
        st = enable_kernel_smem();
        if (st != PB200_OK) return st;
        c = make_consts(p);
        c.multi = 0;
        c.fixed = 1;
        c.max_iter = 1LL << 60;
        c.trace_len = 0;
        st = f.allocate(g);
        if (st != PB200_OK) return st;
        CK(cudaMalloc(&Ap, f.elems * sizeof(double)));
        for (double *q : {f.a, f.b, f.D, f.w, f.r, f.p[0], f.p[1], Ap}) CK(cudaMemset(q, 0, f.elems * sizeof(double)));
        wn = (ny + K1_OWN - 1) / K1_OWN;
        // small pieces on purpose, so that the tests cross many piece and window boundaries: 3-row bands over the first
        // two thirds of the rows, 2-row bands over the rest
        bands.nxA = nx >= 12 ? nx - nx / 3 : nx;
        bands.nbA = (bands.nxA + 2) / 3;
        bands.nbB = nx > bands.nxA ? (nx - bands.nxA + 1) / 2 : 0;
        k1_grid = dim3((bands.nbA + bands.nbB) * ((wn + K1_WARPS - 1) / K1_WARPS), 1, 1);
        k2_grid = k1_grid;
        init_grid = dim3((ny + INIT_THREADS - 1) / INIT_THREADS, std::min(nx, 1024), 1);
        size_t nb = (size_t)k1_grid.x * k1_grid.y;
        nb = std::max(nb, (size_t)init_grid.x * init_grid.y);
        nb = std::max(nb, (size_t)k2_grid.x * k2_grid.y);
        nb = std::max(nb, (size_t)((ny + 255) / 256) * (size_t)std::min(nx, 512));
        CK(cudaMalloc(&ctl, sizeof(Ctl)));
        CK(cudaMemset(ctl, 0, sizeof(Ctl)));
        CK(cudaMalloc(&part, 2 * nb * sizeof(double)));
        CK(cudaMalloc(&out, 4 * sizeof(double)));
        CK(cudaMemset(out, 0, 4 * sizeof(double)));
        return PB200_OK;
    }
    // caller-supplied coefficients: outside div_const's verified range -> IEEE divides everywhere
    void scan_coefficients(const double *host) {
        const size_t n = (size_t)(f.g.nx + 2) * (f.g.ny + 2);
        for (size_t i = 0; i < n && !safe_div; ++i)
            if (!coef_in_div_range(host[i])) safe_div = true;
    }
    // host (nx+2)x(ny+2) -> device pitched
    pb200_status put(double *dev, const double *host) {
        const Geom &g = f.g;
        CK(cudaMemcpy2D(dev, g.pitch * sizeof(double), host, (g.ny + 2) * sizeof(double), (g.ny + 2) * sizeof(double),
                        g.nx + 2, cudaMemcpyHostToDevice));
        return PB200_OK;
    }
    pb200_status get_all(double *host, const double *dev) {
        const Geom &g = f.g;
        CK(cudaMemcpy2D(host, (g.ny + 2) * sizeof(double), dev, g.pitch * sizeof(double), (g.ny + 2) * sizeof(double),
                        g.nx + 2, cudaMemcpyDeviceToHost));
        return PB200_OK;
    }
    // interior only: the reference's operators leave the ring of their output untouched
    pb200_status get_interior(double *host, const double *dev) {
        const Geom &g = f.g;
        CK(cudaMemcpy2D(host + (g.ny + 2) + 1, (g.ny + 2) * sizeof(double), dev + g.pitch + 1, g.pitch * sizeof(double),
                        g.ny * sizeof(double), g.nx, cudaMemcpyDeviceToHost));
        return PB200_OK;
    }
    pb200_status set_ctl(double beta, double zr_old) {
        Ctl h{};
        h.beta = beta;
        h.zr_old = zr_old;
        CK(cudaMemcpy(ctl, &h, sizeof(Ctl), cudaMemcpyHostToDevice));
        return PB200_OK;
    }
    pb200_status diag() {
        dim3 grid((f.g.pitch + 255) / 256, f.g.nx + 2);
        diag_kernel<<<grid, 256>>>(f.g, c, f.a, f.b, f.D, f.cls);
        CK(cudaGetLastError());
        return PB200_OK;
    }
    pb200_status k1(bool first, const double *p_old, double *p_new) {
        K1Args a{};
        a.r = f.r;
        a.p_old = p_old;
        a.a = f.a;
        a.b = f.b;
        a.D = f.D;
        a.cls = f.cls;
        a.p_new = p_new;
        a.Ap = Ap;
        a.partials = part;
        a.ctl = ctl;
        a.g = f.g;
        a.c = c;
        a.wn = wn;
        a.bands = bands;
        if (first) {
            if (safe_div) k1_kernel<true, true><<<k1_grid, K1_THREADS, K1_SMEM>>>(a);
            else k1_kernel<true, false><<<k1_grid, K1_THREADS, K1_SMEM>>>(a);
        } else {
            if (safe_div) k1_kernel<false, true><<<k1_grid, K1_THREADS, K1_SMEM>>>(a);
            else k1_kernel<false, false><<<k1_grid, K1_THREADS, K1_SMEM>>>(a);
        }
        CK(cudaGetLastError());
        return PB200_OK;
    }
    pb200_status k2(const double *p) {
        K2Args a{};
        a.w = f.w;
        a.r = f.r;
        a.p = p;
        a.a = f.a;
        a.b = f.b;
        a.D = f.D;
        a.cls = f.cls;
        a.partials = part;
        a.ctl = ctl;
        a.g = f.g;
        a.c = c;
        a.wn = wn;
        a.bands = bands;
        if (safe_div) k2_kernel<true><<<k2_grid, K2_THREADS, K2_SMEM>>>(a);
        else k2_kernel<false><<<k2_grid, K2_THREADS, K2_SMEM>>>(a);
        CK(cudaGetLastError());
        return PB200_OK;
    }
};

#define TRY(x)                           \
    do {                                 \
        pb200_status st__ = (x);         \
        if (st__ != PB200_OK) return st__; \
    } while (0)

} // namespace

extern "C" {

int32_t pb200_api_version(void) { return PB200_API_VERSION; }

const char *pb200_status_string(pb200_status s) {
    switch (s) {
    case PB200_OK: return "ok";
    case PB200_ERR_INVALID: return "invalid argument";
    case PB200_ERR_CUDA: return "CUDA error";
    case PB200_ERR_COMM: return "communicator error";
    case PB200_ERR_STATE: return "wrong call order";
    case PB200_ERR_TOO_LARGE: return "request exceeds a fixed limit";
    }
    return "unknown";
}

const char *pb200_last_error(void) { return g_err; }

int32_t pb200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

void pb200_default_params(pb200_params *p) {
    std::memset(p, 0, sizeof(*p));
    p->M = 40; // S4:996
    p->N = 40;
    p->delta = 1e-6; // S4:1006
    p->max_iter = 0; // S4:1007
    p->A1 = -1.0;    // S4:11
    p->B1 = 1.0;
    p->A2 = -0.6; // S4:12
    p->B2 = 0.6;
    p->f_val = 1.0; // S4:13
    p->ell_k = 2.0; // S4:33
    p->ell_ax = 1.0; // S4:33: x*x
    p->rhs_mode = PB200_RHS_INDICATOR; // S4:182-191
    p->eps = 0.0;   // S4:693
    p->norm_mode = PB200_NORM_WEIGHTED;
    p->device = -1;
    p->nranks = 1;
    p->use_graph = 1;
}

pb200_status pb200_choose_process_grid(int32_t size, int32_t *Px, int32_t *Py) {
    if (size < 1 || !Px || !Py) { set_error("bad arguments"); return PB200_ERR_INVALID; }
    int px, py;
    choose_process_grid(size, px, py);
    *Px = px;
    *Py = py;
    return PB200_OK;
}

pb200_status pb200_launch_plan(int32_t nx, int32_t ny, int32_t resident_ctas, int32_t plan[6]) {
    if (nx < 1 || ny < 1 || resident_ctas < 1 || !plan) {
        set_error("bad arguments");
        return PB200_ERR_INVALID;
    }
    const int wn = (ny + K1_OWN - 1) / K1_OWN;
    const MarchPlan m = march_plan(resident_ctas, wn, nx);
    plan[0] = m.bands.nbA;
    plan[1] = m.bands.nxA;
    plan[2] = m.bands.nbB;
    plan[3] = m.ctas;
    plan[4] = wn;
    plan[5] = K1_WARPS;
    return PB200_OK;
}

pb200_status pb200_tile_plan(int32_t nx, int32_t ny, int32_t sms, int64_t smem_bytes, int32_t plan[4]) {
    if (nx < 1 || ny < 1 || sms < 1 || smem_bytes < 1 || !plan) {
        set_error("bad arguments");
        return PB200_ERR_INVALID;
    }
    int R = 0, C = 0, tr = 0, tc = 0;
    if (!tile_plan(nx, ny, std::min<int>(sms, TL_MAX_BLOCKS), static_cast<size_t>(smem_bytes), R, C, tr, tc)) {
        set_error("a block of %d x %d cells does not fit the shared memory of %d SMs", nx, ny, sms);
        return PB200_ERR_TOO_LARGE;
    }
    plan[0] = R;
    plan[1] = C;
    plan[2] = tr;
    plan[3] = tc;
    return PB200_OK;
}

pb200_status pb200_decompose_2d(int32_t M, int32_t N, int32_t Px, int32_t Py, int32_t rank, int32_t *i_start,
                                int32_t *i_end, int32_t *j_start, int32_t *j_end) {
    if (M < 2 || N < 2 || Px < 1 || Py < 1 || rank < 0 || rank >= Px * Py || !i_start || !i_end || !j_start || !j_end) {
        set_error("bad arguments");
        return PB200_ERR_INVALID;
    }
    int a, b, c, d;
    decompose_2d(M, N, Px, Py, rank, a, b, c, d);
    *i_start = a;
    *i_end = b;
    *j_start = c;
    *j_end = d;
    return PB200_OK;
}

pb200_status pb200_create(const pb200_params *p, pb200_solver **out) {
    if (!p || !out) { set_error("null argument"); return PB200_ERR_INVALID; }
    DeviceGuard guard;
    *out = nullptr;
    pb200_solver *s = new (std::nothrow) pb200_solver();
    if (!s) { set_error("out of host memory"); return PB200_ERR_INVALID; }
    pb200_status st = s->impl.create(*p);
    if (st != PB200_OK) {
        delete s;
        return st;
    }
    *out = s;
    return PB200_OK;
}

pb200_status pb200_comm_unique_id(uint8_t id_out[128]) {
    if (!id_out) { set_error("null argument"); return PB200_ERR_INVALID; }
    return comm_unique_id(id_out);
}

pb200_status pb200_attach_comm(pb200_solver *s, const uint8_t unique_id[128]) {
    if (!s || !unique_id) { set_error("null argument"); return PB200_ERR_INVALID; }
    DeviceGuard guard;
    return s->impl.attach_comm(unique_id);
}

pb200_status pb200_peer_export(pb200_solver *s, uint8_t blob_out[256]) {
    if (!s || !blob_out) { set_error("null argument"); return PB200_ERR_INVALID; }
    DeviceGuard guard;
    return s->impl.peer_export(blob_out);
}

pb200_status pb200_peer_attach(pb200_solver *s, const uint8_t *blobs, int32_t nranks) {
    if (!s || !blobs) { set_error("null argument"); return PB200_ERR_INVALID; }
    DeviceGuard guard;
    return s->impl.peer_attach(blobs, nranks);
}

pb200_status pb200_solve(pb200_solver *s, pb200_result *res) {
    if (!s) { set_error("null solver"); return PB200_ERR_INVALID; }
    DeviceGuard guard;
    return s->impl.solve(res);
}

pb200_status pb200_get_solution(pb200_solver *s, double *host, int64_t ld) {
    if (!s) { set_error("null solver"); return PB200_ERR_INVALID; }
    DeviceGuard guard;
    return s->impl.get_solution(host, ld);
}

pb200_status pb200_get_trace(pb200_solver *s, double *host_3n, int32_t n, int32_t *rows) {
    if (!s || (!host_3n && n > 0)) { set_error("null argument"); return PB200_ERR_INVALID; }
    DeviceGuard guard;
    int r = 0;
    pb200_status st = s->impl.get_trace(host_3n, n, &r);
    if (rows) *rows = r;
    return st;
}

pb200_status pb200_get_field(pb200_solver *s, int32_t which, double **dev_ptr, int64_t *pitch, int32_t *nx, int32_t *ny) {
    if (!s || !dev_ptr) { set_error("null argument"); return PB200_ERR_INVALID; }
    Solver &S = s->impl;
    double *tab[7] = {S.f.w, S.f.r, S.f.p[S.last_iter & 1], nullptr, S.f.a, S.f.b, S.f.D};
    if (which < 0 || which > 6 || !tab[which]) { set_error("bad field id (3 = Ap is never stored)"); return PB200_ERR_INVALID; }
    *dev_ptr = tab[which];
    if (pitch) *pitch = S.f.g.pitch;
    if (nx) *nx = S.f.g.nx;
    if (ny) *ny = S.f.g.ny;
    return PB200_OK;
}

pb200_status pb200_profile_iteration(pb200_solver *s, int32_t iters, double ms_out[2]) {
    if (!s || !ms_out) { set_error("null argument"); return PB200_ERR_INVALID; }
    DeviceGuard guard;
    return s->impl.profile_iteration(iters, ms_out);
}

void pb200_destroy(pb200_solver *s) {
    DeviceGuard guard;
    delete s;
}

// pb200_run keeps the handle of its last call (per host thread): an identical call then skips the device allocation
// and the coefficient kernel, which at 8192^2 cost about as much as 15 iterations.
namespace {
struct RunCache {
    std::unique_ptr<pb200_solver> solver;
    pb200_params params{};
};
thread_local RunCache g_run_cache;
} // namespace

void pb200_release_cache(void) {
    DeviceGuard guard;
    g_run_cache.solver.reset();
}

pb200_status pb200_run(const pb200_params *p, pb200_result *res, double *w_host, int64_t ld) {
    if (!p) { set_error("null argument"); return PB200_ERR_INVALID; }
    if (p->nranks != 1) { set_error("pb200_run is single-rank; use create/attach_comm/solve"); return PB200_ERR_INVALID; }
    TRY(validate(*p)); // argument errors come before any device error
    DeviceGuard guard;
    RunCache &rc = g_run_cache;
    // the cache key names the device explicitly: "the current device" (device < 0) is resolved first, so a caller that
    // switched devices between two otherwise identical calls gets a workspace on the new device
    pb200_params key = *p;
    if (key.device < 0) {
        int cur = 0;
        if (cudaGetDevice(&cur) != cudaSuccess) {
            cudaGetLastError();
            set_error("no CUDA device: libpoisson_b200 has no CPU path");
            return PB200_ERR_CUDA;
        }
        key.device = cur;
    }
    const bool reuse = rc.solver && std::memcmp(&rc.params, &key, sizeof(key)) == 0 && !std::getenv("PB200_NO_CACHE");
    double t_setup = 0.0;
    if (!reuse) {
        rc.solver.reset(); // free the previous workspace before allocating the next
        pb200_solver *s = nullptr;
        TRY(pb200_create(&key, &s));
        rc.solver.reset(s);
        rc.params = key;
        t_setup = s->impl.t_setup;
    }
    Solver &S = rc.solver->impl;
    pb200_status st = S.solve(res);
    if (st == PB200_OK && w_host) st = S.get_solution(w_host, ld);
    if (st != PB200_OK) {
        rc.solver.reset();
        return st;
    }
    if (res) res->t_setup_s = t_setup; // 0 when the cached workspace was reused
    return PB200_OK;
}

// ---- operator entry points ------------------------------------------------------------------

pb200_status pb200_op_setup(const pb200_params *p, int32_t i0, int32_t i1, int32_t j0, int32_t j1, double *a, double *b,
                            double *B, double *D) {
    if (!p) { set_error("null argument"); return PB200_ERR_INVALID; }
    if (i0 < 1 || i1 > p->M - 1 || j0 < 1 || j1 > p->N - 1 || i1 < i0 || j1 < j0) { set_error("bad block"); return PB200_ERR_INVALID; }
    OpBlock ob;
    TRY(ob.init(*p, i1 - i0 + 1, j1 - j0 + 1, i0 - 1, j0 - 1));
    dim3 grid((ob.f.g.pitch + 255) / 256, ob.f.g.nx + 2);
    setup_kernel<<<grid, 256>>>(ob.f.g, ob.c, ob.f.a, ob.f.b, ob.f.D, ob.f.r, nullptr, nullptr, nullptr, nullptr, ob.f.cls);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    if (a) TRY(ob.get_all(a, ob.f.a));
    if (b) TRY(ob.get_all(b, ob.f.b));
    if (B) TRY(ob.get_interior(B, ob.f.r));
    if (D) TRY(ob.get_interior(D, ob.f.D));
    return PB200_OK;
}

pb200_status pb200_op_apply_A(const pb200_params *p, const double *pvec, const double *a, const double *b, double *Ap,
                              int32_t nx, int32_t ny) {
    if (!p || !pvec || !a || !b || !Ap) { set_error("null argument"); return PB200_ERR_INVALID; }
    OpBlock ob;
    TRY(ob.init(*p, nx, ny));
    TRY(ob.put(ob.f.p[0], pvec));
    TRY(ob.put(ob.f.a, a));
    TRY(ob.put(ob.f.b, b));
    ob.scan_coefficients(a);
    ob.scan_coefficients(b);
    TRY(ob.diag()); // classifies the caller's coefficients
    // r = 0, beta = 1: K1 forms p' = 0/D + 1*p = p exactly, then Ap = A p'
    TRY(ob.set_ctl(1.0, 0.0));
    TRY(ob.k1(false, ob.f.p[0], ob.f.p[1]));
    CK(cudaDeviceSynchronize());
    TRY(ob.get_interior(Ap, ob.Ap));
    return PB200_OK;
}

pb200_status pb200_op_apply_Dinv(const pb200_params *p, const double *r, const double *a, const double *b, double *z,
                                 int32_t nx, int32_t ny) {
    if (!p || !r || !a || !b || !z) { set_error("null argument"); return PB200_ERR_INVALID; }
    OpBlock ob;
    TRY(ob.init(*p, nx, ny));
    TRY(ob.put(ob.f.r, r));
    TRY(ob.put(ob.f.a, a));
    TRY(ob.put(ob.f.b, b));
    ob.scan_coefficients(a);
    ob.scan_coefficients(b);
    TRY(ob.diag());
    TRY(ob.set_ctl(0.0, 0.0));
    TRY(ob.k1(true, ob.f.p[0], ob.f.p[1])); // iteration-1 form of K1: p' = D^-1 r
    CK(cudaDeviceSynchronize());
    TRY(ob.get_interior(z, ob.f.p[1]));
    return PB200_OK;
}

pb200_status pb200_op_iteration(const pb200_params *p, double *w, double *r, double *pvec, double *Ap, const double *a,
                                const double *b, int32_t nx, int32_t ny, double beta, double zr_old, int32_t first,
                                double out[4]) {
    if (!p || !w || !r || !pvec || !a || !b || !out) { set_error("null argument"); return PB200_ERR_INVALID; }
    OpBlock ob;
    TRY(ob.init(*p, nx, ny));
    TRY(ob.put(ob.f.w, w));
    TRY(ob.put(ob.f.r, r));
    TRY(ob.put(ob.f.p[0], pvec));
    TRY(ob.put(ob.f.a, a));
    TRY(ob.put(ob.f.b, b));
    ob.scan_coefficients(a);
    ob.scan_coefficients(b);
    TRY(ob.diag());
    TRY(ob.set_ctl(beta, zr_old));
    TRY(ob.k1(first != 0, ob.f.p[0], ob.f.p[1]));
    TRY(ob.k2(ob.f.p[1]));
    CK(cudaDeviceSynchronize());
    Ctl h{};
    CK(cudaMemcpy(&h, ob.ctl, sizeof(Ctl), cudaMemcpyDeviceToHost));
    out[0] = h.denom;
    out[1] = h.alpha;
    out[2] = h.red[1];
    out[3] = h.red[2];
    TRY(ob.get_interior(w, ob.f.w));
    TRY(ob.get_interior(r, ob.f.r));
    TRY(ob.get_interior(pvec, ob.f.p[1]));
    if (Ap) TRY(ob.get_interior(Ap, ob.Ap));
    return PB200_OK;
}

pb200_status pb200_op_dot(const pb200_params *p, const double *u, const double *v, int32_t nx, int32_t ny, double *out) {
    if (!p || !u || !v || !out) { set_error("null argument"); return PB200_ERR_INVALID; }
    OpBlock ob;
    TRY(ob.init(*p, nx, ny));
    TRY(ob.put(ob.f.w, u));
    TRY(ob.put(ob.f.r, v));
    dim3 grid((ny + 255) / 256, std::min(nx, 512));
    dot_kernel<<<grid, 256>>>(ob.f.g, ob.f.w, ob.f.r, ob.part, &ob.ctl->ticket[0], ob.out);
    CK(cudaGetLastError());
    double s = 0.0;
    CK(cudaMemcpy(&s, ob.out, sizeof(double), cudaMemcpyDeviceToHost));
    *out = s * ob.c.h1 * ob.c.h2; // S3:191
    return PB200_OK;
}

pb200_status pb200_selftest_div(double h, int64_t n, uint64_t seed, int64_t *mismatches) {
    if (!mismatches || n < 1 || !(h > 0x1p-100 && h < 0x1p100)) { set_error("bad arguments"); return PB200_ERR_INVALID; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        set_error("no CUDA device: libpoisson_b200 has no CPU path");
        return PB200_ERR_CUDA;
    }
    unsigned long long *d = nullptr;
    CK(cudaMalloc(&d, sizeof(unsigned long long)));
    unsigned long long bad = 0;
    cudaError_t e = cudaMemset(d, 0, sizeof(unsigned long long));
    if (e == cudaSuccess) {
        selftest_div_kernel<<<296, 256>>>(h, 1.0 / h, n, seed, d);
        e = cudaMemcpy(&bad, d, sizeof(bad), cudaMemcpyDeviceToHost);
    }
    cudaFree(d);
    CK(e);
    *mismatches = static_cast<int64_t>(bad);
    return PB200_OK;
}

} // extern "C"
