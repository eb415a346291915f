// pb200_solver.cu -- host side of libpoisson_b200.so: the solver object behind the C ABI of
// include/poisson_b200.h. Replaces gradient_solver_mpi (stage4-mpi+cuda/poisson_mpi_cuda_f.cu,
// "S4", lines 688-983) and the parts of main that feed it (S4:996-1007).
//
// Differences in structure from the reference (by design):
//   * coefficients, RHS and preconditioner diagonal are computed by a kernel, not on the host
//     (S4:717 + 723-760);
//   * one iteration is two kernels with the scalar logic on the device; the host never waits for
//     a scalar (the reference blocks ~20 times per iteration). The host only polls
//     a device-side stop flag every `check_every` iterations, one batch behind the GPU;
//   * the iteration is replayed as a CUDA graph (two iterations per graph: p ping-pongs).
#include "pb200_solver.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace pb200 {

thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// S4:82-86
void choose_process_grid(int size, int &Px, int &Py) {
    Px = static_cast<int>(std::sqrt(static_cast<double>(size)));
    while (Px > 1 && size % Px != 0) --Px;
    Py = size / Px;
}

// S4:97-133
void decompose_2d(int M, int N, int Px, int Py, int rank, int &i0, int &i1, int &j0, int &j1) {
    const int px = rank % Px, py = rank / Px;
    auto split = [](int total, int parts, int coord, int &lo, int &hi) {
        const int base = total / parts, rem = total % parts;
        lo = 1 + coord * base + std::min(coord, rem); // 1 + sum_{k<coord} (base + (k<rem))
        hi = lo + base + (coord < rem ? 1 : 0) - 1;
    };
    split(M - 1, Px, px, i0, i1);
    split(N - 1, Py, py, j0, j1);
}

pb200_status validate(const pb200_params &p) {
    if (p.M < 2 || p.N < 2) { set_error("M and N must be >= 2 (got %d, %d)", p.M, p.N); return PB200_ERR_INVALID; }
    if (!(p.B1 > p.A1) || !(p.B2 > p.A2)) { set_error("empty rectangle"); return PB200_ERR_INVALID; }
    if (!(p.ell_k > 0.0)) { set_error("ell_k must be > 0"); return PB200_ERR_INVALID; }
    if (p.ell_ax < 0.0 || p.ell_ax != p.ell_ax) { set_error("ell_ax must be > 0 (or 0 for the reference's 1)"); return PB200_ERR_INVALID; }
    if (p.rhs_mode != PB200_RHS_INDICATOR && p.rhs_mode != PB200_RHS_AREA) { set_error("bad rhs_mode"); return PB200_ERR_INVALID; }
    if (!(p.delta >= 0.0)) { set_error("delta must be >= 0"); return PB200_ERR_INVALID; }
    if (p.norm_mode != PB200_NORM_WEIGHTED && p.norm_mode != PB200_NORM_STAGE0) { set_error("bad norm_mode"); return PB200_ERR_INVALID; }
    if (p.nranks < 1 || p.rank < 0 || p.rank >= p.nranks) { set_error("bad rank %d of %d", p.rank, p.nranks); return PB200_ERR_INVALID; }
    if (p.Px > 0 && p.Py > 0 && p.Px * p.Py != p.nranks) { set_error("Px*Py != nranks"); return PB200_ERR_INVALID; }
    const double h1 = (p.B1 - p.A1) / p.M, h2 = (p.B2 - p.A2) / p.N;
    if (!(h1 > 0x1p-100 && h1 < 0x1p100 && h2 > 0x1p-100 && h2 < 0x1p100)) { set_error("grid step out of range"); return PB200_ERR_INVALID; }
    return PB200_OK;
}

Consts make_consts(const pb200_params &p) {
    Consts c{};
    c.A1 = p.A1;
    c.A2 = p.A2;
    c.h1 = (p.B1 - p.A1) / p.M; // S4:691
    c.h2 = (p.B2 - p.A2) / p.N; // S4:692
    c.rh1 = 1.0 / c.h1;
    c.rh2 = 1.0 / c.h2;
    int e1 = 0, e2 = 0;
    c.pow2 = (std::frexp(c.h1, &e1) == 0.5 ? 1 : 0) | (std::frexp(c.h2, &e2) == 0.5 ? 2 : 0);
    c.mih1 = -1.0 / c.h1; // S4:526
    c.mih2 = -1.0 / c.h2; // S4:530
    c.h1h1 = c.h1 * c.h1; // S4:556
    c.h2h2 = c.h2 * c.h2; // S4:557
    c.eps = p.eps > 0.0 ? p.eps : std::max(c.h1, c.h2) * std::max(c.h1, c.h2); // S4:693
    c.inv_eps = 1.0 / c.eps;
    c.ell_k = p.ell_k;
    c.ell_c = p.ell_k * p.ell_k;
    c.ell_ax = p.ell_ax > 0.0 ? p.ell_ax : 1.0; // S4:32-34: the reference's ellipse has x semi-axis 1
    c.inv_ax2 = 1.0 / (c.ell_ax * c.ell_ax);
    c.rhs_mode = p.rhs_mode;
    c.u_scale = p.f_val / (2.0 * c.inv_ax2 + 2.0 * c.ell_c); // -laplace(u) = f for u = u_scale (1 - (x/ax)^2 - (k y)^2)
    c.f_val = p.f_val;
    c.delta = p.delta;
    c.norm_mode = p.norm_mode;
    c.fixed = p.fixed_iters > 0;
    c.multi = p.nranks > 1;
    c.trace_len = std::max(0, p.trace_len);
    // diagonal classes: (a_ip + a_c) / (h1*h1) + (b_jp + b_c) / (h2*h2) with each coefficient 1.0 or 1/eps (S4:556-557)
    const double pair[3] = {1.0 + 1.0, c.inv_eps + 1.0, c.inv_eps + c.inv_eps};
    for (int na = 0; na < 3; ++na)
        for (int nb = 0; nb < 3; ++nb) {
            const double D = pair[na] / c.h1h1 + pair[nb] / c.h2h2;
            c.lutD[3 * na + nb] = D;
            c.lutR[3 * na + nb] = 1.0 / D;
        }
    for (int k = 9; k < 16; ++k) c.lutD[k] = c.lutR[k] = 1.0;
    const long long dflt = static_cast<long long>(p.M - 1) * static_cast<long long>(p.N - 1); // S4:1007
    c.max_iter = p.fixed_iters > 0 ? p.fixed_iters : (p.max_iter > 0 ? p.max_iter : dflt);
    return c;
}

int pitch_for(int ny) { return ((ny + 2 + 15) / 16) * 16; }

// bytes per row of the class map: one 128-byte record per warp window, whole CTAs of windows
int cpitch_for(int ny) {
    const int wcols = (ny + K1_OWN - 1) / K1_OWN;
    return ((wcols + K1_WARPS - 1) / K1_WARPS) * K1_WARPS * CLS_WIN;
}

// the iteration kernels stage their operands in dynamic shared memory (more than the 48 KB default for K2)
pb200_status enable_kernel_smem() {
    static thread_local int done_for_device = -1;
    int dev = -1;
    CK(cudaGetDevice(&dev));
    if (done_for_device == dev) return PB200_OK;
    CK(cudaFuncSetAttribute(k1_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, K1_SMEM));
    CK(cudaFuncSetAttribute(k1_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, K1_SMEM));
    CK(cudaFuncSetAttribute(k1_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, K1_SMEM));
    CK(cudaFuncSetAttribute(k1_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, K1_SMEM));
    CK(cudaFuncSetAttribute(k2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, K2_SMEM));
    CK(cudaFuncSetAttribute(k2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, K2_SMEM));
    done_for_device = dev;
    return PB200_OK;
}

// coefficient window for the fast divisions (pb200_kernels.cuh: in_div_range)
bool coef_in_div_range(double v) {
    const double a = std::fabs(v);
    return a == 0.0 || (a >= 0x1p-50 && a <= 0x1p300);
}

// ---------------------------------------------------------------------------------------------

Fields::~Fields() { release(); }

void Fields::release() {
    double **all[] = {&a, &b, &D, &w, &r, &p[0], &p[1]};
    for (double **q : all) {
        if (*q) cudaFree(*q);
        *q = nullptr;
    }
    if (cls) cudaFree(cls);
    cls = nullptr;
}

pb200_status Fields::allocate(const Geom &geom) {
    g = geom;
    elems = static_cast<size_t>(g.nx + 2) * static_cast<size_t>(g.pitch);
    // The last warp window of a row may reach past the pitch: the bulk copies then read on into the next row, and
    // past the last row into this zeroed slack.
    const size_t slack = 2 * K1_WIN;
    double **all[] = {&a, &b, &D, &w, &r, &p[0], &p[1]};
    for (double **q : all) {
        CK(cudaMalloc(q, (elems + slack) * sizeof(double)));
        CK(cudaMemset(*q + elems, 0, slack * sizeof(double)));
    }
    cls_bytes = static_cast<size_t>(g.nx + 2) * static_cast<size_t>(g.cpitch);
    CK(cudaMalloc(&cls, cls_bytes));
    CK(cudaMemset(cls, 0, cls_bytes));
    return PB200_OK;
}

// ---------------------------------------------------------------------------------------------

Solver::~Solver() {
    if (dev >= 0) cudaSetDevice(dev);
    peer_release();
    if (mail) cudaFree(mail);
    destroy_comm();
    if (graph_exec) cudaGraphExecDestroy(graph_exec);
    for (auto &e : ev_poll) if (e) cudaEventDestroy(e);
    if (ev_start) cudaEventDestroy(ev_start);
    if (ev_stop) cudaEventDestroy(ev_stop);
    if (ev_call0) cudaEventDestroy(ev_call0);
    if (ev_call1) cudaEventDestroy(ev_call1);
    if (ctl) cudaFree(ctl);
    if (part1) cudaFree(part1);
    if (part2) cudaFree(part2);
    if (trace) cudaFree(trace);
    if (norms) cudaFree(norms);
    if (tile_slots) cudaFree(tile_slots);
    if (tile_r2) cudaFree(tile_r2);
    if (h_ctl) cudaFreeHost(h_ctl);
    f.release();
    if (stream) cudaStreamDestroy(stream);
}

// Grid of a marching kernel: the rows are cut into bands of `rows` rows and every warp takes one 124-column window of
// one band (make_part). The band height is chosen with a small cost model fitted to B200 measurements at 8192^2:
//   * a piece costs rows + 2 row steps (two extra rows start every piece);
//   * the kernel lasts ceil(blocks / resident blocks) waves;
//   * per-warp speed grows with the number of waves in flight behind each other (46 / 60 / 64.5 / 67 / 71 / 72.5
//     G pt*it/s at 1 / 2 / 3 / 4 / 6 / 8 waves) -- many short pieces beat one resident wave.
// PB200_ROWS overrides the choice (tuning). The plan also carries the lower bands that end the sweep (below).
static double wave_gain(double w) {
    static const double x[] = {0.0, 0.5, 1.0, 2.0, 3.0, 4.0, 6.0, 8.0}, y[] = {0.0, 0.42, 0.63, 0.83, 0.89, 0.925, 0.98, 1.0};
    if (w >= 8.0) return 1.0;
    for (int i = 1; i < 8; ++i)
        if (w <= x[i]) return y[i - 1] + (y[i] - y[i - 1]) * (w - x[i - 1]) / (x[i] - x[i - 1]);
    return 1.0;
}

MarchPlan march_plan(int resident_blocks, int wn, int nx) {
    const int wcn = (wn + K1_WARPS - 1) / K1_WARPS;
    auto blocks_for = [&](int rows) { // bands x CTA-windows (make_part)
        return ((nx + rows - 1) / rows) * wcn;
    };
    const double resident = std::max(1, resident_blocks);
    int best_rows = 20;
    if (const char *e = std::getenv("PB200_ROWS")) best_rows = std::min(nx, std::max(1, atoi(e)));
    else {
        double best_cost = 1e300;
        for (int rows = 2; rows <= 24; ++rows) {
            const double waves = blocks_for(rows) / resident;
            const double cost = std::ceil(waves) * (rows + 2) / wave_gain(waves);
            if (cost < best_cost * 0.999) {
                best_cost = cost;
                best_rows = rows;
            }
        }
    }
    // Tail: the last CTAs to start get bands half as high, so that the SMs run dry within a few row steps of each other
    // instead of within a whole band: the last wave and a half of a long kernel, three quarters of a wave of a short one.
    // Measured: 8192^2 78.0 -> 80.7 G pt*it/s, 4096^2 291 -> 270 us, 2400x3200 166 -> 159 us per iteration; bands a third
    // or a sixth as high gain less. PB200_TAIL=0 switches it off, PB200_TAIL_ROWS / PB200_TAIL_WAVES override (tuning).
    MarchPlan m{};
    const double waves = blocks_for(best_rows) / resident;
    int tail_rows = std::max(2, best_rows / 2), nxB = 0;
    double tail_waves = waves >= 4.0 ? 1.5 : 0.75;
    if (const char *e = std::getenv("PB200_TAIL_WAVES")) tail_waves = std::max(0.1, atof(e));
    if (const char *e = std::getenv("PB200_TAIL_ROWS")) tail_rows = std::max(1, atoi(e));
    bool tail = best_rows >= 4 && waves >= 1.0;
    if (const char *e = std::getenv("PB200_TAIL")) tail = tail && atoi(e) != 0;
    if (tail) {
        const long long tail_bands = (long long)std::ceil(tail_waves * resident / wcn);
        nxB = (int)std::min<long long>(nx / 5, tail_bands * tail_rows);
    }
    m.bands.nxA = nx - nxB;
    m.bands.nbA = (m.bands.nxA + best_rows - 1) / best_rows;
    m.bands.nbB = nxB > 0 ? (nxB + tail_rows - 1) / tail_rows : 0;
    m.ctas = (m.bands.nbA + m.bands.nbB) * wcn;
    return m;
}

// Tiles of tile_loop_kernel: tr x tc tiles of at most R x C cells, one per CTA, at most one CTA per SM. Small blocks use
// fewer CTAs (a barrier among few is cheaper than idle threads are useful): about one cell per thread at least.
bool tile_plan(int nx, int ny, int sms, size_t smem_max, int &R, int &C, int &tr, int &tc, bool lean) {
    const long long cells = static_cast<long long>(nx) * ny;
    int want = static_cast<int>(std::min<long long>(sms, (cells + TL_THREADS - 1) / TL_THREADS));
    want = std::max(1, std::min(want, TL_MAX_BLOCKS));
    bool found = false;
    const char *force = std::getenv("PB200_TILE_TC"); // tuning: tiles across, instead of the cost model's choice
    struct Cand {
        double cost;
        int R, C, tr, tc;
    };
    std::vector<Cand> cands;
    for (int c_try = 1; c_try <= want; ++c_try) {
        if (force && atoi(force) > 0 && c_try != atoi(force)) continue;
        const int r_try = std::min(want / c_try, nx);
        if (r_try < 1 || c_try > ny) continue;
        const int Rq = (nx + r_try - 1) / r_try, Cq = (ny + c_try - 1) / c_try;
        if (Cq > TL_MAXC || 2 * (Rq + Cq) > TL_NH * TL_THREADS) continue;
        if (tile_smem_bytes(Rq, Cq, lean) > smem_max) continue;
        // cost: what the slowest warp does per phase -- a thread owns a pair of columns and marches down its share of the
        // rows (pb200_tile.cuh) -- times the warps that share the SM's issue slots, plus the halo ring (one p' each)
        const int npair = (Cq + 1) / 2, txp = ((npair + 31) / 32) * 32, ty = std::max(1, std::min(TL_THREADS / txp, Rq));
        const int rows_per_thread = (Rq + ty - 1) / ty, warps = ty * txp / 32;
        const double cost = static_cast<double>(rows_per_thread) * std::max(4, warps) + (2.0 * Cq + 2.0 * Rq) / 64.0;
        cands.push_back({cost, Rq, Cq, (nx + Rq - 1) / Rq, (ny + Cq - 1) / Cq});
    }
    if (cands.empty()) return false;
    size_t pick = 0;
    for (size_t i = 1; i < cands.size(); ++i)
        if (cands[i].cost < cands[pick].cost) pick = i;
    if (lean) {
        // The lean form's time depends on the cut in ways the SM-side cost above does not capture (its r, z, A p' go through the
        // L2): among the cuts within 15 % of the cheapest, the one that keeps the most SMs busy. Measured at
        // 1200 x 1600: 148 tiles of 33 x 400 cells 46 us per iteration, 145 tiles of 42 x 320 (the cheapest by the model) 63.
        const double limit = cands[pick].cost * 1.15;
        for (size_t i = 0; i < cands.size(); ++i)
            if (cands[i].cost <= limit && cands[i].tr * cands[i].tc > cands[pick].tr * cands[pick].tc) pick = i;
    }
    R = cands[pick].R;
    C = cands[pick].C;
    tr = cands[pick].tr;
    tc = cands[pick].tc;
    found = true;
    return found;
}

// Could this decomposition run the resident loop kernel? Decided on the LARGEST block of the decomposition, so that every
// rank gets the same answer (a rank in that kernel and its neighbour in the marching kernels would wait for each other
// for ever).
bool Solver::tile_feasible(int sms, size_t smem_max, int &R, int &C, bool lean) const {
    const int nx_max = (prm.M - 1 + Px - 1) / Px, ny_max = (prm.N - 1 + Py - 1) / Py;
    int tr = 0, tc = 0;
    return tile_plan(nx_max, ny_max, std::min(sms, TL_MAX_BLOCKS), smem_max, R, C, tr, tc, lean);
}

void Solver::plan_tile(int sms) {
    use_tile = false;
    const char *env = std::getenv("PB200_TILE"), *env_loop = std::getenv("PB200_LOOP");
    if ((env && atoi(env) == 0) || (env_loop && atoi(env_loop) == 0)) return; // tuning / tests: the other kernels
    if (c.multi == 1) return; // NCCL mode exchanges between kernels: the marching kernels
    // One GPU, tiny block: the register-resident loop kernel with a thread per cell is quicker (40 x 40: 7.8 against 10.1 us
    // per iteration, 100 x 100: 7.7 against 9.6; 400 x 600: 16.3 against 13.6). PB200_TILE=2 keeps this kernel (tests).
    const long long cells = static_cast<long long>(f.g.nx) * f.g.ny;
    if (c.multi == 0 && use_loop && loop_cells == 1 && cells <= 60000 && !(env && atoi(env) == 2)) return;
    int coop = 0, smem_optin = 0;
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
    cudaDeviceGetAttribute(&smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    cudaFuncAttributes fa{};
    if (!coop || cudaFuncGetAttributes(&fa, tile_loop_kernel<false>) != cudaSuccess) {
        cudaGetLastError();
        return;
    }
    const size_t smem_max = static_cast<size_t>(smem_optin) - fa.sharedSizeBytes;
    int R = 0, C = 0, tr = 0, tc = 0;
    if (!tile_r2) return;
    // the form with everything in shared memory if the largest block fits it, else the lean one (PB200_TILE_LEAN=0: never,
    // =1: always -- tuning / tests)
    const char *env_lean = std::getenv("PB200_TILE_LEAN");
    bool lean = env_lean && atoi(env_lean) == 1;
    if (lean || !tile_feasible(sms, smem_max, R, C, false)) {
        if (env_lean && atoi(env_lean) == 0) return;
        if (!tile_feasible(sms, smem_max, R, C, true)) return;
        lean = true;
    }
    int R2 = 0, C2 = 0, tr2 = 0, tc2 = 0; // this rank's own block: its own best cut, else the largest block's tile size
    if (tile_plan(f.g.nx, f.g.ny, std::min(sms, TL_MAX_BLOCKS), smem_max, R2, C2, tr2, tc2, lean)) {
        R = R2;
        C = C2;
    }
    tr = (f.g.nx + R - 1) / R;
    tc = (f.g.ny + C - 1) / C;
    const size_t smem = tile_smem_bytes(R, C, lean);
    if (cudaFuncSetAttribute(tile_loop_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_max)) != cudaSuccess ||
        cudaFuncSetAttribute(tile_loop_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_max)) != cudaSuccess) {
        cudaGetLastError();
        return;
    }
    int occ = 0;
    const cudaError_t e_occ = lean ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, tile_loop_kernel<true>, TL_THREADS, smem)
                                   : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, tile_loop_kernel<false>, TL_THREADS, smem);
    if (e_occ != cudaSuccess || occ < 1) {
        cudaGetLastError();
        return;
    }
    if (static_cast<long long>(tr) * tc > static_cast<long long>(occ) * sms) return;
    use_tile = true;
    tile_lean = lean;
    tile_R = R;
    tile_C = C;
    tile_tr = tr;
    tile_tc = tc;
    tile_smem = smem;
}

void Solver::plan_launch() {
    wn = (f.g.ny + K1_OWN - 1) / K1_OWN;
    int sms = 148, occ1 = 1, occ2 = 1;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ1, k1_kernel<false, false>, K1_THREADS, K1_SMEM);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ2, k2_kernel<false>, K2_THREADS, K2_SMEM);
    k1_plan = march_plan(std::max(1, occ1) * sms, wn, f.g.nx);
    k2_plan = march_plan(std::max(1, occ2) * sms, wn, f.g.nx);
    k1_grid = dim3(k1_plan.ctas, 1, 1);
    k2_grid = dim3(k2_plan.ctas, 1, 1);
    // Small single-GPU blocks: the whole loop as one cooperative launch with a resident thread per cell (cell_loop_kernel),
    // if the GPU can hold a thread for every cell at once. PB200_LOOP=0 switches it off (tuning, tests).
    use_loop = false;
    int coop = 0;
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
    const char *env_loop = std::getenv("PB200_LOOP");
    if (c.multi == 0 && coop && !(env_loop && atoi(env_loop) == 0)) {
        for (int nc : {1, 2, 4}) { // cells per thread: the smallest for which every block is resident
            int occL = 0;
            if (nc == 1) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occL, cell_loop_kernel<1>, CL_THREADS, 0);
            if (nc == 2) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occL, cell_loop_kernel<2>, CL_THREADS, 0);
            if (nc == 4) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occL, cell_loop_kernel<4>, CL_THREADS, 0);
            const dim3 grid((f.g.ny + CL_BX - 1) / CL_BX, (f.g.nx + CL_BY * nc - 1) / (CL_BY * nc), 1);
            if (static_cast<long long>(grid.x) * grid.y <= static_cast<long long>(occL) * sms) {
                use_loop = true;
                loop_cells = nc;
                loop_grid = grid;
                break;
            }
        }
        // too large for registers, small enough for the L2: the streaming form of the loop (loop_cells = 0)
        const double pts_local = static_cast<double>(f.g.nx) * f.g.ny;
        double stream_max = LOOP_STREAM_MAX_POINTS;
        if (const char *e = std::getenv("PB200_LOOP_STREAM_MAX")) stream_max = atof(e);
        if (!use_loop && pts_local <= stream_max) {
            int occS = 0;
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occS, cell_loop_stream_kernel, CL_THREADS, 0);
            const long long tiles = static_cast<long long>((f.g.ny + CL_BX - 1) / CL_BX) * ((f.g.nx + CL_BY - 1) / CL_BY);
            if (occS > 0) {
                use_loop = true;
                loop_cells = 0;
                loop_grid = dim3(static_cast<unsigned>(std::min<long long>(tiles, static_cast<long long>(occS) * sms)), 1, 1);
            }
        }
    }
    plan_tile(sms);
    norms_grid = dim3((f.g.ny + 255) / 256, std::min(f.g.nx, 512), 1);
    init_grid = dim3((f.g.ny + INIT_THREADS - 1) / INIT_THREADS, std::min(f.g.nx, 1024), 1);
}

pb200_status Solver::create(const pb200_params &params) {
    const double t0 = now_s();
    prm = params;
    pb200_status st = validate(prm);
    if (st != PB200_OK) return st;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        set_error("no CUDA device: libpoisson_b200 has no CPU path");
        return PB200_ERR_CUDA;
    }
    if (prm.device >= 0) CK(cudaSetDevice(prm.device));
    CK(cudaGetDevice(&dev));
    Px = prm.Px;
    Py = prm.Py;
    if (Px <= 0 || Py <= 0) choose_process_grid(prm.nranks, Px, Py);
    if (Px > prm.M - 1 || Py > prm.N - 1) { set_error("more blocks than interior nodes"); return PB200_ERR_INVALID; }
    decompose_2d(prm.M, prm.N, Px, Py, prm.rank, i0, i1, j0, j1);
    Geom g{};
    g.nx = i1 - i0 + 1;
    g.ny = j1 - j0 + 1;
    g.pitch = pitch_for(g.ny);
    g.cpitch = cpitch_for(g.ny);
    g.gi0 = i0 - 1;
    g.gj0 = j0 - 1;
    g.M = prm.M;
    g.N = prm.N;
    c = make_consts(prm);
    // the coefficients are 1, 1/eps or a blend of the two (S4:167-177): div_const's fast path needs them in range
    safe_div = !(coef_in_div_range(1.0 / c.eps) && c.h1 <= 0x1p10 && c.h2 <= 0x1p10);
    CK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    st = enable_kernel_smem();
    if (st != PB200_OK) return st;
    st = f.allocate(g);
    if (st != PB200_OK) return st;
    plan_launch();
    const size_t nbL = use_loop ? static_cast<size_t>(loop_grid.x) * loop_grid.y : 0; // cell_loop_kernel: 3 partials per block
    const size_t nb1 = static_cast<size_t>(k1_grid.x) * k1_grid.y, nb2 = std::max(static_cast<size_t>(k2_grid.x) * k2_grid.y, (3 * nbL + 1) / 2);
    CK(cudaMalloc(&ctl, sizeof(Ctl)));
    CK(cudaMalloc(&part1, nb1 * sizeof(double)));
    const size_t nb3 = std::max(static_cast<size_t>(norms_grid.x) * norms_grid.y, static_cast<size_t>(init_grid.x) * init_grid.y);
    CK(cudaMalloc(&part2, std::max(2 * nb2, 3 * nb3) * sizeof(double))); // norms_kernel: 3 partials per block
    CK(cudaMalloc(&norms, 8 * sizeof(double)));
    CK(cudaMalloc(&tile_slots, TL_SLOT_WORDS * sizeof(unsigned long long)));
    { // the tagged copy of r, for decompositions whose blocks fit the resident loop kernel (before any peer export)
        int sms = 148, smem_optin = 0, R = 0, C = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaDeviceGetAttribute(&smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        cudaFuncAttributes fa{};
        if (cudaFuncGetAttributes(&fa, tile_loop_kernel<false>) == cudaSuccess && smem_optin > static_cast<int>(fa.sharedSizeBytes) &&
            tile_feasible(sms, static_cast<size_t>(smem_optin) - fa.sharedSizeBytes, R, C, true)) {
            CK(cudaMalloc(&tile_r2, 2 * f.elems * sizeof(unsigned long long)));
            plan_launch(); // again, now that the kernel's halo buffer exists
        } else
            cudaGetLastError();
    }
    if (!mail) { // the mailbox of the in-kernel all-reduce (peer_export hands it to the other ranks)
        CK(cudaMalloc(&mail, MAIL_DOUBLES * sizeof(double)));
        CK(cudaMemset(mail, 0, MAIL_DOUBLES * sizeof(double)));
    }
    if (c.trace_len > 0) {
        CK(cudaMalloc(&trace, static_cast<size_t>(c.trace_len) * 3 * sizeof(double)));
        CK(cudaMemsetAsync(trace, 0, static_cast<size_t>(c.trace_len) * 3 * sizeof(double), stream));
    }
    CK(cudaMallocHost(&h_ctl, 2 * sizeof(Ctl)));
    CK(cudaEventCreate(&ev_start));
    CK(cudaEventCreate(&ev_stop));
    CK(cudaEventCreate(&ev_call0));
    CK(cudaEventCreate(&ev_call1));
    for (auto &e : ev_poll) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    st = run_setup();
    if (st != PB200_OK) return st;
    CK(cudaStreamSynchronize(stream));
    t_setup = now_s() - t0;
    return PB200_OK;
}

pb200_status Solver::run_setup() {
    dim3 grid((f.g.pitch + 255) / 256, f.g.nx + 2);
    if (coefficients_ready) reset_kernel<<<grid, 256, 0, stream>>>(f.g, c, f.r, f.w, f.p[0], f.p[1]); // a, b, D, classes stay
    else setup_kernel<<<grid, 256, 0, stream>>>(f.g, c, f.a, f.b, f.D, f.r, f.w, f.p[0], f.p[1], nullptr, f.cls);
    coefficients_ready = true;
    ++launches;
    CK(cudaGetLastError());
    return PB200_OK;
}

K1Args Solver::k1_args(long long k) const {
    K1Args a{};
    a.r = f.r;
    a.p_old = f.p[(k - 1) & 1];
    a.a = f.a;
    a.b = f.b;
    a.D = f.D;
    a.cls = f.cls;
    a.p_new = f.p[k & 1];
    a.Ap = nullptr;
    a.partials = part1;
    a.ctl = ctl;
    a.g = f.g;
    a.c = c;
    a.wn = wn;
    a.bands = k1_plan.bands;
    a.pv = pv;
    return a;
}

K2Args Solver::k2_args(long long k) const {
    K2Args a{};
    a.w = f.w;
    a.r = f.r;
    a.p = f.p[k & 1];
    a.a = f.a;
    a.b = f.b;
    a.D = f.D;
    a.cls = f.cls;
    a.partials = part2;
    a.trace = trace;
    a.ctl = ctl;
    a.g = f.g;
    a.c = c;
    a.wn = wn;
    a.bands = k2_plan.bands;
    a.pv = pv;
    return a;
}

void Solver::launch_k1(const K1Args &a, bool first) {
    if (first) {
        if (safe_div) k1_kernel<true, true><<<k1_grid, K1_THREADS, K1_SMEM, stream>>>(a);
        else k1_kernel<true, false><<<k1_grid, K1_THREADS, K1_SMEM, stream>>>(a);
    } else {
        if (safe_div) k1_kernel<false, true><<<k1_grid, K1_THREADS, K1_SMEM, stream>>>(a);
        else k1_kernel<false, false><<<k1_grid, K1_THREADS, K1_SMEM, stream>>>(a);
    }
}

void Solver::launch_k2(const K2Args &a) {
    if (safe_div) k2_kernel<true><<<k2_grid, K2_THREADS, K2_SMEM, stream>>>(a);
    else k2_kernel<false><<<k2_grid, K2_THREADS, K2_SMEM, stream>>>(a);
}

void Solver::launch_init() {
    InitArgs ia{};
    ia.r = f.r;
    ia.D = f.D;
    ia.partials = part2;
    ia.ctl = ctl;
    ia.g = f.g;
    ia.c = c;
    init_kernel<<<init_grid, INIT_THREADS, 0, stream>>>(ia);
}

void Solver::launch_iteration(long long k, bool first) {
    launch_k1(k1_args(k), first);
    if (c.multi == 1) allreduce(0, 1);
    launch_k2(k2_args(k));
    if (c.multi == 1) { // NCCL mode; in peer-memory mode (2) the two kernels exchange everything themselves
        exchange_r_halo();
        allreduce(1, 2);
        finalize_kernel<<<1, 1, 0, stream>>>(ctl, c, trace, 0);
    }
}

int Solver::kernels_per_iteration() const { return c.multi == 1 ? 3 + halo_kernels_per_iteration() : 2; }

pb200_status Solver::solve(pb200_result *res) {
    const double t0 = now_s();
    CK(cudaSetDevice(dev));
    if (c.multi && !comm_ready()) { set_error("nranks > 1 but no communicator attached (pb200_attach_comm)"); return PB200_ERR_COMM; }
    const long long launches0 = launches;
    CK(cudaEventRecord(ev_call0, stream));
    if (solved) { // a second solve on the same handle starts from scratch (S4:803-834)
        pb200_status st = run_setup();
        if (st != PB200_OK) return st;
    }
    // ---- init (S4:803-843): w = 0 and r = B were written by setup; zr_old = (D^-1 r, r)
    CK(cudaMemsetAsync(ctl, 0, sizeof(Ctl), stream));
    CK(cudaMemsetAsync(norms, 0, 8 * sizeof(double), stream));
    // peer-memory mode: clear the mailbox; the NCCL all-reduce of the start-up (z,r) below doubles as the barrier that
    // keeps any rank from posting into a mailbox that has not been cleared yet
    if (c.multi == 2 || use_tile) CK(cudaMemsetAsync(mail, 0, MAIL_DOUBLES * sizeof(double), stream));
    launch_init();
    ++launches;
    if (c.multi) {
        allreduce(2, 1);
        finalize_kernel<<<1, 1, 0, stream>>>(ctl, c, trace, 1);
        ++launches;
    }
    CK(cudaGetLastError());

    // ---- loop (S4:845-943)
    // The poll interval must be the same on every rank: in NCCL mode a rank that leaves the loop after a different
    // number of launched iterations would issue collectives the others do not match. So it is derived from the global
    // grid and the rank count, never from this rank's own block (blocks differ by a row or a column); an explicit
    // check_every has to be the same number on all ranks.
    const double pts_rank = static_cast<double>(prm.M - 1) * (prm.N - 1) / prm.nranks;
    int check = prm.check_every;
    if (check <= 0) { // poll roughly every 250 us of device work
        const double t_it = std::max(6e-6, pts_rank * 112.0 / 6.0e12);
        check = static_cast<int>(std::min(64.0, std::max(2.0, 250e-6 / t_it)));
    }
    check += check & 1; // even: graph replays cover two iterations
    // No graph replay in NCCL mode: with NCCL collectives captured in the graph the replay hung on this stack
    // (NCCL 2.27.3 / driver 580, 2 ranks). In peer-memory mode the iteration is two plain kernels and is replayed.
    const bool use_graph = prm.use_graph != 0 && c.multi != 1;
    CK(cudaEventRecord(ev_start, stream));
    long long k = 1;
    if (use_tile) { // one cooperative launch per rank runs every iteration, halo and sums included (pb200_tile.cuh)
        CK(cudaMemsetAsync(tile_slots, 0, TL_SLOT_WORDS * sizeof(unsigned long long), stream));
        CK(cudaMemsetAsync(tile_r2, 0, 2 * f.elems * sizeof(unsigned long long), stream));
        TileArgs ta{};
        ta.w = f.w;
        ta.r = f.r;
        ta.p[0] = f.p[0];
        ta.p[1] = f.p[1];
        ta.a = f.a;
        ta.b = f.b;
        ta.D = f.D;
        ta.cls = f.cls;
        ta.slots = tile_slots;
        ta.r2 = tile_r2;
        ta.gz = f.p[0];
        ta.gap = f.p[1];
        ta.trace = trace;
        ta.ctl = ctl;
        ta.g = f.g;
        ta.c = c;
        ta.R = tile_R;
        ta.C = tile_C;
        ta.tc = tile_tc;
        ta.safe = safe_div ? 1 : 0;
        ta.pv = pv;
        if (c.multi == 0) { // one GPU: the all-reduce still goes through the (own) mailbox
            ta.pv = PeerView{};
            ta.pv.nranks = 1;
            ta.pv.rank = 0;
            ta.pv.mail[0] = mail;
        }
        void *args[] = {&ta};
        CK(cudaLaunchCooperativeKernel(tile_lean ? reinterpret_cast<const void *>(tile_loop_kernel<true>)
                                                 : reinterpret_cast<const void *>(tile_loop_kernel<false>),
                                       dim3(tile_tr * tile_tc, 1, 1),
                                       dim3(TL_THREADS, 1, 1), args, tile_smem, stream));
        ++launches;
        k = c.max_iter + 1;
    } else if (use_loop) { // one cooperative launch runs every iteration; delta and max_iter are enforced on the device
        CellLoopArgs la{};
        la.w = f.w;
        la.r = f.r;
        la.p[0] = f.p[0];
        la.p[1] = f.p[1];
        la.a = f.a;
        la.b = f.b;
        la.D = f.D;
        la.partials = part2;
        la.trace = trace;
        la.ctl = ctl;
        la.g = f.g;
        la.c = c;
        void *args[] = {&la};
        const void *fn = loop_cells == 0   ? reinterpret_cast<const void *>(cell_loop_stream_kernel)
                         : loop_cells == 1 ? reinterpret_cast<const void *>(cell_loop_kernel<1>)
                         : loop_cells == 2 ? reinterpret_cast<const void *>(cell_loop_kernel<2>)
                                           : reinterpret_cast<const void *>(cell_loop_kernel<4>);
        CK(cudaLaunchCooperativeKernel(fn, loop_grid, dim3(CL_THREADS, 1, 1), args, 0, stream));
        ++launches;
        k = c.max_iter + 1;
    } else {
        launch_iteration(k, true); // iteration 1: p = z (S4:830)
        launches += kernels_per_iteration();
        ++k;
    }
    CK(cudaGetLastError());
    if (!use_loop && !use_tile && use_graph && !graph_exec && c.max_iter > 1) {
        cudaGraph_t graph = nullptr;
        CK(cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal));
        launch_iteration(2, false);
        launch_iteration(3, false);
        // whatever happened in between, the capture is ended here, so an error never leaves the stream capturing
        const cudaError_t e_end = cudaStreamEndCapture(stream, &graph);
        cudaError_t e_inst = cudaSuccess;
        if (e_end == cudaSuccess) e_inst = cudaGraphInstantiate(&graph_exec, graph, 0);
        if (graph) cudaGraphDestroy(graph);
        if (e_end != cudaSuccess || e_inst != cudaSuccess) graph_exec = nullptr;
        CK(e_end);
        CK(e_inst);
    }
    int slot = 0, pending = -1;
    bool done = false;
    long long since_poll = 1;
    while (!done && k <= c.max_iter) {
        if (comm_failed()) break; // NCCL mode: a collective could not be enqueued; reduce_norms_host reports it
        if (use_graph && graph_exec) {
            CK(cudaGraphLaunch(graph_exec, stream)); // iterations k, k+1 (the second is a no-op past the end)
            launches += 2 * kernels_per_iteration();
            k += 2;
            since_poll += 2;
        } else {
            launch_iteration(k, false);
            launches += kernels_per_iteration();
            ++k;
            ++since_poll;
        }
        if (since_poll >= check || k > c.max_iter) {
            since_poll = 0;
            CK(cudaMemcpyAsync(&h_ctl[slot], ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, stream));
            CK(cudaEventRecord(ev_poll[slot], stream));
            if (pending >= 0) { // look at the batch BEFORE the one just queued: the GPU never idles
                CK(cudaEventSynchronize(ev_poll[pending]));
                done = h_ctl[pending].done != 0;
            }
            pending = slot;
            slot ^= 1;
        }
    }
    CK(cudaEventRecord(ev_stop, stream));
    CK(cudaGetLastError());
    // ---- results
    {
        norms_kernel<<<norms_grid, 256, 0, stream>>>(f.g, c, f.w, f.r, part2, &ctl->ticket[1], norms);
        ++launches;
    }
    double h_norms[5] = {0, 0, 0, 0, 0}; // sum w^2, sum r^2, sum (w-u)^2 in D | max |w|, max |w-u| in D
    if (c.multi) { // sums and max over all ranks, in stream
        pb200_status st = reduce_norms_host(h_norms);
        if (st != PB200_OK) return st;
    }
    CK(cudaMemcpyAsync(&h_ctl[0], ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, stream));
    CK(cudaMemcpyAsync(h_norms, norms, 5 * sizeof(double), cudaMemcpyDeviceToHost, stream));
    CK(cudaEventRecord(ev_call1, stream));
    CK(cudaStreamSynchronize(stream));
    CK(cudaGetLastError());
    float ms = 0.f, ms_call = 0.f;
    CK(cudaEventElapsedTime(&ms, ev_start, ev_stop));
    CK(cudaEventElapsedTime(&ms_call, ev_call0, ev_call1));
    const Ctl &hc = h_ctl[0];
    solved = true;
    last_iter = hc.iter;
    if (hc.status == STOP_COMM_TIMEOUT) { set_error("a peer rank did not answer within the in-kernel time limit (peer-memory mode)"); return PB200_ERR_COMM; }
    if (res) {
        std::memset(res, 0, sizeof(*res));
        res->iterations = hc.iter;
        res->stop_reason = hc.status;
        res->first_converged_iter = hc.conv_iter;
        res->diff_norm = hc.diff;
        res->zr = hc.zr_old;
        res->alpha = hc.alpha;
        res->beta = hc.beta;
        res->w_norm_E = std::sqrt(h_norms[0] * c.h1 * c.h2);
        res->r_norm_E = std::sqrt(h_norms[1] * c.h1 * c.h2);
        res->w_max = h_norms[3];
        res->err_max_in_D = h_norms[4];
        res->err_l2_in_D = std::sqrt(h_norms[2] * c.h1 * c.h2);
        res->t_setup_s = t_setup;
        res->t_solve_s = ms * 1e-3;
        res->t_total_s = now_s() - t0;
        const double pts = static_cast<double>(prm.M - 1) * (prm.N - 1);
        res->pt_it_per_s = res->t_solve_s > 0 ? pts * static_cast<double>(hc.iter) / res->t_solve_s : 0.0;
        res->algo_gbps = res->pt_it_per_s * 104.0 * 1e-9;
        res->kernel_launches = launches - launches0;
        res->t_device_s = ms_call * 1e-3;
        res->safe_mode = safe_div ? 1.0 : 0.0;
    }
    return PB200_OK;
}

// Measurement aid: mean duration of K1 and of K2 over `iters` iterations from a fresh start,
// CUDA events on the solver stream around every launch (plain launches, no graph).
pb200_status Solver::profile_iteration(int iters, double ms_out[2]) {
    if (c.multi) { set_error("profile_iteration is single-rank"); return PB200_ERR_INVALID; }
    if (iters < 1) { set_error("iters must be >= 1"); return PB200_ERR_INVALID; }
    CK(cudaSetDevice(dev));
    pb200_status st = run_setup();
    if (st != PB200_OK) return st;
    CK(cudaMemsetAsync(ctl, 0, sizeof(Ctl), stream));
    const Consts saved = c;
    c.fixed = 1;
    c.max_iter = 1LL << 60;
    launch_init();
    std::vector<cudaEvent_t> ev(3 * (size_t)iters);
    for (auto &e : ev) CK(cudaEventCreate(&e));
    for (int k = 1; k <= iters; ++k) {
        K2Args a2 = k2_args(k);
        a2.trace = nullptr;
        cudaEventRecord(ev[3 * (k - 1) + 0], stream);
        launch_k1(k1_args(k), k == 1);
        cudaEventRecord(ev[3 * (k - 1) + 1], stream);
        launch_k2(a2);
        cudaEventRecord(ev[3 * (k - 1) + 2], stream);
    }
    c = saved;
    launches += 2 * iters + 2;
    CK(cudaStreamSynchronize(stream));
    CK(cudaGetLastError());
    double t1 = 0.0, t2 = 0.0;
    for (int k = 0; k < iters; ++k) {
        float a = 0.f, b = 0.f;
        cudaEventElapsedTime(&a, ev[3 * k], ev[3 * k + 1]);
        cudaEventElapsedTime(&b, ev[3 * k + 1], ev[3 * k + 2]);
        t1 += a;
        t2 += b;
    }
    for (auto &e : ev) cudaEventDestroy(e);
    ms_out[0] = t1 / iters;
    ms_out[1] = t2 / iters;
    solved = true; // the fields now hold `iters` iterations of a fixed-count run
    last_iter = iters;
    return PB200_OK;
}

pb200_status Solver::get_solution(double *host, int64_t ld) {
    if (!solved) { set_error("get_solution before solve"); return PB200_ERR_STATE; }
    if (!host || ld < prm.N + 1) { set_error("bad host array / ld"); return PB200_ERR_INVALID; }
    CK(cudaSetDevice(dev));
    CK(cudaMemcpy2DAsync(host + static_cast<size_t>(i0) * ld + j0, ld * sizeof(double), f.w + f.g.pitch + 1,
                         f.g.pitch * sizeof(double), f.g.ny * sizeof(double), f.g.nx, cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    return PB200_OK;
}

pb200_status Solver::get_trace(double *host, int n, int *rows) {
    if (!solved) { set_error("get_trace before solve"); return PB200_ERR_STATE; }
    const int have = static_cast<int>(std::min<long long>(std::min(n, c.trace_len), last_iter));
    if (rows) *rows = have;
    if (have > 0) {
        CK(cudaSetDevice(dev));
        CK(cudaMemcpy(host, trace, static_cast<size_t>(have) * 3 * sizeof(double), cudaMemcpyDeviceToHost));
    }
    return PB200_OK;
}

} // namespace pb200
