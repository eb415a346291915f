// pb200_kernels.cuh -- device code of the fused PCG iteration for sm_100a.
//
// Replaces the seven kernels of the reference (stage4-mpi+cuda/poisson_mpi_cuda_f.cu, "S4"):
//   setup_kernel  <- fictitious_regions_setup_local (host loop, S4:146-192) + zero_kernel (S4:599-609)
//   init_kernel   <- apply_Dinv_kernel + dot_kernel at start-up (S4:820-843)
//   k1_kernel     <- apply_Dinv_kernel (S4:541-562) + update_p_kernel (S4:663-676) / copy_kernel (S4:612-623)
//                    + apply_A_kernel (S4:507-536) + dot_kernel for (Ap,p) (S4:574-598)
//   k2_kernel     <- apply_A_kernel again (Ap is recomputed, never stored) + update_w_r_kernel (S4:626-660)
//                    + apply_Dinv_kernel + dot_kernel for (z,r) + the host-side scalar logic (S4:872-906, 926-932)
//
// Arithmetic contract: every element-wise expression keeps the reference's expression tree with
// one rounding per operation (this file is compiled with -fmad=false; the only fused operations
// are the explicit __fma_rn inside div_const_nocheck, which implements a correctly rounded division).
//
// Layout of a block (nx x ny interior cells, 1-cell ring): rows li = 0..nx+1, columns lj = 0..ny+1,
// element (li,lj) at base[li*pitch + lj]; pitch is a multiple of 16 doubles so every row starts on
// a 128-byte line; columns ny+2..pitch-1 are padding.
#pragma once
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace pb200 {

struct Geom {
    int nx, ny;   // interior cells of this block
    int pitch;    // doubles per row, multiple of 16
    int gi0, gj0; // global index of local (0,0): i_start-1, j_start-1 (S4:156-157)
    int M, N;
    int cpitch;   // bytes per row of the class map (one 128-byte record per 124-column warp window)
};

// The class map is stored per warp window so that a warp's 128 class bytes of a row are one aligned
// 128-byte record (bulk-copyable): byte k of record w of row li belongs to column 124*w + k.
constexpr int CLS_OWN = 124, CLS_WIN = 128;
__host__ __device__ __forceinline__ size_t cls_index(int cpitch, int li, int w, int k) {
    return (size_t)li * (size_t)cpitch + (size_t)w * CLS_WIN + (size_t)k;
}
// store the class byte of column lj of row li into every record that covers it (the first 4 columns of a
// window are also the last 4 of the previous one)
__device__ __forceinline__ void cls_store(unsigned char *cls, int cpitch, int li, int lj, unsigned v) {
    const int w = lj / CLS_OWN, k = lj - w * CLS_OWN;
    if ((w + 1) * CLS_WIN <= cpitch) cls[cls_index(cpitch, li, w, k)] = (unsigned char)v;
    if (w > 0 && k < CLS_WIN - CLS_OWN) cls[cls_index(cpitch, li, w - 1, k + CLS_OWN)] = (unsigned char)v;
}

struct Consts {
    double A1, A2;             // rectangle origin (S4:11-12)
    double h1, h2;             // S4:691-692
    double rh1, rh2;           // RN(1/h1), RN(1/h2) for div_const
    double mih1, mih2;         // -1.0/h1, -1.0/h2 (S4:526, 530)
    int pow2;                  // bit 0: h1 is a power of two, bit 1: h2 is (x / h is then the exact product x * rh)
    double h1h1, h2h2;         // h1*h1, h2*h2 (S4:556-557)
    double eps, inv_eps;       // S4:693; 1.0/eps (S4:169)
    double ell_k, ell_c;       // ellipse (x/ax)^2 + (k y)^2 < 1; c = k*k (S4:33: 4.0)
    double ell_ax, inv_ax2;    // x semi-axis (the reference: 1) and 1/(ax*ax); with ax = 1 every product below is exact
    int rhs_mode;              // 0: point indicator (S4:182-191); 1: area fraction of the cell (no reference counterpart)
    double u_scale;            // f / (2/ax^2 + 2 k^2): the analytic solution on D is u_scale * (1 - (x/ax)^2 - (k y)^2)
    double f_val;              // S4:13
    double delta;              // S4:1006
    int norm_mode;             // 0 weighted (S4:897), 1 stage0 (S0:154)
    int fixed;                 // 1: delta test does not stop the loop
    int multi;                 // 0: one GPU. 1: scalars all-reduced by NCCL between kernels, finalize_kernel runs the tail.
                               // 2: scalars and halos travel through peer memory inside K1/K2 (PeerView)
    int trace_len;
    long long max_iter;        // S4:1007 (or fixed_iters): enforced on the device
    // Coefficient classes: a, b are 1.0 (class 0) or 1/eps (class 1) everywhere except on the
    // O(M+N) cells whose edge is cut by the ellipse (class 2: value kept in the full array). The diagonal
    // D = (a+a)/(h1*h1) + (b+b)/(h2*h2) then takes one of 9 values: lutD[3*na+nb], na/nb = how many of the
    // two a's / b's are 1/eps; lutR = RN(1/lutD). Class 15 = anything else (value kept in the full array).
    double lutD[16], lutR[16];
};

constexpr unsigned CLS_GENERAL = 2u;  // a / b class: read the full array
constexpr unsigned DCLS_GENERAL = 15u; // D class: read the full array
constexpr unsigned DCLS_RING = 9u;     // cells outside the global interior: D = 1 (their r is 0)

// class byte of one cell: bits 0-1 a class, bits 2-3 b class, bits 4-7 D class
__device__ __forceinline__ unsigned coef_class(double v, double inv_eps) {
    return (v == 1.0) ? 0u : ((v == inv_eps) ? 1u : CLS_GENERAL);
}
__device__ __forceinline__ unsigned diag_class(const Consts &c, unsigned ca_ip, unsigned ca, unsigned cb_jp, unsigned cb, double vD) {
    if (ca_ip < 2u && ca < 2u && cb_jp < 2u && cb < 2u) {
        const unsigned idx = (ca_ip + ca) * 3u + (cb_jp + cb);
        if (c.lutD[idx] == vD) return idx;
    }
    return DCLS_GENERAL;
}

// Device-resident loop state: what the reference keeps in host variables (S4:836-848, 872-932).
struct Ctl {
    double zr_old;   // (z,r) of the previous iteration
    double denom;    // (Ap,p)
    double alpha, beta;
    double diff;     // last ||w(k+1)-w(k)||
    double red[4];   // [0] denom (this rank) [1] sum dw^2 (this rank, unweighted) [2] (z,r) (this rank) -- all-reduced in place when multi
    long long iter;  // S4:848
    int done;        // loop has ended; later kernels return immediately
    int status;      // PB200_STOP_*
    int conv_iter;   // first k with diff < delta
    unsigned int ticket[2];
};

enum { STOP_MAX_ITER = 0, STOP_CONVERGED = 1, STOP_DENOM = 2, STOP_COMM_TIMEOUT = 3 };

// ------------------------------------------------------------------------------------------
// Peer-memory communication (multi == 2): every rank maps every other rank's mailbox and its neighbours' r arrays
// (CUDA IPC between processes, peer access between threads) and the iteration kernels do the exchange themselves:
//   * the last block of K1 / K2 stores this rank's partial sums into EVERY rank's mailbox, fences, raises a flag, then
//     waits until all flags of its own mailbox show the current iteration and adds the partials in rank order -- an
//     all-reduce made of NVLink stores, one thread per GPU spinning on local memory, no NCCL call, graph-replayable;
//   * K2 stores the new r of its boundary rows / columns straight into the neighbours' ring cells.
// Replaces MPI_Allreduce (S4:872, 893, 926) and exchange_halos_2d_gpu (S4:331-500).
constexpr int MAX_RANKS = 8;            // one NVSwitch box
constexpr int MAIL_DOUBLES = 64;        // den[2][8] | dif[2][8] | zr[2][8] | flag_k1[8] | flag_k2[8]
constexpr int MAIL_DEN = 0, MAIL_DIF = 16, MAIL_ZR = 32, MAIL_FLAG1 = 48, MAIL_FLAG2 = 56;

struct PeerView {
    int nranks, rank;
    double *mail[MAX_RANKS];                    // every rank's mailbox (own included)
    double *r_left, *r_right, *r_down, *r_up;   // neighbours' r arrays; null = no neighbour on that side
    int nx_left;                                // the left neighbour's ring row nx_left+1 receives my row 1
    int ny_down, pitch_down, pitch_up;          // the lower neighbour's ring column ny_down+1 receives my column 1
    // the neighbours' tagged copies of r (pb200_tile.cuh: two 64-bit words per cell, same indexing as r); null as r_*
    unsigned long long *t_left, *t_right, *t_down, *t_up;
};

// One thread: publish nv values (slots at offsets off[v]) for iteration k to every rank, then wait for everyone's and
// return the rank-ordered sums. False if a peer does not show up within ~2 s (the caller stops the solve).
__device__ __forceinline__ bool peer_allreduce(const PeerView &pv, const int flag_off, const unsigned long long k, const int nv,
                                               const int *off, const double *vals, double *sums) {
    const int par = (int)(k & 1ull) * MAX_RANKS;
    for (int q = 0; q < pv.nranks; ++q)
        for (int v = 0; v < nv; ++v) pv.mail[q][off[v] + par + pv.rank] = vals[v];
    __threadfence_system();
    for (int q = 0; q < pv.nranks; ++q)
        *reinterpret_cast<volatile unsigned long long *>(pv.mail[q] + flag_off + pv.rank) = k;
    const volatile unsigned long long *flags = reinterpret_cast<const volatile unsigned long long *>(pv.mail[pv.rank] + flag_off);
    const long long t0 = clock64();
    for (int q = 0; q < pv.nranks; ++q)
        while (flags[q] < k) {
            if (clock64() - t0 > 4000000000ll) return false;
            __nanosleep(64);
        }
    __threadfence_system();
    const volatile double *mine = pv.mail[pv.rank];
    for (int v = 0; v < nv; ++v) {
        double s = 0.0;
        for (int q = 0; q < pv.nranks; ++q) s += mine[off[v] + par + q]; // rank order, like the reference's MPI_SUM
        sums[v] = s;
    }
    return true;
}

// ------------------------------------------------------------------------------------------
// small helpers

__device__ __forceinline__ double dmax_std(double a, double b) { return (a < b) ? b : a; } // std::max
__device__ __forceinline__ double dmin_std(double a, double b) { return (b < a) ? b : a; } // std::min

__device__ __forceinline__ double shfl_up1(double v) { return __shfl_up_sync(0xffffffffu, v, 1); }
__device__ __forceinline__ double shfl_dn1(double v) { return __shfl_down_sync(0xffffffffu, v, 1); }

// Fixed-shape warp sum: lane 0 ends with ((..(v0+v16)+(v8+v24)..)) -- same tree every run.
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

// Block-level deterministic reduction of up to NV values per thread, followed by the
// "last block sums all block partials in index order" pass. Returns true in thread 0 of the
// last block with the grid totals in out[0..NV). partials: [NV][nblocks].
template <int NV, int THREADS>
__device__ __forceinline__ bool grid_reduce(double (&v)[NV], double *partials, unsigned int *ticket, double (&out)[NV],
                                            const bool system_scope = false) {
    __shared__ double s_w[NV][THREADS / 32];
    __shared__ bool s_last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned nblocks = gridDim.x * gridDim.y;
    const unsigned bid = blockIdx.y * gridDim.x + blockIdx.x;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        double w = warp_sum(v[k]);
        if (lane == 0) s_w[k][warp] = w;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            double s = 0.0;
#pragma unroll
            for (int w = 0; w < THREADS / 32; ++w) s += s_w[k][w];
            __stcg(&partials[(size_t)k * nblocks + bid], s);
        }
        if (system_scope) __threadfence_system(); // this block's boundary stores may have gone to a peer GPU
        else __threadfence();
        unsigned t = atomicAdd(ticket, 1u);
        s_last = (t == nblocks - 1);
    }
    __syncthreads();
    if (!s_last) return false;
    __threadfence();
    // last block: thread t sums partials t, t+THREADS, ... in that order (loads issued eight at a time so that their
    // latencies overlap; the additions keep the sequential order); then the fixed tree above.
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        const double *pk = partials + (size_t)k * nblocks;
        double s = 0.0;
        unsigned i = threadIdx.x;
        for (; i + 7u * THREADS < nblocks; i += 8u * THREADS) {
            double v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = __ldcg(pk + i + u * THREADS);
#pragma unroll
            for (int u = 0; u < 8; ++u) s += v[u];
        }
        for (; i < nblocks; i += THREADS) s += __ldcg(pk + i);
        double w = warp_sum(s);
        if (lane == 0) s_w[k][warp] = w;
    }
    __syncthreads();
    if (threadIdx.x != 0) return false;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < THREADS / 32; ++w) s += s_w[k][w];
        out[k] = s;
    }
    *ticket = 0u; // ready for the next launch
    return true;
}

// ------------------------------------------------------------------------------------------
// geometry (S4:32-76, 159-177). Expression trees follow the reference exactly.

__device__ __forceinline__ double seg_len(const Consts &c, double cc, double s, double e, bool vertical) {
    double l;
    if (vertical) {
        const double x0 = cc;
        if (fabs(x0) >= c.ell_ax) l = 0.0; // S4:51
        else {
            const double root = sqrt(dmax_std(0.0, (1.0 - x0 * x0 * c.inv_ax2) / c.ell_c)); // S4:53-54
            l = dmax_std(0.0, dmin_std(e, root) - dmax_std(s, -root));                      // S4:55
        }
    } else {
        const double y0 = cc;
        if (fabs(c.ell_k * y0) >= 1.0) l = 0.0; // S4:59
        else {
            const double root = c.ell_ax * sqrt(dmax_std(0.0, 1.0 - c.ell_c * y0 * y0)); // S4:61-62
            l = dmax_std(0.0, dmin_std(e, root) - dmax_std(s, -root));                   // S4:63
        }
    }
    return l;
}

// length inside D of the vertical edge left of node (gi,gj) / of the horizontal edge below it (S4:159-165)
__device__ __forceinline__ double len_a(const Consts &c, int gi, int gj) {
    const double x = c.A1 + gi * c.h1, y = c.A2 + gj * c.h2;
    return seg_len(c, x - 0.5 * c.h1, y - 0.5 * c.h2, y + 0.5 * c.h2, true);
}
__device__ __forceinline__ double len_b(const Consts &c, int gi, int gj) {
    const double x = c.A1 + gi * c.h1, y = c.A2 + gj * c.h2;
    return seg_len(c, y - 0.5 * c.h2, x - 0.5 * c.h1, x + 0.5 * c.h1, false);
}
// S4:167-171 / 173-177: the coefficient of an edge of length h of which l lies inside D
__device__ __forceinline__ double coef_of(const Consts &c, double l, double h) {
    return (fabs(l - h) < 1e-9) ? 1.0 : (l < 1e-9 ? 1.0 / c.eps : (l / h) + (1.0 - l / h) / c.eps);
}
__device__ __forceinline__ double coef_a(const Consts &c, int gi, int gj) { return coef_of(c, len_a(c, gi, gj), c.h2); }
__device__ __forceinline__ double coef_b(const Consts &c, int gi, int gj) { return coef_of(c, len_b(c, gi, gj), c.h1); }

// S4:32-34
__device__ __forceinline__ bool in_D(const Consts &c, double x, double y) { return x * x * c.inv_ax2 + c.ell_c * y * y < 1.0; }

// The part [lo, hi] of a cell edge that lies in D: the two clamped end points whose difference seg_len returns
// (S4:46-76, same expressions). False when the edge misses D.
__device__ __forceinline__ bool seg_ends(const Consts &c, double cc, double s, double e, bool vertical, double &lo, double &hi) {
    double root;
    if (vertical) {
        if (fabs(cc) >= c.ell_ax) return false;                                // S4:51
        root = sqrt(dmax_std(0.0, (1.0 - cc * cc * c.inv_ax2) / c.ell_c));     // S4:53-54
    } else {
        if (fabs(c.ell_k * cc) >= 1.0) return false;                           // S4:59
        root = c.ell_ax * sqrt(dmax_std(0.0, 1.0 - c.ell_c * cc * cc));        // S4:61-62
    }
    lo = dmax_std(s, -root); // S4:55, 63
    hi = dmin_std(e, root);
    return hi > lo;
}

// running shoelace sum over the vertices of a polygon given in order
struct Shoelace {
    int n = 0;
    double x0 = 0.0, y0 = 0.0, xp = 0.0, yp = 0.0, sum = 0.0;
    __device__ __forceinline__ void add(double x, double y) {
        if (n == 0) {
            x0 = x;
            y0 = y;
        } else
            sum += xp * y - x * yp;
        xp = x;
        yp = y;
        ++n;
    }
};

// Right-hand side of node (gi,gj). Mode 0 is the reference's point indicator (S4:182-191). Mode 1 (north_star's "area
// fractions", not in the reference) weighs f by the share of the node's cell [x_i-h1/2, x_{i+1}-h1/2] x [y_j-h2/2,
// y_{j+1}-h2/2] that lies in D: the area of the polygon whose vertices are the end points of the in-D parts of the cell's
// four edges -- the edge intersections S4:46-76 computes for a and b -- walked counter-clockwise in coordinates relative
// to the node (the cell clipped by the chords of the ellipse, second-order accurate), snapped to 0 / 1 with the
// reference's 1e-9 threshold so that cells wholly inside get exactly f.
__device__ __forceinline__ double rhs_value(const Consts &c, int gi, int gj) {
    const double x_i = c.A1 + gi * c.h1, y_j = c.A2 + gj * c.h2; // S4:159-160
    if (c.rhs_mode == 0) return in_D(c, x_i, y_j) ? c.f_val : 0.0; // S4:189
    const double x_ip = c.A1 + (gi + 1) * c.h1, y_jp = c.A2 + (gj + 1) * c.h2;
    const double xl = x_i - 0.5 * c.h1, xr = x_ip - 0.5 * c.h1, yb = y_j - 0.5 * c.h2, yt = y_jp - 0.5 * c.h2;
    double lo, hi;
    Shoelace q;
    if (seg_ends(c, yb, xl, xr, false, lo, hi)) { // bottom edge, left to right
        q.add(lo - x_i, yb - y_j);
        q.add(hi - x_i, yb - y_j);
    }
    if (seg_ends(c, xr, yb, yt, true, lo, hi)) { // right edge, upwards
        q.add(xr - x_i, lo - y_j);
        q.add(xr - x_i, hi - y_j);
    }
    if (seg_ends(c, yt, xl, xr, false, lo, hi)) { // top edge, right to left
        q.add(hi - x_i, yt - y_j);
        q.add(lo - x_i, yt - y_j);
    }
    if (seg_ends(c, xl, yb, yt, true, lo, hi)) { // left edge, downwards
        q.add(xl - x_i, hi - y_j);
        q.add(xl - x_i, lo - y_j);
    }
    if (q.n == 0) return 0.0;
    q.sum += q.xp * q.y0 - q.x0 * q.yp;
    double frac = (0.5 * q.sum) / (c.h1 * c.h2);
    frac = (fabs(frac - 1.0) < 1e-9) ? 1.0 : (frac < 1e-9 ? 0.0 : frac);
    return c.f_val * frac;
}

// One thread per stored cell (ring and padding included). Writes a, b (block + ring, S4:154-179),
// r = B (S4:182-191, 813) on every cell that is a global interior node -- ring cells owned by a
// neighbour block get the neighbour's values, so no halo exchange is needed before iteration 1 --
// D (S4:556-557) likewise, and w = p = p' = Ap = 0 (S4:805; zero ring = Dirichlet condition S4:444-492).
__global__ void __launch_bounds__(256) setup_kernel(Geom g, Consts c, double *__restrict__ a, double *__restrict__ b,
                                                    double *__restrict__ D, double *__restrict__ r,
                                                    double *__restrict__ w, double *__restrict__ p0,
                                                    double *__restrict__ p1, double *__restrict__ Ap,
                                                    unsigned char *__restrict__ cls) {
    const int lj = blockIdx.x * blockDim.x + threadIdx.x;
    const int li = blockIdx.y;
    if (lj >= g.pitch) return;
    const size_t o = (size_t)li * g.pitch + lj;
    double va = 0.0, vb = 0.0, vD = 1.0, vr = 0.0;
    unsigned vc = CLS_GENERAL | (CLS_GENERAL << 2) | (DCLS_RING << 4);
    if (lj <= g.ny + 1) {
        const int gi = g.gi0 + li, gj = g.gj0 + lj;
        const double la = len_a(c, gi, gj), lb = len_b(c, gi, gj);
        va = coef_of(c, la, c.h2);
        vb = coef_of(c, lb, c.h1);
        const unsigned ca = coef_class(va, c.inv_eps), cb = coef_class(vb, c.inv_eps);
        unsigned cD = DCLS_RING;
        if (gi >= 1 && gi <= g.M - 1 && gj >= 1 && gj <= g.N - 1) {
            const double la_ip = len_a(c, gi + 1, gj), lb_jp = len_b(c, gi, gj + 1);
            vr = rhs_value(c, gi, gj); // S4:32-34, 189
            const double a_ip = coef_of(c, la_ip, c.h2), b_jp = coef_of(c, lb_jp, c.h1);
            vD = (a_ip + va) / c.h1h1 + (b_jp + vb) / c.h2h2; // S4:556-557
            cD = diag_class(c, coef_class(a_ip, c.inv_eps), ca, coef_class(b_jp, c.inv_eps), cb, vD);
        }
        vc = ca | (cb << 2) | (cD << 4);
    }
    if (cls && lj <= g.ny + 1) cls_store(cls, g.cpitch, li, lj, vc);
    a[o] = va;
    b[o] = vb;
    D[o] = vD;
    r[o] = vr;
    if (w) w[o] = 0.0;
    if (p0) p0[o] = 0.0;
    if (p1) p1[o] = 0.0;
    if (Ap) Ap[o] = 0.0;
}

// D from caller-supplied a, b (operator entry points only: the solver's D comes from setup_kernel).
__global__ void __launch_bounds__(256) diag_kernel(Geom g, Consts c, const double *__restrict__ a,
                                                   const double *__restrict__ b, double *__restrict__ D,
                                                   unsigned char *__restrict__ cls) {
    const int lj = blockIdx.x * blockDim.x + threadIdx.x;
    const int li = blockIdx.y;
    if (lj >= g.pitch) return;
    const size_t o = (size_t)li * g.pitch + lj;
    double vD = 1.0;
    unsigned vc = CLS_GENERAL | (CLS_GENERAL << 2) | (DCLS_RING << 4);
    if (lj <= g.ny + 1) {
        const unsigned ca = coef_class(a[o], c.inv_eps), cb = coef_class(b[o], c.inv_eps);
        unsigned cD = DCLS_RING;
        if (li >= 1 && li <= g.nx && lj >= 1 && lj <= g.ny) {
            const double a_ip = a[o + g.pitch], b_jp = b[o + 1];
            vD = (a_ip + a[o]) / c.h1h1 + (b_jp + b[o]) / c.h2h2;
            cD = diag_class(c, coef_class(a_ip, c.inv_eps), ca, coef_class(b_jp, c.inv_eps), cb, vD);
        }
        vc = ca | (cb << 2) | (cD << 4);
    }
    D[o] = vD;
    if (lj <= g.ny + 1) cls_store(cls, g.cpitch, li, lj, vc);
}

// ------------------------------------------------------------------------------------------
// Division helpers for the iteration kernels.

// True when v is zero or 2^-840 <= |v| < 2^601. For such p' and r, with coefficients in 2^-50..2^300 and grid
// steps in 2^-100..2^10 (checked on the host, else the SAFE kernels run), every numerator x of a fast division is 0
// or 2^-943 <= |x| <= 2^901: the residual x - h*q of div_const is then exactly representable (needs |x| >= 2^-970)
// and no intermediate overflows, so the two-correction quotient is the IEEE quotient. Values outside the window
// DO occur in this problem -- the iterates decay by about a factor 4 per cell into the penalised exterior until they
// underflow -- so rows that contain one take the IEEE-divide path (a thin ring of the grid). Integer pipe only.
__device__ __forceinline__ bool in_div_range(double v) {
    const unsigned hi = (unsigned)__double2hiint(v) & 0x7fffffffu;
    const unsigned e = hi >> 20;
    return (e - 183u <= 1440u) || ((hi | (unsigned)__double2loint(v)) == 0u);
}
// The same test as a 0/1 word, written without short-circuit operators so that four of them combine into straight-line
// predicate logic (the marching kernels test every p' and r they produce).
__device__ __forceinline__ unsigned out_of_div_range(double v) {
    const unsigned hi = (unsigned)__double2hiint(v) & 0x7fffffffu;
    const unsigned outside = ((hi >> 20) - 183u > 1440u) ? 1u : 0u;
    const unsigned nonzero = ((hi | (unsigned)__double2loint(v)) != 0u) ? 1u : 0u;
    return outside & nonzero;
}

// Correctly rounded x / h for a divisor h that is the same for many x, given rh = RN(1/h) (computed on the host).
// q0 = RN(x*rh) is within 1.5 ulp of x/h; one residual step makes it faithful (<= 0.5 ulp + 2^-53 ulp);
// a second residual step on a faithful quotient with a correctly rounded reciprocal yields RN(x/h)
// (Markstein's theorem). Both residuals x - h*q are exact in an FMA provided they are representable, which is
// what the caller's range test (in_div_range) guarantees; there is no test in here. 5 FP64 ops instead of the
// ~10 + MUFU + branch of the generic divide. pb200_selftest_div checks it against x/h on the device.
__device__ __forceinline__ double div_const_nocheck(double x, double h, double rh) {
    double q = __dmul_rn(x, rh);
    double e = __fma_rn(-h, q, x);
    q = __fma_rn(e, rh, q);
    e = __fma_rn(-h, q, x);
    return __fma_rn(e, rh, q);
}

// r / D as the reference writes it (S4:559), without the generic divide's slow path for r == 0: the
// divide sees 1.0 instead of a zero numerator and the zero (with its sign; D > 0 in practice, and the
// sign of a zero never reaches a non-zero result) is put back afterwards.
__device__ __forceinline__ double precond_div(double r, double D) {
    const bool zero = (r == 0.0);
    const double q = __ddiv_rn(zero ? 1.0 : r, D);
    return (D != 0.0) ? (zero ? ((D < 0.0) ? -r : r) : q) : 0.0;
}

// ------------------------------------------------------------------------------------------
// init: zr_old = (D^-1 r, r) over the interior (S4:820-843). z itself is never stored.
constexpr int INIT_THREADS = 256;

struct InitArgs {
    const double *r, *D;
    double *partials;
    Ctl *ctl;
    Geom g;
    Consts c;
};

__global__ void __launch_bounds__(INIT_THREADS) init_kernel(const InitArgs A) {
    const Geom g = A.g;
    const int lj = 1 + blockIdx.x * blockDim.x + threadIdx.x;
    double acc[1] = {0.0};
    if (lj <= g.ny)
        for (int li = 1 + blockIdx.y; li <= g.nx; li += gridDim.y) {
            const size_t o = (size_t)li * g.pitch + lj;
            const double r = A.r[o];
            acc[0] += precond_div(r, A.D[o]) * r; // S4:559, 594
        }
    double tot[1];
    if (grid_reduce<1, INIT_THREADS>(acc, A.partials, &A.ctl->ticket[1], tot)) {
        const double zr = tot[0] * A.c.h1 * A.c.h2; // S3:191
        A.ctl->red[2] = zr;
        if (A.c.multi == 0) A.ctl->zr_old = zr; // S4:838-843 (multi-GPU: after the NCCL all-reduce, finalize_kernel)
    }
}

// ------------------------------------------------------------------------------------------
// The two kernels of an iteration. Both march down the grid with the same skeleton:
//
// A warp owns 124 adjacent columns and a segment of rows. Lane l holds the four columns s+4l..s+4l+3 of a
// 128-column window [s, s+128), s = 124*warp_column (so every lane's 32 bytes are 32-byte aligned and move
// with one 256-bit load/store). Column s is the left neighbour of the owned range s+1..s+124, column s+125
// the right neighbour; they are recomputed instead of exchanged, so warps never synchronise with each other.
// p' of the current row and the flux through its lower face live in registers; each step brings in row i+1.
//
// The operator is evaluated in flux form, which is the reference's expression tree term by term
// (S4:526-535): Ax = -1/h1 * (Fx(i+1/2) - Fx(i-1/2)) with Fx(i+1/2) = a(i+1,j)*(p(i+1,j)-p(i,j))/h1.
// The flux through a face is the same floating-point expression for the two cells that share the
// face, so it is computed once: Fx(i+1/2) is carried to the next row in a register, Fy(j+1/2) is
// shared inside the lane and passed to the neighbouring lane by one shuffle.
//
//   K1 (k1_kernel): p' = D^-1 r + beta p  (S4:559, 675), stored; Ap = A p' stays in registers and only
//                   its inner product with p' leaves the kernel (S4:866). 25 B per cell: r, p, class in; p' out.
//   K2 (k2_kernel): recomputes Ap = A p' from the stored p' (bit-identical: same expressions), then
//                   w += alpha p', r -= alpha Ap, sum dw^2, z = D^-1 r, sum z r (S4:640-656, 559, 920) and the scalar
//                   tail of the iteration. 41 B per cell: p', w, r, class in; w, r out.
// Neither Ap nor z is ever written to memory; a, b, D are decoded from one class byte per cell, with
// warp-uniform fast paths for rows that lie wholly inside or wholly outside the ellipse.
//
// Divisions: the fast kernels (SAFE = false) use the exact-residual scheme of div_const without per-call
// range tests; instead every p' and every r is range-tested once per row (integer pipe, see in_div_range) and
// a warp whose row holds a value outside the proven window computes that row with IEEE divides.

constexpr int LW = 4;                  // columns per lane
constexpr int K1_WIN = 32 * LW;        // window per warp
constexpr int K1_OWN = K1_WIN - 4;     // owned columns per warp (124: keeps lane offsets 32-byte aligned)
static_assert(K1_OWN == CLS_OWN && K1_WIN == CLS_WIN, "class map records follow the warp windows");
#ifndef PB200_K1_WARPS
#define PB200_K1_WARPS 4
#endif
constexpr int K1_WARPS = PB200_K1_WARPS;
constexpr int K1_THREADS = K1_WARPS * 32;
#ifndef PB200_K1_MINB
#define PB200_K1_MINB 4 // resident CTAs per SM the register allocation is held to (tuned: see DESIGN.md)
#endif
#ifndef PB200_K2_MINB
#define PB200_K2_MINB 4
#endif
#ifndef PB200_STAGES
#define PB200_STAGES 4
#endif
constexpr int K2_THREADS = K1_THREADS;
#ifndef PB200_ROW_UNROLL
#define PB200_ROW_UNROLL 1
#endif
constexpr int ROW_UNROLL = PB200_ROW_UNROLL; // row steps per trip of the marching loops
// Operand rows reach the SM through the bulk-copy engine (cp.async.bulk -> shared memory, completion on an
// mbarrier), STAGES rows ahead of the row being computed, one private ring per warp: the bytes in flight no
// longer depend on registers, scoreboard slots or occupancy.
constexpr int STAGES = PB200_STAGES;
constexpr int ROW_BYTES = K1_WIN * 8;                      // one warp window of one row of doubles
constexpr int K1_STAGE_BYTES = 2 * ROW_BYTES + CLS_WIN;     // r, p, classes of row n
constexpr int K2_STAGE_BYTES = 3 * ROW_BYTES + CLS_WIN;     // p'(i+1), w(i), r(i), classes of row i+1
constexpr int K1_SMEM = K1_WARPS * STAGES * K1_STAGE_BYTES;
constexpr int K2_SMEM = K1_WARPS * STAGES * K2_STAGE_BYTES;

// How the rows are cut into bands (see Part below): nbA bands over rows 1..nxA, then nbB bands over the rest.
struct Bands {
    int nbA, nxA, nbB;
};

struct K1Args {
    const double *r, *p_old;
    const double *a, *b, *D; // full arrays, read only where the class says "general"
    const unsigned char *cls;
    double *p_new;
    double *Ap; // operator entry points only (null in the solver): store A p'
    double *partials; // [nblocks]
    Ctl *ctl;
    Geom g;
    Consts c;
    int wn; // warp windows per row: ceil(ny / 124)
    Bands bands;
    PeerView pv; // multi == 2
};

struct K2Args {
    double *w, *r;
    const double *p;
    const double *a, *b, *D;
    const unsigned char *cls;
    double *partials; // [2][nblocks]
    double *trace;    // [trace_len][3] or null
    Ctl *ctl;
    Geom g;
    Consts c;
    int wn;
    Bands bands;
    PeerView pv; // multi == 2
};

struct V4 {
    double v[LW];
};

// The warp's index in its CTA, taken from lane 0 so that the compiler knows it is the same in all lanes: everything
// derived from it (ring and barrier addresses, the rows and window of the warp's piece, the source addresses of the
// bulk copies) can then live in uniform registers.
__device__ __forceinline__ int uniform_warp_id() { return __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0); }

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void *src, unsigned bytes, unsigned bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
// One pipeline unit: arm the stage's mbarrier with the byte count and start the bulk copies of up to three operand
// rows plus the class record, all from ONE elected lane of a converged warp. Everything here is warp-uniform (see
// uniform_warp_id), so the addresses stay in uniform registers and the whole unit is a dozen instructions.
template <int NROWS>
__device__ __forceinline__ void issue_unit(unsigned dst, unsigned bar, const void *s0, const void *s1, const void *s2,
                                           const void *cls) {
    static_assert(NROWS >= 1 && NROWS <= 3, "one to three operand rows per unit");
    if (NROWS == 1)
        asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\t"
                     "@p mbarrier.arrive.expect_tx.shared::cta.b64 _, [%1], %2;\n\t"
                     "@p cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%3], %5, [%1];\n\t"
                     "@p cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%6], [%4], %7, [%1];\n\t}" ::"r"(dst),
                     "r"(bar), "r"(ROW_BYTES + CLS_WIN), "l"(s0), "l"(cls), "r"(ROW_BYTES), "r"(dst + 2 * ROW_BYTES),
                     "r"(CLS_WIN)
                     : "memory");
    if (NROWS == 2)
        asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\t"
                     "@p mbarrier.arrive.expect_tx.shared::cta.b64 _, [%1], %2;\n\t"
                     "@p cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%3], %6, [%1];\n\t"
                     "@p cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%7], [%4], %6, [%1];\n\t"
                     "@p cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%8], [%5], %9, [%1];\n\t}" ::"r"(dst),
                     "r"(bar), "r"(2 * ROW_BYTES + CLS_WIN), "l"(s0), "l"(s1), "l"(cls), "r"(ROW_BYTES), "r"(dst + ROW_BYTES),
                     "r"(dst + 2 * ROW_BYTES), "r"(CLS_WIN)
                     : "memory");
    if (NROWS == 3)
        asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\t"
                     "@p mbarrier.arrive.expect_tx.shared::cta.b64 _, [%1], %2;\n\t"
                     "@p cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%3], %7, [%1];\n\t"
                     "@p cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%8], [%4], %7, [%1];\n\t"
                     "@p cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%9], [%5], %7, [%1];\n\t"
                     "@p cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%10], [%6], %11, [%1];\n\t}" ::"r"(dst),
                     "r"(bar), "r"(3 * ROW_BYTES + CLS_WIN), "l"(s0), "l"(s1), "l"(s2), "l"(cls), "r"(ROW_BYTES),
                     "r"(dst + ROW_BYTES), "r"(dst + 2 * ROW_BYTES), "r"(dst + 3 * ROW_BYTES), "r"(CLS_WIN)
                     : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_LOOP:\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
                 "@p bra WAIT_DONE;\n\tbra WAIT_LOOP;\n\tWAIT_DONE:\n\t}" ::"r"(bar),
                 "r"(parity)
                 : "memory");
}
// a lane's four doubles of a staged row (two 16-byte shared-memory reads)
__device__ __forceinline__ V4 lds4(const unsigned char *row, int lane) {
    const double2 *q = reinterpret_cast<const double2 *>(row + 32 * lane);
    const double2 lo = q[0], hi = q[1];
    V4 x;
    x.v[0] = lo.x; x.v[1] = lo.y; x.v[2] = hi.x; x.v[3] = hi.y;
    return x;
}

__device__ __forceinline__ V4 ldg4_nc(const double *p, bool ok) { // read-only for the whole kernel
    V4 x;
    x.v[0] = x.v[1] = x.v[2] = x.v[3] = 0.0;
    if (ok) asm("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(x.v[0]), "=d"(x.v[1]), "=d"(x.v[2]), "=d"(x.v[3]) : "l"(p));
    return x;
}
// Store of a lane's four columns under a per-column mask: one 256-bit store when all four are writable (every lane
// but the few on a window or block edge; the mask is loop-invariant, so the test is hoisted), else 8-byte stores.
__device__ __forceinline__ void st4_mask(double *p, const V4 &x, const unsigned mask) {
    if (mask == 15u) {
        asm volatile("st.global.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(p), "d"(x.v[0]), "d"(x.v[1]), "d"(x.v[2]), "d"(x.v[3])
                     : "memory");
    } else {
#pragma unroll
        for (int k = 0; k < LW; ++k)
            if (mask & (1u << k)) p[k] = x.v[k];
    }
}

__device__ __forceinline__ double class_value(unsigned k, double inv_eps) { return (k & 1u) ? inv_eps : 1.0; }

// true when one of the four values is outside the window in which the fast divisions are proven (in_div_range)
__device__ __forceinline__ bool range_bad4(const V4 &x) {
    return (out_of_div_range(x.v[0]) | out_of_div_range(x.v[1]) | out_of_div_range(x.v[2]) | out_of_div_range(x.v[3])) != 0u;
}

// What a lane knows about its place in the march (computed once per kernel).
struct Lane {
    int lane, warp, wcol, s, c0, i_lo, i_hi;
    unsigned own;  // bit k: column c0+k is an interior cell this lane owns (stencil outputs, w, r)
    unsigned pw;   // bit k: column c0+k gets this lane's p' (owned interior cells + ring columns)
};

// Work partition. The rows are cut into bands -- nbA bands of equal height (+-1 row) over rows 1..nxA, then nbB lower
// bands over the rest (Bands) -- and the 124-column warp windows into CTA-windows of K1_WARPS adjacent ones. CTA c takes
// CTA-window c % wcn of band c / wcn, wcn = ceil(windows / K1_WARPS): warps that run side by side read side-by-side
// windows of the same rows (DRAM page locality). CTAs start in index order, so the lower bands come last and keep the
// end of the kernel -- when the SMs run out of CTAs one by one -- short. The host picks the bands (solver.cu:
// march_plan). `descending`: the same bands counted from the last row upwards. K1 sweeps up the rows and K2 sweeps
// back down, so each kernel starts on the ~100 MB the previous one touched last and finds them in the 126 MB L2.
struct Part {
    int nsub;                  // 0 or 1 sub-segments
    int wcol[1], lo[1], hi[1]; // window and inclusive row range
};

__device__ __forceinline__ Part make_part(int cta, int warp, int wn, int nx, const Bands &B, bool descending = false) {
    Part p;
    const int wcn = (wn + K1_WARPS - 1) / K1_WARPS;
    const int b = cta / wcn;
    p.wcol[0] = (cta - b * wcn) * K1_WARPS + warp;
    int lo = 1, hi = 0;
    if (b < B.nbA) {
        lo = 1 + (int)(((long long)B.nxA * b) / B.nbA);
        hi = (int)(((long long)B.nxA * (b + 1)) / B.nbA);
    } else if (b < B.nbA + B.nbB) {
        const int k = b - B.nbA, nxB = nx - B.nxA;
        lo = 1 + B.nxA + (int)(((long long)nxB * k) / B.nbB);
        hi = B.nxA + (int)(((long long)nxB * (k + 1)) / B.nbB);
    }
    p.lo[0] = descending ? nx + 1 - hi : lo;
    p.hi[0] = descending ? nx + 1 - lo : hi;
    p.nsub = (lo <= hi && p.wcol[0] < wn) ? 1 : 0;
    return p;
}

__device__ __forceinline__ Lane make_lane(const Geom &g, int lane, int warp, int wcol, int i_lo, int i_hi) {
    Lane L;
    L.lane = lane;
    L.warp = warp;
    L.wcol = wcol;
    L.s = L.wcol * K1_OWN;
    L.c0 = L.s + LW * L.lane;
    L.i_lo = i_lo;
    L.i_hi = i_hi;
    L.own = 0u;
    L.pw = 0u;
#pragma unroll
    for (int k = 0; k < LW; ++k) {
        const int col = L.c0 + k;
        const bool in_own_range = (col >= L.s + 1) && (col <= L.s + K1_OWN);
        if (in_own_range && col >= 1 && col <= g.ny) L.own |= 1u << k;
        const bool ring = (L.s == 0 && col == 0) || (col == g.ny + 1 && col == L.s + K1_OWN + 1);
        if (col <= g.ny + 1 && (in_own_range || ring)) L.pw |= 1u << k;
    }
    return L;
}

// Exact coefficients of the "general" cells (cut by the ellipse) a row step needs. They are the only operands that do
// not come through the shared-memory ring, so all of them are requested together, first thing in the step, behind one
// warp vote: one exposed memory round trip on the few rows that have such cells, none elsewhere.
struct Exc {
    V4 a;  // a(row_a, c0+k) where its class is general
    V4 bn; // b(row_b, c0+k+1): the cell right of face k (k = 3: the next lane's first column)
    V4 D;  // D(row_d, c0+k)
};

__device__ __forceinline__ bool load_exceptions(const unsigned cls_a, const unsigned cls_b, const unsigned cls_r,
                                                const unsigned cls_d, const int lane, const double *a_row,
                                                const double *b_row, const double *D_row, Exc &e) {
    const unsigned ga = cls_a & 0x02020202u;
    const unsigned gb = (cls_b & 0x08080800u) | ((lane < 31) ? (cls_r & 0x08u) : 0u);
    const unsigned kd = (cls_d >> 4) & 0x0f0f0f0fu;
    const unsigned gd = kd & (kd >> 1) & (kd >> 2) & (kd >> 3) & 0x01010101u; // nibble == DCLS_GENERAL (15)
    if (!__any_sync(0xffffffffu, (ga | gb | gd) != 0u)) return false;
#pragma unroll
    for (int k = 0; k < LW; ++k) {
        e.a.v[k] = (ga & (0x02u << (8 * k))) ? __ldg(a_row + k) : 0.0;
        e.D.v[k] = (gd & (0x01u << (8 * k))) ? __ldg(D_row + k) : 1.0;
    }
#pragma unroll
    for (int k = 0; k < LW - 1; ++k) e.bn.v[k] = (gb & (0x08u << (8 * (k + 1)))) ? __ldg(b_row + k + 1) : 0.0;
    e.bn.v[LW - 1] = (gb & 0x08u) ? __ldg(b_row + LW) : 0.0;
    return true;
}

// Rows of a warp window that lie wholly inside the ellipse (every class byte 0x00: a = b = 1, D = lutD[0]) or wholly
// outside it (every byte 0x85: a = b = 1/eps, D = lutD[8]) are the bulk of the grid. One pair of warp reductions per
// row (row_kind) recognises them, and the row step then skips the class decoding, the search for cut cells and the
// per-cell table look-ups altogether.
constexpr unsigned CLS_ROW_OUTSIDE = 0x85858585u;
enum { ROW_GENERAL = 0, ROW_INSIDE = 1, ROW_OUTSIDE = 2 };

struct ClsSummary {
    unsigned any, all; // OR and AND of the class words of the warp's 32 lanes (warp-uniform)
};
__device__ __forceinline__ ClsSummary cls_summary(unsigned cls) {
    ClsSummary s;
    s.any = __reduce_or_sync(0xffffffffu, cls);
    s.all = __reduce_and_sync(0xffffffffu, cls);
    return s;
}
// kind of the row pair (i, i+1) a row step works on
__device__ __forceinline__ int row_kind(const ClsSummary &c, const ClsSummary &n) {
    const unsigned any = c.any | n.any, all = c.all & n.all;
    return (any == 0u) ? ROW_INSIDE : ((all == CLS_ROW_OUTSIDE && any == CLS_ROW_OUTSIDE) ? ROW_OUTSIDE : ROW_GENERAL);
}

// The numerators of the eight face fluxes of a lane in one row, S4:527/531:
//   nx[k] = a(i+1, c0+k) * (p(i+1, c0+k) - p(i, c0+k))          (x faces above the lane's cells)
//   ny[k] = b(i, c0+k+1) * (p(i, c0+k+1) - p(i, c0+k))          (y faces right of the lane's cells)
// a, b come from the class bytes: cls_n = classes of row i+1, cls_c = row i, cls_r = next lane's row i (general rows
// only). Inside the ellipse a = b = 1: the product with 1.0 is exact and skipped.
__device__ __forceinline__ void flux_numerators(const V4 &pn, const V4 &pc, const double pr, const unsigned cls_n,
                                                const unsigned cls_c, const unsigned cls_r, const int lane,
                                                const double inv_eps, const int kind, const bool have_exc,
                                                const Exc &exc, V4 &nx, V4 &ny) {
#pragma unroll
    for (int k = 0; k < LW; ++k) nx.v[k] = pn.v[k] - pc.v[k];
#pragma unroll
    for (int k = 0; k < LW - 1; ++k) ny.v[k] = pc.v[k + 1] - pc.v[k];
    ny.v[LW - 1] = pr - pc.v[LW - 1];
    if (kind == ROW_INSIDE) return; // a = b = 1
    if (kind == ROW_OUTSIDE) {
#pragma unroll
        for (int k = 0; k < LW; ++k) {
            nx.v[k] = inv_eps * nx.v[k];
            ny.v[k] = inv_eps * ny.v[k];
        }
        return;
    }
    V4 an, bn; // bn[k] = b of the cell right of face k
#pragma unroll
    for (int k = 0; k < LW; ++k) an.v[k] = class_value(cls_n >> (8 * k), inv_eps);
#pragma unroll
    for (int k = 0; k < LW - 1; ++k) bn.v[k] = class_value(cls_c >> (8 * (k + 1) + 2), inv_eps);
    bn.v[LW - 1] = class_value(cls_r >> 2, inv_eps);
    if (have_exc) { // cells cut by the ellipse: the exact values fetched at the top of the step
#pragma unroll
        for (int k = 0; k < LW; ++k)
            if (cls_n & (0x02u << (8 * k))) an.v[k] = exc.a.v[k];
#pragma unroll
        for (int k = 0; k < LW - 1; ++k)
            if (cls_c & (0x08u << (8 * (k + 1)))) bn.v[k] = exc.bn.v[k];
        if ((lane < 31) && (cls_r & 0x08u)) bn.v[LW - 1] = exc.bn.v[LW - 1];
    }
#pragma unroll
    for (int k = 0; k < LW; ++k) {
        nx.v[k] = an.v[k] * nx.v[k];
        ny.v[k] = bn.v[k] * ny.v[k];
    }
}

// x-face numerators only (start of a segment: the face between rows i_lo-1 and i_lo)
__device__ __forceinline__ V4 xface_numerators(const V4 &pc, const V4 &pm, const unsigned cls_c, const double inv_eps,
                                               const double *a_c) {
    V4 nx;
#pragma unroll
    for (int k = 0; k < LW; ++k) {
        double a = class_value(cls_c >> (8 * k), inv_eps);
        if (cls_c & (0x02u << (8 * k))) a = __ldg(a_c + k);
        nx.v[k] = a * (pc.v[k] - pm.v[k]);
    }
    return nx;
}

// D^-1 r for the lane's four cells of one row (S4:559): D and RN(1/D) from the class table; one broadcast
// table read when the whole warp sits in one diagonal class, per-cell reads otherwise; cells next to the
// ellipse boundary (general class) divide by the stored D, passed in D_exc.
template <bool SAFE>
__device__ __forceinline__ V4 precond4(const V4 &r, const unsigned cls, const double2 *s_lut, const V4 &D_exc,
                                       const int kind = ROW_GENERAL) {
    V4 z;
    const bool slow = SAFE || __any_sync(0xffffffffu, range_bad4(r)); // a value outside the proven window: IEEE divides
    const unsigned kd = (cls >> 4) & 0x0f0f0f0fu;
    unsigned k0 = (kind == ROW_INSIDE) ? 0u : 8u;
    bool uniform = kind != ROW_GENERAL;
    if (!uniform) {
        k0 = __shfl_sync(0xffffffffu, kd, 0) & 0xfu;
        uniform = __all_sync(0xffffffffu, kd == k0 * 0x01010101u) && k0 != DCLS_GENERAL;
    }
    if (uniform) {
        const double2 DR = s_lut[k0];
        if (slow) {
#pragma unroll
            for (int k = 0; k < LW; ++k) z.v[k] = precond_div(r.v[k], DR.x);
        } else {
#pragma unroll
            for (int k = 0; k < LW; ++k) z.v[k] = div_const_nocheck(r.v[k], DR.x, DR.y);
        }
        return z;
    }
#pragma unroll
    for (int k = 0; k < LW; ++k) {
        const unsigned kk = (kd >> (8 * k)) & 0xfu;
        const double2 DR = s_lut[kk];
        z.v[k] = slow ? precond_div(r.v[k], DR.x) : div_const_nocheck(r.v[k], DR.x, DR.y);
        if (kk == DCLS_GENERAL) z.v[k] = precond_div(r.v[k], D_exc.v[k]);
    }
    return z;
}

// Stencil of one row for a lane (S4:526-535), given the face numerators. Updates the carried x flux.
template <bool SAFE>
__device__ __forceinline__ V4 stencil_row(const V4 &nx, const V4 &ny, V4 &fxl, const bool slow, const double h1,
                                          const double rh1, const double h2, const double rh2, const double mih1,
                                          const double mih2, const int pow2) {
    V4 fx, fy, out;
    if (SAFE || slow) { // warp-uniform
#pragma unroll
        for (int k = 0; k < LW; ++k) {
            fx.v[k] = __ddiv_rn(nx.v[k], h1);
            fy.v[k] = __ddiv_rn(ny.v[k], h2);
        }
    } else if (pow2 == 0) {
#pragma unroll
        for (int k = 0; k < LW; ++k) {
            fx.v[k] = div_const_nocheck(nx.v[k], h1, rh1);
            fy.v[k] = div_const_nocheck(ny.v[k], h2, rh2);
        }
    } else {
        // A grid step that is a power of two (h1 = 2/M with M = 8192, 32768, ...): rh is exact and, inside the window
        // the row has been tested for, the product x * rh neither overflows nor falls below the normal range, so it IS
        // the IEEE quotient x / h -- one multiplication instead of five operations.
        if (pow2 == 1) {
#pragma unroll
            for (int k = 0; k < LW; ++k) {
                fx.v[k] = __dmul_rn(nx.v[k], rh1);
                fy.v[k] = div_const_nocheck(ny.v[k], h2, rh2);
            }
        } else if (pow2 == 2) {
#pragma unroll
            for (int k = 0; k < LW; ++k) {
                fx.v[k] = div_const_nocheck(nx.v[k], h1, rh1);
                fy.v[k] = __dmul_rn(ny.v[k], rh2);
            }
        } else {
#pragma unroll
            for (int k = 0; k < LW; ++k) {
                fx.v[k] = __dmul_rn(nx.v[k], rh1);
                fy.v[k] = __dmul_rn(ny.v[k], rh2);
            }
        }
    }
    const double fl = shfl_up1(fy.v[LW - 1]); // Fy left of the lane's first column = previous lane's last face
#pragma unroll
    for (int k = 0; k < LW; ++k) {
        const double fym = (k == 0) ? fl : fy.v[k - 1];
        out.v[k] = mih1 * (fx.v[k] - fxl.v[k]) + mih2 * (fy.v[k] - fym);
        fxl.v[k] = fx.v[k];
    }
    return out;
}

__device__ __forceinline__ bool stopped(const Ctl *ctl) { return *reinterpret_cast<const volatile int *>(&ctl->done) != 0; }

// FIRST: iteration 1, p' = z (S4:830). SAFE: IEEE divides everywhere (coefficients or data outside the range in
// which the exact-residual divisions are proven).
template <bool FIRST, bool SAFE>
__global__ void __launch_bounds__(K1_THREADS, PB200_K1_MINB) k1_kernel(const K1Args A) {
    Ctl *ctl = A.ctl;
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ __align__(8) unsigned long long s_bar[K1_WARPS][STAGES];
    __shared__ double2 s_lut[16];
    if (stopped(ctl)) return;
    const Geom g = A.g;
    const int lane_id = threadIdx.x & 31, warp_id = uniform_warp_id();
    if (threadIdx.x < 16) s_lut[threadIdx.x] = make_double2(A.c.lutD[threadIdx.x], A.c.lutR[threadIdx.x]);
    if (lane_id == 0) {
#pragma unroll
        for (int st = 0; st < STAGES; ++st) mbar_init(smem_u32(&s_bar[warp_id][st]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const double h1 = A.c.h1, h2 = A.c.h2, rh1 = A.c.rh1, rh2 = A.c.rh2, mih1 = A.c.mih1, mih2 = A.c.mih2;
    const double inv_eps = A.c.inv_eps;
    const int pow2 = A.c.pow2;
    const double beta = FIRST ? 0.0 : ctl->beta;
    const Part part = make_part(blockIdx.x, warp_id, A.wn, g.nx, A.bands);

    double acc[1] = {0.0};
    if (part.nsub > 0) {
        const Lane L = make_lane(g, lane_id, warp_id, part.wcol[0], part.lo[0], part.hi[0]);
        const size_t P = (size_t)g.pitch;
        const double *ab = A.a + L.c0, *bb = A.b + L.c0, *Db = A.D + L.c0;
        double *po_row = A.p_new + L.c0 + (size_t)(L.i_lo - 1) * P; // p' row being written (advanced by the stores)
        double *apb = A.Ap ? A.Ap + L.c0 : nullptr;

        struct RowIn {
            V4 r, q;       // row n
            unsigned cls;  // class bytes of the lane's four columns in row n (byte k = column c0+k)
        };
        // Operand pipeline: unit t is row (i_lo-1)+t; lane 0 asks the copy engine for r, p and the class record of
        // the warp's window, STAGES units ahead; the bytes land in this warp's ring and flip the stage's mbarrier.
        // Units are issued and consumed strictly in order, so a refill always goes into the stage just read: one
        // cursor (stage offset, barrier offset, parity) serves both, and the sources are running pointers.
        const int nunits = L.i_hi - L.i_lo + 3;
        unsigned char *ring = dyn_smem + (size_t)warp_id * STAGES * K1_STAGE_BYTES;
        const unsigned ring_s = smem_u32(ring), bar_s = smem_u32(&s_bar[warp_id][0]);
        const size_t row_b = P * sizeof(double), cls_b = (size_t)g.cpitch;
        const char *r_nx = reinterpret_cast<const char *>(A.r + L.s) + (size_t)(L.i_lo - 1) * row_b;
        const char *q_nx = reinterpret_cast<const char *>(A.p_old + L.s) + (size_t)(L.i_lo - 1) * row_b;
        const char *c_nx = reinterpret_cast<const char *>(A.cls + cls_index(g.cpitch, 0, L.wcol, 0)) + (size_t)(L.i_lo - 1) * cls_b;
        int to_issue = nunits;
        unsigned cur_off = 0u, cur_bar = 0u, cur_par = 0u; // stage being consumed: byte offset in the ring, barrier offset, parity
        auto issue = [&](const unsigned off, const unsigned bar_off) { // next unit -> the stage at (off, bar_off)
            if (to_issue > 0) {
                if (FIRST) issue_unit<1>(ring_s + off, bar_s + bar_off, r_nx, nullptr, nullptr, c_nx);
                else issue_unit<2>(ring_s + off, bar_s + bar_off, r_nx, q_nx, nullptr, c_nx);
                r_nx += row_b;
                q_nx += row_b;
                c_nx += cls_b;
                --to_issue;
            }
        };
        auto fetch = [&]() {
            RowIn in;
            mbar_wait(bar_s + cur_bar, cur_par);
            const unsigned char *base = ring + cur_off;
            in.r = lds4(base, lane_id);
            if (!FIRST) in.q = lds4(base + ROW_BYTES, lane_id);
            in.cls = *reinterpret_cast<const unsigned *>(base + 2 * ROW_BYTES + 4 * lane_id);
            // Every lane has its copy; the stage may be refilled. The refill is written by the async proxy: without
            // the proxy fence it may overtake the generic-proxy reads above (observed: wrong, run-to-run varying
            // results on long solves).
            __syncwarp();
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            issue(cur_off, cur_bar);
            const bool wrap = cur_bar == 8u * (STAGES - 1);
            cur_off = wrap ? 0u : cur_off + (unsigned)K1_STAGE_BYTES;
            cur_bar = wrap ? 0u : cur_bar + 8u;
            cur_par ^= wrap ? 1u : 0u;
            return in;
        };
#pragma unroll
        for (int t = 0; t < STAGES; ++t) issue((unsigned)(t * K1_STAGE_BYTES), 8u * t);
        auto p_update = [&](const RowIn &in, const V4 &D_exc, const int kind = ROW_GENERAL) {
            V4 z = precond4<SAFE>(in.r, in.cls, s_lut, D_exc, kind); // S4:559
            if (!FIRST) {
#pragma unroll
                for (int k = 0; k < LW; ++k) z.v[k] = z.v[k] + beta * in.q.v[k]; // S4:675 (FIRST: p = z, S4:830)
            }
            return z;
        };

        // rows i_lo-1 and i_lo: p' of both, and the flux through the face between them
        Exc exc;
        RowIn in = fetch();
        load_exceptions(0u, 0u, 0u, in.cls, lane_id, ab, bb, Db + (size_t)(L.i_lo - 1) * P, exc);
        const V4 pm = p_update(in, exc.D);
        if (L.i_lo == 1) st4_mask(po_row, pm, L.pw); // ring row 0
        po_row += P;
        in = fetch();
        load_exceptions(0u, 0u, 0u, in.cls, lane_id, ab, bb, Db + (size_t)L.i_lo * P, exc);
        V4 pc = p_update(in, exc.D);
        unsigned cls_c = in.cls;
        ClsSummary sum_c = cls_summary(cls_c);
        st4_mask(po_row, pc, L.pw);
        bool bad_c = !SAFE && (range_bad4(pm) || range_bad4(pc));
        V4 fxl;
        {
            const V4 nx = xface_numerators(pc, pm, cls_c, inv_eps, ab + (size_t)L.i_lo * P);
            if (SAFE || __any_sync(0xffffffffu, bad_c)) {
#pragma unroll
                for (int k = 0; k < LW; ++k) fxl.v[k] = __ddiv_rn(nx.v[k], h1);
            } else {
#pragma unroll
                for (int k = 0; k < LW; ++k) fxl.v[k] = div_const_nocheck(nx.v[k], h1, rh1);
            }
        }

        // one row: consumes the operands of row i+1, leaves p'(i+1), Fx(i+1/2) and the classes of row i+1
        auto row_step = [&](const int i, const RowIn &cur, const unsigned st_mask) {
            const ClsSummary sum_n = cls_summary(cur.cls);
            const int kind = row_kind(sum_c, sum_n);
            sum_c = sum_n;
            unsigned cls_r = 0u;
            bool have_exc = false;
            if (kind == ROW_GENERAL) {
                cls_r = __shfl_down_sync(0xffffffffu, cls_c, 1);
                have_exc = load_exceptions(cur.cls, cls_c, cls_r, cur.cls, lane_id, ab + (size_t)(i + 1) * P,
                                           bb + (size_t)i * P, Db + (size_t)(i + 1) * P, exc);
            }
            const V4 pn = p_update(cur, exc.D, kind);
            po_row += P;
            st4_mask(po_row, pn, st_mask);
            const bool bad_n = !SAFE && range_bad4(pn);
            const bool slow = __any_sync(0xffffffffu, bad_c | bad_n);
            bad_c = bad_n;
            const double pr = shfl_dn1(pc.v[0]); // right neighbour of the lane's last column
            V4 nx, ny;
            flux_numerators(pn, pc, pr, cur.cls, cls_c, cls_r, lane_id, inv_eps, kind, have_exc, exc, nx, ny);
            const V4 out = stencil_row<SAFE>(nx, ny, fxl, slow, h1, rh1, h2, rh2, mih1, mih2, pow2);
            if (apb) st4_mask(apb + (size_t)i * P, out, L.own);
#pragma unroll
            for (int k = 0; k < LW; ++k) acc[0] += ((L.own >> k) & 1u) ? out.v[k] * pc.v[k] : 0.0; // S4:594
            pc = pn;
            cls_c = cur.cls;
        };

        const unsigned last = (L.i_hi == g.nx) ? L.pw : 0u; // row i_hi+1 belongs to the next segment unless it is ring row nx+1
#pragma unroll ROW_UNROLL
        for (int i = L.i_lo; i < L.i_hi; ++i) row_step(i, fetch(), L.pw);
        row_step(L.i_hi, fetch(), last);
    }

    double tot[1];
    if (grid_reduce<1, K1_THREADS>(acc, A.partials, &ctl->ticket[0], tot)) {
        ctl->red[0] = tot[0] * h1 * h2; // S3:191: sum * h1 * h2
        ctl->iter = ctl->iter + 1;      // S4:847-848: k
        if (A.c.multi == 0) ctl->denom = ctl->red[0];
        if (A.c.multi == 2) { // S4:872: the sum over ranks, through peer memory
            const int off[1] = {MAIL_DEN};
            double sum[1];
            if (peer_allreduce(A.pv, MAIL_FLAG1, (unsigned long long)ctl->iter, 1, off, &ctl->red[0], sum)) ctl->denom = sum[0];
            else {
                ctl->status = STOP_COMM_TIMEOUT;
                __threadfence();
                ctl->done = 1;
            }
        }
    }
}

// The scalar tail of one iteration (S4:893-906, 926-932), run by exactly one thread.
__device__ __forceinline__ void finalize_iteration(Ctl *ctl, const Consts &c, double *trace, double alpha, double diff2,
                                                   double zr_new) {
    const long long k = ctl->iter;
    const double diff = (c.norm_mode == 0) ? sqrt(diff2 * c.h1 * c.h2) : sqrt(diff2); // S3:449 | S0:154
    ctl->diff = diff;
    ctl->alpha = alpha;
    const bool conv = diff < c.delta; // S4:899
    if (trace && k <= c.trace_len) {
        trace[3 * (k - 1) + 0] = alpha;
        trace[3 * (k - 1) + 1] = 0.0;
        trace[3 * (k - 1) + 2] = diff;
    }
    if (conv && ctl->conv_iter == 0) ctl->conv_iter = (int)k;
    if (conv && !c.fixed) {
        ctl->status = STOP_CONVERGED;
        __threadfence();
        ctl->done = 1;
        return;
    }
    const double beta = zr_new / ctl->zr_old; // S4:931
    ctl->beta = beta;
    ctl->zr_old = zr_new; // S4:932
    if (trace && k <= c.trace_len) trace[3 * (k - 1) + 1] = beta;
    if (k >= c.max_iter) { // S4:847: loop bound
        ctl->status = STOP_MAX_ITER;
        __threadfence();
        ctl->done = 1;
    }
}

template <bool SAFE>
__global__ void __launch_bounds__(K2_THREADS, PB200_K2_MINB) k2_kernel(const K2Args A) {
    Ctl *ctl = A.ctl;
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ __align__(8) unsigned long long s_bar[K1_WARPS][STAGES];
    __shared__ double2 s_lut[16];
    if (stopped(ctl)) return;
    const Geom g = A.g;
    const int lane_id = threadIdx.x & 31, warp_id = uniform_warp_id();
    if (threadIdx.x < 16) s_lut[threadIdx.x] = make_double2(A.c.lutD[threadIdx.x], A.c.lutR[threadIdx.x]);
    if (lane_id == 0) {
#pragma unroll
        for (int st = 0; st < STAGES; ++st) mbar_init(smem_u32(&s_bar[warp_id][st]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const double denom = (A.c.multi == 1) ? ctl->red[0] : ctl->denom;
    const bool brk = (A.c.norm_mode == 1) ? (denom < 1e-15) : (fabs(denom) < 1e-15); // S0:128 | S4:876
    if (brk) {
        if (blockIdx.x == 0 && threadIdx.x == 0) {
            ctl->denom = denom;
            ctl->status = STOP_DENOM;
            __threadfence();
            ctl->done = 1;
        }
        return;
    }
    const double alpha = ctl->zr_old / denom; // S4:877
    const double h1 = A.c.h1, h2 = A.c.h2, rh1 = A.c.rh1, rh2 = A.c.rh2, mih1 = A.c.mih1, mih2 = A.c.mih2;
    const double inv_eps = A.c.inv_eps;
    const int pow2 = A.c.pow2;

    const Part part = make_part(blockIdx.x, warp_id, A.wn, g.nx, A.bands, true);

    double acc[2] = {0.0, 0.0};
    if (part.nsub > 0) {
        const Lane L = make_lane(g, lane_id, warp_id, part.wcol[0], part.lo[0], part.hi[0]);
        const size_t P = (size_t)g.pitch;
        const double *pb = A.p + L.c0, *ab = A.a + L.c0, *bb = A.b + L.c0, *Db = A.D + L.c0;
        double *w_row = A.w + L.c0 + (size_t)L.i_lo * P, *r_row = A.r + L.c0 + (size_t)L.i_lo * P; // rows being written

        struct RowIn {
            V4 pn;         // p' of row i+1
            unsigned cls;  // classes of row i+1
            V4 w, r;       // row i
        };
        // Operand pipeline (see k1_kernel): unit t = row i_lo+t brings p'(i+1), w(i), r(i) and the classes of row i+1.
        const int nunits = L.i_hi - L.i_lo + 1;
        unsigned char *ring = dyn_smem + (size_t)warp_id * STAGES * K2_STAGE_BYTES;
        const unsigned ring_s = smem_u32(ring), bar_s = smem_u32(&s_bar[warp_id][0]);
        const size_t row_b = P * sizeof(double), cls_b = (size_t)g.cpitch;
        const char *p_nx = reinterpret_cast<const char *>(A.p + L.s) + (size_t)(L.i_lo + 1) * row_b;
        const char *w_nx = reinterpret_cast<const char *>(A.w + L.s) + (size_t)L.i_lo * row_b;
        const char *r_nx = reinterpret_cast<const char *>(A.r + L.s) + (size_t)L.i_lo * row_b;
        const char *c_nx = reinterpret_cast<const char *>(A.cls + cls_index(g.cpitch, 0, L.wcol, 0)) + (size_t)(L.i_lo + 1) * cls_b;
        int to_issue = nunits;
        unsigned cur_off = 0u, cur_bar = 0u, cur_par = 0u;
        auto issue = [&](const unsigned off, const unsigned bar_off) {
            if (to_issue > 0) {
                issue_unit<3>(ring_s + off, bar_s + bar_off, p_nx, w_nx, r_nx, c_nx);
                p_nx += row_b;
                w_nx += row_b;
                r_nx += row_b;
                c_nx += cls_b;
                --to_issue;
            }
        };
        auto fetch = [&]() {
            RowIn in;
            mbar_wait(bar_s + cur_bar, cur_par);
            const unsigned char *base = ring + cur_off;
            in.pn = lds4(base, lane_id);
            in.w = lds4(base + ROW_BYTES, lane_id);
            in.r = lds4(base + 2 * ROW_BYTES, lane_id);
            in.cls = *reinterpret_cast<const unsigned *>(base + 3 * ROW_BYTES + 4 * lane_id);
            __syncwarp(); // see k1_kernel: reads first, proxy fence, then the refill
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            issue(cur_off, cur_bar);
            const bool wrap = cur_bar == 8u * (STAGES - 1);
            cur_off = wrap ? 0u : cur_off + (unsigned)K2_STAGE_BYTES;
            cur_bar = wrap ? 0u : cur_bar + 8u;
            cur_par ^= wrap ? 1u : 0u;
            return in;
        };
#pragma unroll
        for (int t = 0; t < STAGES; ++t) issue((unsigned)(t * K2_STAGE_BYTES), 8u * t);

        // rows i_lo-1 and i_lo of p', the classes of row i_lo, and the flux through the face between them
        const V4 pm = ldg4_nc(pb + (size_t)(L.i_lo - 1) * P, true);
        V4 pc = ldg4_nc(pb + (size_t)L.i_lo * P, true);
        unsigned cls_c = __ldg(reinterpret_cast<const unsigned *>(A.cls + cls_index(g.cpitch, L.i_lo, L.wcol, 4 * lane_id)));
        ClsSummary sum_c = cls_summary(cls_c);
        bool bad_c = !SAFE && (range_bad4(pm) || range_bad4(pc));
        V4 fxl;
        {
            const V4 nx = xface_numerators(pc, pm, cls_c, inv_eps, ab + (size_t)L.i_lo * P);
            if (SAFE || __any_sync(0xffffffffu, bad_c)) {
#pragma unroll
                for (int k = 0; k < LW; ++k) fxl.v[k] = __ddiv_rn(nx.v[k], h1);
            } else {
#pragma unroll
                for (int k = 0; k < LW; ++k) fxl.v[k] = div_const_nocheck(nx.v[k], h1, rh1);
            }
        }

        // peer-memory mode: does this warp's band / window touch a side that has a neighbour block? (warp-uniform)
        const bool halo_out = A.c.multi == 2 &&
                              ((L.i_lo == 1 && A.pv.r_left) || (L.i_hi == g.nx && A.pv.r_right) || (L.wcol == 0 && A.pv.r_down) ||
                               (L.s + K1_OWN >= g.ny && A.pv.r_up));
        auto row_step = [&](const int i, const RowIn &cur) {
            const bool bad_n = !SAFE && range_bad4(cur.pn);
            const bool slow = __any_sync(0xffffffffu, bad_c | bad_n);
            bad_c = bad_n;
            const double pr = shfl_dn1(pc.v[0]);
            const ClsSummary sum_n = cls_summary(cur.cls);
            const int kind = row_kind(sum_c, sum_n);
            sum_c = sum_n;
            unsigned cls_r = 0u;
            Exc exc;
            bool have_exc = false;
            if (kind == ROW_GENERAL) {
                cls_r = __shfl_down_sync(0xffffffffu, cls_c, 1);
                have_exc = load_exceptions(cur.cls, cls_c, cls_r, cls_c, lane_id, ab + (size_t)(i + 1) * P, bb + (size_t)i * P,
                                           Db + (size_t)i * P, exc);
            }
            V4 nx, ny;
            flux_numerators(cur.pn, pc, pr, cur.cls, cls_c, cls_r, lane_id, inv_eps, kind, have_exc, exc, nx, ny);
            const V4 Ap = stencil_row<SAFE>(nx, ny, fxl, slow, h1, rh1, h2, rh2, mih1, mih2, pow2); // the expressions K1 used for (Ap,p)
            V4 wn, rn;
#pragma unroll
            for (int k = 0; k < LW; ++k) {
                wn.v[k] = cur.w.v[k] + alpha * pc.v[k]; // S4:647
                rn.v[k] = cur.r.v[k] - alpha * Ap.v[k]; // S4:648
            }
            const V4 zn = precond4<SAFE>(rn, cls_c, s_lut, exc.D, kind); // S4:559
#pragma unroll
            for (int k = 0; k < LW; ++k) {
                const bool own = (L.own >> k) & 1u;
                const double d = wn.v[k] - cur.w.v[k]; // S4:653
                acc[0] += own ? d * d : 0.0;
                acc[1] += own ? zn.v[k] * rn.v[k] : 0.0; // S4:594
            }
            st4_mask(w_row, wn, L.own);
            st4_mask(r_row, rn, L.own);
            w_row += P;
            r_row += P;
            if (halo_out) { // halo of r (S4:361-398): boundary cells go straight into the neighbours' rings
                if (i == 1 && A.pv.r_left) st4_mask(A.pv.r_left + (size_t)(A.pv.nx_left + 1) * P + L.c0, rn, L.own);
                if (i == g.nx && A.pv.r_right) st4_mask(A.pv.r_right + L.c0, rn, L.own);
#pragma unroll
                for (int k = 0; k < LW; ++k) {
                    if (A.pv.r_down && L.c0 + k == 1 && ((L.own >> k) & 1u))
                        A.pv.r_down[(size_t)i * A.pv.pitch_down + A.pv.ny_down + 1] = rn.v[k];
                    if (A.pv.r_up && L.c0 + k == g.ny && ((L.own >> k) & 1u)) A.pv.r_up[(size_t)i * A.pv.pitch_up] = rn.v[k];
                }
            }
            pc = cur.pn;
            cls_c = cur.cls;
        };

#pragma unroll ROW_UNROLL
        for (int i = L.i_lo; i <= L.i_hi; ++i) row_step(i, fetch());
    }

    double tot[2];
    if (grid_reduce<2, K2_THREADS>(acc, A.partials, &ctl->ticket[1], tot, A.c.multi == 2)) {
        const double zr = tot[1] * A.c.h1 * A.c.h2; // S3:191
        ctl->red[1] = tot[0];
        ctl->red[2] = zr;
        if (A.c.multi == 0) finalize_iteration(ctl, A.c, A.trace, alpha, tot[0], zr);
        if (A.c.multi == 2) { // S4:893, 926: both sums over ranks, through peer memory; every rank then runs the same tail
            const int off[2] = {MAIL_DIF, MAIL_ZR};
            double sum[2];
            if (peer_allreduce(A.pv, MAIL_FLAG2, (unsigned long long)ctl->iter, 2, off, &ctl->red[1], sum))
                finalize_iteration(ctl, A.c, A.trace, alpha, sum[0], sum[1]);
            else {
                ctl->status = STOP_COMM_TIMEOUT;
                __threadfence();
                ctl->done = 1;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// The whole PCG loop (S4:845-943) of a SMALL single-GPU block in one cooperative launch, one thread per cell.
//
// The marching kernels above are built for throughput: a warp walks down its rows, a row step is a dependent sequence of
// some 400 instructions (about 1.6 us for a warp that runs alone), and two dependent kernels per iteration each spend
// about 14 us on launch gap, control-block loads, fence + ticket and the last block's pass. On a block of a few hundred
// rows nothing hides any of that: an iteration costs 27-35 us whatever the size (measured; per-cell kernels in the same
// two-launch scheme: no better). Here every cell has a resident thread that keeps its w, r, p and Ap in registers for
// the whole solve, the two reductions of an iteration are a block sum, one grid-wide barrier and a pass over the block
// partials that EVERY block makes for itself (same values, same order: bit-identical scalars in all threads, so the
// stop decision needs no broadcast), and nothing goes through the control block until the end.
//
// Arithmetic: term by term that of the marching kernels -- p' = r/D + beta p on the cell and its four neighbours
// (S4:559, 675), the four face fluxes and the stencil (S4:526-535) with IEEE divisions (which the constant-divisor
// scheme there reproduces), the updates (S4:640-656) -- so p', w, r are bit-identical to the two-kernel path given the
// same scalars; only the grouping of the partial sums differs. Coefficients come from the full a, b, D arrays.
// Neighbours' r and p are read from global memory past L1 (__ldcg): they were written by other SMs in this launch.
constexpr int CL_BX = 64, CL_BY = 4, CL_THREADS = CL_BX * CL_BY;

struct CellLoopArgs {
    double *w, *r, *p[2];
    const double *a, *b, *D;
    double *partials; // [3][blocks]: (Ap,p) | dw^2 | (z,r)
    double *trace;
    Ctl *ctl;
    Geom g;
    Consts c;
};

// Sum of v over the block, the same value in every thread: fixed warp tree, then the warps in index order.
template <int NV>
__device__ __forceinline__ void block_sum_all(double (&v)[NV], double (*s_w)[CL_THREADS / 32]) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads(); // s_w may still be read from the previous use
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        const double w = warp_sum(v[k]);
        if (lane == 0) s_w[k][warp] = w;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < CL_THREADS / 32; ++w) s += s_w[k][w];
        v[k] = s;
    }
}

// Grid-wide sums: every block leaves its partial, the grid meets, every block adds all partials up in index order.
template <int NV>
__device__ __forceinline__ void grid_sum_all(double (&v)[NV], double *partials, const unsigned nblocks, const unsigned bid,
                                             double (*s_w)[CL_THREADS / 32], cooperative_groups::grid_group &grid) {
    block_sum_all<NV>(v, s_w);
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < NV; ++k) __stcg(&partials[(size_t)k * nblocks + bid], v[k]);
    }
    grid.sync();
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        double s = 0.0;
        for (unsigned i = threadIdx.x; i < nblocks; i += CL_THREADS) s += __ldcg(&partials[(size_t)k * nblocks + i]);
        v[k] = s;
    }
    block_sum_all<NV>(v, s_w);
}

// NC cells per thread (rows li, li + CL_BY, ...): their w, r, p and Ap stay in registers, coefficients and divisors are
// re-read every iteration through the read-only path (they sit in L1). The host picks the smallest NC for which the
// grid is resident (solver.cu: plan_launch).
template <int NC>
__global__ void __launch_bounds__(CL_THREADS, NC <= 2 ? 4 : 3) cell_loop_kernel(const CellLoopArgs A) {
    __shared__ double s_w[2][CL_THREADS / 32];
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    const Geom g = A.g;
    const Consts &c = A.c;
    const unsigned nblocks = gridDim.x * gridDim.y, bid = blockIdx.y * gridDim.x + blockIdx.x;
    const int lj = 1 + (int)blockIdx.x * CL_BX + (int)(threadIdx.x % CL_BX);
    const int li0 = 1 + (int)blockIdx.y * (CL_BY * NC) + (int)(threadIdx.x / CL_BX);
    const size_t P = (size_t)g.pitch;
    bool own[NC];
    size_t o[NC];
    double w[NC], r[NC], pc[NC], Ap[NC];
#pragma unroll
    for (int q = 0; q < NC; ++q) {
        const int li = li0 + q * CL_BY;
        own[q] = li <= g.nx && lj <= g.ny;
        o[q] = own[q] ? (size_t)li * P + lj : P + 1; // idle slots look at cell (1,1), harmlessly, and store nothing
        w[q] = A.w[o[q]]; // S4:805-813: w = 0, r = B, p = 0 (setup_kernel / reset_kernel)
        r[q] = A.r[o[q]];
        pc[q] = 0.0;
        Ap[q] = 0.0;
    }
    double zr_old = __ldcg(&A.ctl->zr_old); // init_kernel (S4:820-843)
    double alpha = 0.0, beta = 0.0, diff = 0.0, denom = 0.0;
    long long k = 0;
    int status = STOP_MAX_ITER, conv_iter = 0;
    for (;;) {
        ++k; // S4:847-848
        const double *p_old = A.p[(k - 1) & 1];
        double *p_new = A.p[k & 1];
        const bool first = k == 1;
        // ---- K1: p' on the cell and its neighbours, Ap, (Ap, p')
        double s1[1] = {0.0};
#pragma unroll
        for (int q = 0; q < NC; ++q) {
            const size_t oc = o[q];
            const double zl = precond_div(__ldcg(&A.r[oc - 1]), __ldg(&A.D[oc - 1]));
            const double zrr = precond_div(__ldcg(&A.r[oc + 1]), __ldg(&A.D[oc + 1]));
            const double zd = precond_div(__ldcg(&A.r[oc - P]), __ldg(&A.D[oc - P]));
            const double zu = precond_div(__ldcg(&A.r[oc + P]), __ldg(&A.D[oc + P]));
            const double zc = precond_div(r[q], __ldg(&A.D[oc])); // S4:559
            const double pl = first ? zl : zl + beta * __ldcg(&p_old[oc - 1]);
            const double pr = first ? zrr : zrr + beta * __ldcg(&p_old[oc + 1]);
            const double pd = first ? zd : zd + beta * __ldcg(&p_old[oc - P]);
            const double pu = first ? zu : zu + beta * __ldcg(&p_old[oc + P]);
            const double pq = first ? zc : zc + beta * pc[q]; // S4:675 (first: p = z, S4:830)
            pc[q] = pq;
            if (own[q]) __stcg(&p_new[oc], pq);
            const double fx_hi = __ddiv_rn(__ldg(&A.a[oc + P]) * (pu - pq), c.h1), fx_lo = __ddiv_rn(__ldg(&A.a[oc]) * (pq - pd), c.h1); // S4:527
            const double fy_hi = __ddiv_rn(__ldg(&A.b[oc + 1]) * (pr - pq), c.h2), fy_lo = __ddiv_rn(__ldg(&A.b[oc]) * (pq - pl), c.h2); // S4:531
            Ap[q] = c.mih1 * (fx_hi - fx_lo) + c.mih2 * (fy_hi - fy_lo); // S4:526-535
            s1[0] += own[q] ? Ap[q] * pq : 0.0;                          // S4:594
        }
        grid_sum_all<1>(s1, A.partials, nblocks, bid, s_w, grid);
        denom = s1[0] * c.h1 * c.h2; // S3:191
        const bool brk = (c.norm_mode == 1) ? (denom < 1e-15) : (fabs(denom) < 1e-15); // S0:128 | S4:876
        if (brk) {
            status = STOP_DENOM;
            break;
        }
        // ---- K2: the updates and the two sums
        alpha = zr_old / denom; // S4:877
        double s2[2] = {0.0, 0.0};
#pragma unroll
        for (int q = 0; q < NC; ++q) {
            const double wn = w[q] + alpha * pc[q]; // S4:647
            const double rn = r[q] - alpha * Ap[q]; // S4:648
            const double d = wn - w[q];             // S4:653
            w[q] = wn;
            r[q] = rn;
            if (own[q]) __stcg(&A.r[o[q]], rn); // the neighbours' next p' needs it
            s2[0] += own[q] ? d * d : 0.0;
            s2[1] += own[q] ? precond_div(rn, __ldg(&A.D[o[q]])) * rn : 0.0; // S4:559, 594
        }
        grid_sum_all<2>(s2, A.partials + nblocks, nblocks, bid, s_w, grid);
        // ---- the scalar tail (finalize_iteration), in every thread
        const double zr_new = s2[1] * c.h1 * c.h2;
        diff = (c.norm_mode == 0) ? sqrt(s2[0] * c.h1 * c.h2) : sqrt(s2[0]); // S3:449 | S0:154
        const bool conv = diff < c.delta;                                     // S4:899
        const bool tr = A.trace && k <= c.trace_len && bid == 0 && threadIdx.x == 0;
        if (tr) {
            A.trace[3 * (k - 1) + 0] = alpha;
            A.trace[3 * (k - 1) + 1] = 0.0;
            A.trace[3 * (k - 1) + 2] = diff;
        }
        if (conv && conv_iter == 0) conv_iter = (int)k;
        if (conv && !c.fixed) {
            status = STOP_CONVERGED;
            break;
        }
        beta = zr_new / zr_old; // S4:931
        zr_old = zr_new;        // S4:932
        if (tr) A.trace[3 * (k - 1) + 1] = beta;
        if (k >= c.max_iter) { // S4:847: loop bound
            status = STOP_MAX_ITER;
            break;
        }
    }
#pragma unroll
    for (int q = 0; q < NC; ++q)
        if (own[q]) A.w[o[q]] = w[q];
    if (bid == 0 && threadIdx.x == 0) {
        Ctl *ctl = A.ctl;
        ctl->zr_old = zr_old;
        ctl->denom = denom;
        ctl->alpha = alpha;
        ctl->beta = beta;
        ctl->diff = diff;
        ctl->iter = k;
        ctl->status = status;
        ctl->conv_iter = conv_iter;
        __threadfence();
        ctl->done = 1;
    }
}

// The same loop for blocks too large to keep in registers but small enough to live in the 126 MB L2: every block walks
// over its tiles of CL_BX x CL_BY cells (tile t of block b: b, b + blocks, ...) twice per iteration, state in global
// memory, Ap recomputed in the second pass from the stored p' (the expressions of the first). Same arithmetic, same
// fixed summation order (tile order inside a thread, then the block and grid sums above).
__device__ __forceinline__ double cell_Ap(const CellLoopArgs &A, const double *p, const size_t o, const size_t P, const double pc) {
    const Consts &c = A.c;
    const double pl = __ldcg(&p[o - 1]), pr = __ldcg(&p[o + 1]), pd = __ldcg(&p[o - P]), pu = __ldcg(&p[o + P]);
    const double fx_hi = __ddiv_rn(__ldg(&A.a[o + P]) * (pu - pc), c.h1), fx_lo = __ddiv_rn(__ldg(&A.a[o]) * (pc - pd), c.h1); // S4:527
    const double fy_hi = __ddiv_rn(__ldg(&A.b[o + 1]) * (pr - pc), c.h2), fy_lo = __ddiv_rn(__ldg(&A.b[o]) * (pc - pl), c.h2); // S4:531
    return c.mih1 * (fx_hi - fx_lo) + c.mih2 * (fy_hi - fy_lo);                                                                  // S4:526-535
}

__global__ void __launch_bounds__(CL_THREADS, 4) cell_loop_stream_kernel(const CellLoopArgs A) {
    __shared__ double s_w[2][CL_THREADS / 32];
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    const Geom g = A.g;
    const Consts &c = A.c;
    const unsigned nblocks = gridDim.x, bid = blockIdx.x;
    const int tiles_x = (g.ny + CL_BX - 1) / CL_BX, tiles_y = (g.nx + CL_BY - 1) / CL_BY, ntiles = tiles_x * tiles_y;
    const int tx = (int)(threadIdx.x % CL_BX), ty = (int)(threadIdx.x / CL_BX);
    const size_t P = (size_t)g.pitch;
    double zr_old = __ldcg(&A.ctl->zr_old); // init_kernel (S4:820-843)
    double alpha = 0.0, beta = 0.0, diff = 0.0, denom = 0.0;
    long long k = 0;
    int status = STOP_MAX_ITER, conv_iter = 0;
    for (;;) {
        ++k; // S4:847-848
        const double *p_old = A.p[(k - 1) & 1];
        double *p_new = A.p[k & 1];
        const bool first = k == 1;
        // ---- K1: p' = D^-1 r + beta p (S4:559, 675; first: p = z, S4:830), all cells of the block's tiles
        for (int t = (int)bid; t < ntiles; t += (int)nblocks) {
            const int li = 1 + (t / tiles_x) * CL_BY + ty, lj = 1 + (t % tiles_x) * CL_BX + tx;
            if (li <= g.nx && lj <= g.ny) {
                const size_t o = (size_t)li * P + lj;
                const double z = precond_div(__ldcg(&A.r[o]), __ldg(&A.D[o]));
                __stcg(&p_new[o], first ? z : z + beta * __ldcg(&p_old[o]));
            }
        }
        grid.sync(); // the stencil needs the neighbours' p'
        double s1[1] = {0.0};
        for (int t = (int)bid; t < ntiles; t += (int)nblocks) {
            const int li = 1 + (t / tiles_x) * CL_BY + ty, lj = 1 + (t % tiles_x) * CL_BX + tx;
            if (li <= g.nx && lj <= g.ny) {
                const size_t o = (size_t)li * P + lj;
                const double pc = __ldcg(&p_new[o]);
                s1[0] += cell_Ap(A, p_new, o, P, pc) * pc; // S4:594
            }
        }
        grid_sum_all<1>(s1, A.partials, nblocks, bid, s_w, grid);
        denom = s1[0] * c.h1 * c.h2; // S3:191
        const bool brk = (c.norm_mode == 1) ? (denom < 1e-15) : (fabs(denom) < 1e-15); // S0:128 | S4:876
        if (brk) {
            status = STOP_DENOM;
            break;
        }
        // ---- K2: the updates and the two sums
        alpha = zr_old / denom; // S4:877
        double s2[2] = {0.0, 0.0};
        for (int t = (int)bid; t < ntiles; t += (int)nblocks) {
            const int li = 1 + (t / tiles_x) * CL_BY + ty, lj = 1 + (t % tiles_x) * CL_BX + tx;
            if (li <= g.nx && lj <= g.ny) {
                const size_t o = (size_t)li * P + lj;
                const double pc = __ldcg(&p_new[o]);
                const double Ap = cell_Ap(A, p_new, o, P, pc);
                const double w0 = __ldcg(&A.w[o]);
                const double wn = w0 + alpha * pc;               // S4:647
                const double rn = __ldcg(&A.r[o]) - alpha * Ap;  // S4:648
                const double d = wn - w0;                        // S4:653
                __stcg(&A.w[o], wn);
                __stcg(&A.r[o], rn);
                s2[0] += d * d;
                s2[1] += precond_div(rn, __ldg(&A.D[o])) * rn; // S4:559, 594
            }
        }
        grid_sum_all<2>(s2, A.partials + nblocks, nblocks, bid, s_w, grid);
        // ---- the scalar tail (finalize_iteration), in every thread
        const double zr_new = s2[1] * c.h1 * c.h2;
        diff = (c.norm_mode == 0) ? sqrt(s2[0] * c.h1 * c.h2) : sqrt(s2[0]); // S3:449 | S0:154
        const bool conv = diff < c.delta;                                     // S4:899
        const bool tr = A.trace && k <= c.trace_len && bid == 0 && threadIdx.x == 0;
        if (tr) {
            A.trace[3 * (k - 1) + 0] = alpha;
            A.trace[3 * (k - 1) + 1] = 0.0;
            A.trace[3 * (k - 1) + 2] = diff;
        }
        if (conv && conv_iter == 0) conv_iter = (int)k;
        if (conv && !c.fixed) {
            status = STOP_CONVERGED;
            break;
        }
        beta = zr_new / zr_old; // S4:931
        zr_old = zr_new;        // S4:932
        if (tr) A.trace[3 * (k - 1) + 1] = beta;
        if (k >= c.max_iter) { // S4:847: loop bound
            status = STOP_MAX_ITER;
            break;
        }
    }
    if (bid == 0 && threadIdx.x == 0) {
        Ctl *ctl = A.ctl;
        ctl->zr_old = zr_old;
        ctl->denom = denom;
        ctl->alpha = alpha;
        ctl->beta = beta;
        ctl->diff = diff;
        ctl->iter = k;
        ctl->status = status;
        ctl->conv_iter = conv_iter;
        __threadfence();
        ctl->done = 1;
    }
}

// Multi-GPU: after the all-reduce of red[1..2] (or red[2] at init) every rank runs the same scalar tail.
__global__ void finalize_kernel(Ctl *ctl, Consts c, double *trace, int init) {
    if (stopped(ctl)) return;
    if (init) {
        ctl->zr_old = ctl->red[2];
        return;
    }
    const double denom = ctl->red[0];
    ctl->denom = denom;
    finalize_iteration(ctl, c, trace, ctl->zr_old / denom, ctl->red[1], ctl->red[2]);
}

// ------------------------------------------------------------------------------------------
// norms of the result (reporting only): sum w^2, sum r^2, max |w| over the interior, and the distance to the analytic
// solution u = f (1 - (x/ax)^2 - (k y)^2) / (2/ax^2 + 2 k^2) of the problem on D (R4 p.1-2: u = (1 - x^2 - 4y^2)/10) over the
// nodes inside D: sum (w-u)^2 and max |w-u|. The reference states u but frees w unseen (S4:946-954).
// out5 = { sum w^2, sum r^2, sum (w-u)^2 | max |w|, max |w-u| }: the first three add over ranks, the last two take the max.
__global__ void __launch_bounds__(256) norms_kernel(Geom g, Consts c, const double *__restrict__ w, const double *__restrict__ r,
                                                    double *partials, unsigned int *ticket, double *out5) {
    __shared__ double s_max[2][8];
    double acc[3] = {0.0, 0.0, 0.0};
    double mx = 0.0, ex = 0.0;
    const int lj = 1 + blockIdx.x * blockDim.x + threadIdx.x;
    if (lj <= g.ny) {
        const double y = c.A2 + (g.gj0 + lj) * c.h2; // S4:160
        for (int li = 1 + blockIdx.y; li <= g.nx; li += gridDim.y) {
            const size_t o = (size_t)li * g.pitch + lj;
            const double vw = w[o], vr = r[o];
            acc[0] += vw * vw;
            acc[1] += vr * vr;
            mx = fmax(mx, fabs(vw));
            const double x = c.A1 + (g.gi0 + li) * c.h1; // S4:159
            if (in_D(c, x, y)) {
                const double e = fabs(vw - c.u_scale * (1.0 - x * x * c.inv_ax2 - c.ell_c * y * y));
                acc[2] += e * e;
                ex = fmax(ex, e);
            }
        }
    }
    // block maxima, then ONE atomic per block and value (bit pattern of a non-negative double orders like an integer)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mx = fmax(mx, __shfl_down_sync(0xffffffffu, mx, o));
        ex = fmax(ex, __shfl_down_sync(0xffffffffu, ex, o));
    }
    if ((threadIdx.x & 31) == 0) {
        s_max[0][threadIdx.x >> 5] = mx;
        s_max[1][threadIdx.x >> 5] = ex;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < 8; ++k) {
            mx = fmax(mx, s_max[0][k]);
            ex = fmax(ex, s_max[1][k]);
        }
        atomicMax(reinterpret_cast<unsigned long long *>(&out5[3]), (unsigned long long)__double_as_longlong(mx));
        atomicMax(reinterpret_cast<unsigned long long *>(&out5[4]), (unsigned long long)__double_as_longlong(ex));
    }
    double tot[3];
    if (grid_reduce<3, 256>(acc, partials, ticket, tot)) {
        out5[0] = tot[0];
        out5[1] = tot[1];
        out5[2] = tot[2];
    }
}

// Back to the initial state of a solve without recomputing the coefficients: w = p = 0, r = B (S4:805, 813, 182-191).
__global__ void __launch_bounds__(256) reset_kernel(Geom g, Consts c, double *__restrict__ r, double *__restrict__ w,
                                                    double *__restrict__ p0, double *__restrict__ p1) {
    const int lj = blockIdx.x * blockDim.x + threadIdx.x;
    const int li = blockIdx.y;
    if (lj >= g.pitch) return;
    const size_t o = (size_t)li * g.pitch + lj;
    double vr = 0.0;
    const int gi = g.gi0 + li, gj = g.gj0 + lj;
    if (lj <= g.ny + 1 && gi >= 1 && gi <= g.M - 1 && gj >= 1 && gj <= g.N - 1) {
        vr = rhs_value(c, gi, gj); // S4:32-34, 189
    }
    r[o] = vr;
    w[o] = 0.0;
    p0[o] = 0.0;
    p1[o] = 0.0;
}

// Generic weighted inner product (operator entry point; the solver's dots are fused into K1/K2).
__global__ void __launch_bounds__(256) dot_kernel(Geom g, const double *__restrict__ u, const double *__restrict__ v,
                                                  double *partials, unsigned int *ticket, double *out) {
    double acc[1] = {0.0};
    const int lj = 1 + blockIdx.x * blockDim.x + threadIdx.x;
    if (lj <= g.ny)
        for (int li = 1 + blockIdx.y; li <= g.nx; li += gridDim.y) {
            const size_t o = (size_t)li * g.pitch + lj;
            acc[0] += u[o] * v[o];
        }
    double tot[1];
    if (grid_reduce<1, 256>(acc, partials, ticket, tot)) out[0] = tot[0];
}

// ------------------------------------------------------------------------------------------
// self-test of div_const_nocheck against the IEEE divide.
__global__ void selftest_div_kernel(double h, double rh, long long n, unsigned long long seed,
                                    unsigned long long *mismatches) {
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    unsigned long long bad = 0;
    for (long long i = tid; i < n; i += stride) {
        unsigned long long z = seed + 0x9E3779B97F4A7C15ull * (unsigned long long)(i + 1);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        z ^= z >> 31;
        // (1) random mantissa, exponent anywhere in the window the kernels use the fast division in (2^-940 .. 2^900)
        const int ex = (int)((z >> 52) % 1841) - 940;
        double x = ldexp(1.0 + (double)(z & ((1ull << 52) - 1)) * 0x1p-52, ex);
        if (z >> 63) x = -x;
        if (__double_as_longlong(div_const_nocheck(x, h, rh)) != __double_as_longlong(__ddiv_rn(x, h))) ++bad;
        // (2) adversarial: x next to h*m, so that x/h sits next to a representable value or a midpoint
        const double m = 1.0 + (double)((z >> 7) & ((1ull << 52) - 1)) * 0x1p-52;
        const double hm = __dmul_rn(h, m);
        const long long bits = __double_as_longlong(hm);
#pragma unroll
        for (int d = -2; d <= 2; ++d) {
            const double xa = __longlong_as_double(bits + d);
            if (__double_as_longlong(div_const_nocheck(xa, h, rh)) != __double_as_longlong(__ddiv_rn(xa, h))) ++bad;
        }
        // (3) x/h near a midpoint: x = h*(m + half an ulp) rounded either way
        const double mh = __fma_rn(h, m, __dmul_rn(h, 0x1p-53));
        const long long b2 = __double_as_longlong(mh);
#pragma unroll
        for (int d = -1; d <= 1; ++d) {
            const double xa = __longlong_as_double(b2 + d);
            if (__double_as_longlong(div_const_nocheck(xa, h, rh)) != __double_as_longlong(__ddiv_rn(xa, h))) ++bad;
        }
    }
    if (bad) atomicAdd(mismatches, bad);
}

} // namespace pb200
