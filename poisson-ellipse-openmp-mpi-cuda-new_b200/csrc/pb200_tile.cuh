// pb200_tile.cuh -- the whole PCG loop (S4:845-943) of a block that fits in the GPU's shared memory, in ONE cooperative
// launch per rank, on one GPU or on several ("device-resident persistent solver").
//
// One CTA per SM owns a rectangular tile of the block for the whole solve. The tile's p' (with a one-cell halo), r and
// A p' live in the CTA's shared memory (up to 227 KB on sm_100a: 1.3 M cells per GPU), the class bytes of its cells too;
// only w stays in global memory (it is touched once per iteration, by its owner alone). Per iteration:
//
//   phase A  p' = D^-1 r + beta p' (S4:559, 675) on the tile AND on its halo ring: the halo cells' r is read from global
//            memory (written there by the neighbouring tile, or by the neighbouring GPU through NVLink, during the previous
//            iteration), their p' is this CTA's own previous halo value -- bit-identical to the owner's, so p' is never
//            exchanged, only r (as in the marching kernels);
//   phase B  A p' from shared memory (S4:526-535), kept in shared memory; (A p', p')     -> all-reduce 1
//   phase C  w += alpha p', r -= alpha A p', sum dw^2, z = D^-1 r, sum z r (S4:640-656, 920); the tile's edge cells' r go to
//            global memory and, on a block edge with a neighbour GPU, straight into that GPU's ring    -> all-reduce 2
//   the scalar tail of the iteration (S4:893-906, 926-932) in every thread, identically.
//
// A thread owns a column of the tile and marches down its rows: p' of the rows above and below and the flux through the
// upper face stay in registers; the columns left and right come from shared memory.
//
// An all-reduce is the grid barrier, built from "tagged words": a double travels as two 64-bit words, each carrying half
// of its bits and the number of the reduction, so a reader that sees the tag has the payload -- no flag, no fence and no
// atomic on the value path, one memory round trip per hop. Every CTA posts its partial sums as tagged words; CTA 0 gathers
// them (one polling thread per word), adds them in CTA order and posts the GPU's totals, again as tagged words, into the
// mailbox of EVERY GPU (NVLink stores into peer memory; its own included); one warp of every CTA of every GPU polls its own
// GPU's mailbox and adds the totals in rank order -- the same bits everywhere. Replaces exchange_halos_2d_gpu (S4:331-500)
// and the three MPI_Allreduce per iteration (S4:872, 893, 926) without a kernel boundary anywhere in the loop.
//
// Arithmetic: term by term that of the other kernels -- the expression trees of S4:526-535, 559, 647-656, 675, one rounding
// per operation, divisions by the grid steps and by the class-table diagonals through the exact-residual scheme when every
// operand of the warp lies in the window it is proven for (in_div_range), IEEE divisions otherwise.
#pragma once
#include "pb200_kernels.cuh"

namespace pb200 {

constexpr int TL_THREADS = 1024;
constexpr int TL_NH = 2;           // halo cells per thread at most: 2 (R + C) <= TL_NH * TL_THREADS
constexpr int TL_MAX_BLOCKS = 192; // one CTA per SM
constexpr int TL_MAXC = 1024;      // a tile is at most this wide: one thread per column
constexpr int TL_SLOT_WORDS = 4 * TL_MAX_BLOCKS; // CTA partials: [value k < 2][half][CTA]
constexpr int TL_MAIL_WORDS = 64;  // per GPU: [parity][rank < 8][value k < 2][half]  (= MAIL_DOUBLES: the same 512 bytes)
static_assert(TL_MAIL_WORDS == MAIL_DOUBLES, "the resident loop kernel uses the marching kernels' mailbox");

struct TileArgs {
    double *w, *r, *p[2];
    const double *a, *b, *D;    // full arrays: read only where the class byte says "general"
    const unsigned char *cls;
    unsigned long long *slots;  // TL_SLOT_WORDS tagged words, zeroed before every solve
    double *trace;
    Ctl *ctl;
    Geom g;
    Consts c;
    int R, C, tc;               // tile height and width (the last tiles may be smaller), tiles per row of tiles
    int safe;                   // IEEE divisions everywhere (coefficients outside the proven window)
    PeerView pv;                // mail[q]: every rank's mailbox (one GPU: nranks = 1, its own); r_*: c.multi == 2
};

__host__ __device__ inline size_t tile_smem_bytes(int R, int C) {
    const size_t halo = (size_t)(R + 2) * (size_t)(C + 2), own = (size_t)R * (size_t)C;
    return (halo + 2 * own) * sizeof(double) + ((halo + 15) / 16) * 16 + ((own + 15) / 16) * 16;
}

__device__ __forceinline__ double tl_coef(unsigned k, double inv_eps, const double *cell) {
    return (k == 0u) ? 1.0 : ((k == 1u) ? inv_eps : __ldg(cell));
}

// class byte of stored cell (li, lj) from the per-window class map (cls_store)
__device__ __forceinline__ unsigned tl_class(const unsigned char *cls, int cpitch, int li, int lj) {
    int w = lj / CLS_OWN, k = lj - w * CLS_OWN;
    if ((w + 1) * CLS_WIN > cpitch) { // the last columns of the last window live at the end of the previous record
        --w;
        k += CLS_OWN;
    }
    return cls[cls_index(cpitch, li, w, k)];
}

// ---- tagged words
__device__ __forceinline__ unsigned long long tl_word(double v, int half, unsigned tag) {
    const unsigned bits = half ? (unsigned)__double2loint(v) : (unsigned)__double2hiint(v);
    return ((unsigned long long)bits << 32) | (unsigned long long)tag;
}
// wait until the word at p carries `tag`; returns its payload. *fail is set if it does not arrive within ~3 s.
__device__ __forceinline__ unsigned tl_poll(const unsigned long long *p, unsigned tag, int *fail) {
    const volatile unsigned long long *q = reinterpret_cast<const volatile unsigned long long *>(p);
    unsigned long long wv = *q;
    if ((unsigned)wv != tag) {
        const long long t0 = clock64();
        int spins = 0;
        while ((unsigned)(wv = *q) != tag) {
            if ((++spins & 63) == 0 && clock64() - t0 > 6000000000ll) {
                *fail = 1;
                break;
            }
        }
    }
    return (unsigned)(wv >> 32);
}

// Sum over the whole grid and over all GPUs of NV <= 2 values per thread; every thread of every CTA of every GPU returns
// with the same totals in v. wmask bit k: total k is an inner product, weighted by h1*h2 (S3:191) before it leaves the GPU.
// publish: the CTA wrote global / peer memory that others read after this reduction (the r of its edge cells).
// False: a CTA or a peer GPU did not show up within the time limit -- the caller ends the solve.
// (Measured: letting EVERY CTA gather all partials instead of CTA 0 saves a hop on paper and costs 4 us per iteration at
// 148 CTAs -- 148 x 148 sector reads per polling round on the same few cache lines.)
template <int NV>
__device__ __forceinline__ bool tile_allreduce(double (&v)[NV], const TileArgs &A, const unsigned gen, const unsigned wmask,
                                               const bool publish, double (*s_red)[32], unsigned (*s_words)[TL_MAX_BLOCKS],
                                               double *s_bc, int *s_fail) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nblocks = gridDim.x, bid = blockIdx.x;
    const bool sys = A.c.multi == 2;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        const double s = warp_sum(v[k]);
        if (lane == 0) s_red[k][warp] = s;
    }
    __syncthreads();
    if (tid < 2 * NV) { // one thread per word: the CTA's sum of value tid / 2, warps in index order
        const int k = tid >> 1;
        double s = 0.0;
#pragma unroll
        for (int q = 0; q < TL_THREADS / 32; ++q) s += s_red[k][q];
        if (publish) {
            if (sys) __threadfence_system(); // edge cells may have gone to a peer GPU
            else __threadfence();
        }
        *reinterpret_cast<volatile unsigned long long *>(&A.slots[(2 * k + (tid & 1)) * TL_MAX_BLOCKS + bid]) = tl_word(s, tid & 1, gen);
    }
    if (bid == 0) { // gather the CTAs' partials, add them in CTA order, post the GPU's totals to every GPU
        for (int i = tid; i < 2 * NV * nblocks; i += TL_THREADS) {
            const int kh = i / nblocks, b = i - kh * nblocks;
            s_words[kh][b] = tl_poll(&A.slots[kh * TL_MAX_BLOCKS + b], gen, s_fail);
        }
        __syncthreads();
        if (warp == 0) {
            double tot[NV];
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                double s = 0.0;
                for (int b = lane; b < nblocks; b += 32) s += __hiloint2double((int)s_words[2 * k][b], (int)s_words[2 * k + 1][b]);
                s = warp_sum(s);
                s = ((wmask >> k) & 1u) ? s * A.c.h1 * A.c.h2 : s; // S3:191: sum * h1 * h2
                tot[k] = __shfl_sync(0xffffffffu, s, 0);
            }
            if (publish) {
                if (sys) __threadfence_system();
                else __threadfence();
            }
            const int nw = A.pv.nranks * 2 * NV; // lane -> (destination rank, value, half)
            if (lane < nw) {
                const int q = lane / (2 * NV), kh = lane - q * (2 * NV);
                const double val = (NV > 1 && (kh >> 1)) ? tot[NV - 1] : tot[0];
                unsigned long long *dst = reinterpret_cast<unsigned long long *>(A.pv.mail[q]) + (gen & 1u) * 32 + A.pv.rank * 4 + kh;
                *reinterpret_cast<volatile unsigned long long *>(dst) = tl_word(val, kh & 1, gen);
            }
        }
    }
    if (warp == 0) { // every CTA: this GPU's mailbox, all ranks' totals, added in rank order
        const int nw = A.pv.nranks * 2 * NV;
        unsigned pay = 0u;
        if (lane < nw) {
            const int q = lane / (2 * NV), kh = lane - q * (2 * NV);
            const unsigned long long *src = reinterpret_cast<const unsigned long long *>(A.pv.mail[A.pv.rank]) + (gen & 1u) * 32 + q * 4 + kh;
            pay = tl_poll(src, gen, s_fail);
        }
        __syncwarp();
        if (publish) {
            if (sys) __threadfence_system();
            else __threadfence();
        }
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            double s = 0.0;
            for (int q = 0; q < A.pv.nranks; ++q) { // rank order, like the reference's MPI_SUM
                const unsigned hi = __shfl_sync(0xffffffffu, pay, q * 2 * NV + 2 * k), lo = __shfl_sync(0xffffffffu, pay, q * 2 * NV + 2 * k + 1);
                s += __hiloint2double((int)hi, (int)lo);
            }
            if (lane == 0) s_bc[k] = s;
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < NV; ++k) v[k] = s_bc[k];
    return *reinterpret_cast<volatile int *>(s_fail) == 0;
}

__global__ void __launch_bounds__(TL_THREADS, 1) tile_loop_kernel(const TileArgs A) {
    extern __shared__ __align__(16) unsigned char tl_smem[];
    __shared__ double2 s_lut[16];
    __shared__ double s_red[2][32];
    __shared__ unsigned s_words[4][TL_MAX_BLOCKS];
    __shared__ double s_bc[2];
    __shared__ int s_fail;
    const Geom g = A.g;
    const Consts &c = A.c;
    const int tid = threadIdx.x, lane = tid & 31;
    const unsigned bid = blockIdx.x;
    const int ti = (int)bid / A.tc, tj = (int)bid - ti * A.tc;
    const int i_lo = 1 + ti * A.R, i_hi = min(g.nx, i_lo + A.R - 1);
    const int j_lo = 1 + tj * A.C, j_hi = min(g.ny, j_lo + A.C - 1);
    const int Rt = i_hi - i_lo + 1, Ct = j_hi - j_lo + 1;
    const int PC = A.C + 2; // pitch of the haloed arrays
    const size_t n_halo = (size_t)(A.R + 2) * PC, n_own = (size_t)A.R * A.C;
    double *sP = reinterpret_cast<double *>(tl_smem); // p', (R+2) x (C+2): own cell (lr, lc) at [(lr+1) * PC + lc + 1]
    double *sR = sP + n_halo;                         // r of the own cells, [lr * Ct + lc]
    double *sAp = sR + n_own;                         // A p' of the own cells
    unsigned char *sD = reinterpret_cast<unsigned char *>(sAp + n_own); // diagonal class, indexed like sP
    unsigned char *sK = sD + ((n_halo + 15) / 16) * 16; // coefficient classes of a cell's stencil: a(c) | a(up) << 2 | b(c) << 4 | b(right) << 6
    const int P = g.pitch;
    const size_t o00 = (size_t)(i_lo - 1) * P + (size_t)(j_lo - 1); // global offset of tile-local (0, 0)
    const bool safe = A.safe != 0;
    const bool pw1 = (c.pow2 & 1) != 0, pw2 = (c.pow2 & 2) != 0;
    const double h1 = c.h1, h2 = c.h2, rh1 = c.rh1, rh2 = c.rh2, mih1 = c.mih1, mih2 = c.mih2, inv_eps = c.inv_eps;

    if (tid < 16) s_lut[tid] = make_double2(c.lutD[tid], c.lutR[tid]);
    if (tid == 0) s_fail = 0;
    for (int q = tid; q < (Rt + 2) * (Ct + 2); q += TL_THREADS) {
        const int hr = q / (Ct + 2), hc = q - hr * (Ct + 2);
        sP[hr * PC + hc] = 0.0; // S4:805-834: p starts from z, the ring from zero
        sD[hr * PC + hc] = (unsigned char)(tl_class(A.cls, g.cpitch, i_lo - 1 + hr, j_lo - 1 + hc) >> 4);
    }
    for (int q = tid; q < Rt * Ct; q += TL_THREADS) {
        const int lr = q / Ct, lc = q - lr * Ct, li = i_lo + lr, lj = j_lo + lc;
        const unsigned kc = tl_class(A.cls, g.cpitch, li, lj), ku = tl_class(A.cls, g.cpitch, li + 1, lj),
                       kr = tl_class(A.cls, g.cpitch, li, lj + 1);
        sK[q] = (unsigned char)((kc & 3u) | ((ku & 3u) << 2) | (((kc >> 2) & 3u) << 4) | (((kr >> 2) & 3u) << 6));
        sR[q] = A.r[(size_t)li * P + lj]; // S4:813: r = B (setup / reset kernel)
    }
    // Thread layout: one thread per column of the tile; when the tile is narrower than half the CTA, TY groups of
    // threads share the rows. Lanes past the last column shadow it (same reads, no stores, nothing summed), so that
    // every warp that works is whole and its votes are warp-wide.
    const int TXp = ((Ct + 31) / 32) * 32, TY = max(1, min(TL_THREADS / TXp, Rt));
    const int ty = tid / TXp, tx = tid - ty * TXp;
    const bool work = ty < TY;                 // warp-uniform: TXp is a multiple of 32
    const bool own = work && tx < Ct;
    const int txc = min(tx, Ct - 1);
    const int RS = (Rt + TY - 1) / TY, r0 = min(Rt, ty * RS), r1 = min(Rt, r0 + RS);
    // the thread's halo cells (loop-invariant): h < Ct: row 0; < 2 Ct: row Rt+1; < 2 Ct + Rt: column 0; else column Ct+1
    const int nh = 2 * Ct + 2 * Rt;
    int h_sm[TL_NH], h_gl[TL_NH];
#pragma unroll
    for (int n = 0; n < TL_NH; ++n) {
        const int h = (TL_THREADS - 1 - tid) + n * TL_THREADS; // the threads without a column come first
        int hr = 0, hc = 0;
        if (h < Ct) { hr = 0; hc = 1 + h; }
        else if (h < 2 * Ct) { hr = Rt + 1; hc = 1 + h - Ct; }
        else if (h < 2 * Ct + Rt) { hr = 1 + h - 2 * Ct; hc = 0; }
        else { hr = 1 + h - 2 * Ct - Rt; hc = Ct + 1; }
        h_sm[n] = (h < nh) ? hr * PC + hc : -1;
        h_gl[n] = hr * P + hc;
    }
    __syncthreads();

    double zr_old = __ldcg(&A.ctl->zr_old); // init_kernel (S4:820-843)
    double alpha = 0.0, beta = 0.0, diff = 0.0, denom = 0.0;
    long long k = 0;
    int status = STOP_MAX_ITER, conv_iter = 0;
    for (;;) {
        ++k; // S4:847-848
        const bool first = k == 1;
        // ---- phase A: p' = D^-1 r + beta p' (S4:559, 675; first: p = z, S4:830) on the halo ring -- its r comes from global
        // memory: ask for it first -- and on the own cells
        double r_halo[TL_NH];
#pragma unroll
        for (int n = 0; n < TL_NH; ++n) r_halo[n] = (h_sm[n] >= 0) ? __ldcg(&A.r[o00 + h_gl[n]]) : 0.0;
        if (work) {
            int ip = (r0 + 1) * PC + txc + 1, ic = r0 * Ct + txc;
            for (int r = r0; r < r1; ++r, ip += PC, ic += Ct) {
                const double rv = sR[ic];
                const unsigned kd = sD[ip];
                const double2 DR = s_lut[kd];
                double z;
                if (__any_sync(0xffffffffu, safe || kd == DCLS_GENERAL || !in_div_range(rv)))
                    z = precond_div(rv, kd == DCLS_GENERAL ? __ldg(A.D + o00 + (size_t)(r + 1) * P + txc + 1) : DR.x);
                else
                    z = div_const_nocheck(rv, DR.x, DR.y);
                const double pn = first ? z : z + beta * sP[ip];
                if (own) sP[ip] = pn;
            }
        }
#pragma unroll
        for (int n = 0; n < TL_NH; ++n)
            if (h_sm[n] >= 0) {
                const int ip = h_sm[n];
                const unsigned kd = sD[ip];
                const double2 DR = s_lut[kd];
                const double rv = r_halo[n];
                double z;
                if (safe || kd == DCLS_GENERAL || !in_div_range(rv)) z = precond_div(rv, kd == DCLS_GENERAL ? __ldg(A.D + o00 + h_gl[n]) : DR.x);
                else z = div_const_nocheck(rv, DR.x, DR.y);
                sP[ip] = first ? z : z + beta * sP[ip];
            }
        __syncthreads();
        // ---- phase B: A p' (S4:526-535) marching down the thread's rows: p' of the rows above and below and the flux through
        // the upper face stay in registers (the flux through a face is one expression for both cells that share it)
        double s1[1] = {0.0};
        if (work && r0 < r1) {
            int ip = (r0 + 1) * PC + txc + 1, ic = r0 * Ct + txc;
            const size_t o_col = o00 + txc + 1;
            double pd = sP[ip - PC], pc = sP[ip];
            bool bad_c = out_of_div_range(pc) != 0u;
            double fx_lo;
            {
                const unsigned K = sK[ic];
                const double a_c = tl_coef(K & 3u, inv_eps, A.a + o_col + (size_t)(r0 + 1) * P);
                const double num = a_c * (pc - pd);
                if (__any_sync(0xffffffffu, safe || bad_c || out_of_div_range(pd) != 0u)) fx_lo = __ddiv_rn(num, h1);
                else fx_lo = pw1 ? __dmul_rn(num, rh1) : div_const_nocheck(num, h1, rh1);
            }
            for (int r = r0; r < r1; ++r, ip += PC, ic += Ct) {
                const double pu = sP[ip + PC], pl = sP[ip - 1], pr = sP[ip + 1];
                const unsigned K = sK[ic];
                const bool bad_u = out_of_div_range(pu) != 0u;
                const bool bad_e = out_of_div_range(lane == 0 ? pl : pr) != 0u; // the neighbours outside this warp (inside: bad_c of the lanes)
                const bool slow = __any_sync(0xffffffffu, safe || bad_c || bad_u || bad_e);
                double a_u, b_c, b_r;
                if (__any_sync(0xffffffffu, (K & 0xa8u) != 0u)) { // a cell cut by the ellipse: its exact coefficient
                    const size_t o = o_col + (size_t)(r + 1) * P;
                    a_u = tl_coef((K >> 2) & 3u, inv_eps, A.a + o + P);
                    b_c = tl_coef((K >> 4) & 3u, inv_eps, A.b + o);
                    b_r = tl_coef((K >> 6) & 3u, inv_eps, A.b + o + 1);
                } else {
                    a_u = (K & 0x04u) ? inv_eps : 1.0;
                    b_c = (K & 0x10u) ? inv_eps : 1.0;
                    b_r = (K & 0x40u) ? inv_eps : 1.0;
                }
                const double nxu = a_u * (pu - pc), nyh = b_r * (pr - pc), nyl = b_c * (pc - pl); // S4:527, 531
                double fx_hi, fy_hi, fy_lo;
                if (slow) {
                    fx_hi = __ddiv_rn(nxu, h1);
                    fy_hi = __ddiv_rn(nyh, h2);
                    fy_lo = __ddiv_rn(nyl, h2);
                } else {
                    fx_hi = pw1 ? __dmul_rn(nxu, rh1) : div_const_nocheck(nxu, h1, rh1);
                    fy_hi = pw2 ? __dmul_rn(nyh, rh2) : div_const_nocheck(nyh, h2, rh2);
                    fy_lo = pw2 ? __dmul_rn(nyl, rh2) : div_const_nocheck(nyl, h2, rh2);
                }
                const double Ap = mih1 * (fx_hi - fx_lo) + mih2 * (fy_hi - fy_lo); // S4:526-535
                if (own) {
                    sAp[ic] = Ap;
                    s1[0] += Ap * pc; // S4:594
                }
                pd = pc;
                pc = pu;
                bad_c = bad_u;
                fx_lo = fx_hi;
            }
        }
        if (!tile_allreduce<1>(s1, A, (unsigned)(2 * k - 1), 1u, false, s_red, s_words, s_bc, &s_fail)) {
            status = STOP_COMM_TIMEOUT;
            break;
        }
        denom = s1[0]; // S4:866-872
        const bool brk = (c.norm_mode == 1) ? (denom < 1e-15) : (fabs(denom) < 1e-15); // S0:128 | S4:876
        if (brk) {
            status = STOP_DENOM;
            break;
        }
        // ---- phase C: the updates and the two sums
        alpha = zr_old / denom; // S4:877
        double s2[2] = {0.0, 0.0};
        if (work && r0 < r1) {
            int ip = (r0 + 1) * PC + txc + 1, ic = r0 * Ct + txc;
            size_t o = o00 + (size_t)(r0 + 1) * P + txc + 1;
            const int lj = j_lo + txc;
            for (int r = r0; r < r1; ++r, ip += PC, ic += Ct, o += P) {
                const double w0 = __ldcg(&A.w[o]);
                const double pc = sP[ip];
                const double wn = w0 + alpha * pc;          // S4:647
                const double rn = sR[ic] - alpha * sAp[ic]; // S4:648
                const double d = wn - w0;                   // S4:653
                const unsigned kd = sD[ip];
                const double2 DR = s_lut[kd];
                double z;
                if (__any_sync(0xffffffffu, safe || kd == DCLS_GENERAL || !in_div_range(rn)))
                    z = precond_div(rn, kd == DCLS_GENERAL ? __ldg(A.D + o) : DR.x); // S4:559
                else
                    z = div_const_nocheck(rn, DR.x, DR.y);
                if (own) {
                    __stcg(&A.w[o], wn);
                    sR[ic] = rn;
                    s2[0] += d * d;
                    s2[1] += z * rn; // S4:594
                    if (r == 0 || r == Rt - 1 || txc == 0 || txc == Ct - 1) { // the neighbouring tiles' halo (S4:361-398)
                        __stcg(&A.r[o], rn);
                        if (A.c.multi == 2) { // ... and the neighbouring GPUs' rings, through NVLink
                            const int li = i_lo + r;
                            if (li == 1 && A.pv.r_left) A.pv.r_left[(size_t)(A.pv.nx_left + 1) * P + lj] = rn;
                            if (li == g.nx && A.pv.r_right) A.pv.r_right[lj] = rn;
                            if (lj == 1 && A.pv.r_down) A.pv.r_down[(size_t)li * A.pv.pitch_down + A.pv.ny_down + 1] = rn;
                            if (lj == g.ny && A.pv.r_up) A.pv.r_up[(size_t)li * A.pv.pitch_up] = rn;
                        }
                    }
                }
            }
        }
        if (!tile_allreduce<2>(s2, A, (unsigned)(2 * k), 2u, true, s_red, s_words, s_bc, &s_fail)) {
            status = STOP_COMM_TIMEOUT;
            break;
        }
        // ---- the scalar tail (finalize_iteration), in every thread
        const double zr_new = s2[1]; // S4:920-926
        diff = (c.norm_mode == 0) ? sqrt(s2[0] * h1 * h2) : sqrt(s2[0]); // S3:449 | S0:154
        const bool conv = diff < c.delta;                                 // S4:899
        const bool tr = A.trace && k <= c.trace_len && bid == 0 && tid == 0;
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
    // ---- the state goes back to global memory: r and the last p' of the own cells (w is there already)
    double *p_out = A.p[k & 1];
    for (int q = tid; q < Rt * Ct; q += TL_THREADS) {
        const int lr = q / Ct, lc = q - lr * Ct;
        const size_t o = o00 + (size_t)(lr + 1) * P + (lc + 1);
        A.r[o] = sR[q];
        p_out[o] = sP[(lr + 1) * PC + (lc + 1)];
    }
    if (bid == 0 && tid == 0) {
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

} // namespace pb200
