// pb200_tile.cuh -- the whole PCG loop (S4:845-943) of a block that fits in the GPU's shared memory, in ONE cooperative
// launch per rank, on one GPU or on several ("device-resident persistent solver").
//
// One CTA per SM owns a rectangular tile of the block for the whole solve. The tile's p' (with a one-cell halo), r and
// A p' live in the CTA's shared memory (up to 227 KB on sm_100a: 1.3 M cells per GPU), the class bytes of its cells too;
// only w stays in global memory (it is touched once per iteration, by its owner alone). Per iteration:
//
//   phase A  p' = z + beta p' (S4:675) on the tile AND on its halo ring, z = D^-1 r (S4:559). The own cells' z was left in
//            shared memory by phase C of the previous iteration; the halo cells' z arrives as tagged words from the tile (or,
//            through NVLink, the GPU) that owns them, and their p' is this CTA's own previous halo value -- bit-identical to
//            the owner's, so p' is never exchanged (as in the marching kernels). Only iteration 1 divides here.
//   phase B  A p' from shared memory (S4:526-535), kept in shared memory; (A p', p')     -> all-reduce 1
//   phase C  w += alpha p', r -= alpha A p', sum dw^2, z = D^-1 r, sum z r (S4:640-656, 920); z takes A p's place in shared
//            memory; the tile's edge cells' z go out as tagged words: to the neighbouring tiles and, on a block edge with a
//            neighbour GPU, straight into that GPU's halo copy    -> all-reduce 2
//   the scalar tail of the iteration (S4:893-906, 926-932) in every thread, identically.
//
// A thread owns a pair of adjacent columns of the tile and marches down its rows: p' of the current row and the fluxes
// through its upper faces stay in registers (each face flux is computed once), the columns left and right of the pair
// come from shared memory.
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

constexpr int TL_THREADS = 512;    // one CTA per SM, up to 128 registers per thread
constexpr int TL_NH = 4;           // halo cells per thread at most: 2 (R + C) <= TL_NH * TL_THREADS
constexpr int TL_MAX_BLOCKS = 192; // one CTA per SM
constexpr int TL_MAXC = 1024;      // a tile is at most this wide: one thread per pair of columns
constexpr int TL_SLOT_WORDS = 4 * TL_MAX_BLOCKS; // CTA partials: [value k < 2][half][CTA]
constexpr int TL_MAIL_WORDS = 64;  // per GPU: [parity][rank < 8][value k < 2][half]  (= MAIL_DOUBLES: the same 512 bytes)
static_assert(TL_MAIL_WORDS == MAIL_DOUBLES, "the resident loop kernel uses the marching kernels' mailbox");

struct TileArgs {
    double *w, *r, *p[2];
    const double *a, *b, *D;    // full arrays: read only where the class byte says "general"
    const unsigned char *cls;
    unsigned long long *slots;  // TL_SLOT_WORDS tagged words, zeroed before every solve
    unsigned long long *r2;     // tagged copy of r, two words per stored cell (index 2 * (li * pitch + lj)), zeroed before every
                                // solve: the tiles' edge cells travel through it (and through the neighbour GPUs' copies)
    double *gz, *gap;           // lean form only: z and A p' of the own cells in global memory (the two p arrays, idle here)
    double *trace;
    Ctl *ctl;
    Geom g;
    Consts c;
    int R, C, tc;               // tile height and width (the last tiles may be smaller), tiles per row of tiles
    int safe;                   // IEEE divisions everywhere (coefficients outside the proven window)
    PeerView pv;                // mail[q]: every rank's mailbox (one GPU: nranks = 1, its own); r_*: c.multi == 2
};

// Shared-memory layout of a tile of R x C cells (all pitches even, so that a thread's pair of columns is one 16-byte access):
//   sP   (R+2) rows of PC = C+4 (rounded up to even) doubles: p' with its halo; own cell (lr, lc) at [(lr+1)*PC + lc + 2],
//        the left halo column at index 1, the right one at Ct + 2
//   sR, sAp   R rows of Cp = C rounded up to even doubles: r and A p' of the own cells, [lr*Cp + lc]
//   sD   bytes indexed like sP: diagonal class;   sK   bytes indexed like sR: the coefficient classes of the cell's stencil
__host__ __device__ inline int tile_pc(int C) { return (C + 4 + 1) & ~1; }
__host__ __device__ inline int tile_cp(int C) { return (C + 1) & ~1; }
//   sHv, sHk   a hash table of TL_HASH exact values: the a, b, D of the cells cut by the ellipse ("general" class), so that the
//        loop never waits for global memory (the gpu-scope fences of the all-reduce empty the L1 every iteration)
constexpr int TL_HASH = 2048, TL_HASH_BITS = 11, TL_PROBES = 24;
// lean: r and A p' / z stay in global memory (they are touched by their owner only and live in the L2); shared memory keeps
// p' and the class bytes: 10 instead of 26 bytes per cell, for blocks of 1.1-2.9 M cells per GPU.
__host__ __device__ inline size_t tile_smem_bytes(int R, int C, bool lean = false) {
    const size_t halo = (size_t)(R + 2) * (size_t)tile_pc(C), own = (size_t)R * (size_t)tile_cp(C);
    return (halo + (lean ? 0 : 2 * own)) * sizeof(double) + ((halo + 15) / 16) * 16 + ((own + 15) / 16) * 16 +
           TL_HASH * (sizeof(double) + sizeof(unsigned));
}

// key of an exact value: field 0 = D of the cell at haloed index i, 1 = a of the cell above compact index i,
// 2 = b of the cell itself, 3 = b of the cell right of it
__device__ __forceinline__ unsigned tl_key(int index, unsigned field) { return ((unsigned)(index + 1) << 3) | field; }
__device__ __forceinline__ unsigned tl_hash(unsigned key) { return (key * 2654435761u) >> (32 - TL_HASH_BITS); }
__device__ __forceinline__ void tl_insert(unsigned *keys, double *vals, unsigned key, double v) {
    unsigned slot = tl_hash(key);
    for (int probe = 0; probe < TL_PROBES; ++probe) {
        const unsigned old = atomicCAS(&keys[slot], 0u, key);
        if (old == 0u || old == key) {
            vals[slot] = v;
            return;
        }
        slot = (slot + 1u) & (TL_HASH - 1);
    }
    // table full around here: the value stays in global memory only (tl_lookup falls back to it)
}
__device__ __forceinline__ double tl_lookup(const unsigned *keys, const double *vals, unsigned key, const double *global_cell) {
    unsigned slot = tl_hash(key);
    for (int probe = 0; probe < TL_PROBES; ++probe) {
        const unsigned kk = keys[slot];
        if (kk == key) return vals[slot];
        if (kk == 0u) break;
        slot = (slot + 1u) & (TL_HASH - 1);
    }
    return __ldg(global_cell);
}
// z = D^-1 r (S4:559) for a thread's pair of cells: the exact-residual division with (D, RN(1/D)) from the class table when
// every r of the warp lies in the window it is proven for (in_div_range) and no cell of the warp sits next to the ellipse
// boundary; otherwise the whole warp divides as the reference does, the boundary cells by their exact D from the hash
// table. (Measured: dividing the boundary cells by a cached reciprocal instead saved instructions and cost 5 us per
// iteration at 1200 x 800 -- two more table look-ups on the slowest warps' path.)
__device__ __forceinline__ void tl_precond2(double r0, double r1, unsigned kd0, unsigned kd1, const double2 *lut,
                                            const unsigned *keys, const double *vals, int ip, const double *D_cell,
                                            bool safe, double &z0, double &z1) {
    const double2 DR0 = lut[kd0], DR1 = lut[kd1];
    if (__any_sync(0xffffffffu, safe || !in_div_range(r0) || !in_div_range(r1) || kd0 == DCLS_GENERAL || kd1 == DCLS_GENERAL)) {
        z0 = precond_div(r0, kd0 == DCLS_GENERAL ? tl_lookup(keys, vals, tl_key(ip, 0u), D_cell) : DR0.x);
        z1 = precond_div(r1, kd1 == DCLS_GENERAL ? tl_lookup(keys, vals, tl_key(ip + 1, 0u), D_cell + 1) : DR1.x);
    } else {
        z0 = div_const_nocheck(r0, DR0.x, DR0.y);
        z1 = div_const_nocheck(r1, DR1.x, DR1.y);
    }
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
    auto ld = [q]() { return *q; };
    unsigned long long wv = ld();
    if ((unsigned)wv != tag) {
        const long long t0 = clock64();
        int spins = 0;
        while ((unsigned)(wv = ld()) != tag) {
            if ((++spins & 63) == 0 && clock64() - t0 > 6000000000ll) {
                *fail = 1;
                break;
            }
        }
    }
    return (unsigned)(wv >> 32);
}

// A double as a pair of tagged words in one 16-byte access: the halo values of r. No fence orders them against anything:
// whoever finds both tags has the value.
__device__ __forceinline__ void tl_store_tagged(unsigned long long *p, double v, unsigned tag) {
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(tl_word(v, 0, tag)), "l"(tl_word(v, 1, tag)) : "memory");
}
__device__ __forceinline__ double tl_poll_tagged(const unsigned long long *p, unsigned tag, int *fail) {
    unsigned long long a, b;
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
    if ((unsigned)a != tag || (unsigned)b != tag) {
        const long long t0 = clock64();
        int spins = 0;
        for (;;) {
            asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
            if ((unsigned)a == tag && (unsigned)b == tag) break;
            if ((++spins & 63) == 0 && clock64() - t0 > 6000000000ll) {
                *fail = 1;
                break;
            }
        }
    }
    return __hiloint2double((int)(a >> 32), (int)(b >> 32));
}

// Sum over the whole grid and over all GPUs of NV <= 2 values per thread; every thread of every CTA of every GPU returns
// with the same totals in v. wmask bit k: total k is an inner product, weighted by h1*h2 (S3:191) before it leaves the GPU.
// Nothing else is ordered by it: the halo values of r are tagged words themselves (tl_store_tagged), so no fence is needed.
// False: a CTA or a peer GPU did not show up within the time limit -- the caller ends the solve.
// (Measured: letting EVERY CTA gather all partials instead of CTA 0 saves a hop on paper and costs 4 us per iteration at
// 148 CTAs -- 148 x 148 sector reads per polling round on the same few cache lines.)
template <int NV>
__device__ __forceinline__ bool tile_allreduce(double (&v)[NV], const TileArgs &A, const unsigned gen, const unsigned wmask,
                                               double (*s_red)[TL_THREADS / 32], unsigned (*s_words)[TL_MAX_BLOCKS],
                                               double *s_bc, int *s_fail, long long *stamp = nullptr) {
#ifdef PB200_TILE_TIMING
#define AR_STAMP(i) if (stamp) stamp[i] = clock64()
#else
#define AR_STAMP(i)
#endif
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nblocks = gridDim.x, bid = blockIdx.x;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        const double s = warp_sum(v[k]);
        if (lane == 0) s_red[k][warp] = s;
    }
    AR_STAMP(0);
    __syncthreads();
    AR_STAMP(1);
    if (warp == 0) { // the CTA's sums: a fixed tree over its 16 warps, one half-warp per value, then one lane per word
        static_assert(TL_THREADS / 32 == 16 && NV <= 2, "a half-warp adds the 16 warps' partial sums of one value");
        const int k = lane >> 4, j = lane & 15;
        double s = (k < NV) ? s_red[k < NV ? k : 0][j] : 0.0;
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o, 16);
        s = __shfl_sync(0xffffffffu, s, 0, 16);
        if (k < NV && j < 2)
            *reinterpret_cast<volatile unsigned long long *>(&A.slots[(2 * k + j) * TL_MAX_BLOCKS + bid]) = tl_word(s, j, gen);
    }
    AR_STAMP(2);
    if (bid == 0) { // gather the CTAs' partials, add them in CTA order, post the GPU's totals to every GPU
        for (int i = tid; i < 2 * NV * nblocks; i += TL_THREADS) {
            const int kh = i / nblocks, b = i - kh * nblocks;
            s_words[kh][b] = tl_poll(&A.slots[kh * TL_MAX_BLOCKS + b], gen, s_fail);
        }
        AR_STAMP(3);
        __syncthreads();
        AR_STAMP(4);
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
            const int nw = A.pv.nranks * 2 * NV; // lane -> (destination rank, value, half)
            if (lane < nw) {
                const int q = lane / (2 * NV), kh = lane - q * (2 * NV);
                const double val = (NV > 1 && (kh >> 1)) ? tot[NV - 1] : tot[0];
                unsigned long long *dst = reinterpret_cast<unsigned long long *>(A.pv.mail[q]) + (gen & 1u) * 32 + A.pv.rank * 4 + kh;
                *reinterpret_cast<volatile unsigned long long *>(dst) = tl_word(val, kh & 1, gen);
            }
        }
    }
    AR_STAMP(5);
    if (warp == 0) { // every CTA: this GPU's mailbox, all ranks' totals, added in rank order
        const int nw = A.pv.nranks * 2 * NV;
        unsigned pay = 0u;
        if (lane < nw) {
            const int q = lane / (2 * NV), kh = lane - q * (2 * NV);
            const unsigned long long *src = reinterpret_cast<const unsigned long long *>(A.pv.mail[A.pv.rank]) + (gen & 1u) * 32 + q * 4 + kh;
            pay = tl_poll(src, gen, s_fail);
        }
        __syncwarp();
        AR_STAMP(6);
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
    AR_STAMP(7);
#pragma unroll
    for (int k = 0; k < NV; ++k) v[k] = s_bc[k];
    return *reinterpret_cast<volatile int *>(s_fail) == 0;
}

// a thread's pair of adjacent doubles in shared memory (16-byte aligned by construction of the layout)
__device__ __forceinline__ double2 lds2(const double *p) { return *reinterpret_cast<const double2 *>(p); }
__device__ __forceinline__ void sts2(double *p, double a, double b, bool both) {
    if (both) *reinterpret_cast<double2 *>(p) = make_double2(a, b);
    else *p = a;
}

// a thread's pair in global memory (lean form): two 8-byte accesses past L1, the second cell only where it is owned
__device__ __forceinline__ double2 ldg2(const double *p, bool both) {
    return make_double2(__ldcg(p), both ? __ldcg(p + 1) : 0.0);
}
__device__ __forceinline__ void stg2(double *p, double a, double b, bool both) {
    __stcg(p, a);
    if (both) __stcg(p + 1, b);
}

template <bool LEAN>
__global__ void __launch_bounds__(TL_THREADS, 1) tile_loop_kernel(const TileArgs A) {
    extern __shared__ __align__(16) unsigned char tl_smem[];
    __shared__ double2 s_lut[16];
    __shared__ double s_red[2][TL_THREADS / 32];
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
    const int PC = tile_pc(A.C), Cp = tile_cp(A.C);
    const size_t n_halo = (size_t)(A.R + 2) * PC, n_own = (size_t)A.R * Cp;
    double *sP = reinterpret_cast<double *>(tl_smem);
    double *sR = sP + n_halo;                       // (sR, sAp: not in the lean form)
    double *sAp = sR + n_own;
    unsigned char *sD = reinterpret_cast<unsigned char *>(LEAN ? sP + n_halo : sAp + n_own);
    unsigned char *sK = sD + ((n_halo + 15) / 16) * 16; // a(c) | a(up) << 2 | b(c) << 4 | b(right) << 6
    double *sHv = reinterpret_cast<double *>(sK + ((n_own + 15) / 16) * 16);
    unsigned *sHk = reinterpret_cast<unsigned *>(sHv + TL_HASH);
    const int P = g.pitch;
    const size_t o00 = (size_t)(i_lo - 1) * P + (size_t)(j_lo - 1); // global offset of the tile's halo corner (row i_lo-1, column j_lo-1)
    const bool safe = A.safe != 0;
    const bool pw1 = (c.pow2 & 1) != 0, pw2 = (c.pow2 & 2) != 0;
    const double h1 = c.h1, h2 = c.h2, rh1 = c.rh1, rh2 = c.rh2, mih1 = c.mih1, mih2 = c.mih2, inv_eps = c.inv_eps;

    if (tid < 16) s_lut[tid] = make_double2(c.lutD[tid], c.lutR[tid]);
    if (tid == 0) s_fail = 0;
    for (int q = tid; q < (int)n_halo; q += TL_THREADS) { // p starts from zero (S4:805-834); pads hold zeros for ever
        sP[q] = 0.0;
        sD[q] = (unsigned char)DCLS_RING;
    }
    for (int q = tid; q < (int)n_own; q += TL_THREADS) {
        if (!LEAN) {
            sR[q] = 0.0;
            sAp[q] = 0.0;
        }
        sK[q] = 0;
    }
    for (int q = tid; q < TL_HASH; q += TL_THREADS) sHk[q] = 0u;
    __syncthreads();
    for (int q = tid; q < (Rt + 2) * (Ct + 2); q += TL_THREADS) { // halo coordinates (hr, hc): row i_lo-1+hr, column j_lo-1+hc
        const int hr = q / (Ct + 2), hc = q - hr * (Ct + 2);
        const unsigned kd = tl_class(A.cls, g.cpitch, i_lo - 1 + hr, j_lo - 1 + hc) >> 4;
        sD[hr * PC + hc + 1] = (unsigned char)kd;
        if (kd == DCLS_GENERAL) tl_insert(sHk, sHv, tl_key(hr * PC + hc + 1, 0u), A.D[o00 + (size_t)hr * P + hc]);
    }
    for (int q = tid; q < Rt * Ct; q += TL_THREADS) {
        const int lr = q / Ct, lc = q - lr * Ct, li = i_lo + lr, lj = j_lo + lc;
        const unsigned kc = tl_class(A.cls, g.cpitch, li, lj), ku = tl_class(A.cls, g.cpitch, li + 1, lj),
                       kr = tl_class(A.cls, g.cpitch, li, lj + 1);
        sK[lr * Cp + lc] = (unsigned char)((kc & 3u) | ((ku & 3u) << 2) | (((kc >> 2) & 3u) << 4) | (((kr >> 2) & 3u) << 6));
        const size_t o = (size_t)li * P + lj;
        if ((ku & 3u) == CLS_GENERAL) tl_insert(sHk, sHv, tl_key(lr * Cp + lc, 1u), A.a[o + P]);
        if (((kc >> 2) & 3u) == CLS_GENERAL) tl_insert(sHk, sHv, tl_key(lr * Cp + lc, 2u), A.b[o]);
        if (((kr >> 2) & 3u) == CLS_GENERAL) tl_insert(sHk, sHv, tl_key(lr * Cp + lc, 3u), A.b[o + 1]);
        if (!LEAN) sR[lr * Cp + lc] = A.r[o]; // S4:813: r = B (setup / reset kernel)
    }
    // Thread layout: a thread owns a PAIR of adjacent columns and marches down its rows; when the tile is narrow, TY groups
    // of threads share the rows. Lanes past the last pair shadow it (same reads, no stores, nothing summed), so that every
    // warp that works is whole and its votes are warp-wide. An odd last column is a pair whose second cell is not owned.
    const int NPAIR = (Ct + 1) / 2, TXp = ((NPAIR + 31) / 32) * 32, TY = max(1, min(TL_THREADS / TXp, Rt));
    const int ty = tid / TXp, tx = tid - ty * TXp;
    const bool work = ty < TY;                  // warp-uniform: TXp is a multiple of 32
    const int c0 = 2 * min(tx, NPAIR - 1);      // first column of the pair (tile-local, even)
    const bool own0 = work && tx < NPAIR, own1 = own0 && c0 + 1 < Ct;
    const int RS = (Rt + TY - 1) / TY, r0 = min(Rt, ty * RS), r1 = min(Rt, r0 + RS);
    // the thread's halo cells (loop-invariant): h < Ct: row above; < 2 Ct: row below; < 2 Ct + Rt: left column; else right column
    const int nh = 2 * Ct + 2 * Rt;
    int h_sm[TL_NH], h_gl[TL_NH];
    unsigned h_live = 0u; // bit n: somebody (a neighbouring tile, a neighbouring GPU) rewrites that cell's r every iteration
#pragma unroll
    for (int n = 0; n < TL_NH; ++n) {
        const int h = (TL_THREADS - 1 - tid) + n * TL_THREADS; // the threads without columns come first
        int hr = 0, hc = 0;
        if (h < Ct) { hr = 0; hc = 1 + h; }
        else if (h < 2 * Ct) { hr = Rt + 1; hc = 1 + h - Ct; }
        else if (h < 2 * Ct + Rt) { hr = 1 + h - 2 * Ct; hc = 0; }
        else { hr = 1 + h - 2 * Ct - Rt; hc = Ct + 1; }
        h_sm[n] = (h < nh) ? hr * PC + hc + 1 : -1;
        h_gl[n] = hr * P + hc;
        const int li = i_lo - 1 + hr, lj = j_lo - 1 + hc; // block-local: 0 and nx+1 / ny+1 are the block's ring
        const bool live = (li == 0) ? A.pv.t_left != nullptr
                        : (li == g.nx + 1) ? A.pv.t_right != nullptr
                        : (lj == 0) ? A.pv.t_down != nullptr
                        : (lj == g.ny + 1) ? A.pv.t_up != nullptr : true;
        if (h < nh && live) h_live |= 1u << n;
    }
    __syncthreads();

    double zr_old = __ldcg(&A.ctl->zr_old); // init_kernel (S4:820-843)
    double alpha = 0.0, beta = 0.0, diff = 0.0, denom = 0.0;
    long long k = 0;
    int status = STOP_MAX_ITER, conv_iter = 0;
    const int ip0 = (r0 + 1) * PC + c0 + 2, ic0 = r0 * Cp + c0;  // the pair's first cell in the haloed / compact arrays
    const size_t og0 = o00 + (size_t)(r0 + 1) * P + c0 + 1;      // ... and in global memory
#ifdef PB200_TILE_TIMING // variant builds only (make variant): cycle stamps of one CTA's thread 0 instead of the scalar trace
    const bool t_rec = A.trace && tid == 0 && (PB200_TILE_TIMING == 5 || bid == ((PB200_TILE_TIMING & 1) ? gridDim.x - 1u : 0u));
    long long t_s[7], t_a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#define TL_STAMP(i) t_s[i] = clock64()
#else
#define TL_STAMP(i)
#endif
    for (;;) {
        ++k; // S4:847-848
        const bool first = k == 1;
        TL_STAMP(0);
        // ---- phase A: p' = z + beta p' (S4:675; first: p = z, S4:830) on the halo ring and on the own cells. z = D^-1 r (S4:559)
        // of the own cells was left in sAp by phase C of the previous iteration (the same expression on the same r), the halo
        // cells' z comes as tagged words from the tile / GPU that owns them; only iteration 1 divides here, r = B from setup.
        double v_halo[TL_NH];
        unsigned h_isz = 0u; // bit n: v_halo[n] is z already
#pragma unroll
        for (int n = 0; n < TL_NH; ++n) {
            v_halo[n] = 0.0;
            if (h_sm[n] >= 0) {
                if (first || !((h_live >> n) & 1u)) v_halo[n] = __ldcg(&A.r[o00 + h_gl[n]]); // r = B (setup kernel); a physical boundary: 0
                else {
                    v_halo[n] = tl_poll_tagged(A.r2 + 2 * (o00 + h_gl[n]), (unsigned)(k - 1), &s_fail);
                    h_isz |= 1u << n;
                }
            }
        }
        if (work) {
            int ip = ip0, ic = ic0;
            size_t og = og0;
            if (first) {
                for (int r = r0; r < r1; ++r, ip += PC, ic += Cp, og += P) {
                    const double2 rv = LEAN ? ldg2(A.r + og, own1) : lds2(sR + ic);
                    const unsigned kd2 = *reinterpret_cast<const unsigned short *>(sD + ip), kd0 = kd2 & 0xffu, kd1 = kd2 >> 8;
                    double z0, z1;
                    tl_precond2(rv.x, rv.y, kd0, kd1, s_lut, sHk, sHv, ip, A.D + og, safe, z0, z1); // S4:559
                    if (own0) sts2(sP + ip, z0, z1, own1);
                }
            } else {
                for (int r = r0; r < r1; ++r, ip += PC, ic += Cp, og += P) {
                    const double2 z = LEAN ? ldg2(A.gz + og, own1) : lds2(sAp + ic), po = lds2(sP + ip);
                    if (own0) sts2(sP + ip, z.x + beta * po.x, z.y + beta * po.y, own1);
                }
            }
        }
#pragma unroll
        for (int n = 0; n < TL_NH; ++n)
            if (h_sm[n] >= 0) {
                const int ip = h_sm[n];
                double z = v_halo[n];
                if (!((h_isz >> n) & 1u)) {
                    const unsigned kd = sD[ip];
                    const double2 DR = s_lut[kd];
                    if (safe || kd == DCLS_GENERAL || !in_div_range(z))
                        z = precond_div(z, kd == DCLS_GENERAL ? tl_lookup(sHk, sHv, tl_key(ip, 0u), A.D + o00 + h_gl[n]) : DR.x);
                    else z = div_const_nocheck(z, DR.x, DR.y);
                }
                sP[ip] = first ? z : z + beta * sP[ip];
            }
        __syncthreads();
        TL_STAMP(1);
        // ---- phase B: A p' (S4:526-535) marching down the thread's rows: p' of the current row and the fluxes through its
        // upper faces stay in registers; the flux through a face is one expression for both cells that share it, so the
        // face between the pair's two cells is computed once
        double s1[1] = {0.0};
        if (work && r0 < r1) {
            int ip = ip0, ic = ic0;
            size_t og = og0;
            const double2 pd = lds2(sP + ip - PC);
            double2 pc = lds2(sP + ip);
            bool bad_c = (out_of_div_range(pc.x) | out_of_div_range(pc.y)) != 0u;
            double fxl0, fxl1;
            {
                const unsigned K2 = *reinterpret_cast<const unsigned short *>(sK + ic);
                const double a0 = tl_coef(K2 & 3u, inv_eps, A.a + og), a1 = tl_coef((K2 >> 8) & 3u, inv_eps, A.a + og + 1);
                const double n0 = a0 * (pc.x - pd.x), n1 = a1 * (pc.y - pd.y);
                if (__any_sync(0xffffffffu, safe || bad_c || (out_of_div_range(pd.x) | out_of_div_range(pd.y)) != 0u)) {
                    fxl0 = __ddiv_rn(n0, h1);
                    fxl1 = __ddiv_rn(n1, h1);
                } else {
                    fxl0 = pw1 ? __dmul_rn(n0, rh1) : div_const_nocheck(n0, h1, rh1);
                    fxl1 = pw1 ? __dmul_rn(n1, rh1) : div_const_nocheck(n1, h1, rh1);
                }
            }
            for (int r = r0; r < r1; ++r, ip += PC, ic += Cp, og += P) {
                const double2 pu = lds2(sP + ip + PC);
                const double pl = sP[ip - 1], pr = sP[ip + 2];
                const unsigned K2 = *reinterpret_cast<const unsigned short *>(sK + ic), K0 = K2 & 0xffu, K1 = K2 >> 8;
                const bool bad_u = (out_of_div_range(pu.x) | out_of_div_range(pu.y)) != 0u;
                const bool bad_e = out_of_div_range(lane == 0 ? pl : pr) != 0u; // the neighbours outside this warp (inside: the lanes' own cells)
                const bool slow = __any_sync(0xffffffffu, safe || bad_c || bad_u || bad_e);
                double au0, au1, bc0, bm, br1; // a above the two cells; b left of the pair, between its cells, right of it
                if (__any_sync(0xffffffffu, ((K0 | K1) & 0xa8u) != 0u)) { // a cell cut by the ellipse: its exact coefficient
                    au0 = (K0 & 0x08u) ? tl_lookup(sHk, sHv, tl_key(ic, 1u), A.a + og + P) : ((K0 & 0x04u) ? inv_eps : 1.0);
                    au1 = (K1 & 0x08u) ? tl_lookup(sHk, sHv, tl_key(ic + 1, 1u), A.a + og + P + 1) : ((K1 & 0x04u) ? inv_eps : 1.0);
                    bc0 = (K0 & 0x20u) ? tl_lookup(sHk, sHv, tl_key(ic, 2u), A.b + og) : ((K0 & 0x10u) ? inv_eps : 1.0);
                    bm = (K0 & 0x80u) ? tl_lookup(sHk, sHv, tl_key(ic, 3u), A.b + og + 1) : ((K0 & 0x40u) ? inv_eps : 1.0);
                    br1 = (K1 & 0x80u) ? tl_lookup(sHk, sHv, tl_key(ic + 1, 3u), A.b + og + 2) : ((K1 & 0x40u) ? inv_eps : 1.0);
                } else {
                    au0 = (K0 & 0x04u) ? inv_eps : 1.0;
                    au1 = (K1 & 0x04u) ? inv_eps : 1.0;
                    bc0 = (K0 & 0x10u) ? inv_eps : 1.0;
                    bm = (K0 & 0x40u) ? inv_eps : 1.0;
                    br1 = (K1 & 0x40u) ? inv_eps : 1.0;
                }
                const double nx0 = au0 * (pu.x - pc.x), nx1 = au1 * (pu.y - pc.y);                                  // S4:527
                const double nl = bc0 * (pc.x - pl), nm = bm * (pc.y - pc.x), nr = br1 * (pr - pc.y);                // S4:531
                double fxh0, fxh1, fl, fm, fr;
                if (slow) {
                    fxh0 = __ddiv_rn(nx0, h1);
                    fxh1 = __ddiv_rn(nx1, h1);
                    fl = __ddiv_rn(nl, h2);
                    fm = __ddiv_rn(nm, h2);
                    fr = __ddiv_rn(nr, h2);
                } else {
                    fxh0 = pw1 ? __dmul_rn(nx0, rh1) : div_const_nocheck(nx0, h1, rh1);
                    fxh1 = pw1 ? __dmul_rn(nx1, rh1) : div_const_nocheck(nx1, h1, rh1);
                    fl = pw2 ? __dmul_rn(nl, rh2) : div_const_nocheck(nl, h2, rh2);
                    fm = pw2 ? __dmul_rn(nm, rh2) : div_const_nocheck(nm, h2, rh2);
                    fr = pw2 ? __dmul_rn(nr, rh2) : div_const_nocheck(nr, h2, rh2);
                }
                const double Ap0 = mih1 * (fxh0 - fxl0) + mih2 * (fm - fl); // S4:526-535
                const double Ap1 = mih1 * (fxh1 - fxl1) + mih2 * (fr - fm);
                if (own0) {
                    if (LEAN) stg2(A.gap + og, Ap0, Ap1, own1);
                    else sts2(sAp + ic, Ap0, Ap1, own1);
                    s1[0] += Ap0 * pc.x; // S4:594
                    if (own1) s1[0] += Ap1 * pc.y;
                }
                pc = pu;
                bad_c = bad_u;
                fxl0 = fxh0;
                fxl1 = fxh1;
            }
        }
        TL_STAMP(2);
#ifdef PB200_TILE_TIMING
        long long t_a1[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (!tile_allreduce<1>(s1, A, (unsigned)(2 * k - 1), 1u, s_red, s_words, s_bc, &s_fail, t_rec ? t_a1 : nullptr)) {
#else
        if (!tile_allreduce<1>(s1, A, (unsigned)(2 * k - 1), 1u, s_red, s_words, s_bc, &s_fail)) {
#endif
            status = STOP_COMM_TIMEOUT;
            break;
        }
        TL_STAMP(3);
        denom = s1[0]; // S4:866-872
        const bool brk = (c.norm_mode == 1) ? (denom < 1e-15) : (fabs(denom) < 1e-15); // S0:128 | S4:876
        if (brk) {
            status = STOP_DENOM;
            break;
        }
        // ---- phase C: the updates and the two sums; w of the next row is requested before this row is worked on
        alpha = zr_old / denom; // S4:877
        double s2[2] = {0.0, 0.0};
        if (work && r0 < r1) {
            int ip = ip0, ic = ic0;
            size_t og = og0;
            const int lj = j_lo + c0;
            const bool col_edge0 = c0 == 0 || c0 == Ct - 1, col_edge1 = c0 + 1 == Ct - 1; // first / last column of the tile
            double wa = own0 ? __ldcg(&A.w[og]) : 0.0, wb = own1 ? __ldcg(&A.w[og + 1]) : 0.0;
            double2 g_r = make_double2(0.0, 0.0), g_ap = g_r; // lean form: r and A p' of the row come from global memory, a row ahead
            if (LEAN && own0) {
                g_r = ldg2(A.r + og, own1);
                g_ap = ldg2(A.gap + og, own1);
            }
            for (int r = r0; r < r1; ++r, ip += PC, ic += Cp, og += P) {
                const double w0 = wa, w1 = wb;
                const double2 rv = LEAN ? g_r : lds2(sR + ic), Ap = LEAN ? g_ap : lds2(sAp + ic);
                if (r + 1 < r1) {
                    wa = own0 ? __ldcg(&A.w[og + P]) : 0.0;
                    wb = own1 ? __ldcg(&A.w[og + P + 1]) : 0.0;
                    if (LEAN && own0) {
                        g_r = ldg2(A.r + og + P, own1);
                        g_ap = ldg2(A.gap + og + P, own1);
                    }
                }
                const double2 pc = lds2(sP + ip);
                const double wn0 = w0 + alpha * pc.x, wn1 = w1 + alpha * pc.y;       // S4:647
                const double rn0 = rv.x - alpha * Ap.x, rn1 = rv.y - alpha * Ap.y;   // S4:648
                const double d0 = wn0 - w0, d1 = wn1 - w1;                           // S4:653
                const unsigned kd2 = *reinterpret_cast<const unsigned short *>(sD + ip), kd0 = kd2 & 0xffu, kd1 = kd2 >> 8;
                double z0, z1;
                tl_precond2(rn0, rn1, kd0, kd1, s_lut, sHk, sHv, ip, A.D + og, safe, z0, z1); // S4:559
                if (own0) {
                    __stcg(&A.w[og], wn0);
                    if (own1) __stcg(&A.w[og + 1], wn1);
                    if (LEAN) {
                        stg2(A.r + og, rn0, rn1, own1);
                        stg2(A.gz + og, z0, z1, own1);
                    } else {
                        sts2(sR + ic, rn0, rn1, own1);
                        sts2(sAp + ic, z0, z1, own1); // A p' is used up: its place keeps z for the next iteration's p'
                    }
                    s2[0] += d0 * d0;
                    s2[1] += z0 * rn0; // S4:594
                    if (own1) {
                        s2[0] += d1 * d1;
                        s2[1] += z1 * rn1;
                    }
                    const bool row_edge = r == 0 || r == Rt - 1;
                    // the neighbouring tiles' halo (S4:361-398; what they need of r is z = D^-1 r) and, on a block edge with a neighbour
                    // GPU, that GPU's ring (NVLink)
                    if (row_edge || col_edge0) {
                        tl_store_tagged(A.r2 + 2 * og, z0, (unsigned)k);
                        if (A.c.multi == 2) {
                            const int li = i_lo + r;
                            if (li == 1 && A.pv.t_left) tl_store_tagged(A.pv.t_left + 2 * ((size_t)(A.pv.nx_left + 1) * P + lj), z0, (unsigned)k);
                            if (li == g.nx && A.pv.t_right) tl_store_tagged(A.pv.t_right + 2 * (size_t)lj, z0, (unsigned)k);
                            if (lj == 1 && A.pv.t_down) tl_store_tagged(A.pv.t_down + 2 * ((size_t)li * A.pv.pitch_down + A.pv.ny_down + 1), z0, (unsigned)k);
                            if (lj == g.ny && A.pv.t_up) tl_store_tagged(A.pv.t_up + 2 * ((size_t)li * A.pv.pitch_up), z0, (unsigned)k);
                        }
                    }
                    if (own1 && (row_edge || col_edge1)) {
                        tl_store_tagged(A.r2 + 2 * (og + 1), z1, (unsigned)k);
                        if (A.c.multi == 2) {
                            const int li = i_lo + r;
                            if (li == 1 && A.pv.t_left) tl_store_tagged(A.pv.t_left + 2 * ((size_t)(A.pv.nx_left + 1) * P + lj + 1), z1, (unsigned)k);
                            if (li == g.nx && A.pv.t_right) tl_store_tagged(A.pv.t_right + 2 * (size_t)(lj + 1), z1, (unsigned)k);
                            if (lj + 1 == g.ny && A.pv.t_up) tl_store_tagged(A.pv.t_up + 2 * ((size_t)li * A.pv.pitch_up), z1, (unsigned)k);
                        }
                    }
                }
            }
        }
        TL_STAMP(4);
#ifdef PB200_TILE_TIMING
        if (!tile_allreduce<2>(s2, A, (unsigned)(2 * k), 2u, s_red, s_words, s_bc, &s_fail, t_rec ? t_a : nullptr)) {
#else
        if (!tile_allreduce<2>(s2, A, (unsigned)(2 * k), 2u, s_red, s_words, s_bc, &s_fail)) {
#endif
            status = STOP_COMM_TIMEOUT;
            break;
        }
        TL_STAMP(5);
        // ---- the scalar tail (finalize_iteration), in every thread
        const double zr_new = s2[1]; // S4:920-926
        diff = (c.norm_mode == 0) ? sqrt(s2[0] * h1 * h2) : sqrt(s2[0]); // S3:449 | S0:154
        const bool conv = diff < c.delta;                                 // S4:899
#ifdef PB200_TILE_TIMING
        const bool tr = false;
        if (t_rec && 2 * k <= c.trace_len) { // rows 2k-2, 2k-1: A, B, all-reduce 1 | C, all-reduce 2, (unused)
            if (PB200_TILE_TIMING == 5) { // every CTA, iteration 30 only: row bid = A, B, all-reduce 1, C, all-reduce 2, tile cells
                if (k == 30 && (int)bid < c.trace_len / 2) {
                    A.trace[6 * bid + 0] = (double)(t_s[1] - t_s[0]);   // phase A, whole CTA
                    A.trace[6 * bid + 1] = (double)(t_a1[1] - t_s[1]);  // phase B, whole CTA (to the all-reduce's first barrier)
                    A.trace[6 * bid + 2] = (double)(t_s[3] - t_a1[1]);  // rest of all-reduce 1
                    A.trace[6 * bid + 3] = (double)(t_a[1] - t_s[3]);   // phase C, whole CTA
                    A.trace[6 * bid + 4] = (double)(t_s[5] - t_a[1]);   // rest of all-reduce 2
                    A.trace[6 * bid + 5] = (double)(Rt * 10000 + Ct);
                }
            } else if (PB200_TILE_TIMING <= 2) {
                for (int j = 0; j < 5; ++j) A.trace[6 * (k - 1) + j] = (double)(t_s[j + 1] - t_s[j]);
                A.trace[6 * (k - 1) + 5] = (double)t_s[0];
            } else { // inside all-reduce 2: entry->sync | ->posted | gather | ->all gathered | sum+mail | mailbox poll+exit
                A.trace[6 * (k - 1) + 0] = (double)(t_a[1] - t_a[0]);
                A.trace[6 * (k - 1) + 1] = (double)(t_a[2] - t_a[1]);
                A.trace[6 * (k - 1) + 2] = (double)(t_a[3] - t_a[2]);
                A.trace[6 * (k - 1) + 3] = (double)(t_a[4] - t_a[3]);
                A.trace[6 * (k - 1) + 4] = (double)(t_a[5] - t_a[4]);
                A.trace[6 * (k - 1) + 5] = (double)(t_a[7] - t_a[5]);
            }
        }
#else
        const bool tr = A.trace && k <= c.trace_len && bid == 0 && tid == 0;
#endif
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
        if (!LEAN) A.r[o] = sR[lr * Cp + lc];
        p_out[o] = sP[(lr + 1) * PC + lc + 2];
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
