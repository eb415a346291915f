/*
 * poisson_b200.h -- C ABI of libpoisson_b200.so: the B200-native (sm_100a) replacement for the
 * reference's MPI+CUDA solver path, stage4-mpi+cuda/poisson_mpi_cuda_f.cu ("S4" below;
 * "S3" = stage3-openmp+mpi/hybrid.cpp, "S0" = stage0/Withoutopenmp1.cpp).
 *
 * The reference has no FFI: its boundary is `gradient_solver_mpi` (S4:688) called from `main`
 * (S4:986-1039). Every entry point below names the reference interface it replaces.
 * Plain C types only; the library never calls exit() (S4:20-27 did) -- errors come back as
 * pb200_status plus pb200_last_error(). There is NO CPU fallback: without a CUDA device every
 * compute entry point fails with PB200_ERR_CUDA.
 *
 * Threading: one solver handle per host thread at a time; handles are independent.
 */
#ifndef POISSON_B200_H
#define POISSON_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PB200_API_VERSION 2 /* 2: rhs_mode, ell_ax (taken from v1's reserved words: a zeroed v1 struct means the same), analytic error in the result */

typedef enum pb200_status {
    PB200_OK = 0,
    PB200_ERR_INVALID = 1,     /* bad argument / parameter combination */
    PB200_ERR_CUDA = 2,        /* CUDA runtime error, or no device */
    PB200_ERR_COMM = 3,        /* NCCL error / communicator missing */
    PB200_ERR_STATE = 4,       /* call order (e.g. get_solution before solve) */
    PB200_ERR_TOO_LARGE = 5    /* request exceeds a documented fixed limit */
} pb200_status;

/* Which stopping norm / denominator guard the loop uses. */
enum {
    PB200_NORM_WEIGHTED = 0, /* stages 1-4: sqrt(sum dw^2 * h1*h2) (S4:659,897; S3:449); guard |denom| < 1e-15 (S4:876) */
    PB200_NORM_STAGE0 = 1    /* stage0: sqrt(sum dw^2) (S0:149-154); guard denom < 1e-15 (S0:128) */
};

/* Right-hand side. */
enum {
    PB200_RHS_INDICATOR = 0, /* the reference: B_ij = f if node (x_i,y_j) lies in D, else 0 (S4:182-191) */
    PB200_RHS_AREA = 1       /* no reference counterpart (BASELINE north_star's "area fractions"): B_ij = f * share of the cell
                                [x_i-h1/2, x_i+h1/2] x [y_j-h2/2, y_j+h2/2] inside D, computed as the area of the polygon
                                spanned by the end points of the in-D parts of the cell's four edges (the edge intersections of
                                S4:46-76): the cell clipped by the chords of the ellipse. Changes every number of the solve,
                                so it is never the default. */
};

/* Why the loop ended. */
enum {
    PB200_STOP_MAX_ITER = 0,  /* ran max_iter (or fixed_iters) iterations (S4:847) */
    PB200_STOP_CONVERGED = 1, /* ||w(k+1)-w(k)|| < delta (S4:899-906) */
    PB200_STOP_DENOM = 2,     /* (Ap,p) guard fired (S4:876) */
    PB200_STOP_COMM_TIMEOUT = 3 /* peer-memory mode: a rank did not answer in time (pb200_solve returns PB200_ERR_COMM) */
};

/*
 * Everything the reference fixes at compile time (S4:11-13, 32-34, 693, 1006-1007) plus the
 * decomposition its MPI launcher provides (S4:987-991). pb200_default_params() fills the
 * reference's values; `prog M N` (S4:996-1000) then only sets M and N.
 */
typedef struct pb200_params {
    int32_t M, N;          /* grid cells per side; unknowns are i=1..M-1, j=1..N-1 (S4:996-1000; default 40,40) */
    double delta;          /* stopping tolerance (S4:1006: 1e-6) */
    int64_t max_iter;      /* <= 0: (M-1)*(N-1) (S4:1007) */
    double A1, B1, A2, B2; /* rectangle (S4:11-12: -1,1,-0.6,0.6) */
    double f_val;          /* right-hand side inside D (S4:13: 1.0); RHS is the point indicator (S4:182-191) */
    double ell_k;          /* domain D = { (x/ell_ax)^2 + (ell_k*y)^2 < 1 } (S4:32-34: ell_k = 2, ell_ax = 1) */
    double eps;            /* fictitious-domain penalty; <= 0: max(h1,h2)^2 (S4:693) */
    int32_t norm_mode;     /* PB200_NORM_* */
    int32_t fixed_iters;   /* > 0: run exactly this many iterations; the delta test is evaluated and recorded
                              (result.first_converged_iter) but does not stop the loop. Throughput runs. */
    int32_t device;        /* CUDA device ordinal; < 0: the current device */
    int32_t check_every;   /* host polls the device-side stop flag every this many iterations; <= 0: auto */
    int32_t trace_len;     /* keep alpha_k, beta_k, ||dw||_k for the first trace_len iterations (pb200_get_trace) */
    /* decomposition: this handle owns block `rank` of a Px x Py grid of blocks (S4:82-133) */
    int32_t nranks;        /* 1 = whole grid on this GPU */
    int32_t rank;          /* 0..nranks-1; rank -> (px,py) = (rank % Px, rank / Px) (S4:100-101) */
    int32_t Px, Py;        /* <= 0: pb200_choose_process_grid(nranks) */
    int32_t use_graph;     /* 1: replay the iteration as a CUDA graph (default); 0: plain launches */
    int32_t rhs_mode;      /* PB200_RHS_*; 0 = the reference's indicator */
    int32_t reserved0;
    double ell_ax;         /* x semi-axis of the ellipse; <= 0: 1.0 (the reference, S4:32-34) */
    int32_t reserved[4];
} pb200_params;

/* What the reference prints (S4:901-903, 969-980, 1027-1035) plus what it never exposed. */
typedef struct pb200_result {
    int64_t iterations;          /* `iter` as S4:982 returns it */
    int32_t stop_reason;         /* PB200_STOP_* */
    int32_t first_converged_iter;/* first k with ||dw|| < delta (0 if never); == iterations on a normal stop */
    double diff_norm;            /* last ||w(k+1)-w(k)|| in the selected norm */
    double zr;                   /* (z,r) carried out of the loop (zr_old, S4:932) */
    double alpha, beta;          /* last step scalars */
    double w_norm_E;             /* sqrt(sum w^2 h1 h2) over all ranks */
    double r_norm_E;             /* sqrt(sum r^2 h1 h2) */
    double w_max;                /* max |w| */
    double t_setup_s;            /* allocation + coefficient/RHS kernel (reference: host loop S4:717 + H2D S4:753-758) */
    double t_solve_s;            /* PCG loop, CUDA events on the solver stream */
    double t_total_s;            /* wall clock of the call */
    double pt_it_per_s;          /* (M-1)(N-1) * iterations / t_solve_s, whole job */
    double algo_gbps;            /* pt_it_per_s * 104 B */
    int64_t kernel_launches;     /* launches of this library's kernels during the call */
    double t_device_s;           /* device time of the whole pb200_solve call (re-setup, init, loop, norms), CUDA events */
    double safe_mode;            /* 1.0 if the run used IEEE divides in every row: the coefficients (1/eps) lie outside
                                    2^-50..2^300 or a grid step exceeds 2^10, the window in which the fast exact divisions
                                    are proven. 0.0 otherwise; single rows whose data leave 2^-840..2^601 still fall back
                                    to IEEE divides on their own. */
    double err_max_in_D;         /* max |w - u| over the nodes inside D, u = f (1 - (x/ax)^2 - (k y)^2) / (2/ax^2 + 2 k^2) the
                                    analytic solution on D (R4 p.1-2: (1 - x^2 - 4 y^2)/10); all ranks. The reference states u
                                    but never compares: it frees w unseen (S4:946-954). */
    double err_l2_in_D;          /* sqrt(sum_{nodes in D} (w - u)^2 h1 h2) */
    double reserved[2];
} pb200_result;

typedef struct pb200_solver pb200_solver;

/* ---- housekeeping ------------------------------------------------------------------ */
int32_t pb200_api_version(void);
const char *pb200_status_string(pb200_status s);
const char *pb200_last_error(void);      /* thread-local text of the last failure */
int32_t pb200_device_count(void);        /* 0 when no usable CUDA device */

void pb200_default_params(pb200_params *p);                   /* S4:11-13, 1006-1007 */
/* replaces choose_process_grid (S4:82-86): Px = largest divisor of size <= sqrt(size) */
pb200_status pb200_choose_process_grid(int32_t size, int32_t *Px, int32_t *Py);
/* replaces decompose_2d (S4:97-133): inclusive global index ranges of block `rank` */
pb200_status pb200_decompose_2d(int32_t M, int32_t N, int32_t Px, int32_t Py, int32_t rank, int32_t *i_start,
                                int32_t *i_end, int32_t *j_start, int32_t *j_end);

/* Host-only introspection, no reference counterpart: how the two iteration kernels would cut a block of nx x ny
 * interior cells into work when `resident_ctas` of their CTAs fit on the GPU at once (592 on a B200: 148 SMs x 4).
 * plan[0] = bands of the main part, plan[1] = rows the main part covers (1..plan[1]), plan[2] = lower bands over the
 * remaining rows, plan[3] = CTAs launched, plan[4] = 124-column warp windows per row, plan[5] = warps per CTA.
 * Band b < plan[0] holds rows 1 + plan[1]*b/plan[0] .. plan[1]*(b+1)/plan[0] (integer division), the lower bands
 * split the rest the same way. Needs no GPU; used by the CPU tests to check that every row is covered once. */
pb200_status pb200_launch_plan(int32_t nx, int32_t ny, int32_t resident_ctas, int32_t plan[6]);

/* Host-only introspection, no reference counterpart: how the resident loop kernel (one cooperative launch for the whole
 * solve, the block's state in shared memory) would cut a block of nx x ny cells into tiles for `sms` CTAs with
 * `smem_bytes` of shared memory each (148 and 231000 on a B200): plan = { tile rows R, tile columns C, tiles down tr, tiles
 * across tc }; tile (ti, tj) holds rows 1+ti*R .. min(nx,(ti+1)*R) and columns 1+tj*C .. min(ny,(tj+1)*C).
 * PB200_ERR_TOO_LARGE: the block does not fit (it runs the two marching kernels instead). */
pb200_status pb200_tile_plan(int32_t nx, int32_t ny, int32_t sms, int64_t smem_bytes, int32_t plan[4]);

/* ---- the solver (replaces gradient_solver_mpi, S4:688-983) ---------------------------- */
/* Allocates the block on the device and runs the coefficient/RHS kernel
 * (replaces fictitious_regions_setup_local S4:146-192 + flatten/upload S4:723-760). */
pb200_status pb200_create(const pb200_params *p, pb200_solver **out);
/* Multi-GPU: hand the handle its communicator. `unique_id` is the 128-byte ncclUniqueId that
 * rank 0 obtained from pb200_comm_unique_id() and the host distributed to every rank
 * (replaces MPI_Init/Comm_rank/Comm_size, S4:987-991). Collective over all ranks. */
pb200_status pb200_comm_unique_id(uint8_t id_out[128]);
pb200_status pb200_attach_comm(pb200_solver *s, const uint8_t unique_id[128]);
/* Optional, after pb200_attach_comm: peer-memory mode. Each rank exports 256 bytes (where its mailbox, its residual
 * array and the tagged halo copy of the resident loop kernel live), the host gathers the exports of all ranks in rank order (replaces nothing in the reference: it would be
 * an MPI_Allgather there) and every rank attaches them. From then on the two iteration kernels exchange halos and
 * partial sums themselves through NVLink stores into peer memory (CUDA IPC between processes, peer access between
 * threads) -- no NCCL call inside the loop; blocks of up to ~2.9 M cells per GPU run the whole loop as one cooperative
 * launch per rank, larger ones replay two kernels per iteration as a CUDA graph. Up to 8 ranks on one node.
 * Replaces exchange_halos_2d_gpu (S4:331-500) and the three MPI_Allreduce per iteration (S4:872, 893, 926). */
pb200_status pb200_peer_export(pb200_solver *s, uint8_t blob_out[256]);
pb200_status pb200_peer_attach(pb200_solver *s, const uint8_t *blobs_in_rank_order, int32_t nranks);
/* Init (S4:803-843) + PCG loop (S4:845-943). Collective when nranks > 1. */
pb200_status pb200_solve(pb200_solver *s, pb200_result *res);
/* Copies this rank's block of w into a host array laid out as the WHOLE grid, (M+1) rows of
 * `ld` doubles (ld >= N+1): element (i,j) at host[i*ld + j]. Only the block's interior cells are
 * written (the reference frees w unseen, S4:946-954; north_star asks for the dump). */
pb200_status pb200_get_solution(pb200_solver *s, double *host, int64_t ld);
/* Copies min(n, trace_len, iterations) rows of (alpha_k, beta_k, ||dw||_k); returns rows in *rows. */
pb200_status pb200_get_trace(pb200_solver *s, double *host_3n, int32_t n, int32_t *rows);
/* Device pointer + pitch (in doubles) of a field of this block, for callers that stay on the GPU.
 * which: 0 w, 1 r, 2 p, 4 a, 5 b, 6 D (3 was Ap, which is never stored: invalid). Rows li=0..nx+1, columns lj=0..ny+1 at ptr[li*pitch+lj]. */
pb200_status pb200_get_field(pb200_solver *s, int32_t which, double **dev_ptr, int64_t *pitch, int32_t *nx, int32_t *ny);
/* Measurement aid (bench.py's roofline): restarts from the initial state, runs `iters` iterations as plain
 * launches with CUDA events on the solver stream around each kernel, and returns the mean duration in
 * milliseconds of K1 (ms_out[0]: p-update + A p + (Ap,p)) and K2 (ms_out[1]: w,r update + norms). Single rank. */
pb200_status pb200_profile_iteration(pb200_solver *s, int32_t iters, double ms_out[2]);
void pb200_destroy(pb200_solver *s);

/* One call = the reference binary's whole solver phase: create + solve (+ solution to host when
 * w_host != NULL, ld as above). Single rank. The handle is kept (one per host thread) so that an identical
 * next call reuses the device workspace and the coefficients instead of allocating and recomputing them;
 * any other call replaces it, pb200_release_cache() or the end of the thread frees it. Call pb200_release_cache() yourself
 * before the thread (or the process) ends: a workspace freed by the thread's own exit handlers may find the CUDA context
 * already gone. */
pb200_status pb200_run(const pb200_params *p, pb200_result *res, double *w_host, int64_t ld);
void pb200_release_cache(void);

/* ---- operator-level entry points (parity tests; same kernels as the solver) ------------- */
/* All arrays are HOST pointers in the reference's layout: (nx+2)*(ny+2) doubles, idx = li*(ny+2)+lj
 * (S4:723-736). Each call stages to the device, runs the production kernel(s), and copies back. */
/* a, b (block + halo ring), B (interior), D (S4:556-557 divisor, interior) for block [i0..i1]x[j0..j1];
 * any output may be NULL. Replaces fictitious_regions_setup_local (S4:146-192). */
pb200_status pb200_op_setup(const pb200_params *p, int32_t i0, int32_t i1, int32_t j0, int32_t j1, double *a, double *b,
                            double *B, double *D);
/* Ap = A p (replaces apply_A_kernel, S4:507-536). */
pb200_status pb200_op_apply_A(const pb200_params *p, const double *pvec, const double *a, const double *b, double *Ap,
                              int32_t nx, int32_t ny);
/* z = D^-1 r (replaces apply_Dinv_kernel, S4:541-562). */
pb200_status pb200_op_apply_Dinv(const pb200_params *p, const double *r, const double *a, const double *b, double *z,
                                 int32_t nx, int32_t ny);
/* One fused iteration on caller data: given r, p_old, w, a, b and the scalars beta, zr_old it performs
 * p = D^-1 r + beta*p_old (S4:559,675), Ap = A p (S4:526-535), denom = (Ap,p) (S4:866), alpha = zr_old/denom,
 * w += alpha p, r -= alpha Ap, diff2 = sum dw^2 (S4:640-656), zr = (D^-1 r, r) (S4:920). In/out: w, r, p.
 * out[0..3] = denom, alpha, diff2 (unweighted sum), zr (weighted). first != 0: p = D^-1 r (S4:830). */
pb200_status pb200_op_iteration(const pb200_params *p, double *w, double *r, double *pvec, double *Ap, const double *a,
                                const double *b, int32_t nx, int32_t ny, double beta, double zr_old, int32_t first,
                                double out[4]);
/* h1*h2-weighted inner product over the interior (replaces dot_kernel + dot_gpu, S4:574-598, 772-787). */
pb200_status pb200_op_dot(const pb200_params *p, const double *u, const double *v, int32_t nx, int32_t ny, double *out);
/* No counterpart in the reference. The same weighted inner product for callers that cannot allocate device
 * memory: programs whose GPU memory is wholly owned by their own pool or framework allocator (a cudaMalloc from
 * this library would fail or fragment it) and latency-critical loops that cannot afford cudaMalloc/cudaFree,
 * which synchronise the device. Inputs and output are DEVICE pointers supplied by the caller: u, v are blocks in
 * the solver's layout (rows li = 0..nx+1 of `pitch` doubles, columns lj = 0..ny+1, pitch >= ny+2), *dev_out
 * receives sum_{interior} u*v * h1 * h2; the launch goes to `cuda_stream` (a cudaStream_t, NULL = default stream)
 * and the call returns without waiting for it. Nothing is allocated or freed: the block partials live in a
 * 64 KiB + 256 B array that is part of the library image on every GPU that loads it (held until the process exits).
 * Limit: ceil(ny/256) * min(nx,512) <= 8192 reduction blocks (e.g. ny <= 4096 when nx >= 512); beyond that the
 * call fails with PB200_ERR_TOO_LARGE. The array is shared by all calls on a GPU and keeps the last call's
 * contents (the call re-zeroes what must start at zero): calls that follow one another in a stream are fine,
 * calls whose kernels overlap on the GPU race. The first call on each device resolves the array's device
 * address (with that device current); make that first call before any stream capture. */
pb200_status pb200_dot_noalloc(const double *dev_u, const double *dev_v, int32_t nx, int32_t ny, int64_t pitch,
                               double h1, double h2, double *dev_out, void *cuda_stream);
/* Device self-test of the correctly-rounded division by a grid-uniform constant used inside the
 * stencil: compares it with IEEE x/h on n pseudo-random and n adversarial x; *mismatches must be 0. */
pb200_status pb200_selftest_div(double h, int64_t n, uint64_t seed, int64_t *mismatches);

#ifdef __cplusplus
}
#endif
#endif /* POISSON_B200_H */
