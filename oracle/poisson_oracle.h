/*
 * poisson_oracle.h -- CPU restatement (plain C) of the reference's fictitious-domain PCG
 * Poisson solver. TEST INFRASTRUCTURE ONLY: only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this; the product library
 * (libpoisson_b200.so) never links or calls it.
 *
 * Parity status: PINNED. Every function is checked bit-for-bit (operators) or to the
 * stated tolerance (whole solves) against the reference's own code compiled from
 * /root/reference (oracle/_ref, see oracle/Makefile) and against the committed fixtures
 * in tests/golden/ that were generated from it (tests/golden/make_golden.py).
 *
 * Reference files: S4 = stage4-mpi+cuda/poisson_mpi_cuda_f.cu, S3 = stage3-openmp+mpi/hybrid.cpp,
 * S0 = stage0/Withoutopenmp1.cpp.
 *
 * Array layout everywhere: flat, i-major, 1-cell halo, idx = li*(ny+2)+lj (S4:723-736).
 */
#ifndef POISSON_ORACLE_H
#define POISSON_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

/* Problem constants; defaults are the reference's compile-time constants (S4:11-13, 32-34). */
typedef struct po_problem {
    double A1, B1, A2, B2; /* rectangle, S4:11-12: -1,1,-0.6,0.6 */
    double f_val;          /* right-hand side inside D, S4:13: 1.0 */
    double ell_k;          /* ellipse x^2 + (k*y)^2 < 1; reference k = 2 (S4:32-34: 4.0*y*y) */
    double eps;            /* <= 0: max(h1,h2)^2 (S4:693) */
    /* Extensions beyond the reference; the defaults reproduce it bit for bit. */
    double ell_ax;         /* x semi-axis: ellipse (x/ax)^2 + (k*y)^2 < 1; <= 0: 1.0 (S4:32-34) */
    int rhs_mode;          /* 0: point indicator (S4:182-191); 1: area fraction of the node's cell (polygon through its edge/ellipse intersections) */
} po_problem;

void po_default_problem(po_problem *pb);

enum { PO_NORM_WEIGHTED = 0 /* stages 1-4: S3:449, S4:659+897 */, PO_NORM_STAGE0 = 1 /* S0:149-154, guard S0:128 */ };

typedef struct po_result {
    int iterations;   /* `iter` as returned by the reference (S4:847-848, 982) */
    int converged;    /* 1 if the delta test fired (S4:899) */
    int denom_break;  /* 1 if the |denom|<1e-15 guard fired (S4:876) */
    double diff_norm; /* last ||w(k+1)-w(k)|| in the selected norm */
    double zr;        /* zr_old as the loop left it */
    double alpha;     /* last alpha */
    double w_norm_E;  /* sqrt(sum w^2 h1 h2) */
    double r_norm_E;  /* sqrt(sum r^2 h1 h2) */
    double w_max;     /* max |w| */
    double err_max_in_D; /* max |w - (1-x^2-4y^2)/10| over nodes in D (analytic solution, reference report) */
    double err_l2_in_D;  /* sqrt(sum over nodes in D of (w-u)^2 h1 h2) */
} po_result;

int po_is_in_D(const po_problem *pb, double x, double y);                          /* S4:32-34 */
double po_segment_length(const po_problem *pb, double c, double s, double e, int vertical); /* S4:46-76 */
void po_choose_process_grid(int size, int *Px, int *Py);                            /* S4:82-86 */
void po_decompose_2d(int M, int N, int Px, int Py, int rank, int *i0, int *i1, int *j0, int *j1); /* S4:97-133 */

/* a, b on the block incl. halo ring; B on the interior, 0 on the ring (S4:146-192). */
void po_setup_local(const po_problem *pb, int M, int N, int i0, int i1, int j0, int j1, double *a, double *b, double *B);
/* D_ij of S4:556-557 on the interior (0 on the ring): what apply_Dinv divides by. */
void po_diag(const double *a, const double *b, double *D, int nx, int ny, double h1, double h2);

void po_apply_A(const double *p, const double *a, const double *b, double *Ap, int nx, int ny, double h1, double h2); /* S4:507-536 */
void po_apply_Dinv(const double *r, const double *a, const double *b, double *z, int nx, int ny, double h1, double h2); /* S4:541-562 */
double po_dot(const double *u, const double *v, int nx, int ny, double h1, double h2); /* S3:181-192 */
/* w += alpha p; r -= alpha Ap; returns sum (w_new-w_old)^2, UNweighted (S3:426-436). */
double po_update_w_r(double *w, double *r, const double *p, const double *Ap, double alpha, int nx, int ny);
void po_update_p(double *p, const double *z, double beta, int nx, int ny); /* S4:663-676 */

/*
 * Whole solve on one rank (S4:688-983 / S3:364-470 / S0:106-172 by norm_mode).
 * w_out: NULL or (M+1)*(N+1) doubles, boundary zero. trace: NULL or 3*trace_len doubles
 * (alpha_k, beta_k, diff_k). fixed_iters > 0: run exactly that many iterations, the delta
 * test is evaluated but does not stop the loop (throughput runs).
 */
int po_solve(const po_problem *pb, int M, int N, double delta, long max_iter, int norm_mode, int fixed_iters,
             po_result *res, double *w_out, double *trace, int trace_len);

/*
 * The same solve on P = Px*Py emulated ranks in one process: per-rank blocks from
 * po_decompose_2d, halo exchange of p as S3:249-355, scalar sums combined in rank order
 * (what MPI_Allreduce does in oracle/shim). Px<=0: po_choose_process_grid(P).
 */
int po_solve_blocks(const po_problem *pb, int M, int N, double delta, long max_iter, int norm_mode, int P, int Px,
                    int Py, po_result *res, double *w_out);

#ifdef __cplusplus
}
#endif
#endif
