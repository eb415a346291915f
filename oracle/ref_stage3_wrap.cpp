/*
 * ref_stage3_wrap.cpp -- exposes the UNMODIFIED reference stage3 translation unit
 * (/root/reference/stage3-openmp+mpi/hybrid.cpp) as a shared library + launcher.
 *
 * TEST INFRASTRUCTURE ONLY. No reference source is copied: the file is #include'd from
 * where it lies (the Makefile passes -I/root/reference/stage3-openmp+mpi) with its
 * `main` renamed, and compiled against oracle/shim/mpi.h. Output goes to oracle/_ref/.
 *
 * What is exported (all extern "C", flat row-major arrays of (nx+2)*(ny+2) doubles,
 * pitch ny+2, exactly the layout of poisson_mpi_cuda_f.cu:723-736):
 *   ref3_setup        -> fic_reg_local      (hybrid.cpp:126-175)
 *   ref3_apply_A      -> mat_A_local        (hybrid.cpp:200-220)
 *   ref3_apply_Dinv   -> mat_D              (hybrid.cpp:226-240)
 *   ref3_dot          -> dot_local          (hybrid.cpp:181-192)
 *   ref3_grid/decomp  -> choose_process_grid / decompose_2d (hybrid.cpp:61-112)
 *   ref3_solve        -> solve_mpi          (hybrid.cpp:364-470), P ranks as threads
 *   ref3_solve_traced -> the same loop (hybrid.cpp:398-467) driven from here through the
 *                        reference's own operators so that w, the norms and the
 *                        alpha/beta/diff traces can be read out (solve_mpi keeps them
 *                        private and frees them). 1 rank only.
 *   ref3_main         -> the reference `main` (hybrid.cpp:473-512) on P thread-ranks.
 */
#include <mpi.h>

#define main reference_stage3_main
#include "hybrid.cpp"
#undef main

#include <cstdint>

namespace {

Matrix to_matrix(const double *flat, int nx, int ny) {
    Matrix m(nx + 2, std::vector<double>(ny + 2, 0.0));
    if (flat)
        for (int li = 0; li <= nx + 1; ++li)
            for (int lj = 0; lj <= ny + 1; ++lj) m[li][lj] = flat[(size_t)li * (ny + 2) + lj];
    return m;
}
void to_flat(const Matrix &m, double *flat, int nx, int ny) {
    for (int li = 0; li <= nx + 1; ++li)
        for (int lj = 0; lj <= ny + 1; ++lj) flat[(size_t)li * (ny + 2) + lj] = m[li][lj];
}

} // namespace

extern "C" {

void ref3_grid(int size, int *Px, int *Py) { choose_process_grid(size, *Px, *Py); }

void ref3_decomp(int M, int N, int Px, int Py, int rank, int *i0, int *i1, int *j0, int *j1) {
    decompose_2d(M, N, Px, Py, rank, *i0, *i1, *j0, *j1);
}

void ref3_setup(int M, int N, double h1, double h2, double eps, int i0, int i1, int j0, int j1, double *a, double *b,
                double *B) {
    int nx = i1 - i0 + 1, ny = j1 - j0 + 1;
    Matrix ma = to_matrix(nullptr, nx, ny), mb = ma, mB = ma;
    fic_reg_local(ma, mb, mB, M, N, h1, h2, eps, i0, i1, j0, j1, nx, ny);
    to_flat(ma, a, nx, ny);
    to_flat(mb, b, nx, ny);
    to_flat(mB, B, nx, ny);
}

void ref3_apply_A(const double *p, const double *a, const double *b, double *Ap, int nx, int ny, double h1, double h2) {
    Matrix mp = to_matrix(p, nx, ny), ma = to_matrix(a, nx, ny), mb = to_matrix(b, nx, ny);
    Matrix mAp = to_matrix(nullptr, nx, ny);
    mat_A_local(mp, ma, mb, mAp, nx, ny, h1, h2);
    to_flat(mAp, Ap, nx, ny);
}

void ref3_apply_Dinv(const double *r, const double *a, const double *b, double *z, int nx, int ny, double h1,
                     double h2) {
    Matrix mr = to_matrix(r, nx, ny), ma = to_matrix(a, nx, ny), mb = to_matrix(b, nx, ny);
    Matrix mz = to_matrix(nullptr, nx, ny);
    mat_D(mr, ma, mb, mz, nx, ny, h1, h2);
    to_flat(mz, z, nx, ny);
}

double ref3_dot(const double *u, const double *v, int nx, int ny, double h1, double h2) {
    Matrix mu = to_matrix(u, nx, ny), mv = to_matrix(v, nx, ny);
    return dot_local(mu, mv, nx, ny, h1, h2);
}

/* Unmodified solve_mpi on `ranks` thread-ranks; returns rank 0's iteration count. */
int ref3_solve(int M, int N, double delta, int max_iter, int ranks) {
    int iters = 0;
    if (ranks <= 1) return solve_mpi(M, N, delta, max_iter, 0, 1, MPI_COMM_WORLD);
    shim_mpi_launch(ranks, [&](int rank) {
        int it = solve_mpi(M, N, delta, max_iter, rank, ranks, MPI_COMM_WORLD);
        if (rank == 0) iters = it;
    });
    return iters;
}

/*
 * 1-rank PCG driven through the reference operators, statement for statement as
 * hybrid.cpp:367-469, with read-outs. out[0..7] = iterations, ||w||_E, max|w|,
 * last ||dw||_E, (z,r) at exit (zr_old as left by the loop), ||r||_E, last alpha, converged.
 * w_out (may be NULL): (M+1)*(N+1) doubles. trace (may be NULL): 3*trace_len doubles
 * (alpha_k, beta_k, diff_k for k = 1..trace_len; beta_k is 0 on the converging iteration).
 */
int ref3_solve_traced(int M, int N, double delta, int max_iter, double *out, double *w_out, double *trace,
                      int trace_len) {
    double h1 = (B1 - A1) / M;
    double h2 = (B2 - A2) / N;
    double eps = std::max(h1, h2) * std::max(h1, h2);
    int nx = M - 1, ny = N - 1;
    Matrix a = to_matrix(nullptr, nx, ny), b = a, B = a;
    fic_reg_local(a, b, B, M, N, h1, h2, eps, 1, nx, 1, ny, nx, ny);
    Matrix w = to_matrix(nullptr, nx, ny), r = B, z = w, p = w, Ap = w;
    mat_D(r, a, b, z, nx, ny, h1, h2);
    p = z;
    double zr_old = dot_local(z, r, nx, ny, h1, h2);
    int iter = 0, converged = 0;
    double last_diff = 0.0, last_alpha = 0.0;
    for (int k = 1; k <= max_iter; ++k) {
        iter = k;
        exchange_halos_2d(p, nx, ny, 1, 1, 0, MPI_COMM_WORLD);
        mat_A_local(p, a, b, Ap, nx, ny, h1, h2);
        double denom = dot_local(Ap, p, nx, ny, h1, h2);
        if (std::abs(denom) < 1e-15) break;
        double alpha = zr_old / denom;
        last_alpha = alpha;
        double local_diff = 0.0;
#pragma omp parallel for collapse(2) reduction(+ : local_diff)
        for (int li = 1; li <= nx; ++li) {
            for (int lj = 1; lj <= ny; ++lj) {
                double w_old = w[li][lj];
                w[li][lj] = w_old + alpha * p[li][lj];
                r[li][lj] -= alpha * Ap[li][lj];
                double diff = w[li][lj] - w_old;
                local_diff += diff * diff;
            }
        }
        mat_D(r, a, b, z, nx, ny, h1, h2);
        double zr_new = dot_local(z, r, nx, ny, h1, h2);
        double global_diff = std::sqrt(local_diff * h1 * h2);
        last_diff = global_diff;
        if (trace && k <= trace_len) {
            trace[3 * (k - 1) + 0] = alpha;
            trace[3 * (k - 1) + 1] = 0.0;
            trace[3 * (k - 1) + 2] = global_diff;
        }
        if (global_diff < delta) {
            converged = 1;
            break;
        }
        double beta = zr_new / zr_old;
        zr_old = zr_new;
        if (trace && k <= trace_len) trace[3 * (k - 1) + 1] = beta;
#pragma omp parallel for collapse(2)
        for (int li = 1; li <= nx; ++li)
            for (int lj = 1; lj <= ny; ++lj) p[li][lj] = z[li][lj] + beta * p[li][lj];
    }
    double wmax = 0.0;
    for (int li = 1; li <= nx; ++li)
        for (int lj = 1; lj <= ny; ++lj) wmax = std::max(wmax, std::abs(w[li][lj]));
    out[0] = iter;
    out[1] = std::sqrt(dot_local(w, w, nx, ny, h1, h2));
    out[2] = wmax;
    out[3] = last_diff;
    out[4] = zr_old;
    out[5] = std::sqrt(dot_local(r, r, nx, ny, h1, h2));
    out[6] = last_alpha;
    out[7] = converged;
    if (w_out) to_flat(w, w_out, nx, ny);
    return iter;
}

/* The reference's own main() on `ranks` thread-ranks (what `mpirun -np ranks` would do). */
int ref3_main(int M, int N, int ranks) {
    std::string sm = std::to_string(M), sn = std::to_string(N);
    int rc = 0;
    auto run = [&](int) {
        char prog[] = "hybrid";
        std::vector<char> am(sm.begin(), sm.end()), an(sn.begin(), sn.end());
        am.push_back(0);
        an.push_back(0);
        char *argv[] = {prog, am.data(), an.data(), nullptr};
        int r = reference_stage3_main(3, argv);
        if (r) rc = r;
    };
    if (ranks <= 1) run(0);
    else shim_mpi_launch(ranks, run);
    return rc;
}

} // extern "C"

#ifdef REF_STAGE3_MAIN
/* Launcher: `ref_stage3 [M N]`, ranks from $SHIM_MPI_RANKS (default 1), threads from $OMP_NUM_THREADS. */
int main(int argc, char **argv) {
    const char *e = std::getenv("SHIM_MPI_RANKS");
    int ranks = e ? std::atoi(e) : 1;
    if (ranks <= 1) return reference_stage3_main(argc, argv);
    int rc = 0;
    shim_mpi_launch(ranks, [&](int) { rc |= reference_stage3_main(argc, argv); });
    return rc;
}
#endif
