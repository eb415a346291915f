/*
 * poisson_oracle.c -- CPU restatement of the reference algorithm. See poisson_oracle.h
 * for status ("PINNED" against oracle/_ref and tests/golden) and the layout convention.
 *
 * TEST INFRASTRUCTURE ONLY -- never linked into or called by the product library.
 *
 * Built with -ffp-contract=off and no -mfma: every a*b+c below is two roundings, as in the
 * reference's CPU stages built with g++ for x86-64.
 * Point-wise loops carry OpenMP pragmas; reductions are summed per row (in j order) and the
 * row sums are then added in i order, so results do not depend on the thread count.
 */
#include "poisson_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define IDX(li, lj, ny) ((size_t)(li) * (size_t)((ny) + 2) + (size_t)(lj))

static double dmax(double a, double b) { return (a < b) ? b : a; } /* std::max */
static double dmin(double a, double b) { return (b < a) ? b : a; } /* std::min */

void po_default_problem(po_problem *pb) {
    pb->A1 = -1.0; /* S4:11 */
    pb->B1 = 1.0;
    pb->A2 = -0.6; /* S4:12 */
    pb->B2 = 0.6;
    pb->f_val = 1.0; /* S4:13 */
    pb->ell_k = 2.0; /* S4:33: x*x + 4.0*y*y */
    pb->eps = 0.0;   /* S4:693 */
    pb->ell_ax = 1.0; /* S4:33: x*x */
    pb->rhs_mode = 0; /* S4:182-191 */
}

static double ell_ax(const po_problem *pb) { return pb->ell_ax > 0.0 ? pb->ell_ax : 1.0; }

/* S4:32-34 (S3:19-21): x*x + 4.0*y*y < 1.0; 4.0 = k*k, product evaluated left to right. With the x semi-axis ax of
 * the generalised ellipse, x*x is scaled by 1/(ax*ax): 1.0 for the reference's ax = 1, an exact product. */
int po_is_in_D(const po_problem *pb, double x, double y) {
    double c = pb->ell_k * pb->ell_k;
    double inv_ax2 = 1.0 / (ell_ax(pb) * ell_ax(pb));
    return (x * x * inv_ax2 + c * y * y < 1.0);
}

/* S4:46-76 (S3:33-55). */
double po_segment_length(const po_problem *pb, double cc, double s, double e, int vertical) {
    double c = pb->ell_k * pb->ell_k;
    double ax = ell_ax(pb), inv_ax2 = 1.0 / (ax * ax); /* reference: 1, 1 */
    double lij;
    if (vertical) {
        double x0 = cc;
        if (fabs(x0) >= ax) lij = 0.0; /* S4:51: abs(x0) >= 1.0 */
        else {
            double root = sqrt(dmax(0.0, (1.0 - x0 * x0 * inv_ax2) / c)); /* S4:53-54 */
            double y_min = -root, y_max = root;
            lij = dmax(0.0, dmin(e, y_max) - dmax(s, y_min)); /* S4:55 */
        }
    } else {
        double y0 = cc;
        if (fabs(pb->ell_k * y0) >= 1.0) lij = 0.0; /* S4:59: abs(2.0*y0) >= 1.0 */
        else {
            double root = ax * sqrt(dmax(0.0, 1.0 - c * y0 * y0)); /* S4:61-62 */
            double x_min = -root, x_max = root;
            lij = dmax(0.0, dmin(e, x_max) - dmax(s, x_min)); /* S4:63 */
        }
    }
    return lij;
}

/* S4:82-86: Px = largest divisor of size not above sqrt(size). */
void po_choose_process_grid(int size, int *Px, int *Py) {
    int px = (int)sqrt((double)size);
    while (px > 1 && size % px != 0) --px;
    *Px = px;
    *Py = size / px;
}

/* S4:97-133: balanced ranges of interior indices, remainder to the low coordinates. */
void po_decompose_2d(int M, int N, int Px, int Py, int rank, int *i0, int *i1, int *j0, int *j1) {
    int px = rank % Px, py = rank / Px;
    int base_i = (M - 1) / Px, rem_i = (M - 1) % Px;
    int base_j = (N - 1) / Py, rem_j = (N - 1) % Py;
    int off_i = 1, off_j = 1, k;
    for (k = 0; k < px; ++k) off_i += base_i + (k < rem_i ? 1 : 0);
    for (k = 0; k < py; ++k) off_j += base_j + (k < rem_j ? 1 : 0);
    *i0 = off_i;
    *i1 = off_i + base_i + (px < rem_i ? 1 : 0) - 1;
    *j0 = off_j;
    *j1 = off_j + base_j + (py < rem_j ? 1 : 0) - 1;
}

static void grid_steps(const po_problem *pb, int M, int N, double *h1, double *h2, double *eps) {
    *h1 = (pb->B1 - pb->A1) / M; /* S4:691 */
    *h2 = (pb->B2 - pb->A2) / N; /* S4:692 */
    *eps = pb->eps > 0.0 ? pb->eps : dmax(*h1, *h2) * dmax(*h1, *h2); /* S4:693 */
}

/* The part [*lo, *hi] of a cell edge that lies in D: the two clamped end points whose difference
 * po_segment_length returns (S4:46-76, same expressions). Returns 0 when the edge misses D. */
static int segment_ends(const po_problem *pb, double cc, double s, double e, int vertical, double *lo, double *hi) {
    double c = pb->ell_k * pb->ell_k;
    double ax = ell_ax(pb), inv_ax2 = 1.0 / (ax * ax);
    double root;
    if (vertical) {
        if (fabs(cc) >= ax) return 0;                               /* S4:51 */
        root = sqrt(dmax(0.0, (1.0 - cc * cc * inv_ax2) / c));      /* S4:53-54 */
    } else {
        if (fabs(pb->ell_k * cc) >= 1.0) return 0;                  /* S4:59 */
        root = ax * sqrt(dmax(0.0, 1.0 - c * cc * cc));             /* S4:61-62 */
    }
    *lo = dmax(s, -root); /* S4:55, 63 */
    *hi = dmin(e, root);
    return *hi > *lo;
}

/* Running shoelace sum over the vertices of a polygon given in order. */
typedef struct shoelace {
    int n;
    double x0, y0, xp, yp, sum;
} shoelace;
static void shoelace_add(shoelace *q, double x, double y) {
    if (q->n == 0) {
        q->x0 = x;
        q->y0 = y;
    } else
        q->sum += q->xp * y - x * q->yp;
    q->xp = x;
    q->yp = y;
    q->n++;
}

/* Area-fraction right-hand side (rhs_mode 1; no reference counterpart): f times the share of the node's cell
 * [x_i - h1/2, x_{i+1} - h1/2] x [y_j - h2/2, y_{j+1} - h2/2] that lies in D. The share is the area of the polygon whose
 * vertices are the end points of the in-D parts of the cell's four edges (the edge intersections of S4:46-76; the
 * cell's right / upper edges are the left / lower edges of nodes (gi+1,gj) / (gi,gj+1)), walked counter-clockwise,
 * in coordinates relative to the node -- the cell clipped by the chords of the ellipse, second-order accurate.
 * Snapped to 0 / 1 with the reference's 1e-9 threshold: cells wholly inside get exactly f. */
static double rhs_area(const po_problem *pb, int gi, int gj, double h1, double h2) {
    double x_i = pb->A1 + gi * h1, y_j = pb->A2 + gj * h2;           /* S4:159-160 */
    double x_ip = pb->A1 + (gi + 1) * h1, y_jp = pb->A2 + (gj + 1) * h2;
    double xl = x_i - 0.5 * h1, xr = x_ip - 0.5 * h1, yb = y_j - 0.5 * h2, yt = y_jp - 0.5 * h2;
    double lo, hi, frac;
    shoelace q = {0, 0.0, 0.0, 0.0, 0.0, 0.0};
    if (segment_ends(pb, yb, xl, xr, 0, &lo, &hi)) { /* bottom edge, left to right */
        shoelace_add(&q, lo - x_i, yb - y_j);
        shoelace_add(&q, hi - x_i, yb - y_j);
    }
    if (segment_ends(pb, xr, yb, yt, 1, &lo, &hi)) { /* right edge, upwards */
        shoelace_add(&q, xr - x_i, lo - y_j);
        shoelace_add(&q, xr - x_i, hi - y_j);
    }
    if (segment_ends(pb, yt, xl, xr, 0, &lo, &hi)) { /* top edge, right to left */
        shoelace_add(&q, hi - x_i, yt - y_j);
        shoelace_add(&q, lo - x_i, yt - y_j);
    }
    if (segment_ends(pb, xl, yb, yt, 1, &lo, &hi)) { /* left edge, downwards */
        shoelace_add(&q, xl - x_i, hi - y_j);
        shoelace_add(&q, xl - x_i, lo - y_j);
    }
    if (q.n == 0) return 0.0;
    q.sum += q.xp * q.y0 - q.x0 * q.yp;
    frac = (0.5 * q.sum) / (h1 * h2);
    frac = (fabs(frac - 1.0) < 1e-9) ? 1.0 : (frac < 1e-9 ? 0.0 : frac);
    return pb->f_val * frac;
}

/* S4:146-192 (S3:126-175). */
void po_setup_local(const po_problem *pb, int M, int N, int i0, int i1, int j0, int j1, double *a, double *b, double *B) {
    double h1, h2, eps;
    int nx = i1 - i0 + 1, ny = j1 - j0 + 1;
    grid_steps(pb, M, N, &h1, &h2, &eps);
    memset(B, 0, sizeof(double) * (size_t)(nx + 2) * (size_t)(ny + 2)); /* S4:715: B starts at 0 */
#pragma omp parallel for
    for (int gi = i0 - 1; gi <= i1 + 1; ++gi) {
        for (int gj = j0 - 1; gj <= j1 + 1; ++gj) {
            int li = gi - (i0 - 1), lj = gj - (j0 - 1);
            double x_i = pb->A1 + gi * h1; /* S4:159 */
            double y_j = pb->A2 + gj * h2; /* S4:160 */
            double la = po_segment_length(pb, x_i - 0.5 * h1, y_j - 0.5 * h2, y_j + 0.5 * h2, 1); /* S4:162-163 */
            double lb = po_segment_length(pb, y_j - 0.5 * h2, x_i - 0.5 * h1, x_i + 0.5 * h1, 0); /* S4:164-165 */
            a[IDX(li, lj, ny)] = (fabs(la - h2) < 1e-9) ? 1.0
                                 : (la < 1e-9 ? 1.0 / eps : (la / h2) + (1.0 - la / h2) / eps); /* S4:167-171 */
            b[IDX(li, lj, ny)] = (fabs(lb - h1) < 1e-9) ? 1.0
                                 : (lb < 1e-9 ? 1.0 / eps : (lb / h1) + (1.0 - lb / h1) / eps); /* S4:173-177 */
            if (gi >= i0 && gi <= i1 && gj >= j0 && gj <= j1)
                B[IDX(li, lj, ny)] = pb->rhs_mode == 1 ? rhs_area(pb, gi, gj, h1, h2)
                                                       : (po_is_in_D(pb, x_i, y_j) ? pb->f_val : 0.0); /* S4:182-191 */
        }
    }
}

/* The divisor of S4:556-557 (S3:235-236). */
void po_diag(const double *a, const double *b, double *D, int nx, int ny, double h1, double h2) {
    memset(D, 0, sizeof(double) * (size_t)(nx + 2) * (size_t)(ny + 2));
#pragma omp parallel for
    for (int li = 1; li <= nx; ++li)
        for (int lj = 1; lj <= ny; ++lj)
            D[IDX(li, lj, ny)] = (a[IDX(li + 1, lj, ny)] + a[IDX(li, lj, ny)]) / (h1 * h1) +
                                 (b[IDX(li, lj + 1, ny)] + b[IDX(li, lj, ny)]) / (h2 * h2);
}

/* S4:507-536 (S3:200-220): expression tree kept verbatim: -1.0 / h * ( a*(dp)/h - a*(dp)/h ). */
void po_apply_A(const double *p, const double *a, const double *b, double *Ap, int nx, int ny, double h1, double h2) {
#pragma omp parallel for
    for (int li = 1; li <= nx; ++li) {
        for (int lj = 1; lj <= ny; ++lj) {
            size_t c = IDX(li, lj, ny), ip = IDX(li + 1, lj, ny), im = IDX(li - 1, lj, ny);
            size_t jp = c + 1, jm = c - 1;
            double Ax = -1.0 / h1 * (a[ip] * (p[ip] - p[c]) / h1 - a[c] * (p[c] - p[im]) / h1);
            double Ay = -1.0 / h2 * (b[jp] * (p[jp] - p[c]) / h2 - b[c] * (p[c] - p[jm]) / h2);
            Ap[c] = Ax + Ay;
        }
    }
}

/* S4:541-562 (S3:226-240). */
void po_apply_Dinv(const double *r, const double *a, const double *b, double *z, int nx, int ny, double h1, double h2) {
#pragma omp parallel for
    for (int li = 1; li <= nx; ++li) {
        for (int lj = 1; lj <= ny; ++lj) {
            size_t c = IDX(li, lj, ny);
            double D = (a[IDX(li + 1, lj, ny)] + a[c]) / (h1 * h1) + (b[c + 1] + b[c]) / (h2 * h2);
            z[c] = (D != 0.0) ? r[c] / D : 0.0;
        }
    }
}

static double sum_rows(const double *rows, int nx) {
    double s = 0.0;
    for (int li = 1; li <= nx; ++li) s += rows[li];
    return s;
}

/* S3:181-192: sum u*v over the interior, then * h1 * h2 (left to right). */
double po_dot(const double *u, const double *v, int nx, int ny, double h1, double h2) {
    double *rows = (double *)malloc(sizeof(double) * (size_t)(nx + 2));
    double s;
#pragma omp parallel for
    for (int li = 1; li <= nx; ++li) {
        double acc = 0.0;
        for (int lj = 1; lj <= ny; ++lj) acc += u[IDX(li, lj, ny)] * v[IDX(li, lj, ny)];
        rows[li] = acc;
    }
    s = sum_rows(rows, nx);
    free(rows);
    return s * h1 * h2;
}

/* S3:426-436 (S4:626-660 without the per-lane h1h2 factor). */
double po_update_w_r(double *w, double *r, const double *p, const double *Ap, double alpha, int nx, int ny) {
    double *rows = (double *)malloc(sizeof(double) * (size_t)(nx + 2));
    double s;
#pragma omp parallel for
    for (int li = 1; li <= nx; ++li) {
        double acc = 0.0;
        for (int lj = 1; lj <= ny; ++lj) {
            size_t c = IDX(li, lj, ny);
            double w_old = w[c];
            double w_new = w_old + alpha * p[c];
            double diff;
            w[c] = w_new;
            r[c] = r[c] - alpha * Ap[c];
            diff = w_new - w_old;
            acc += diff * diff;
        }
        rows[li] = acc;
    }
    s = sum_rows(rows, nx);
    free(rows);
    return s;
}

/* S4:663-676 (S3:463-466). */
void po_update_p(double *p, const double *z, double beta, int nx, int ny) {
#pragma omp parallel for
    for (int li = 1; li <= nx; ++li)
        for (int lj = 1; lj <= ny; ++lj) p[IDX(li, lj, ny)] = z[IDX(li, lj, ny)] + beta * p[IDX(li, lj, ny)];
}

/* ------------------------------------------------------------------------------------ */

typedef struct block {
    int i0, i1, j0, j1, nx, ny, px, py;
    double *a, *b, *B, *w, *r, *z, *p, *Ap;
} block;

static double *zeros(int nx, int ny) { return (double *)calloc((size_t)(nx + 2) * (size_t)(ny + 2), sizeof(double)); }

static void block_alloc(block *bk) {
    bk->a = zeros(bk->nx, bk->ny);
    bk->b = zeros(bk->nx, bk->ny);
    bk->B = zeros(bk->nx, bk->ny);
    bk->w = zeros(bk->nx, bk->ny);
    bk->r = zeros(bk->nx, bk->ny);
    bk->z = zeros(bk->nx, bk->ny);
    bk->p = zeros(bk->nx, bk->ny);
    bk->Ap = zeros(bk->nx, bk->ny);
}
static void block_free(block *bk) {
    free(bk->a); free(bk->b); free(bk->B); free(bk->w);
    free(bk->r); free(bk->z); free(bk->p); free(bk->Ap);
}

/* S3:249-355 for all ranks at once: neighbour (px+-1, py) shares rows li, (px, py+-1) shares columns lj;
 * whole lines incl. their end cells travel (S3:272-287); physical sides are set to zero (S3:298, 309, 320, 331). */
static void exchange_all(block *bks, int Px, int Py) {
    for (int rnk = 0; rnk < Px * Py; ++rnk) {
        block *me = &bks[rnk];
        int nx = me->nx, ny = me->ny;
        if (me->px > 0) {
            block *nb = &bks[rnk - 1];
            for (int lj = 0; lj <= ny + 1; ++lj) me->p[IDX(0, lj, ny)] = nb->p[IDX(nb->nx, lj, nb->ny)];
        } else
            for (int lj = 0; lj <= ny + 1; ++lj) me->p[IDX(0, lj, ny)] = 0.0;
        if (me->px < Px - 1) {
            block *nb = &bks[rnk + 1];
            for (int lj = 0; lj <= ny + 1; ++lj) me->p[IDX(nx + 1, lj, ny)] = nb->p[IDX(1, lj, nb->ny)];
        } else
            for (int lj = 0; lj <= ny + 1; ++lj) me->p[IDX(nx + 1, lj, ny)] = 0.0;
    }
    /* second sweep for the j-direction so that it sees this iteration's i-halos exactly as the
     * reference does not: the reference packs all four buffers BEFORE any unpack (S3:271-288), so
     * the end cells of a travelling line are the sender's OLD halo values. They are never read
     * by the 5-point stencil (SURVEY.md 5 "race detection" row); only interior cells matter here. */
    for (int rnk = 0; rnk < Px * Py; ++rnk) {
        block *me = &bks[rnk];
        int nx = me->nx, ny = me->ny;
        if (me->py > 0) {
            block *nb = &bks[rnk - Px];
            for (int li = 1; li <= nx; ++li) me->p[IDX(li, 0, ny)] = nb->p[IDX(li, nb->ny, nb->ny)];
        } else
            for (int li = 0; li <= nx + 1; ++li) me->p[IDX(li, 0, ny)] = 0.0;
        if (me->py < Py - 1) {
            block *nb = &bks[rnk + Px];
            for (int li = 1; li <= nx; ++li) me->p[IDX(li, ny + 1, ny)] = nb->p[IDX(li, 1, nb->ny)];
        } else
            for (int li = 0; li <= nx + 1; ++li) me->p[IDX(li, ny + 1, ny)] = 0.0;
    }
}

static void finish(const po_problem *pb, block *bks, int P, int M, int N, double h1, double h2, po_result *res,
                   double *w_out) {
    double ww = 0.0, rr = 0.0, wmax = 0.0, emax = 0.0, e2 = 0.0;
    double kk = pb->ell_k * pb->ell_k;
    double inv_ax2 = 1.0 / (ell_ax(pb) * ell_ax(pb));
    double scale = pb->f_val / (2.0 * inv_ax2 + 2.0 * kk); /* u = (1 - x^2 - 4y^2)/10 for k = 2, ax = 1, f = 1 (R4 p.1-2) */
    if (w_out) memset(w_out, 0, sizeof(double) * (size_t)(M + 1) * (size_t)(N + 1));
    for (int rnk = 0; rnk < P; ++rnk) {
        block *bk = &bks[rnk];
        ww += po_dot(bk->w, bk->w, bk->nx, bk->ny, h1, h2);
        rr += po_dot(bk->r, bk->r, bk->nx, bk->ny, h1, h2);
        for (int li = 1; li <= bk->nx; ++li)
            for (int lj = 1; lj <= bk->ny; ++lj) {
                int gi = bk->i0 + li - 1, gj = bk->j0 + lj - 1;
                double v = bk->w[IDX(li, lj, bk->ny)];
                double x = pb->A1 + gi * h1, y = pb->A2 + gj * h2;
                if (fabs(v) > wmax) wmax = fabs(v);
                if (po_is_in_D(pb, x, y)) {
                    double e = fabs(v - scale * (1.0 - x * x * inv_ax2 - kk * y * y));
                    if (e > emax) emax = e;
                    e2 += e * e;
                }
                if (w_out) w_out[(size_t)gi * (size_t)(N + 1) + (size_t)gj] = v;
            }
    }
    res->w_norm_E = sqrt(ww);
    res->r_norm_E = sqrt(rr);
    res->w_max = wmax;
    res->err_max_in_D = emax;
    res->err_l2_in_D = sqrt(e2 * h1 * h2);
}

static int solve_impl(const po_problem *pb, int M, int N, double delta, long max_iter, int norm_mode, int fixed_iters,
                      int P, int Px, int Py, po_result *res, double *w_out, double *trace, int trace_len) {
    double h1, h2, eps;
    block *bks;
    double zr_old = 0.0;
    int iter = 0;
    grid_steps(pb, M, N, &h1, &h2, &eps);
    if (max_iter <= 0) max_iter = (long)(M - 1) * (long)(N - 1); /* S4:1007 */
    if (fixed_iters > 0) max_iter = fixed_iters;
    memset(res, 0, sizeof(*res));
    bks = (block *)calloc((size_t)P, sizeof(block));
    for (int rnk = 0; rnk < P; ++rnk) {
        block *bk = &bks[rnk];
        po_decompose_2d(M, N, Px, Py, rnk, &bk->i0, &bk->i1, &bk->j0, &bk->j1);
        bk->nx = bk->i1 - bk->i0 + 1;
        bk->ny = bk->j1 - bk->j0 + 1;
        bk->px = rnk % Px;
        bk->py = rnk / Px;
        block_alloc(bk);
        po_setup_local(pb, M, N, bk->i0, bk->i1, bk->j0, bk->j1, bk->a, bk->b, bk->B);
        memcpy(bk->r, bk->B, sizeof(double) * (size_t)(bk->nx + 2) * (size_t)(bk->ny + 2)); /* S4:813: r = B */
        po_apply_Dinv(bk->r, bk->a, bk->b, bk->z, bk->nx, bk->ny, h1, h2);                  /* S4:820 */
        for (int li = 1; li <= bk->nx; ++li)                                                 /* S4:830: p = z */
            memcpy(&bk->p[IDX(li, 1, bk->ny)], &bk->z[IDX(li, 1, bk->ny)], sizeof(double) * (size_t)bk->ny);
        zr_old += po_dot(bk->z, bk->r, bk->nx, bk->ny, h1, h2);                              /* S4:838-843 */
    }
    for (long k = 1; k <= max_iter; ++k) {
        double denom = 0.0, diff2 = 0.0, zr_new = 0.0, alpha, beta, diff;
        iter = (int)k; /* S4:848 */
        exchange_all(bks, Px, Py); /* S4:851 */
        for (int rnk = 0; rnk < P; ++rnk) {
            block *bk = &bks[rnk];
            po_apply_A(bk->p, bk->a, bk->b, bk->Ap, bk->nx, bk->ny, h1, h2); /* S4:857 */
            denom += po_dot(bk->Ap, bk->p, bk->nx, bk->ny, h1, h2);         /* S4:866-872 */
        }
        if (norm_mode == PO_NORM_STAGE0 ? (denom < 1e-15) : (fabs(denom) < 1e-15)) { /* S0:128 | S4:876 */
            res->denom_break = 1;
            break;
        }
        alpha = zr_old / denom; /* S4:877 */
        for (int rnk = 0; rnk < P; ++rnk) {
            block *bk = &bks[rnk];
            diff2 += po_update_w_r(bk->w, bk->r, bk->p, bk->Ap, alpha, bk->nx, bk->ny); /* S3:426-436 */
            po_apply_Dinv(bk->r, bk->a, bk->b, bk->z, bk->nx, bk->ny, h1, h2);          /* S3:439 */
            zr_new += po_dot(bk->z, bk->r, bk->nx, bk->ny, h1, h2);                     /* S3:442-444 */
        }
        diff = (norm_mode == PO_NORM_STAGE0) ? sqrt(diff2) : sqrt(diff2 * h1 * h2); /* S0:154 | S3:449 */
        res->diff_norm = diff;
        res->alpha = alpha;
        if (trace && k <= trace_len) {
            trace[3 * (k - 1) + 0] = alpha;
            trace[3 * (k - 1) + 1] = 0.0;
            trace[3 * (k - 1) + 2] = diff;
        }
        if (diff < delta && fixed_iters <= 0) { /* S4:899 */
            res->converged = 1;
            break;
        }
        beta = zr_new / zr_old; /* S4:931 */
        zr_old = zr_new;
        if (trace && k <= trace_len) trace[3 * (k - 1) + 1] = beta;
        for (int rnk = 0; rnk < P; ++rnk) po_update_p(bks[rnk].p, bks[rnk].z, beta, bks[rnk].nx, bks[rnk].ny); /* S4:936 */
    }
    res->iterations = iter;
    res->zr = zr_old;
    finish(pb, bks, P, M, N, h1, h2, res, w_out);
    for (int rnk = 0; rnk < P; ++rnk) block_free(&bks[rnk]);
    free(bks);
    return iter;
}

int po_solve(const po_problem *pb, int M, int N, double delta, long max_iter, int norm_mode, int fixed_iters,
             po_result *res, double *w_out, double *trace, int trace_len) {
    return solve_impl(pb, M, N, delta, max_iter, norm_mode, fixed_iters, 1, 1, 1, res, w_out, trace, trace_len);
}

int po_solve_blocks(const po_problem *pb, int M, int N, double delta, long max_iter, int norm_mode, int P, int Px,
                    int Py, po_result *res, double *w_out) {
    if (Px <= 0 || Py <= 0) po_choose_process_grid(P, &Px, &Py);
    return solve_impl(pb, M, N, delta, max_iter, norm_mode, 0, Px * Py, Px, Py, res, w_out, NULL, 0);
}
