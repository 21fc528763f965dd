/* C/OpenMP restatement of Open3D's correspondence RANSAC.  TEST INFRASTRUCTURE ONLY.
 *
 * PARITY UNPINNED (see oracle/ransac_np.py for the statement of the algorithm and why):
 * restates open3d::pipelines::registration::RegistrationRANSACBasedOnCorrespondence as the
 * reference calls it at utils/eval_pose.py:95-97 (ransac_n=10, point-to-point, no scaling,
 * no checkers, max_iteration 100000, confidence 0.999), driven by EXPLICIT minimal sets.
 *
 * Hypotheses are evaluated in waves across OpenMP threads and folded in hypothesis order, so
 * the result equals the sequential scan regardless of the thread count (the reference's own
 * fold is thread-order dependent, README.md:260).
 *
 * Build: gcc -O3 -fopenmp -shared -fPIC -o oracle/_build/liboracle_ransac.so oracle/ransac_c.c -lm
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* One-sided (Hestenes) Jacobi SVD of a 3x3 matrix held column-wise in a[col][row].
 * On exit a[:,j] = sigma_j u_j and v[:,j] = v_j (v[col][row]). */
static void jacobi3(double a[3][3], double v[3][3]) {
    memset(v, 0, 9 * sizeof(double));
    v[0][0] = v[1][1] = v[2][2] = 1.0;
    for (int sweep = 0; sweep < 30; ++sweep) {
        int rotated = 0;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                double al = 0, be = 0, ga = 0;
                for (int r = 0; r < 3; ++r) {
                    al += a[p][r] * a[p][r];
                    be += a[q][r] * a[q][r];
                    ga += a[p][r] * a[q][r];
                }
                if (ga == 0.0 || fabs(ga) <= 1e-300 + 2.3e-16 * sqrt(al * be)) continue;
                rotated = 1;
                double zeta = (be - al) / (2.0 * ga);
                double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
                for (int r = 0; r < 3; ++r) {
                    double x = a[p][r], y = a[q][r];
                    a[p][r] = c * x - s * y;
                    a[q][r] = s * x + c * y;
                    x = v[p][r]; y = v[q][r];
                    v[p][r] = c * x - s * y;
                    v[q][r] = s * x + c * y;
                }
            }
        if (!rotated) break;
    }
}

static void cross3(const double* x, const double* y, double* o) {
    o[0] = x[1] * y[2] - x[2] * y[1];
    o[1] = x[2] * y[0] - x[0] * y[2];
    o[2] = x[0] * y[1] - x[1] * y[0];
}

/* Umeyama without scale on n gathered pairs; T row-major 4x4. */
static void fit(const double* src, const double* tgt, const int32_t* idx, int n, double* T) {
    double ms[3] = {0, 0, 0}, mt[3] = {0, 0, 0};
    for (int i = 0; i < n; ++i)
        for (int r = 0; r < 3; ++r) {
            ms[r] += src[3 * (size_t)idx[i] + r];
            mt[r] += tgt[3 * (size_t)idx[i] + r];
        }
    for (int r = 0; r < 3; ++r) { ms[r] /= n; mt[r] /= n; }
    double a[3][3] = {{0}}; /* a[col][row] of Sigma = 1/n sum (t-mt)(s-ms)^T */
    for (int i = 0; i < n; ++i) {
        double ds[3], dt[3];
        for (int r = 0; r < 3; ++r) {
            ds[r] = src[3 * (size_t)idx[i] + r] - ms[r];
            dt[r] = tgt[3 * (size_t)idx[i] + r] - mt[r];
        }
        for (int c = 0; c < 3; ++c)
            for (int r = 0; r < 3; ++r) a[c][r] += dt[r] * ds[c];
    }
    for (int c = 0; c < 3; ++c)
        for (int r = 0; r < 3; ++r) a[c][r] /= n;
    double v[3][3];
    jacobi3(a, v);
    /* order columns by decreasing norm */
    double nr[3];
    int o[3] = {0, 1, 2};
    for (int c = 0; c < 3; ++c) nr[c] = sqrt(a[c][0] * a[c][0] + a[c][1] * a[c][1] + a[c][2] * a[c][2]);
    for (int i = 0; i < 2; ++i)
        for (int j = i + 1; j < 3; ++j)
            if (nr[o[j]] > nr[o[i]]) { int tmp = o[i]; o[i] = o[j]; o[j] = tmp; }
    double u1[3], u2[3], u3[3], v1[3], v2[3], v3[3], vx[3];
    for (int r = 0; r < 3; ++r) { v1[r] = v[o[0]][r]; v2[r] = v[o[1]][r]; v3[r] = v[o[2]][r]; }
    double n1 = nr[o[0]], n2 = nr[o[1]];
    if (n1 > 0) for (int r = 0; r < 3; ++r) u1[r] = a[o[0]][r] / n1;
    else { u1[0] = 1; u1[1] = 0; u1[2] = 0; }
    if (n2 > 1e-300 && n2 > 1e-14 * n1) {
        for (int r = 0; r < 3; ++r) u2[r] = a[o[1]][r] / n2;
        /* re-orthogonalise against u1 (columns are orthogonal to ~eps already) */
        double d = u1[0] * u2[0] + u1[1] * u2[1] + u1[2] * u2[2];
        for (int r = 0; r < 3; ++r) u2[r] -= d * u1[r];
    } else { /* rank <= 1: any unit vector orthogonal to u1 */
        int k = fabs(u1[0]) < fabs(u1[1]) ? (fabs(u1[0]) < fabs(u1[2]) ? 0 : 2)
                                          : (fabs(u1[1]) < fabs(u1[2]) ? 1 : 2);
        double e[3] = {0, 0, 0};
        e[k] = 1;
        cross3(u1, e, u2);
    }
    double l2 = sqrt(u2[0] * u2[0] + u2[1] * u2[1] + u2[2] * u2[2]);
    for (int r = 0; r < 3; ++r) u2[r] /= l2;
    cross3(u1, u2, u3);
    cross3(v1, v2, vx);
    double detv = vx[0] * v3[0] + vx[1] * v3[1] + vx[2] * v3[2]; /* +-1 */
    double sg = detv < 0 ? -1.0 : 1.0;
    /* R = u1 v1^T + u2 v2^T + det(V) (u1 x u2) v3^T  ==  U diag(1,1,det U det V) V^T */
    double R[3][3];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) R[r][c] = u1[r] * v1[c] + u2[r] * v2[c] + sg * u3[r] * v3[c];
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) T[4 * r + c] = R[r][c];
        T[4 * r + 3] = mt[r] - (R[r][0] * ms[0] + R[r][1] * ms[1] + R[r][2] * ms[2]);
    }
    T[12] = T[13] = T[14] = 0;
    T[15] = 1;
}

static void score(const double* T, const double* src, const double* tgt, int M, double thr2,
                  int32_t* good, double* sum) {
    int32_t g = 0;
    double s = 0;
    for (int i = 0; i < M; ++i) {
        const double* p = src + 3 * (size_t)i;
        const double* q = tgt + 3 * (size_t)i;
        double e0 = ((T[0] * p[0] + T[1] * p[1]) + T[2] * p[2]) + T[3] - q[0];
        double e1 = ((T[4] * p[0] + T[5] * p[1]) + T[6] * p[2]) + T[7] - q[1];
        double e2 = ((T[8] * p[0] + T[9] * p[1]) + T[10] * p[2]) + T[11] - q[2];
        double d2 = e0 * e0 + e1 * e1 + e2 * e2;
        if (d2 < thr2) { ++g; s += d2; }
    }
    *good = g;
    *sum = s;
}

static int budget_from_count(int good, int M, int n, double conf, int cap) {
    if (conf >= 1.0 || good <= 0) return cap;
    double w = (double)good / (double)M;
    double denom = log(1.0 - pow(w, (double)n));
    if (denom == 0.0) return cap; /* Open3D divides by zero here (UB int cast); we keep the budget */
    double k = log(1.0 - conf) / denom;
    return k < cap ? (int)ceil(k) : cap;
}

/* Fit only: T_all [H,16]. */
int orc_fit_all(const double* src, const double* tgt, const int32_t* ms, int H, int n, double* T_all) {
#pragma omp parallel for schedule(static)
    for (int h = 0; h < H; ++h) fit(src, tgt, ms + (size_t)h * n, n, T_all + 16 * (size_t)h);
    return 0;
}

/* Full RANSAC.  counts/sums (length H) may be NULL.  Returns number of hypotheses evaluated
 * in sequential order (h_stop). */
int orc_ransac(const double* src, const double* tgt, int M, const int32_t* ms, int H, int n,
               double max_dist, int adaptive, double confidence, double* T_out, int32_t* best_h,
               int32_t* inliers, double* sum_d2, int32_t* counts, double* sums) {
    for (int i = 0; i < 16; ++i) T_out[i] = (i % 5 == 0) ? 1.0 : 0.0;
    *best_h = -1;
    *inliers = 0;
    *sum_d2 = 0.0;
    if (n < 3 || M < n || max_dist <= 0.0) return 0;
    const double thr2 = max_dist * max_dist;
    int wave = 64;
#ifdef _OPENMP
    wave = 8 * omp_get_max_threads();
#endif
    double* Tw = (double*)malloc(sizeof(double) * 16 * wave);
    int32_t* gw = (int32_t*)malloc(sizeof(int32_t) * wave);
    double* sw = (double*)malloc(sizeof(double) * wave);
    double bfit = 0.0, brmse = 0.0;
    int budget = H, h0 = 0, stop = 0, n_eval = -1;
    while (h0 < budget && !stop) {
        int w = wave < budget - h0 ? wave : budget - h0;
#pragma omp parallel for schedule(dynamic, 1)
        for (int j = 0; j < w; ++j) {
            fit(src, tgt, ms + (size_t)(h0 + j) * n, n, Tw + 16 * j);
            score(Tw + 16 * j, src, tgt, M, thr2, gw + j, sw + j);
        }
        for (int j = 0; j < w; ++j) {
            int h = h0 + j;
            if (h >= budget) { stop = 1; n_eval = h; break; }
            if (counts) counts[h] = gw[j];
            if (sums) sums[h] = sw[j];
            double f = (double)gw[j] / (double)M;
            double rm = gw[j] > 0 ? sqrt(sw[j] / (double)gw[j]) : 0.0;
            if (f > bfit || (f == bfit && rm < brmse)) {
                bfit = f;
                brmse = rm;
                memcpy(T_out, Tw + 16 * j, 16 * sizeof(double));
                *best_h = h;
                *inliers = gw[j];
                *sum_d2 = sw[j];
                if (adaptive) {
                    int b = budget_from_count(gw[j], M, n, confidence, budget);
                    if (b < budget) budget = b;
                }
            }
        }
        h0 += w;
    }
    free(Tw);
    free(gw);
    free(sw);
    if (n_eval >= 0) return n_eval; /* first hypothesis index at or beyond the budget */
    return h0 < budget ? h0 : budget;
}

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
