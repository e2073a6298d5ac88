/* TEST INFRASTRUCTURE ONLY -- plain C restatement (CPU oracle / CPU baseline "port") of
 * the LM refinement hot path of pvphan/camera-calibration.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * this; the product path (libcalib_b200.so) never links or calls it.
 *
 * Parity status: PINNED -- tests/test_oracle.py checks every entry point against
 * tests/golden/*.npz (generated from the unmodified reference by oracle/make_golden.py).
 *
 * Follows (reference file:line):
 *   loop semantics        src/calibrate.py:137-171  (strict '<' accept, lambda/10, lambda*10,
 *                         termination after the update, stale sse)
 *   Marquardt damping     src/calibrate.py:147,152  (lambda * diag(JTJ))
 *   projection            src/calibrate.py:190-197, src/mathutils.py:36-51,174-208,
 *                         src/distortion.py:42-59,78-108 (radtan),198-220 (fisheye)
 *   Jacobian layout       src/jacobian.py:48-85,175-186 (u/v rows interleaved; columns
 *                         alpha,beta,gamma,uc,vc,k...,then rho_x,rho_y,rho_z[deg],tx,ty,tz / view)
 * The reference forms a dense J and inverts the dense JTJ; that cannot exist at the large
 * configurations (154 GB), so this restatement keeps JTJ in its block form (A, B_v, D_v)
 * and solves with a Schur complement over the views.  Same minimiser, same loop rules.
 *
 * Build: gcc -O2 -fopenmp -shared -fPIC -o oracle/liboracle.so oracle/calib_oracle.c -lm
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define LMAX 10
#define DEG 0.017453292519943295

typedef struct {
    double R[9], dRa[9], dRb[9], dRg[9], t[3];
} view_t;

static void mat3(const double* a, const double* b, double* c) {
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            c[3 * i + j] = a[3 * i] * b[j] + a[3 * i + 1] * b[3 + j] + a[3 * i + 2] * b[6 + j];
}

/* R = Rz.Ry.Rx, angles in degrees (mathutils.py:36-51); derivatives carry pi/180 */
static void view_setup(const double* e, view_t* v) {
    double a = e[0] * DEG, b = e[1] * DEG, g = e[2] * DEG;
    double ca = cos(a), sa = sin(a), cb = cos(b), sb = sin(b), cg = cos(g), sg = sin(g);
    double Rx[9] = {1, 0, 0, 0, ca, -sa, 0, sa, ca};
    double Ry[9] = {cb, 0, sb, 0, 1, 0, -sb, 0, cb};
    double Rz[9] = {cg, -sg, 0, sg, cg, 0, 0, 0, 1};
    double dRx[9] = {0, 0, 0, 0, -sa, -ca, 0, ca, -sa};
    double dRy[9] = {-sb, 0, cb, 0, 0, 0, -cb, 0, -sb};
    double dRz[9] = {-sg, -cg, 0, cg, -sg, 0, 0, 0, 0};
    double T[9];
    mat3(Rz, Ry, T); mat3(T, Rx, v->R); mat3(T, dRx, v->dRa);
    mat3(Rz, dRy, T); mat3(T, Rx, v->dRb);
    mat3(dRz, Ry, T); mat3(T, Rx, v->dRg);
    for (int i = 0; i < 9; ++i) { v->dRa[i] *= DEG; v->dRb[i] *= DEG; v->dRg[i] *= DEG; }
    v->t[0] = e[3]; v->t[1] = e[4]; v->t[2] = e[5];
}

/* One point.  uv[2]; if Jc != NULL also Jc[2*L] (row u then row v) and Jv[2*6]. */
static void point_eval(int model, const double* c, const view_t* w, const double* X,
                       double* uv, double* Jc, double* Jv) {
    const int nd = model == 0 ? 5 : 4, L = 5 + nd;
    const double al = c[0], be = c[1], ga = c[2], uc = c[3], vc = c[4];
    const double* k = c + 5;
    double Pc[3];
    for (int i = 0; i < 3; ++i)
        Pc[i] = w->R[3 * i] * X[0] + w->R[3 * i + 1] * X[1] + w->R[3 * i + 2] * X[2] + w->t[i];
    double iz = 1.0 / Pc[2], xn = Pc[0] * iz, yn = Pc[1] * iz;
    double r2 = xn * xn + yn * yn;
    double xd, yd, dxx, dxy, dyy, dxk[5], dyk[5];
    if (model == 0) {
        double r4 = r2 * r2, r6 = r4 * r2;
        double rad = 1 + k[0] * r2 + k[1] * r4 + k[4] * r6;
        double drad = k[0] + 2 * k[1] * r2 + 3 * k[4] * r4;
        double p1 = k[2], p2 = k[3];
        xd = rad * xn + 2 * p1 * xn * yn + p2 * (r2 + 2 * xn * xn);
        yd = rad * yn + p1 * (r2 + 2 * yn * yn) + 2 * p2 * xn * yn;
        dxx = rad + 2 * xn * xn * drad + 2 * p1 * yn + 6 * p2 * xn;
        dxy = 2 * xn * yn * drad + 2 * p1 * xn + 2 * p2 * yn;
        dyy = rad + 2 * yn * yn * drad + 6 * p1 * yn + 2 * p2 * xn;
        dxk[0] = xn * r2; dxk[1] = xn * r4; dxk[2] = 2 * xn * yn; dxk[3] = r2 + 2 * xn * xn; dxk[4] = xn * r6;
        dyk[0] = yn * r2; dyk[1] = yn * r4; dyk[2] = r2 + 2 * yn * yn; dyk[3] = 2 * xn * yn; dyk[4] = yn * r6;
    } else {
        double r = sqrt(r2), th = atan(r), t2 = th * th;
        double t4 = t2 * t2, t6 = t4 * t2, t8 = t4 * t4;
        double q = 1 + k[0] * t2 + k[1] * t4 + k[2] * t6 + k[3] * t8;
        double s = th * q / r;   /* r == 0 -> NaN, as distortion.py:215 */
        double dq = th * (2 * k[0] + 4 * k[1] * t2 + 6 * k[2] * t4 + 8 * k[3] * t6);
        double dsdr = ((q + th * dq) * r / (1 + r2) - th * q) / r2;
        double wv = dsdr / r, tr = th / r;
        xd = s * xn; yd = s * yn;
        dxx = s + xn * xn * wv; dxy = xn * yn * wv; dyy = s + yn * yn * wv;
        dxk[0] = xn * tr * t2; dxk[1] = xn * tr * t4; dxk[2] = xn * tr * t6; dxk[3] = xn * tr * t8;
        dyk[0] = yn * tr * t2; dyk[1] = yn * tr * t4; dyk[2] = yn * tr * t6; dyk[3] = yn * tr * t8;
    }
    uv[0] = al * xd + ga * yd + uc;
    uv[1] = be * yd + vc;
    if (!Jc) return;
    double* Ju = Jc; double* Jw = Jc + L;
    memset(Jc, 0, sizeof(double) * 2 * L);
    Ju[0] = xd; Ju[2] = yd; Ju[3] = 1; Jw[1] = yd; Jw[4] = 1;
    for (int i = 0; i < nd; ++i) { Ju[5 + i] = al * dxk[i] + ga * dyk[i]; Jw[5 + i] = be * dyk[i]; }
    double cu0 = al * dxx + ga * dxy, cu1 = al * dxy + ga * dyy, cv0 = be * dxy, cv1 = be * dyy;
    double gu[3] = {cu0 * iz, cu1 * iz, -(cu0 * xn + cu1 * yn) * iz};
    double gv[3] = {cv0 * iz, cv1 * iz, -(cv0 * xn + cv1 * yn) * iz};
    const double* dR[3] = {w->dRa, w->dRb, w->dRg};
    for (int j = 0; j < 3; ++j) {
        double d[3];
        for (int i = 0; i < 3; ++i)
            d[i] = dR[j][3 * i] * X[0] + dR[j][3 * i + 1] * X[1] + dR[j][3 * i + 2] * X[2];
        Jv[j] = gu[0] * d[0] + gu[1] * d[1] + gu[2] * d[2];
        Jv[6 + j] = gv[0] * d[0] + gv[1] * d[1] + gv[2] * d[2];
    }
    for (int i = 0; i < 3; ++i) { Jv[3 + i] = gu[i]; Jv[9 + i] = gv[i]; }
}

static int nintr(int model) { return model == 0 ? 10 : 9; }

int oracle_project(int model, int64_t M, const int64_t* offs, const double* pts,
                   const double* P, double* y) {
    const int L = nintr(model);
#pragma omp parallel for schedule(dynamic, 4)
    for (int64_t v = 0; v < M; ++v) {
        view_t w; view_setup(P + L + 6 * v, &w);
        for (int64_t i = offs[v]; i < offs[v + 1]; ++i)
            point_eval(model, P, &w, pts + 3 * i, y + 2 * i, NULL, NULL);
    }
    return 0;
}

int oracle_jacobian_blocks(int model, int64_t M, const int64_t* offs, const double* pts,
                           const double* P, double* Jc, double* Jv) {
    const int L = nintr(model);
#pragma omp parallel for schedule(dynamic, 4)
    for (int64_t v = 0; v < M; ++v) {
        view_t w; view_setup(P + L + 6 * v, &w);
        double uv[2];
        for (int64_t i = offs[v]; i < offs[v + 1]; ++i)
            point_eval(model, P, &w, pts + 3 * i, uv, Jc + 2 * L * i, Jv + 12 * i);
    }
    return 0;
}

/* cost: per-view partial sums, then a sequential sum over views (deterministic) */
int oracle_cost(int model, int64_t M, const int64_t* offs, const double* pts,
                const double* sens, const double* P, double* E) {
    const int L = nintr(model);
    double* part = (double*)malloc(sizeof(double) * (M > 0 ? M : 1));
#pragma omp parallel for schedule(dynamic, 4)
    for (int64_t v = 0; v < M; ++v) {
        view_t w; view_setup(P + L + 6 * v, &w);
        double acc = 0, uv[2];
        for (int64_t i = offs[v]; i < offs[v + 1]; ++i) {
            point_eval(model, P, &w, pts + 3 * i, uv, NULL, NULL);
            double du = sens[2 * i] - uv[0], dv = sens[2 * i + 1] - uv[1];
            acc += du * du + dv * dv;
        }
        part[v] = acc;
    }
    double s = 0;
    for (int64_t v = 0; v < M; ++v) s += part[v];
    free(part);
    *E = s;
    return 0;
}

/* A (L*L), B (M*L*6), D (M*36), g (L+6M), E.  Av is scratch M*L*L (per-view A). */
static void normal_eq(int model, int64_t M, const int64_t* offs, const double* pts,
                      const double* sens, const double* P, double* A, double* B, double* D,
                      double* g, double* E, double* Av, double* gcv) {
    const int L = nintr(model);
#pragma omp parallel for schedule(dynamic, 4)
    for (int64_t v = 0; v < M; ++v) {
        view_t w; view_setup(P + L + 6 * v, &w);
        double a[LMAX * LMAX] = {0}, b[LMAX * 6] = {0}, d[36] = {0}, gc[LMAX + 1] = {0}, gv[6] = {0};
        double Jc[2 * LMAX], Jv[12], uv[2];
        for (int64_t i = offs[v]; i < offs[v + 1]; ++i) {
            point_eval(model, P, &w, pts + 3 * i, uv, Jc, Jv);
            double r[2] = {sens[2 * i] - uv[0], sens[2 * i + 1] - uv[1]};
            for (int row = 0; row < 2; ++row) {
                const double* jc = Jc + row * L; const double* jv = Jv + row * 6;
                for (int p = 0; p < L; ++p) {
                    for (int q = 0; q < L; ++q) a[p * L + q] += jc[p] * jc[q];
                    for (int q = 0; q < 6; ++q) b[p * 6 + q] += jc[p] * jv[q];
                    gc[p] += jc[p] * r[row];
                }
                for (int p = 0; p < 6; ++p) {
                    for (int q = 0; q < 6; ++q) d[p * 6 + q] += jv[p] * jv[q];
                    gv[p] += jv[p] * r[row];
                }
            }
            gc[L] += r[0] * r[0] + r[1] * r[1];
        }
        memcpy(Av + v * L * L, a, sizeof(double) * L * L);
        memcpy(gcv + v * (L + 1), gc, sizeof(double) * (L + 1));
        memcpy(B + v * L * 6, b, sizeof(double) * L * 6);
        memcpy(D + v * 36, d, sizeof(double) * 36);
        memcpy(g + L + 6 * v, gv, sizeof(double) * 6);
    }
    memset(A, 0, sizeof(double) * L * L);
    memset(g, 0, sizeof(double) * L);
    double e = 0;
    for (int64_t v = 0; v < M; ++v) {
        for (int i = 0; i < L * L; ++i) A[i] += Av[v * L * L + i];
        for (int i = 0; i < L; ++i) g[i] += gcv[v * (L + 1) + i];
        e += gcv[v * (L + 1) + L];
    }
    *E = e;
}

int oracle_normal_equations(int model, int64_t M, const int64_t* offs, const double* pts,
                            const double* sens, const double* P, double* A, double* B,
                            double* D, double* g, double* E) {
    const int L = nintr(model);
    double* Av = (double*)malloc(sizeof(double) * (M + 1) * L * L);
    double* gcv = (double*)malloc(sizeof(double) * (M + 1) * (L + 1));
    normal_eq(model, M, offs, pts, sens, P, A, B, D, g, E, Av, gcv);
    free(Av); free(gcv);
    return 0;
}

/* in-place lower Cholesky of n x n (row-major, ld n); returns 0 ok, 1 not SPD */
static int chol(double* a, int n) {
    for (int j = 0; j < n; ++j) {
        double d = a[j * n + j];
        for (int k = 0; k < j; ++k) d -= a[j * n + k] * a[j * n + k];
        if (!(d > 0)) return 1;
        d = sqrt(d); a[j * n + j] = d;
        for (int i = j + 1; i < n; ++i) {
            double s = a[i * n + j];
            for (int k = 0; k < j; ++k) s -= a[i * n + k] * a[j * n + k];
            a[i * n + j] = s / d;
        }
    }
    return 0;
}
static void chol_solve(const double* l, int n, double* x) {
    for (int i = 0; i < n; ++i) {
        double s = x[i];
        for (int k = 0; k < i; ++k) s -= l[i * n + k] * x[k];
        x[i] = s / l[i * n + i];
    }
    for (int i = n - 1; i >= 0; --i) {
        double s = x[i];
        for (int k = i + 1; k < n; ++k) s -= l[k * n + i] * x[k];
        x[i] = s / l[i * n + i];
    }
}

/* delta (L+6M) from the damped block system; returns number of non-SPD blocks */
static int schur_step(int L, int64_t M, const double* A, const double* B, const double* D,
                      const double* g, double lam, double* delta, double* Sv) {
    /* Sv: scratch M*(L*L+L): per-view B Y and B z */
    int bad = 0;
#pragma omp parallel for schedule(static) reduction(+ : bad)
    for (int64_t v = 0; v < M; ++v) {
        double l[36], Y[6 * (LMAX + 1)];
        memcpy(l, D + 36 * v, sizeof(l));
        for (int i = 0; i < 6; ++i) l[7 * i] += lam * l[7 * i];
        double* out = Sv + v * (L * L + L);
        if (chol(l, 6)) { bad += 1; for (int i = 0; i < L * L + L; ++i) out[i] = NAN; continue; }
        /* columns of Y: D'^-1 B^T (L columns) and D'^-1 g_v */
        for (int c = 0; c <= L; ++c) {
            double x[6];
            for (int i = 0; i < 6; ++i) x[i] = c < L ? B[v * L * 6 + c * 6 + i] : g[L + 6 * v + i];
            chol_solve(l, 6, x);
            for (int i = 0; i < 6; ++i) Y[c * 6 + i] = x[i];
        }
        for (int p = 0; p < L; ++p)
            for (int c = 0; c <= L; ++c) {
                double s = 0;
                for (int i = 0; i < 6; ++i) s += B[v * L * 6 + p * 6 + i] * Y[c * 6 + i];
                if (c < L) out[p * L + c] = s; else out[L * L + p] = s;
            }
    }
    double S[LMAX * LMAX], s[LMAX];
    for (int i = 0; i < L * L; ++i) S[i] = A[i];
    for (int i = 0; i < L; ++i) { S[i * L + i] += lam * A[i * L + i]; s[i] = g[i]; }
    for (int64_t v = 0; v < M; ++v) {
        const double* o = Sv + v * (L * L + L);
        for (int i = 0; i < L * L; ++i) S[i] -= o[i];
        for (int i = 0; i < L; ++i) s[i] -= o[L * L + i];
    }
    if (bad || chol(S, L)) { for (int64_t i = 0; i < L + 6 * M; ++i) delta[i] = NAN; return bad + 1; }
    chol_solve(S, L, s);
    for (int i = 0; i < L; ++i) delta[i] = s[i];
#pragma omp parallel for schedule(static)
    for (int64_t v = 0; v < M; ++v) {
        double l[36], x[6];
        memcpy(l, D + 36 * v, sizeof(l));
        for (int i = 0; i < 6; ++i) l[7 * i] += lam * l[7 * i];
        chol(l, 6);
        for (int i = 0; i < 6; ++i) {
            double t = g[L + 6 * v + i];
            for (int p = 0; p < L; ++p) t -= B[v * L * 6 + p * 6 + i] * s[p];
            x[i] = t;
        }
        chol_solve(l, 6, x);
        for (int i = 0; i < 6; ++i) delta[L + 6 * v + i] = x[i];
    }
    return 0;
}

int oracle_lm_step(int model, int64_t M, const double* A, const double* B, const double* D,
                   const double* g, double lam, double* delta) {
    const int L = nintr(model);
    double* Sv = (double*)malloc(sizeof(double) * (M + 1) * (L * L + L));
    int rc = schur_step(L, M, A, B, D, g, lam, delta, Sv);
    free(Sv);
    return rc;
}

/* calibrate.py:137-171.  trace: maxIters x 4 (E, E', lambda, accepted) or NULL.
 * The normal equations are kept across rejected steps (only lambda changed). */
int oracle_lm(int model, int64_t M, const int64_t* offs, const double* pts, const double* sens,
              double* P, int maxIters, double lam0, double lam_min, double lam_max,
              double err_min, double* sse, int* iters, double* trace, int nthreads) {
    const int L = nintr(model);
    const int64_t K = L + 6 * M;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
    if (maxIters <= 0) return -1; /* the reference raises UnboundLocalError */
    double* A = (double*)malloc(sizeof(double) * L * L);
    double* B = (double*)malloc(sizeof(double) * (M + 1) * L * 6);
    double* D = (double*)malloc(sizeof(double) * (M + 1) * 36);
    double* g = (double*)malloc(sizeof(double) * K);
    double* Av = (double*)malloc(sizeof(double) * (M + 1) * L * L);
    double* gcv = (double*)malloc(sizeof(double) * (M + 1) * (L + 1));
    double* Sv = (double*)malloc(sizeof(double) * (M + 1) * (L * L + L));
    double* delta = (double*)malloc(sizeof(double) * K);
    double* P1 = (double*)malloc(sizeof(double) * K);
    double lam = lam0, E = 0, E1 = 0;
    int it = 0, fresh = 0;
    for (it = 0; it < maxIters;) {
        if (!fresh) { normal_eq(model, M, offs, pts, sens, P, A, B, D, g, &E, Av, gcv); fresh = 1; }
        schur_step(L, M, A, B, D, g, lam, delta, Sv);
        for (int64_t i = 0; i < K; ++i) P1[i] = P[i] + delta[i];
        oracle_cost(model, M, offs, pts, sens, P1, &E1);
        int acc = E1 < E;
        if (trace) { trace[4 * it] = E; trace[4 * it + 1] = E1; trace[4 * it + 2] = lam; trace[4 * it + 3] = acc; }
        ++it;
        if (acc) { memcpy(P, P1, sizeof(double) * K); lam /= 10; fresh = 0; }
        else lam *= 10;
        if (!(lam_min < lam && lam < lam_max) || E < err_min) break;
    }
    *sse = E; *iters = it;
    free(A); free(B); free(D); free(g); free(Av); free(gcv); free(Sv); free(delta); free(P1);
    return 0;
}

int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
