// Per-view homography refinement: Calibrator._refineHomographies / _refineHomography,
// src/calibrate.py:60-115, with the analytic form of HomographyJacobian (src/jacobian.py:88-121).
// The step right before the hot path in Calibrator.estimateCalibrationParameters:
// M independent 9-parameter Levenberg-Marquardt problems, <= 20 iterations each, same damping rule.
//
// One warp per view, the whole loop in one launch.  With s = (X, Y, 1) / w, w = h6 X + h7 Y + h8,
// the two Jacobian rows of a point are Ju = [s, 0, -u s], Jv = [0, s, -v s], so J^T J consists of the
// 3 x 3 matrix s s^T weighted by 1, u, v and u^2 + v^2: 24 sums instead of 45, plus 9 for J^T r and
// the cost.  Lanes stride over the view's points with the sums in registers, a butterfly gives every
// lane the totals, the damped 9 x 9 system is solved by the warp (solve_dense_core), the trial cost
// is a second sweep.  After a rejected step only lambda changes, so the sums are kept.
#pragma once
#include "calib_solve.cuh"

namespace calib {

struct HomDev {
    const double *X, *Y, *U, *V;      // SoA detections of the handle
    const long long* offs;            // M + 1
    int M, max_iters, want_trace;
    double lam0, lam_min, lam_max, err_min;
    double* H;                        // M x 9, in/out
    int* iters;                       // M
    double* trace;                    // M x max_iters x 4 (nullable)
};

__device__ __forceinline__ void hom_project(const double* h, double X, double Y, double& u, double& v, double& iw) {
    const double w = fma(h[6], X, fma(h[7], Y, h[8]));
    iw = 1.0 / w;
    u = fma(h[0], X, fma(h[1], Y, h[2])) * iw;
    v = fma(h[3], X, fma(h[4], Y, h[5])) * iw;
}

__global__ void __launch_bounds__(kBlockThreads) k_homography_lm(HomDev d) {
    constexpr int NS = 34, NE = 90;   // sums per point sweep; dense system 9 x 9 + 9
    __shared__ double sJ[kBlockWarps][NE];      // J^T J (row-major) and J^T r at the current point
    __shared__ double sD[kBlockWarps][NE];      // damped copy, overwritten by the factor
    __shared__ double sh[kBlockWarps][2][12];   // current and trial homography
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int v = blockIdx.x * kBlockWarps + warp; v < d.M; v += gridDim.x * kBlockWarps) {
        const long long p0 = d.offs[v], p1 = d.offs[v + 1];
        double* h = sh[warp][0];
        double* ht = sh[warp][1];
        __syncwarp();
        if (lane < 9) h[lane] = d.H[9ll * v + lane];
        __syncwarp();
        double lam = d.lam0, E = 0.0;
        bool need_eval = true;
        int it = 0;
        for (; it < d.max_iters;) {
            if (need_eval) {
                double acc[NS];
#pragma unroll
                for (int k = 0; k < NS; ++k) acc[k] = 0.0;
                for (long long i = p0 + lane; i < p1; i += 32) {
                    const double X = __ldg(d.X + i), Y = __ldg(d.Y + i);
                    double u, vv, iw;
                    hom_project(h, X, Y, u, vv, iw);
                    const double ru = __ldg(d.U + i) - u, rv = __ldg(d.V + i) - vv;
                    const double s0 = X * iw, s1 = Y * iw, s2 = iw;
                    const double ss[6] = {s0 * s0, s0 * s1, s0 * s2, s1 * s1, s1 * s2, s2 * s2};
                    const double q = fma(u, u, vv * vv), gr = -fma(u, ru, vv * rv);
#pragma unroll
                    for (int k = 0; k < 6; ++k) {
                        acc[k] += ss[k];
                        acc[6 + k] = fma(u, ss[k], acc[6 + k]);
                        acc[12 + k] = fma(vv, ss[k], acc[12 + k]);
                        acc[18 + k] = fma(q, ss[k], acc[18 + k]);
                    }
                    acc[24] = fma(s0, ru, acc[24]); acc[25] = fma(s1, ru, acc[25]); acc[26] = fma(s2, ru, acc[26]);
                    acc[27] = fma(s0, rv, acc[27]); acc[28] = fma(s1, rv, acc[28]); acc[29] = fma(s2, rv, acc[29]);
                    acc[30] = fma(s0, gr, acc[30]); acc[31] = fma(s1, gr, acc[31]); acc[32] = fma(s2, gr, acc[32]);
                    acc[33] = __dadd_rn(acc[33], fma(rv, rv, __dmul_rn(ru, ru)));
                }
#pragma unroll
                for (int k = 0; k < NS; ++k) acc[k] = warp_sum(acc[k]);
                E = acc[33];
                // unpack: entry e = 9 a + b of J^T J, block (a / 3, b / 3), element (a % 3, b % 3) of s s^T
                __syncwarp();
                for (int e = lane; e < NE; e += 32) {
                    double val;
                    if (e < 81) {
                        const int a = e / 9, b = e - 9 * a;
                        const int ga = a / 3, gb = b / 3, ia = a - 3 * ga, ib = b - 3 * gb;
                        const int lo = ia < ib ? ia : ib, hi = ia < ib ? ib : ia;
                        const int t = lo * 3 - lo * (lo - 1) / 2 + (hi - lo);     // index in ss[6]
                        int base;
                        double sign = 1.0;
                        if (ga == gb) base = ga < 2 ? 0 : 18;
                        else if (ga + gb == 1) base = -1;                         // u rows x v rows: zero
                        else { base = (ga == 0 || gb == 0) ? 6 : 12; sign = -1.0; }
                        double pick = 0.0;
#pragma unroll
                        for (int k = 0; k < 24; ++k) pick = (base >= 0 && k == base + t) ? acc[k] : pick;
                        val = sign * pick;
                    } else {
                        double pick = 0.0;
#pragma unroll
                        for (int k = 0; k < 9; ++k) pick = (k == e - 81) ? acc[24 + k] : pick;
                        val = pick;
                    }
                    sJ[warp][e] = val;
                }
                __syncwarp();
            }
            // (J^T J + lam diag J^T J) delta = J^T r, src/calibrate.py:85-93
            for (int e = lane; e < NE; e += 32) {
                double a = sJ[warp][e];
                if (e < 81 && e % 10 == 0) a = fma(lam, a, a);
                sD[warp][e] = a;
            }
            __syncwarp();
            const double dl = solve_dense_core<9>(sD[warp], lane);
            if (lane < 9) ht[lane] = h[lane] + dl;
            __syncwarp();
            double e1 = 0.0;
            for (long long i = p0 + lane; i < p1; i += 32) {
                double u, vv, iw;
                hom_project(ht, __ldg(d.X + i), __ldg(d.Y + i), u, vv, iw);
                const double ru = __ldg(d.U + i) - u, rv = __ldg(d.V + i) - vv;
                e1 = __dadd_rn(e1, fma(rv, rv, __dmul_rn(ru, ru)));
            }
            e1 = warp_sum(e1);
            const bool acc_step = e1 < E;     // strict; NaN rejects (src/calibrate.py:101)
            if (d.want_trace && lane == 0) {
                double* tr = d.trace + ((long long)v * d.max_iters + it) * 4;
                tr[0] = E; tr[1] = e1; tr[2] = lam; tr[3] = acc_step ? 1.0 : 0.0;
            }
            const double Eprev = E;
            __syncwarp();
            if (acc_step) {
                if (lane < 9) h[lane] = ht[lane];
                lam = lam / 10.0;
                need_eval = true;
            } else {
                lam = lam * 10.0;
                need_eval = false;
            }
            __syncwarp();
            ++it;
            if (!(d.lam_min < lam && lam < d.lam_max) || Eprev < d.err_min) break;
        }
        // Href /= Href[2,2], src/calibrate.py:108-109
        if (lane < 9) d.H[9ll * v + lane] = h[lane] / h[8];
        if (lane == 0) d.iters[v] = it;
        __syncwarp();
    }
}

}  // namespace calib
