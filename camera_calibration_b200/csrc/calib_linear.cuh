// Per-point parts of the closed-form initial estimate (Calibrator.estimateCalibrationParameters,
// src/calibrate.py:41-58), on the handle's resident detections.  The reference
// builds these row by row in Python (np.vstack per point: O(N^2)); here each is one warp per view
// with the sums in registers, and what is left for the host is O(M) small dense algebra
// (9 x 9 symmetric eigenvectors, a 2M x 6 SVD, 3 x 3 SVDs, one nd x nd solve).
//
//   k_dlt_gram          linearcalibrate.estimateHomography, src/linearcalibrate.py:24-58: the point
//                       normalisations (computeNormalizationMatrix, :61-91) and the Gram matrix
//                       M^T M of the 2N x 9 DLT matrix M (:46-52).  With p = (X, Y, 1) the two rows of
//                       a point are (-p, 0, u p) and (0, -p, v p), so M^T M is p p^T weighted by
//                       1, u, v, u^2 + v^2: 24 sums.  The homography is the eigenvector of the smallest
//                       eigenvalue (the reference takes the last right-singular vector of M: same
//                       vector, up to sign, which H /= H[2,2] removes).
//   k_distortion_rows   DistortionModel.estimateDistortion, src/distortion.py:110-191 (radial-
//                       tangential) and :222-271 (fisheye): the rows of D k = Ddot are formed per point
//                       and D^T D, D^T Ddot accumulated; the host solves the nd x nd system
//                       (the reference: pinv(D) @ Ddot, the same least-squares solution).
#pragma once
#include "calib_kernels.cuh"

namespace calib {

constexpr int kGramN = 24;      // p p^T (6) x weights {1, u, v, u^2 + v^2}
constexpr int kNormN = 6;       // mean_u, mean_v, scale_a, mean_X, mean_Y, scale_b

struct LinDev {
    const double *X, *Y, *Z, *U, *V;
    const long long* offs;      // M + 1
    int M, model;
    double* gram;               // M x kGramN
    double* norm;               // M x kNormN
    const double* A;            // 9: intrinsic matrix, row-major
    const double* W;            // M x 12: R row-major, then t
    double* drows;              // M x kDistN: per-view sums of D^T D (upper triangle) and D^T Ddot
    double* Hdlt;               // M x 9 or NULL: k_dlt_gram also solves for the homography
};

constexpr int kDistN = 20;      // radtan: 15 + 5; fisheye: 10 + 4 (padded)

// Eigen-decomposition of a symmetric 9 x 9 matrix by one warp: cyclic Jacobi rotations on A (shared memory,
// overwritten: eigenvalues on its diagonal), eigenvectors accumulated in the columns of V.  Lane k < 9 owns
// row k.  A rotation is skipped when |a_pq| <= eps sqrt(|a_pp a_qq|) (the relative criterion that keeps the
// small eigenvalues of a positive semi-definite matrix accurate); done when a whole sweep skipped everything.
// Returns the index of the smallest eigenvalue.
__device__ __forceinline__ int warp_jacobi9(double (*A)[9], double (*V)[9], int lane) {
    if (lane < 9) {
#pragma unroll
        for (int j = 0; j < 9; ++j) V[lane][j] = lane == j ? 1.0 : 0.0;
    }
    __syncwarp();
    for (int sweep = 0; sweep < 30; ++sweep) {
        int rotated = 0;
        for (int p = 0; p < 8; ++p) {
            for (int q = p + 1; q < 9; ++q) {
                const double apq = A[p][q], app = A[p][p], aqq = A[q][q];
                if (!(fabs(apq) > 1e-17 * sqrt(fabs(app * aqq)))) continue;      // warp-uniform
                rotated = 1;
                const double theta = (aqq - app) / (2.0 * apq);
                const double t = copysign(1.0, theta) / (fabs(theta) + sqrt(fma(theta, theta, 1.0)));
                const double c = rsqrt(fma(t, t, 1.0)), sn = t * c;
                __syncwarp();
                if (lane < 9) {
                    const double vp = V[lane][p], vq = V[lane][q];
                    V[lane][p] = fma(c, vp, -sn * vq);
                    V[lane][q] = fma(sn, vp, c * vq);
                    if (lane != p && lane != q) {
                        const double akp = A[lane][p], akq = A[lane][q];
                        const double np_ = fma(c, akp, -sn * akq), nq = fma(sn, akp, c * akq);
                        A[lane][p] = np_; A[p][lane] = np_;
                        A[lane][q] = nq; A[q][lane] = nq;
                    } else if (lane == p) {
                        A[p][p] = fma(-t, apq, app);
                        A[q][q] = fma(t, apq, aqq);
                        A[p][q] = 0.0; A[q][p] = 0.0;
                    }
                }
                __syncwarp();
            }
        }
        if (!rotated) break;
    }
    int best = 0;
    double lo = A[0][0];
#pragma unroll
    for (int j = 1; j < 9; ++j) {
        const double e = A[j][j];
        if (e < lo) { lo = e; best = j; }
    }
    return best;
}

__global__ void __launch_bounds__(kBlockThreads) k_dlt_gram(LinDev d) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __shared__ double sG[kBlockWarps][9][9], sVec[kBlockWarps][9][9];
    for (int v = blockIdx.x * kBlockWarps + warp; v < d.M; v += gridDim.x * kBlockWarps) {
        const long long p0 = d.offs[v], p1 = d.offs[v + 1];
        const double n = (double)(p1 - p0);
        // computeNormalizationMatrix: centroid, then sqrt(2) / mean distance to it
        double su = 0, sv = 0, sx = 0, sy = 0;
        for (long long i = p0 + lane; i < p1; i += 32) {
            su += __ldg(d.U + i); sv += __ldg(d.V + i); sx += __ldg(d.X + i); sy += __ldg(d.Y + i);
        }
        const double mu_ = warp_sum(su) / n, mv = warp_sum(sv) / n, mx = warp_sum(sx) / n, my = warp_sum(sy) / n;
        double da = 0, db = 0;
        for (long long i = p0 + lane; i < p1; i += 32) {
            const double a0 = __ldg(d.U + i) - mu_, a1 = __ldg(d.V + i) - mv;
            const double b0 = __ldg(d.X + i) - mx, b1 = __ldg(d.Y + i) - my;
            da += sqrt(fma(a0, a0, a1 * a1));
            db += sqrt(fma(b0, b0, b1 * b1));
        }
        const double sa = sqrt(2.0) / (warp_sum(da) / n), sb = sqrt(2.0) / (warp_sum(db) / n);
        double acc[kGramN];
#pragma unroll
        for (int k = 0; k < kGramN; ++k) acc[k] = 0.0;
        for (long long i = p0 + lane; i < p1; i += 32) {
            // Na @ hom(x): scale * x + (-scale * mean)
            const double u = fma(sa, __ldg(d.U + i), -sa * mu_), w = fma(sa, __ldg(d.V + i), -sa * mv);
            const double X = fma(sb, __ldg(d.X + i), -sb * mx), Y = fma(sb, __ldg(d.Y + i), -sb * my);
            const double pp[6] = {X * X, X * Y, X, Y * Y, Y, 1.0};
            const double q = fma(u, u, w * w);
#pragma unroll
            for (int k = 0; k < 6; ++k) {
                acc[k] += pp[k];
                acc[6 + k] = fma(u, pp[k], acc[6 + k]);
                acc[12 + k] = fma(w, pp[k], acc[12 + k]);
                acc[18 + k] = fma(q, pp[k], acc[18 + k]);
            }
        }
#pragma unroll
        for (int k = 0; k < kGramN; ++k) acc[k] = warp_sum(acc[k]);
        if (lane == 0) {
#pragma unroll
            for (int k = 0; k < kGramN; ++k) d.gram[(long long)v * kGramN + k] = acc[k];
            double* nm = d.norm + (long long)v * kNormN;
            nm[0] = mu_; nm[1] = mv; nm[2] = sa; nm[3] = mx; nm[4] = my; nm[5] = sb;
        }
        if (d.Hdlt) {
            // M^T M = [[P1, 0, -Pu], [0, P1, -Pv], [-Pu, -Pv, Pq]] with the 3 x 3 blocks P* = sums of w p p^T
            // (acc[6 k ..] = packed 00 01 02 11 12 22); its eigenvector of the smallest eigenvalue is the
            // normalised homography (src/linearcalibrate.py:52-56), then H = Na^-1 H' Nb, H /= H[2,2] (:57-58)
            double (*G)[9] = sG[warp];
            double (*E)[9] = sVec[warp];
            double* sums = &E[0][0];                 // scratch until the rotations start
            __syncwarp();
            if (lane == 0) {
#pragma unroll
                for (int k = 0; k < kGramN; ++k) sums[k] = acc[k];
            }
            __syncwarp();
            if (lane < 9) {
                const int bi = lane / 3, i = lane % 3;
#pragma unroll
                for (int j = 0; j < 9; ++j) {
                    const int bj = j / 3, jj = j % 3;
                    const int a = i < jj ? i : jj, b = i < jj ? jj : i;
                    const int pk = a * 3 - a * (a - 1) / 2 + (b - a);     // packed index of (a, b), a <= b
                    double g = 0.0;
                    if (bi == bj && bi < 2) g = sums[pk];
                    else if (bi == 2 && bj == 2) g = sums[18 + pk];
                    else if (bi + bj == 2 && bi != bj) g = -sums[6 + pk];      // blocks (0,2), (2,0): -Pu
                    else if (bi + bj == 3) g = -sums[12 + pk];                  // blocks (1,2), (2,1): -Pv
                    G[lane][j] = g;
                }
            }
            __syncwarp();
            const int jmin = warp_jacobi9(G, E, lane);
            __syncwarp();
            if (lane == 0) {
                double hp[9];
#pragma unroll
                for (int k = 0; k < 9; ++k) hp[k] = E[k][jmin];
                // Na^-1 = [[1/sa, 0, mu], [0, 1/sa, mv], [0, 0, 1]] ; Nb = [[sb, 0, -sb mx], [0, sb, -sb my], [0, 0, 1]]
                double t[9], Hh[9];
                const double isa = 1.0 / sa;
#pragma unroll
                for (int j = 0; j < 3; ++j) {
                    t[j] = fma(isa, hp[j], mu_ * hp[6 + j]);
                    t[3 + j] = fma(isa, hp[3 + j], mv * hp[6 + j]);
                    t[6 + j] = hp[6 + j];
                }
#pragma unroll
                for (int r = 0; r < 3; ++r) {
                    Hh[3 * r + 0] = t[3 * r + 0] * sb;
                    Hh[3 * r + 1] = t[3 * r + 1] * sb;
                    Hh[3 * r + 2] = fma(-sb * mx, t[3 * r + 0], fma(-sb * my, t[3 * r + 1], t[3 * r + 2]));
                }
                const double ih = 1.0 / Hh[8];
#pragma unroll
                for (int k = 0; k < 9; ++k) d.Hdlt[(long long)v * 9 + k] = Hh[k] * ih;
            }
            __syncwarp();
        }
    }
}

template <int MODEL>
__global__ void __launch_bounds__(kBlockThreads) k_distortion_rows(LinDev d) {
    constexpr int ND = MODEL == 0 ? 5 : 4, NT = ND * (ND + 1) / 2;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const double fx = d.A[0], sk = d.A[1], uc = d.A[2], fy = d.A[4], vc = d.A[5];
    for (int v = blockIdx.x * kBlockWarps + warp; v < d.M; v += gridDim.x * kBlockWarps) {
        const long long p0 = d.offs[v], p1 = d.offs[v + 1];
        const double* w = d.W + 12ll * v;
        double R[9], t[3];
#pragma unroll
        for (int k = 0; k < 9; ++k) R[k] = __ldg(w + k);
#pragma unroll
        for (int k = 0; k < 3; ++k) t[k] = __ldg(w + 9 + k);
        double acc[kDistN];
#pragma unroll
        for (int k = 0; k < kDistN; ++k) acc[k] = 0.0;
        for (long long i = p0 + lane; i < p1; i += 32) {
            const double X = __ldg(d.X + i), Y = __ldg(d.Y + i), Z = __ldg(d.Z + i);
            // cX = cMb bX (mathutils.transform), x = projectStandard(cX), (u, v) = A hom(x)
            const double cx = fma(R[0], X, fma(R[1], Y, fma(R[2], Z, t[0])));
            const double cy = fma(R[3], X, fma(R[4], Y, fma(R[5], Z, t[1])));
            const double cz = fma(R[6], X, fma(R[7], Y, fma(R[8], Z, t[2])));
            const double xn = cx / cz, yn = cy / cz;
            const double r2 = fma(xn, xn, yn * yn);
            const double u = fma(fx, xn, fma(sk, yn, uc)), vv = fma(fy, yn, vc);
            double Du[ND], Dv[ND];
            if (MODEL == 0) {
                const double r4 = r2 * r2, r6 = r4 * r2, du = u - uc, dv = vv - vc;
                Du[0] = du * r2; Du[1] = du * r4; Du[2] = fx * (2.0 * xn * yn); Du[3] = fx * fma(2.0 * xn, xn, r2); Du[4] = du * r6;
                Dv[0] = dv * r2; Dv[1] = dv * r4; Dv[2] = fy * fma(2.0 * yn, yn, r2); Dv[3] = fy * (2.0 * xn * yn); Dv[4] = dv * r6;
            } else {
                const double r = sqrt(r2), th = atan(r), q = th / r, th2 = th * th;
                const double bu = fx * (u - uc) * q, bv = fy * (vv - vc) * q;
                double p = th2;
#pragma unroll
                for (int c = 0; c < ND; ++c) { Du[c] = bu * p; Dv[c] = bv * p; p *= th2; }
            }
            const double eu = __ldg(d.U + i) - u, ev = __ldg(d.V + i) - vv;
            int s = 0;
#pragma unroll
            for (int a = 0; a < ND; ++a) {
#pragma unroll
                for (int b = a; b < ND; ++b) { acc[s] = fma(Du[a], Du[b], fma(Dv[a], Dv[b], acc[s])); ++s; }
            }
#pragma unroll
            for (int a = 0; a < ND; ++a) acc[NT + a] = fma(Du[a], eu, fma(Dv[a], ev, acc[NT + a]));
        }
#pragma unroll
        for (int k = 0; k < kDistN; ++k) acc[k] = warp_sum(acc[k]);
        if (lane == 0) {
#pragma unroll
            for (int k = 0; k < kDistN; ++k) d.drows[(long long)v * kDistN + k] = acc[k];
        }
    }
}

}  // namespace calib
