// Schur complement over the views, reduced solve and back-substitution (device side of
// `np.linalg.inv(JTJ + lam*diagJTJ) @ J.T @ r`, src/calibrate.py:147,152, in block form).
#pragma once
#include "calib_kernels.cuh"

namespace calib {

// slot of A(i,j), i <= j, in the pass-A layout; -1 = structurally zero
template <int ND>
__host__ __device__ inline int a_slot(int i, int j) {
    using S = SlotsA<ND>;
    if (j < 5) {
        const int tab[5][5] = {{0, -1, 1, 2, -1}, {-1, 3, -1, -1, 4}, {1, -1, 3, 4, -1},
                               {2, -1, 4, 5, -1}, {-1, 4, -1, -1, 5}};
        return tab[i][j];
    }
    if (i < 5) {
        const int a = j - 5;
        switch (i) {
            case 0: return S::UK + 3 * a;
            case 2: return S::UK + 3 * a + 1;
            case 3: return S::UK + 3 * a + 2;
            case 1: return S::VK + 2 * a;
            default: return S::VK + 2 * a + 1;
        }
    }
    return S::kk(i - 5, j - 5);
}

constexpr int vtab_items(int L) { return 21 + 6 * (L + 1); }

template <int MODEL>
__global__ void __launch_bounds__(kSchurThreads) k_schur(Dev d) {
    using T = ModelTraits<MODEL>;
    constexpr int L = T::L;
    using S = SlotsBD<L>;
    LmState* st = d.st;
    if (st->done) return;
    __shared__ double sred[2][kNA];
    __shared__ int s_last;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nvb = gridDim.x - 1;
    if ((int)blockIdx.x == nvb) {
        // the extra block: column sums of the pass-A rows, and the cost
        const int col = tid & (kNA - 1), rg = tid >> 6;
        double s = 0.0;
        for (int r = rg; r < d.nblocks; r += 2) s += d.partA[(long long)r * kNA + col];
        sred[rg][col] = s;
        __syncthreads();
        if (tid < kNA) d.red[kRedA + tid] = sred[0][tid] + sred[1][tid];
        if (warp == 3) {
            const double e = warp_sum_array(d.partE, d.nblocks, lane);
            if (lane == 0) {
                d.red[kRedE] = e;
                if (st->need_eval) st->n_eval += 1;
            }
        }
    } else {
        const int v = blockIdx.x * kSchurThreads + tid;
        double wv[kNBD];
#pragma unroll
        for (int i = 0; i < kNBD; ++i) wv[i] = 0.0;
        if (v < d.M) {
            const int t0 = d.view_task_off[v], t1 = d.view_task_off[v + 1];
            const double lam = st->lam;
            double Lc[36];
#pragma unroll
            for (int a = 0; a < 6; ++a)
#pragma unroll
                for (int b = a; b < 6; ++b) {
                    const double* src = d.partBD + (long long)S::dd(a, b) * d.Tpad;
                    double s = 0.0;
                    for (int t = t0; t < t1; ++t) s += src[t];
                    Lc[b * 6 + a] = (a == b) ? fma(lam, s, s) : s;   // D + lam diag(D)
                }
            chol_inplace<6>(Lc);
            double id[6];
#pragma unroll
            for (int a = 0; a < 6; ++a) {
                id[a] = 1.0 / Lc[a * 6 + a];
#pragma unroll
                for (int b = 0; b <= a; ++b) d.vtab[(long long)(a * (a + 1) / 2 + b) * d.Mpad + v] = Lc[a * 6 + b];
            }
            double z[L + 1][6];
#pragma unroll
            for (int c = 0; c <= L; ++c) {
                // forward substitution L z = column c of [B^T | g_v]
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    const int slot = (c < L) ? (S::B + c * 6 + i) : (S::GV + i);
                    const double* src = d.partBD + (long long)slot * d.Tpad;
                    double s = 0.0;
                    for (int t = t0; t < t1; ++t) s += src[t];
#pragma unroll
                    for (int k = 0; k < i; ++k) s -= Lc[i * 6 + k] * z[c][k];
                    z[c][i] = s * id[i];
                    d.vtab[(long long)(21 + c * 6 + i) * d.Mpad + v] = z[c][i];
                }
            }
            // W = Z^T Z (upper triangle) and w = Z^T z_g
#pragma unroll
            for (int p = 0; p < L; ++p) {
#pragma unroll
                for (int c = p; c < L; ++c) {
                    double s = 0.0;
#pragma unroll
                    for (int i = 0; i < 6; ++i) s = fma(z[p][i], z[c][i], s);
                    wv[tri(L, p, c)] = s;
                }
                double s = 0.0;
#pragma unroll
                for (int i = 0; i < 6; ++i) s = fma(z[p][i], z[L][i], s);
                wv[L * (L + 1) / 2 + p] = s;
            }
        }
        warp_reduce_scatter<kNBD>(wv, lane);
        double* dst = d.partW + (long long)(blockIdx.x * (kSchurThreads / 32) + warp) * kNBD;
#pragma unroll
        for (int j = 0; j < kNBD / 32; ++j) dst[lane * (kNBD / 32) + j] = wv[j];
    }
    // last block to finish sums the per-warp W rows in a fixed order
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const unsigned int prev = atomicAdd(&st->cnt_schur, 1u);
        s_last = (prev == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (tid < kNBD) {
        const int nrows = nvb * (kSchurThreads / 32);
        double s = 0.0;
        for (int r = 0; r < nrows; ++r) s += d.partW[(long long)r * kNBD + tid];
        d.red[kRedW + tid] = s;
    }
    if (tid == 0) st->cnt_schur = 0;
}

template <int MODEL>
__global__ void __launch_bounds__(kSchurThreads) k_solve_backsub(Dev d) {
    using T = ModelTraits<MODEL>;
    constexpr int L = T::L, ND = T::ND;
    using SA = SlotsA<ND>;
    LmState* st = d.st;
    if (st->done) return;
    __shared__ double Sm[L * L];
    __shared__ double dc[L];
    const int tid = threadIdx.x;
    const double lam = st->lam;
    const int cur = st->cur;
    const double* P = d.Pbuf + (long long)cur * d.K;
    double* Pt = d.Pbuf + (long long)(cur ^ 1) * d.K;
    if (tid < L * L) {
        const int i = tid / L, j = tid % L;
        const int lo = i < j ? i : j, hi = i < j ? j : i;
        const int sl = a_slot<ND>(lo, hi);
        double a = sl >= 0 ? d.red[kRedA + sl] : 0.0;
        if (i == j) a = fma(lam, a, a);
        Sm[tid] = a - d.red[kRedW + tri(L, lo, hi)];
    }
    if (tid >= 128 - L) {
        const int i = tid - (128 - L);
        dc[i] = d.red[kRedA + SA::G + i] - d.red[kRedW + L * (L + 1) / 2 + i];
    }
    __syncthreads();
    if (tid == 0) {
        // lower Cholesky of S and the two triangular solves, serial (L <= 10)
        for (int j = 0; j < L; ++j) {
            double dd = Sm[j * L + j];
            for (int k = 0; k < j; ++k) dd -= Sm[j * L + k] * Sm[j * L + k];
            dd = dd > 0.0 ? sqrt(dd) : __longlong_as_double(0x7ff8000000000000LL);
            Sm[j * L + j] = dd;
            const double id = 1.0 / dd;
            for (int i = j + 1; i < L; ++i) {
                double s = Sm[i * L + j];
                for (int k = 0; k < j; ++k) s -= Sm[i * L + k] * Sm[j * L + k];
                Sm[i * L + j] = s * id;
            }
        }
        for (int i = 0; i < L; ++i) {
            double s = dc[i];
            for (int k = 0; k < i; ++k) s -= Sm[i * L + k] * dc[k];
            dc[i] = s / Sm[i * L + i];
        }
        for (int i = L - 1; i >= 0; --i) {
            double s = dc[i];
            for (int k = i + 1; k < L; ++k) s -= Sm[k * L + i] * dc[k];
            dc[i] = s / Sm[i * L + i];
        }
        if (blockIdx.x == 0) st->E_cur = d.red[kRedE];
    }
    __syncthreads();
    if (blockIdx.x == 0 && tid < L) {
        Pt[tid] = P[tid] + dc[tid];
        d.delta[tid] = dc[tid];
    }
    const int v = blockIdx.x * kSchurThreads + tid;
    if (v >= d.M) return;
    double x[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        double s = d.vtab[(long long)(21 + L * 6 + i) * d.Mpad + v];
#pragma unroll
        for (int p = 0; p < L; ++p) s = fma(-d.vtab[(long long)(21 + p * 6 + i) * d.Mpad + v], dc[p], s);
        x[i] = s;
    }
    double Lc[21];
#pragma unroll
    for (int i = 0; i < 21; ++i) Lc[i] = d.vtab[(long long)i * d.Mpad + v];
#pragma unroll
    for (int i = 5; i >= 0; --i) {
        double s = x[i];
#pragma unroll
        for (int k = i + 1; k < 6; ++k) s -= Lc[k * (k + 1) / 2 + i] * x[k];
        x[i] = s / Lc[i * (i + 1) / 2 + i];
    }
    double e[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        const long long idx = L + 6ll * v + i;
        e[i] = P[idx] + x[i];
        Pt[idx] = e[i];
        d.delta[idx] = x[i];
    }
    ViewC w;
    make_view(e, w);
    d.views[(long long)(cur ^ 1) * d.M + v] = w;
}

// ---- parity / debug kernels (not on the timed path) ------------------------------------------
// Per point: y (N x 2), r (2N), Jc (2N x L), Jv (2N x 6); any pointer may be null.
template <int MODEL>
__global__ void k_debug_point(Dev d, double* y, double* r, double* Jc, double* Jv) {
    using T = ModelTraits<MODEL>;
    constexpr int L = T::L, ND = T::ND;
    const int lane = threadIdx.x & 31;
    const int t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (t >= d.ntasks) return;
    const Task task = d.tasks[t];
    const int cur = d.st->cur;
    const double* P = d.Pbuf + (long long)cur * d.K;
    double c[L];
#pragma unroll
    for (int i = 0; i < L; ++i) c[i] = P[i];
    const ViewC w = d.views[(long long)cur * d.M + task.view];
    for (int i = lane; i < task.count; i += 32) {
        const long long n = task.start + i;
        double p[5];
        load_point(d, n, p);
        PointOut<MODEL> o;
        point_eval<MODEL, true, true>(c, w, p[0], p[1], p[2], p[3], p[4], o);
        if (y) { y[2 * n] = o.u; y[2 * n + 1] = o.v; }
        if (r) { r[2 * n] = o.ru; r[2 * n + 1] = o.rv; }
        if (Jc) {
            double* ju = Jc + 2 * n * L;
            double* jv = ju + L;
            ju[0] = o.xd; ju[1] = 0; ju[2] = o.yd; ju[3] = 1; ju[4] = 0;
            jv[0] = 0; jv[1] = o.yd; jv[2] = 0; jv[3] = 0; jv[4] = 1;
#pragma unroll
            for (int a = 0; a < ND; ++a) { ju[5 + a] = o.ku[a]; jv[5 + a] = o.kv[a]; }
        }
        if (Jv) {
#pragma unroll
            for (int j = 0; j < 6; ++j) { Jv[12 * n + j] = o.ju[j]; Jv[12 * n + 6 + j] = o.jv[j]; }
        }
    }
}

// Unpacks the pass partials into the reference's block layout: A L x L, B M x L x 6,
// D M x 6 x 6, g K.  red[] must already hold the column sums (k_schur ran).
template <int MODEL>
__global__ void k_debug_unpack(Dev d, double* A, double* B, double* D, double* g) {
    using T = ModelTraits<MODEL>;
    constexpr int L = T::L, ND = T::ND;
    using SA = SlotsA<ND>;
    using S = SlotsBD<L>;
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v == 0) {
        for (int i = 0; i < L; ++i) {
            for (int j = 0; j < L; ++j) {
                const int sl = a_slot<ND>(i < j ? i : j, i < j ? j : i);
                A[i * L + j] = sl >= 0 ? d.red[kRedA + sl] : 0.0;
            }
            g[i] = d.red[kRedA + SA::G + i];
        }
    }
    if (v >= d.M) return;
    const int t0 = d.view_task_off[v], t1 = d.view_task_off[v + 1];
    auto slot_sum = [&](int slot) {
        double s = 0.0;
        for (int t = t0; t < t1; ++t) s += d.partBD[(long long)slot * d.Tpad + t];
        return s;
    };
    for (int p = 0; p < L; ++p)
        for (int j = 0; j < 6; ++j) B[((long long)v * L + p) * 6 + j] = slot_sum(S::B + p * 6 + j);
    for (int a = 0; a < 6; ++a)
        for (int b = a; b < 6; ++b) {
            const double s = slot_sum(S::dd(a, b));
            D[(long long)v * 36 + a * 6 + b] = s;
            D[(long long)v * 36 + b * 6 + a] = s;
        }
    for (int j = 0; j < 6; ++j) g[L + 6ll * v + j] = slot_sum(S::GV + j);
}

// FP64 FMA pipe micro-benchmark: 8 independent chains per thread
__global__ void k_fp64_peak(double* out, int iters, double a, double b) {
    double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; ++i) {
        x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
        x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}

__global__ void k_fill(double* p, long long n, double v) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) p[i] = v;
}

}  // namespace calib
