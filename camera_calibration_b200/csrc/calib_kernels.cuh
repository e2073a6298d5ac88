// Kernels of the single-calibration ("large") path.  One LM iteration =
//   k_eval_A, k_eval_BD   fused projection + residual + analytic Jacobian + normal equations,
//                         one warp per task (<= seg points of ONE view); no J is ever stored
//   k_schur               per view: sum task partials, damp, 6x6 Cholesky, Z = L^-1 [B^T g_v],
//                         W = Z^T Z; cross-view sums; the last block writes the reduced system
//   k_solve_backsub       L x L Cholesky solve, per-view back-substitution, trial parameters and
//                         the trial views' constants
//   k_cost                residual-only pass at the trial point; last block sums (and, on one
//                         GPU, takes the accept/reject decision)
// Replaces the loop body of src/calibrate.py:143-168 (see include/calib_b200.h).
#pragma once
#include "calib_device.cuh"

namespace calib {

constexpr int kEvalThreads = 128;           // 4 warps = 4 tasks per block
constexpr int kWarpsPerBlock = kEvalThreads / 32;
constexpr int kNA = 64;                     // padded slot count of pass A
constexpr int kNBD = 96;                    // padded slot count of pass BD and of W
constexpr int kRedA = 0, kRedW = kNA, kRedE = kNA + kNBD, kRedN = kNA + kNBD + 1;
constexpr int kSchurThreads = 128;
constexpr int kAsumThreads = 256;           // the extra block of k_schur that sums pass-A rows

struct Task { int view, start, count, pad; };

struct LmState {
    double lam, E_cur, E_trial, sse;
    double lam_min, lam_max, err_min;
    int iter, max_iters, done, need_eval, cur, nranks;
    unsigned int cnt_schur, cnt_cost;
    long long n_eval;   // Jacobian evaluations actually executed
};

struct Dev {  // everything the kernels need; passed by value
    int model, L, M, ntasks, nblocks;   // nblocks = ceil(ntasks / 4)
    long long N, K;
    int Mpad, Tpad;                     // padded strides of per-view / per-task tables
    const double *X, *Y, *Z, *U, *V;    // SoA detections
    const Task* tasks;
    const int* view_task_off;           // M+1
    double* Pbuf;                       // 2 x K
    ViewC* views;                       // 2 x M
    double* partA;                      // nblocks x kNA
    double* partE;                      // nblocks
    double* partBD;                     // kNBD x Tpad (slot-major)
    double* partW;                      // nschur_warps x kNBD
    double* asum;                       // kNA
    double* red;                        // kRedN
    double* vtab;                       // (21 + 6*(L+1)) x Mpad : Cholesky factor and Z per view
    double* partC;                      // nblocks (trial cost)
    double* etrial;                     // 1
    double* delta;                      // K
    double* trace;                      // max_iters x 4
    LmState* st;
};

// ---- slot layouts ---------------------------------------------------------------------------
template <int ND> struct SlotsA {
    static constexpr int UK = 6, VK = 6 + 3 * ND, KK = 6 + 5 * ND,
                         G = KK + ND * (ND + 1) / 2, N = G + 5 + ND;
    __host__ __device__ static constexpr int kk(int i, int j) { return KK + i * ND - i * (i - 1) / 2 + (j - i); }
};
template <int L> struct SlotsBD {
    static constexpr int B = 0, D = 6 * L, GV = 6 * L + 21, N = 6 * L + 27;
    __host__ __device__ static constexpr int dd(int a, int b) { return D + a * 6 - a * (a - 1) / 2 + (b - a); }
};
// upper triangle (p <= c) of an L x L symmetric matrix followed by an L vector
__host__ __device__ constexpr int tri(int L, int p, int c) { return p * L - p * (p - 1) / 2 + (c - p); }

__device__ __forceinline__ void load_point(const Dev& d, long long i, double (&p)[5]) {
    p[0] = __ldg(d.X + i); p[1] = __ldg(d.Y + i); p[2] = __ldg(d.Z + i);
    p[3] = __ldg(d.U + i); p[4] = __ldg(d.V + i);
}

// deterministic sum of n partials by one warp: strided per-lane chains, then the butterfly
__device__ __forceinline__ double warp_sum_array(const double* __restrict__ part, int n, int lane) {
    double acc = 0.0;
    for (int i = lane; i < n; i += 32) acc = __dadd_rn(acc, part[i]);
    return warp_sum(acc);
}

// ---- view constants for parameter buffer `which` --------------------------------------------
__global__ void k_make_views(Dev d, int which_is_cur) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= d.M) return;
    const int b = which_is_cur ? d.st->cur : (d.st->cur ^ 1);
    ViewC w;
    make_view(d.Pbuf + (long long)b * d.K + d.L + 6ll * v, w);
    d.views[(long long)b * d.M + v] = w;
}

// ---- pass A: intrinsic block, its right-hand side, and the cost ----------------------------
template <int MODEL>
__global__ void __launch_bounds__(kEvalThreads, 2) k_eval_A(Dev d) {
    using T = ModelTraits<MODEL>;
    constexpr int ND = T::ND, L = T::L;
    using S = SlotsA<ND>;
    const LmState* st = d.st;
    if (st->done || !st->need_eval) return;
    __shared__ double sc[16];
    __shared__ __align__(16) ViewC sview[kWarpsPerBlock];
    __shared__ double sred[kWarpsPerBlock][kNA];
    __shared__ double sE[kWarpsPerBlock];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int cur = st->cur;
    const double* P = d.Pbuf + (long long)cur * d.K;
    if (threadIdx.x < L) sc[threadIdx.x] = P[threadIdx.x];
    const int t = blockIdx.x * kWarpsPerBlock + warp;
    Task task = {0, 0, 0, 0};
    if (t < d.ntasks) {
        task = d.tasks[t];
        const double* src = reinterpret_cast<const double*>(d.views + (long long)cur * d.M + task.view);
        if (lane < kViewDoubles) reinterpret_cast<double*>(&sview[warp])[lane] = src[lane];
    }
    __syncthreads();
    double acc[kNA];
#pragma unroll
    for (int i = 0; i < kNA; ++i) acc[i] = 0.0;
    double e = 0.0;
    const ViewC& w = sview[warp];
    double p[5], pn[5];
    if (lane < task.count) load_point(d, task.start + lane, p);
    for (int i = lane; i < task.count; i += 32) {
        if (i + 32 < task.count) load_point(d, task.start + i + 32, pn);
        PointOut<MODEL> o;
        point_eval<MODEL, true, false>(sc, w, p[0], p[1], p[2], p[3], p[4], o);
        e = __dadd_rn(e, fma(o.rv, o.rv, __dmul_rn(o.ru, o.ru)));
        acc[0] = fma(o.xd, o.xd, acc[0]);
        acc[1] = fma(o.xd, o.yd, acc[1]);
        acc[2] += o.xd;
        acc[3] = fma(o.yd, o.yd, acc[3]);
        acc[4] += o.yd;
        acc[5] += 1.0;
#pragma unroll
        for (int a = 0; a < ND; ++a) {
            acc[S::UK + 3 * a + 0] = fma(o.xd, o.ku[a], acc[S::UK + 3 * a + 0]);
            acc[S::UK + 3 * a + 1] = fma(o.yd, o.ku[a], acc[S::UK + 3 * a + 1]);
            acc[S::UK + 3 * a + 2] += o.ku[a];
            acc[S::VK + 2 * a + 0] = fma(o.yd, o.kv[a], acc[S::VK + 2 * a + 0]);
            acc[S::VK + 2 * a + 1] += o.kv[a];
#pragma unroll
            for (int b = a; b < ND; ++b)
                acc[S::kk(a, b)] = fma(o.ku[a], o.ku[b], fma(o.kv[a], o.kv[b], acc[S::kk(a, b)]));
            acc[S::G + 5 + a] = fma(o.ku[a], o.ru, fma(o.kv[a], o.rv, acc[S::G + 5 + a]));
        }
        acc[S::G + 0] = fma(o.xd, o.ru, acc[S::G + 0]);
        acc[S::G + 1] = fma(o.yd, o.rv, acc[S::G + 1]);
        acc[S::G + 2] = fma(o.yd, o.ru, acc[S::G + 2]);
        acc[S::G + 3] += o.ru;
        acc[S::G + 4] += o.rv;
#pragma unroll
        for (int j = 0; j < 5; ++j) p[j] = pn[j];
    }
    warp_reduce_scatter<kNA>(acc, lane);
    e = warp_sum(e);
    sred[warp][2 * lane] = acc[0];
    sred[warp][2 * lane + 1] = acc[1];
    if (lane == 0) sE[warp] = e;
    __syncthreads();
    if (threadIdx.x < kNA) {
        double s = sred[0][threadIdx.x];
#pragma unroll
        for (int k = 1; k < kWarpsPerBlock; ++k) s += sred[k][threadIdx.x];
        d.partA[(long long)blockIdx.x * kNA + threadIdx.x] = s;
    } else if (threadIdx.x == kNA) {
        double s = sE[0];
#pragma unroll
        for (int k = 1; k < kWarpsPerBlock; ++k) s = __dadd_rn(s, sE[k]);
        d.partE[blockIdx.x] = s;
    }
}

// ---- pass BD: intrinsic x view coupling B_v, view block D_v, view right-hand side ------------
template <int MODEL>
__global__ void __launch_bounds__(kEvalThreads, 2) k_eval_BD(Dev d) {
    using T = ModelTraits<MODEL>;
    constexpr int ND = T::ND, L = T::L;
    using S = SlotsBD<L>;
    const LmState* st = d.st;
    if (st->done || !st->need_eval) return;
    __shared__ double sc[16];
    __shared__ __align__(16) ViewC sview[kWarpsPerBlock];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int cur = st->cur;
    const double* P = d.Pbuf + (long long)cur * d.K;
    if (threadIdx.x < L) sc[threadIdx.x] = P[threadIdx.x];
    const int t = blockIdx.x * kWarpsPerBlock + warp;
    Task task = {0, 0, 0, 0};
    if (t < d.ntasks) {
        task = d.tasks[t];
        const double* src = reinterpret_cast<const double*>(d.views + (long long)cur * d.M + task.view);
        if (lane < kViewDoubles) reinterpret_cast<double*>(&sview[warp])[lane] = src[lane];
    }
    __syncthreads();
    if (t >= d.ntasks) return;
    double acc[kNBD];
#pragma unroll
    for (int i = 0; i < kNBD; ++i) acc[i] = 0.0;
    const ViewC& w = sview[warp];
    double p[5], pn[5];
    if (lane < task.count) load_point(d, task.start + lane, p);
    for (int i = lane; i < task.count; i += 32) {
        if (i + 32 < task.count) load_point(d, task.start + i + 32, pn);
        PointOut<MODEL> o;
        point_eval<MODEL, true, true>(sc, w, p[0], p[1], p[2], p[3], p[4], o);
#pragma unroll
        for (int j = 0; j < 6; ++j) {
            acc[S::B + 0 * 6 + j] = fma(o.xd, o.ju[j], acc[S::B + 0 * 6 + j]);   // alpha
            acc[S::B + 1 * 6 + j] = fma(o.yd, o.jv[j], acc[S::B + 1 * 6 + j]);   // beta
            acc[S::B + 2 * 6 + j] = fma(o.yd, o.ju[j], acc[S::B + 2 * 6 + j]);   // gamma
            acc[S::B + 3 * 6 + j] += o.ju[j];                                     // uc
            acc[S::B + 4 * 6 + j] += o.jv[j];                                     // vc
#pragma unroll
            for (int a = 0; a < ND; ++a)
                acc[S::B + (5 + a) * 6 + j] =
                    fma(o.ku[a], o.ju[j], fma(o.kv[a], o.jv[j], acc[S::B + (5 + a) * 6 + j]));
#pragma unroll
            for (int b = j; b < 6; ++b)
                acc[S::dd(j, b)] = fma(o.ju[j], o.ju[b], fma(o.jv[j], o.jv[b], acc[S::dd(j, b)]));
            acc[S::GV + j] = fma(o.ju[j], o.ru, fma(o.jv[j], o.rv, acc[S::GV + j]));
        }
#pragma unroll
        for (int j = 0; j < 5; ++j) p[j] = pn[j];
    }
    warp_reduce_scatter<kNBD>(acc, lane);
#pragma unroll
    for (int j = 0; j < kNBD / 32; ++j)
        d.partBD[(long long)(lane * (kNBD / 32) + j) * d.Tpad + t] = acc[j];
}

// ---- cost-only pass at the trial point ------------------------------------------------------
__device__ void lm_accept(const Dev& d);

template <int MODEL>
__global__ void __launch_bounds__(kEvalThreads) k_cost(Dev d, int fuse_accept) {
    LmState* st = d.st;
    if (st->done) return;
    constexpr int L = ModelTraits<MODEL>::L;
    __shared__ double sc[16];
    __shared__ __align__(16) ViewC sview[kWarpsPerBlock];
    __shared__ double sE[kWarpsPerBlock];
    __shared__ int s_last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tb = st->cur ^ 1;
    const double* P = d.Pbuf + (long long)tb * d.K;
    if (threadIdx.x < L) sc[threadIdx.x] = P[threadIdx.x];
    const int t = blockIdx.x * kWarpsPerBlock + warp;
    Task task = {0, 0, 0, 0};
    if (t < d.ntasks) {
        task = d.tasks[t];
        const double* src = reinterpret_cast<const double*>(d.views + (long long)tb * d.M + task.view);
        if (lane < kViewDoubles) reinterpret_cast<double*>(&sview[warp])[lane] = src[lane];
    }
    __syncthreads();
    double e = 0.0;
    const ViewC& w = sview[warp];
    double p[5], pn[5];
    if (lane < task.count) load_point(d, task.start + lane, p);
    for (int i = lane; i < task.count; i += 32) {
        if (i + 32 < task.count) load_point(d, task.start + i + 32, pn);
        PointOut<MODEL> o;
        point_eval<MODEL, false, false>(sc, w, p[0], p[1], p[2], p[3], p[4], o);
        e = __dadd_rn(e, fma(o.rv, o.rv, __dmul_rn(o.ru, o.ru)));
#pragma unroll
        for (int j = 0; j < 5; ++j) p[j] = pn[j];
    }
    e = warp_sum(e);
    if (lane == 0) sE[warp] = e;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = sE[0];
#pragma unroll
        for (int k = 1; k < kWarpsPerBlock; ++k) s = __dadd_rn(s, sE[k]);
        d.partC[blockIdx.x] = s;
        __threadfence();
        const unsigned int prev = atomicAdd(&st->cnt_cost, 1u);
        s_last = (prev == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (warp == 0) {
        const double tot = warp_sum_array(d.partC, d.nblocks, lane);
        if (lane == 0) {
            d.etrial[0] = tot;
            st->cnt_cost = 0;
            if (fuse_accept) lm_accept(d);
        }
    }
}

// ---- accept / reject, lambda update, termination: src/calibrate.py:155-168 ------------------
__device__ void lm_accept(const Dev& d) {
    LmState* st = d.st;
    const double E = st->E_cur, E1 = d.etrial[0], lam = st->lam;
    const int it = st->iter;
    const bool acc = E1 < E;   // strict; NaN rejects
    double* tr = d.trace + 4ll * it;
    tr[0] = E; tr[1] = E1; tr[2] = lam; tr[3] = acc ? 1.0 : 0.0;
    double nl;
    if (acc) { st->cur ^= 1; nl = lam / 10.0; st->need_eval = 1; }
    else     { nl = lam * 10.0; st->need_eval = 0; }
    st->lam = nl;
    st->E_trial = E1;
    st->sse = E;
    st->iter = it + 1;
    if (!(st->lam_min < nl && nl < st->lam_max) || E < st->err_min || it + 1 >= st->max_iters) st->done = 1;
}

__global__ void k_accept(Dev d) {
    if (d.st->done) return;
    lm_accept(d);
}

}  // namespace calib
