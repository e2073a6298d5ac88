// Batch of independent calibrations (BASELINE config 4: a multi-camera rig sweep): one thread block
// per problem, the WHOLE LM loop of src/calibrate.py:141-171 inside one kernel launch.  The
// problem's detections, both parameter vectors, the view constants, every block of the normal
// equations and the Schur factors live in shared memory; nothing but the final parameters, the
// cost, the iteration count (and, on request, the trace) goes back to global memory.
// The per-point math and the accumulation are the same device functions the single-problem
// kernels use (point_eval, accum_A, accum_BD); the small dense stages are re-mapped for a block:
// one THREAD per view for the 6x6 Cholesky and forward substitution, one thread per entry of
// W = sum_v Z_v^T Z_v, one warp for the reduced L x L solve.
#pragma once
#include "calib_solve.cuh"

namespace calib {

#ifndef CALIB_BATCH_WARPS
#define CALIB_BATCH_WARPS 4
#endif
#ifndef CALIB_BATCH_BLOCKS
#define CALIB_BATCH_BLOCKS 2
#endif
constexpr int kBatchWarps = CALIB_BATCH_WARPS;        // warps per problem
constexpr int kBatchBlocks = CALIB_BATCH_BLOCKS;      // problems resident per SM (register and shared-memory budget)
constexpr int kBatchThreads = kBatchWarps * 32;
constexpr int kBDStride = kNBD + 1;     // 97: odd stride in doubles -> conflict-free thread-per-view reads
constexpr int kVtStride = kVtab + 1;    // 89

struct BatchDev {
    int model, L, n_problems, max_views, max_points, stage_points, max_iters, want_trace;
    const long long* prob_view;   // n_problems + 1 : first view of problem p
    const long long* view_off;    // total views + 1 : first point of view v
    const double *X, *Y, *Z, *U, *V;
    double* P;                    // concatenated parameter vectors; problem p starts at L p + 6 prob_view[p]
    double* sse;                  // n_problems
    int* iters;                   // n_problems
    double* trace;                // n_problems x max_iters x 4 (if want_trace)
    double lam0, lam_min, lam_max, err_min;
#ifdef CALIB_STAMPS
    unsigned long long* phase;    // kBatchPhases cycle counters summed over the blocks (profiling build only)
#endif
};

// Cycle accounting per phase of the loop (profiling build, -DCALIB_STAMPS): thread 0 reads the SM clock after
// every block barrier (counters in shared memory: the kernel has no register to spare); the sums over all
// blocks go to BatchDev::phase (profiles/batch_phases.py).
constexpr int kBatchPhases = 12;
#ifdef CALIB_STAMPS
#define BPHASE(k) do { if (tid == 0) { const long long now_ = clock64(); ph[k] += now_ - ph[kBatchPhases]; ph[kBatchPhases] = now_; } __syncwarp(); } while (0)
#else
#define BPHASE(k) do {} while (0)
#endif

struct BatchSmem {      // offsets in doubles into the dynamic shared memory
    int P, views, bd, vt, redA, misc, idx, sc, voff, pview, pts, total;
};

__host__ __device__ inline BatchSmem batch_smem_layout(int L, int max_views, int max_points, int stage_points) {
    BatchSmem s;
    const int K = L + 6 * max_views;
    int o = 0;
    s.P = o;     o += 2 * ((K + 1) & ~1);
    s.views = o; o += 2 * max_views * kViewDoubles;
    s.bd = o;    o += max_views * kBDStride;
    s.vt = o;    o += max_views * kVtStride;
    s.redA = o;  o += kBatchWarps * kNA + kNA + kNBD;     // per-warp A rows, summed A, W
    s.misc = o;  o += 64 + 112;                            // per-warp costs, dc, scalars; dense reduced system
#ifdef CALIB_STAMPS
    o += kBatchPhases + 2;                                 // phase counters
#endif
    s.idx = o;   o += 112;                                 // per entry of the dense reduced system: its slots in A and in W
    s.sc = o;    o += 6 * max_views;                       // sines and cosines of the trial Euler angles
    s.voff = o;  o += (max_views + 2) / 2;                 // first point of every view (int), relative to the problem
    s.pview = o; o += stage_points ? (max_points + 7) / 8 : 0;   // view of every point (one byte each), with the staged points
    o = (o + 1) & ~1;
    s.pts = o;   o += stage_points ? 5 * max_points : 0;
    s.total = o;
    return s;
}

// Per-view Schur stage in two steps.  Factor: ONE thread per view (bd = the view's summed
// (B_v | D_v | g_v) row; vt gets the packed strictly-lower factor with 1/diagonal on the diagonal).
template <int MODEL>
__device__ __forceinline__ void view_factor_thread(const double* bd, double lam, double* vt) {
    constexpr int L = ModelTraits<MODEL>::L;
    using S = SlotsBD<L>;
    double Lc[36], id[6];
#pragma unroll
    for (int a = 0; a < 6; ++a)
#pragma unroll
        for (int b = a; b < 6; ++b) {
            const double s = bd[S::dd(a, b)];
            Lc[b * 6 + a] = (a == b) ? fma(lam, s, s) : s;
            if (a != b) Lc[a * 6 + b] = 0.0;
        }
    chol_inplace<6>(Lc, id);
#pragma unroll
    for (int a = 0; a < 6; ++a)
#pragma unroll
        for (int b = 0; b <= a; ++b) vt[a * (a + 1) / 2 + b] = (a == b) ? id[a] : Lc[a * 6 + b];
}

// Forward substitution L z = column c of [B_v^T | g_v]: one thread per (view, column), so the L + 1
// columns of a view are solved side by side instead of one after the other by the factoring thread.
template <int MODEL>
__device__ __forceinline__ void view_column_thread(const double* bd, int c, double* vt) {
    constexpr int L = ModelTraits<MODEL>::L;
    using S = SlotsBD<L>;
    const double* rhs = bd + ((c < L) ? (S::B + c * 6) : S::GV);
    double z[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        double s = rhs[i];
#pragma unroll
        for (int k = 0; k < 6; ++k)
            if (k < i) s = fma(-vt[i * (i + 1) / 2 + k], z[k], s);
        z[i] = s * vt[i * (i + 1) / 2 + i];
        vt[21 + c * 6 + i] = z[i];
    }
}

// Back-substitution of one view by one thread: x = L^-T (z_g - Z^T dc)
template <int MODEL>
__device__ __forceinline__ void view_backsub_thread(const double* vt, const double* dc, double (&x)[6]) {
    constexpr int L = ModelTraits<MODEL>::L;
#pragma unroll
    for (int i = 0; i < 6; ++i) x[i] = vt[21 + L * 6 + i];
#pragma unroll
    for (int p = 0; p < L; ++p) {          // p outer: six independent chains (per element the same order of operations)
        const double dcp = dc[p];
#pragma unroll
        for (int i = 0; i < 6; ++i) x[i] = fma(-vt[21 + p * 6 + i], dcp, x[i]);
    }
#pragma unroll
    for (int i = 5; i >= 0; --i) {
        double s = x[i];
#pragma unroll
        for (int k = 0; k < 6; ++k)
            if (k > i) s = fma(-vt[k * (k + 1) / 2 + i], x[k], s);
        x[i] = s * vt[i * (i + 1) / 2 + i];
    }
}

// The view a point of the problem belongs to when there is no per-point table (detections not staged):
// last v with voff[v] <= i.
__device__ __forceinline__ int batch_view_search(const int* voff, int M, int i) {
    int lo = 0, hi = M - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (voff[mid] <= i) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// Per-view Schur stage of the views v = slot, slot + nslots, ... of a problem by one warp: lane k factors the
// warp's k-th view (6x6 Cholesky, one thread), then the (L + 1) forward substitutions of those views are spread
// over the lanes -- warp-local, no block barrier in between.
template <int MODEL>
__device__ __forceinline__ void batch_schur_views(const double* sBD, double* vtb, int M, int slot, int nslots, int lane, double lam) {
    constexpr int L = ModelTraits<MODEL>::L;
    const int nv = M > slot ? (M - slot + nslots - 1) / nslots : 0;
    for (int k = lane; k < nv; k += 32) {
        const int v = slot + nslots * k;
        view_factor_thread<MODEL>(sBD + v * kBDStride, lam, vtb + v * kVtStride);
    }
    __syncwarp();
    for (int it = lane; it < nv * (L + 1); it += 32) {
        const int k = it / (L + 1), c = it - k * (L + 1), v = slot + nslots * k;
        view_column_thread<MODEL>(sBD + v * kBDStride, c, vtb + v * kVtStride);
    }
}

// W = sum_v Z_v^T Z_v and w = sum_v Z_v^T z_g: one thread per entry (t = t0, t0 + nthreads, ...); the dot
// products of four views are independent of each other, the sum over the views keeps its order.
template <int MODEL>
__device__ __forceinline__ void batch_schur_w(const double* vtb, double* Wb, int M, int t0, int nthreads) {
    constexpr int L = ModelTraits<MODEL>::L, NT = L * (L + 1) / 2, NWE = NT + L;
    for (int t = t0; t < NWE; t += nthreads) {
        int e = t, p, c;
        if (e < NT) { p = 0; while (e >= L - p) { e -= L - p; ++p; } c = p + e; }
        else { p = e - NT; c = L; }
        const double* zp = vtb + 21 + p * 6;
        const double* zc = vtb + 21 + c * 6;
        double s = 0.0;
        int v = 0;
        for (; v + 4 <= M; v += 4) {
            double tt[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                tt[u] = 0.0;
#pragma unroll
                for (int i = 0; i < 6; ++i) tt[u] = fma(zp[(v + u) * kVtStride + i], zc[(v + u) * kVtStride + i], tt[u]);
            }
            s += tt[0]; s += tt[1]; s += tt[2]; s += tt[3];
        }
        for (; v < M; ++v) {
            double tt = 0.0;
#pragma unroll
            for (int i = 0; i < 6; ++i) tt = fma(zp[v * kVtStride + i], zc[v * kVtStride + i], tt);
            s += tt;
        }
        Wb[t] = s;
    }
}

template <int MODEL>
__global__ void __launch_bounds__(kBatchThreads, kBatchBlocks) k_batch_lm(BatchDev d) {
    using T = ModelTraits<MODEL>;
    constexpr int L = T::L, NT = L * (L + 1) / 2, NWE = NT + L;
    extern __shared__ __align__(16) double smem[];
    const BatchSmem lay = batch_smem_layout(L, d.max_views, d.max_points, d.stage_points);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int prob = blockIdx.x;
    const long long v0 = d.prob_view[prob];
    const int M = (int)(d.prob_view[prob + 1] - v0);
    const int K = L + 6 * M;
    const int Kp = (L + 6 * d.max_views + 1) & ~1;
    const long long pt0 = d.view_off[v0];
    const int npts = (int)(d.view_off[v0 + M] - pt0);
    double* gP = d.P + (long long)L * prob + 6 * v0;

    double* sP = smem + lay.P;                                   // [2][Kp]
    ViewC* sViews = reinterpret_cast<ViewC*>(smem + lay.views);  // [2][max_views]
    double* sBD = smem + lay.bd;                                 // [M][97]
    double* sVt = smem + lay.vt;                                 // [M][89]
    double* sRedA = smem + lay.redA;                             // [kBatchWarps][64]
    double* sA = sRedA + kBatchWarps * kNA;                      // [64]
    double* sW = sA + kNA;                                       // [96]
    double* sE = smem + lay.misc;                                // [kBatchWarps]
    double* sdc = sE + 8;                                        // [L]
    double* sScal = sdc + 16;                                    // lam, E_cur, E_trial, sse
    int* sInt = reinterpret_cast<int*>(sScal + 8);               // cur, need_eval, done, iter
    double* sDense = sE + 64;                                    // [L * L + L]

    short2* sIdx = reinterpret_cast<short2*>(smem + lay.idx);    // [L * L]: slot in A (or -1), slot in W
    double* sSC = smem + lay.sc;                                 // [M][6]
    int* sVoff = reinterpret_cast<int*>(smem + lay.voff);        // [M + 1]
    unsigned char* sPV = reinterpret_cast<unsigned char*>(smem + lay.pview);   // [npts]

    // ---- set-up: parameters, view constants, detections, index tables
    for (int i = tid; i < K; i += kBatchThreads) sP[i] = gP[i];
    for (int v = tid; v <= M; v += kBatchThreads) sVoff[v] = (int)(d.view_off[v0 + v] - pt0);
    for (int e = tid; e < L * L; e += kBatchThreads) {
        const int i = e / L, c = e - i * L;
        const int lo = i < c ? i : c, hi = i < c ? c : i;
        sIdx[e] = make_short2((short)a_slot<T::ND>(lo, hi), (short)tri(L, lo, hi));
    }
    const double *pX, *pY, *pZ, *pU, *pV;
    if (d.stage_points) {
        double* sp = smem + lay.pts;
        const int n = d.max_points;
        for (int i = tid; i < npts; i += kBatchThreads) {
            sp[i] = d.X[pt0 + i]; sp[n + i] = d.Y[pt0 + i]; sp[2 * n + i] = d.Z[pt0 + i];
            sp[3 * n + i] = d.U[pt0 + i]; sp[4 * n + i] = d.V[pt0 + i];
        }
        pX = sp; pY = sp + n; pZ = sp + 2 * n; pU = sp + 3 * n; pV = sp + 4 * n;
    } else {
        pX = d.X + pt0; pY = d.Y + pt0; pZ = d.Z + pt0; pU = d.U + pt0; pV = d.V + pt0;
    }
    if (tid == 0) {
        sScal[0] = d.lam0; sScal[1] = 0.0; sScal[2] = 0.0; sScal[3] = 0.0;
        sInt[0] = 0; sInt[1] = 1; sInt[2] = 0; sInt[3] = 0;
    }
    __syncthreads();
    if (d.stage_points)                                          // the view of every point: the flat sweeps look it up
        for (int v = warp; v < M; v += kBatchWarps)
            for (int i = sVoff[v] + lane; i < sVoff[v + 1]; i += 32) sPV[i] = (unsigned char)v;
    for (int v = tid; v < M; v += kBatchThreads) make_view(sP + L + 6 * v, sViews[v]);
    __syncthreads();
#ifdef CALIB_STAMPS
    volatile long long* ph = reinterpret_cast<volatile long long*>(smem + lay.misc + 64 + 112);
    if (tid == 0) {
        for (int i = 0; i < kBatchPhases; ++i) ph[i] = 0;
        ph[kBatchPhases] = clock64();
        ph[kBatchPhases + 1] = ph[kBatchPhases];
    }
    __syncwarp();
#endif
    // views of this warp in the warp-local stages (per-view factor, back-substitution): v = warp + kBatchWarps k
    const int nv_w = M > warp ? (M - warp + kBatchWarps - 1) / kBatchWarps : 0;

    // ---- the loop: src/calibrate.py:143-168
    for (;;) {
        const int cur = sInt[0];
        const double lam = sScal[0];
        const double* Pc = sP + cur * Kp;
        double* Pt = sP + (cur ^ 1) * Kp;
        const ViewC* vc = sViews + cur * d.max_views;
        ViewC* vt_views = sViews + (cur ^ 1) * d.max_views;
        if (sInt[1]) {
            // sweep 1: B_v, D_v, g_v, one warp per view
            for (int v = warp; v < M; v += kBatchWarps) {
                const int s0 = sVoff[v], cnt = sVoff[v + 1] - s0;
                double acc[kNBD];
#pragma unroll
                for (int i = 0; i < kNBD; ++i) acc[i] = 0.0;
                for (int i = lane; i < cnt; i += 32) {
                    PointOut<MODEL> o;
                    point_eval<MODEL, true, true>(Pc, vc[v], pX[s0 + i], pY[s0 + i], pZ[s0 + i], pU[s0 + i], pV[s0 + i], o);
                    accum_BD<MODEL>(o, acc);
                }
                warp_reduce_scatter<kNBD>(acc, lane);
                double* row = sBD + v * kBDStride + lane * (kNBD / 32);
#pragma unroll
                for (int j = 0; j < kNBD / 32; ++j) row[j] = acc[j];
            }
            // sweep 2: A, g_c, E.  Flat over the problem's points (thread t takes points t, t + 128, ...: no idle
            // lanes at the end of a 54-point view; measured against the per-view form: 13.1 against 13.6 ms at
            // cfg4 together with the flat cost sweep); the cost sweep below visits the points in the same order,
            // so E and E' are sums of the same shape.
            {
                double acc[kNA];
#pragma unroll
                for (int i = 0; i < kNA; ++i) acc[i] = 0.0;
                double e = 0.0;
                for (int i = tid; i < npts; i += kBatchThreads) {
                    PointOut<MODEL> o;
                    const int v = d.stage_points ? (int)sPV[i] : batch_view_search(sVoff, M, i);
                    point_eval<MODEL, true, false>(Pc, vc[v], pX[i], pY[i], pZ[i], pU[i], pV[i], o);
                    accum_A<MODEL>(o, acc, e);
                }
                warp_reduce_scatter<kNA>(acc, lane);
                e = warp_sum(e);
                sRedA[warp * kNA + 2 * lane] = acc[0];
                sRedA[warp * kNA + 2 * lane + 1] = acc[1];
                if (lane == 0) sE[warp] = e;
            }
            __syncthreads();
            BPHASE(0);      // the two evaluation sweeps
            for (int t = tid; t <= kNA; t += kBatchThreads) {
                if (t < kNA) {
                    double s = sRedA[t];
#pragma unroll
                    for (int k = 1; k < kBatchWarps; ++k) s += sRedA[k * kNA + t];
                    sA[t] = s;
                } else {
                    double s = sE[0];
#pragma unroll
                    for (int k = 1; k < kBatchWarps; ++k) s = __dadd_rn(s, sE[k]);
                    if (sInt[3] == 0) sScal[1] = s;      // later: the trial cost of the last accepted step
                }
            }
        }
        // Schur stage (warp-local factor + forward substitutions, then W).  Measured and rejected (profiles/README.md):
        // building the stage of the damping a rejected step leads to (10 lam) in the warps that idle during the
        // reduced solve -- the other block of the SM loses more than this one gains, 14.0 against 13.1 ms at cfg4.
        batch_schur_views<MODEL>(sBD, sVt, M, warp, kBatchWarps, lane, lam);
        __syncthreads();
        BPHASE(1);          // sum of A over the warps, 6x6 factors, forward substitutions
        batch_schur_w<MODEL>(sVt, sW, M, tid, kBatchThreads);
        __syncthreads();
        BPHASE(3);          // W
        // reduced system: lane i of warp 0 puts row i of S = A + lam diag(A) - W and s_i = g_c - w in dense form
        // (slots from the table built at set-up), then the warp solves it -> dc, trial shared parameters
        if (warp == 0) {
            using SA = SlotsA<T::ND>;
            if (lane < L) {
#pragma unroll
                for (int c = 0; c < L; ++c) {
                    const short2 ix = sIdx[lane * L + c];
                    double v = ix.x >= 0 ? sA[ix.x] : 0.0;
                    if (c == lane) v = fma(lam, v, v);
                    sDense[lane * L + c] = v - sW[ix.y];
                }
                sDense[L * L + lane] = sA[SA::G + lane] - sW[NT + lane];
            }
            __syncwarp();
            const double dc = solve_dense_core<L>(sDense, lane);
            if (lane < L) { sdc[lane] = dc; Pt[lane] = Pc[lane] + dc; }
        }
        __syncthreads();
        BPHASE(5);          // dense form + reduced solve, one warp
        // back-substitution and trial view constants, warp-local: lane k back-substitutes the warp's k-th view,
        // the 3 sincos of every view go to 3 lanes, lane k assembles the constants
        for (int k = lane; k < nv_w; k += 32) {
            const int v = warp + kBatchWarps * k;
            double x[6];
            view_backsub_thread<MODEL>(sVt + v * kVtStride, sdc, x);
#pragma unroll
            for (int i = 0; i < 6; ++i) Pt[L + 6 * v + i] = Pc[L + 6 * v + i] + x[i];
        }
        __syncwarp();
        for (int t = lane; t < 3 * nv_w; t += 32) {
            const int k = t / 3, a = t - 3 * k, v = warp + kBatchWarps * k;
            double sn, cs;
            sincos(__dmul_rn(Pt[L + 6 * v + a], kDeg), &sn, &cs);
            sSC[6 * v + 2 * a] = sn; sSC[6 * v + 2 * a + 1] = cs;
        }
        __syncwarp();
        for (int k = lane; k < nv_w; k += 32) {
            const int v = warp + kBatchWarps * k;
            const double* sc = sSC + 6 * v;
            make_view_sc(Pt + L + 6 * v, sc[0], sc[1], sc[2], sc[3], sc[4], sc[5], vt_views[v]);
        }
        __syncthreads();
        BPHASE(6);          // back-substitution + view constants
        // trial cost: flat over the points like sweep 2, three points per thread in flight (three independent
        // chains: the sweep is bound by the latency of one)
        {
            double e = 0.0;
            for (int base = tid; base < npts; base += 3 * kBatchThreads) {
                double r[3];
#pragma unroll
                for (int u = 0; u < 3; ++u) {
                    const int i = base + u * kBatchThreads;
                    const bool ok = i < npts;
                    const int ii = ok ? i : base;
                    PointOut<MODEL> o;
                    const int v = d.stage_points ? (int)sPV[ii] : batch_view_search(sVoff, M, ii);
                    point_eval<MODEL, false, false>(Pt, vt_views[v], pX[ii], pY[ii], pZ[ii], pU[ii], pV[ii], o);
                    r[u] = fma(o.rv, o.rv, __dmul_rn(o.ru, o.ru));
                }
                e = __dadd_rn(e, r[0]);
                if (base + kBatchThreads < npts) e = __dadd_rn(e, r[1]);
                if (base + 2 * kBatchThreads < npts) e = __dadd_rn(e, r[2]);
            }
            e = warp_sum(e);
            if (lane == 0) sE[warp] = e;
        }
        __syncthreads();
        BPHASE(7);          // trial cost sweep
        if (tid == 0) {
            double E1 = sE[0];
#pragma unroll
            for (int k = 1; k < kBatchWarps; ++k) E1 = __dadd_rn(E1, sE[k]);
            // accept / reject, lambda update, termination: src/calibrate.py:155-168
            const double E = sScal[1];
            const int it = sInt[3];
            const bool acc = E1 < E;   // strict; NaN rejects
            if (d.want_trace) {
                double* tr = d.trace + ((long long)prob * d.max_iters + it) * 4;
                tr[0] = E; tr[1] = E1; tr[2] = lam; tr[3] = acc ? 1.0 : 0.0;
            }
            double nl;
            if (acc) { sInt[0] = cur ^ 1; nl = lam / 10.0; sInt[1] = 1; sScal[1] = E1; }
            else     { nl = lam * 10.0; sInt[1] = 0; }
            sScal[0] = nl;
            sScal[3] = E;
            sInt[3] = it + 1;
            if (!(d.lam_min < nl && nl < d.lam_max) || E < d.err_min || it + 1 >= d.max_iters) sInt[2] = 1;
        }
        __syncthreads();
        BPHASE(8);          // accept / reject
        if (sInt[2]) break;
    }
#ifdef CALIB_STAMPS
    if (tid == 0 && d.phase) {
        ph[9] = clock64() - ph[kBatchPhases + 1];      // the whole loop
        ph[10] = sInt[3];                 // iterations
        ph[11] = 1;                       // problems
        for (int i = 0; i < kBatchPhases; ++i) atomicAdd(d.phase + i, (unsigned long long)ph[i]);
    }
#endif
    const double* Pf = sP + sInt[0] * Kp;
    for (int i = tid; i < K; i += kBatchThreads) gP[i] = Pf[i];
    if (tid == 0) { d.sse[prob] = sScal[3]; d.iters[prob] = sInt[3]; }
}

}  // namespace calib
