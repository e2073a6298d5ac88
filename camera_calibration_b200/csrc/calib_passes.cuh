// The two sweeps over the detections: the fused evaluation pass and the trial-cost pass.
#pragma once
#include "calib_stream.cuh"

namespace calib {

// How a pass brings the detections in.  Measured on B200 (profiles/README.md, "staging"):
//   k_eval is FP64-bound; coalesced 8-byte loads with one-step-ahead register prefetch
//   (sweep_pieces) give 60 us on cfg3 against 70 us through the bulk-copy ring, so it keeps them.
//   k_cost is bandwidth-bound; when the detections do not fit L2 the bulk-copy (TMA) ring with
//   2 x 5 KB in flight per warp wins (cfg5: 86 vs 99 us per 8.2 M points), when they are
//   L2-resident the plain loads win (cfg3: 15 vs 18 us).  The host picks per problem size.
constexpr int kWinSteps = 4;     // 32-point steps per window: five 1 KB bulk copies per window
constexpr int kCostStages = 2;   // windows in flight per warp: 16 warps x 2 x 5 KB = 160 KB per SM
using CostRing = PointRing<kCostStages, kWinSteps>;
constexpr size_t kCostDynSmem = kBlockWarps * sizeof(CostRing);

extern __shared__ __align__(128) unsigned char dyn_smem[];

template <bool TMA, int D, int WS, class Begin, class Body, class End>
__device__ __forceinline__ void sweep(const Dev& d, const Piece* __restrict__ pieces, const WarpRec& rec,
                                      const ViewC* __restrict__ views, ViewC& sview, int warp, unsigned long long* bars,
                                      int lane, Begin begin, Body body, End end) {
    if constexpr (TMA) {
        PointRing<D, WS>* ring = reinterpret_cast<PointRing<D, WS>*>(dyn_smem);
        sweep_windows<D, WS>(d, pieces, rec, views, sview, ring[warp], bars, lane, begin, body, end);
    } else {
        sweep_pieces(d, pieces, rec, views, sview, lane, begin, body, end);
    }
}

// ---- the fused evaluation pass ----------------------------------------------------------------
template <int MODEL>
__global__ void __launch_bounds__(kBlockThreads, 1) k_eval(Dev d) {
    using T = ModelTraits<MODEL>;
    constexpr int L = T::L;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int w = blockIdx.x * kBlockWarps + warp;
    WarpRec rec;                       // issued together with the state load below: one round trip
    rec.p0 = rec.p1 = 0;
    if (w < d.ev_warps) rec = d.ev_rec[w];
    const LmState* st = d.st;
    if (st->done || !st->need_eval) return;
    __shared__ double sc[16];
    __shared__ __align__(16) ViewC sview[kBlockWarps];
    __shared__ double sred[kBlockWarps][kNA];
    __shared__ double sE[kBlockWarps];
    const int cur = st->cur;
    const double* P = d.Pbuf + (long long)cur * d.K;
    const ViewC* views = d.views + (long long)cur * d.M;
    if (threadIdx.x < L) sc[threadIdx.x] = P[threadIdx.x];
    __syncthreads();
    // sweep 1: B_v, D_v, g_v -- one row of kNBD slots per piece
    {
        double acc[kNBD];
#pragma unroll
        for (int i = 0; i < kNBD; ++i) acc[i] = 0.0;
        sweep<false, 1, 1>(d, d.ev_piece, rec, views, sview[warp], warp, nullptr, lane,
            [&](const Piece&, int) {},
            [&](const Pt& p) {
                PointOut<MODEL> o;
                point_eval<MODEL, true, true>(sc, sview[warp], p.X, p.Y, p.Z, p.U, p.V, o);
                accum_BD<MODEL>(o, acc);
            },
            [&](const Piece&, int pi) {
                warp_reduce_scatter<kNBD>(acc, lane);
                double* row = d.partBD + (long long)pi * kNBD + lane * (kNBD / 32);
#pragma unroll
                for (int j = 0; j < kNBD / 32; ++j) row[j] = acc[j];
#pragma unroll
                for (int i = 0; i < kNBD; ++i) acc[i] = 0.0;
            });
    }
    // sweep 2: A, g_c, E -- accumulated over all pieces of the warp, one row per block
    {
        double acc[kNA];
#pragma unroll
        for (int i = 0; i < kNA; ++i) acc[i] = 0.0;
        double e = 0.0;
        sweep<false, 1, 1>(d, d.ev_piece, rec, views, sview[warp], warp, nullptr, lane,
            [&](const Piece&, int) {},
            [&](const Pt& p) {
                PointOut<MODEL> o;
                point_eval<MODEL, true, false>(sc, sview[warp], p.X, p.Y, p.Z, p.U, p.V, o);
                accum_A<MODEL>(o, acc, e);
            },
            [&](const Piece&, int) {});
        warp_reduce_scatter<kNA>(acc, lane);
        e = warp_sum(e);
        sred[warp][2 * lane] = acc[0];
        sred[warp][2 * lane + 1] = acc[1];
        if (lane == 0) sE[warp] = e;
    }
    __syncthreads();
    if (threadIdx.x < kNA) {
        double s = sred[0][threadIdx.x];
#pragma unroll
        for (int k = 1; k < kBlockWarps; ++k) s += sred[k][threadIdx.x];
        d.partA[(long long)blockIdx.x * kNA + threadIdx.x] = s;
    } else if (threadIdx.x == kNA) {
        double s = sE[0];
#pragma unroll
        for (int k = 1; k < kBlockWarps; ++k) s = __dadd_rn(s, sE[k]);
        d.partE[blockIdx.x] = s;
    }
}

// ---- accept / reject, lambda update, termination: src/calibrate.py:155-168 ------------------
// COND: the launch sits inside the WHILE node of the whole-loop graph and sets its condition.  (Kept
// out of the plain variants: a kernel that calls the device runtime cannot be profiled by ncu.)
template <bool COND>
__device__ __forceinline__ void lm_accept(const Dev& d) {
    LmState* st = d.st;
    const double E = st->E_cur, E1 = d.etrial[0], lam = st->lam;
    const int it = st->iter;
    const bool acc = E1 < E;   // strict; NaN rejects
    double* tr = d.trace + 4ll * it;
    tr[0] = E; tr[1] = E1; tr[2] = lam; tr[3] = acc ? 1.0 : 0.0;
    double nl;
    if (acc) { st->cur ^= 1; nl = lam / 10.0; st->need_eval = 1; st->E_cur = E1; }
    else     { nl = lam * 10.0; st->need_eval = 0; }
    st->lam = nl;
    st->E_trial = E1;
    st->sse = E;
    st->iter = it + 1;
    const int done = (!(st->lam_min < nl && nl < st->lam_max) || E < st->err_min || it + 1 >= st->max_iters) ? 1 : 0;
    if (done) st->done = 1;
    if (COND) cudaGraphSetConditional(d.cond, done ? 0u : 1u);
}

template <bool COND>
__global__ void k_accept(Dev d) {
    if (d.st->done) return;
    lm_accept<COND>(d);
}

// ---- cost-only sweep at the trial point -------------------------------------------------------
template <int MODEL, bool COND, bool TMA>
__global__ void __launch_bounds__(kBlockThreads, 2) k_cost(Dev d, int fuse_accept) {
    constexpr int L = ModelTraits<MODEL>::L;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int w = blockIdx.x * kBlockWarps + warp;
    WarpRec rec;
    rec.p0 = rec.p1 = 0;
    if (w < d.co_warps) rec = d.co_rec[w];
    LmState* st = d.st;
    if (st->done) {   // also after a failed exchange: the WHILE node must not spin
        if (COND) { if (blockIdx.x == 0 && threadIdx.x == 0) cudaGraphSetConditional(d.cond, 0u); }
        return;
    }
    __shared__ double sc[16];
    __shared__ __align__(16) ViewC sview[kBlockWarps];
    __shared__ double sE[kBlockWarps];
    __shared__ double s_tot;
    __shared__ int s_last;
    __shared__ __align__(8) unsigned long long bars[kBlockWarps][kCostStages];
    const int tb = st->cur ^ 1;
    const double* P = d.Pbuf + (long long)tb * d.K;
    const ViewC* views = d.views + (long long)tb * d.M;
    if (threadIdx.x < L) sc[threadIdx.x] = P[threadIdx.x];
    __syncthreads();
    double e = 0.0;
    sweep<TMA, kCostStages, kWinSteps>(d, d.co_piece, rec, views, sview[warp], warp, bars[warp], lane,
        [&](const Piece&, int) {},
        [&](const Pt& p) {
            PointOut<MODEL> o;
            point_eval<MODEL, false, false>(sc, sview[warp], p.X, p.Y, p.Z, p.U, p.V, o);
            e = __dadd_rn(e, fma(o.rv, o.rv, __dmul_rn(o.ru, o.ru)));
        },
        [&](const Piece&, int) {});
    e = warp_sum(e);
    if (lane == 0) sE[warp] = e;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = sE[0];
#pragma unroll
        for (int k = 1; k < kBlockWarps; ++k) s = __dadd_rn(s, sE[k]);
        d.partC[blockIdx.x] = s;
        __threadfence();
        const unsigned int prev = atomicAdd(&st->cnt_cost, 1u);
        s_last = (prev == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (warp == 0) {
        const double tot = warp_sum_array(d.partC, d.co_blocks, lane);
        if (lane == 0) s_tot = tot;
    }
    __syncthreads();
    bool ok = true;
    if (d.peer.nranks > 1) ok = peer_allreduce(d.peer, 1, &s_tot, 1, &s_last);   // views sharded: sum over the ranks
    if (threadIdx.x == 0) {
        d.etrial[0] = s_tot;
        st->cnt_cost = 0;
        if (!ok) {
            st->comm_fail = 1; st->done = 1;
            if (COND) cudaGraphSetConditional(d.cond, 0u);
        } else if (fuse_accept) {
            lm_accept<COND>(d);
        }
    }
}

}  // namespace calib
