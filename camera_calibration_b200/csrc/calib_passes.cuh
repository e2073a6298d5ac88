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

template <bool TMA, int D, int WS, int NV, int PF, class Begin, class Body, class End>
__device__ __forceinline__ void sweep(const Dev& d, const Piece* __restrict__ pieces, const WarpRec& rec,
                                      const ViewC* const (&views)[NV], ViewC (&sview)[NV], int warp,
                                      unsigned long long* bars, int lane, Begin begin, Body body, End end) {
    if constexpr (TMA) {
        PointRing<D, WS>* ring = reinterpret_cast<PointRing<D, WS>*>(dyn_smem);
        sweep_windows<D, WS, NV>(d, pieces, rec, views, sview, ring[warp], bars, lane, begin, body, end);
    } else {
        sweep_pieces<NV, PF>(d, pieces, rec, views, sview, lane, begin, body, end);
    }
}

// What one pass of the iteration works from.  The per-launch kernels read it from the device state; the
// persistent kernel keeps a copy per block (calib_persistent.cuh).
struct PassCtx { int cur, J, need_eval, iter; double lam; };
__device__ __forceinline__ PassCtx pass_ctx(const LmState& st) {
    PassCtx c;
    c.cur = st.cur; c.J = st.J; c.need_eval = st.need_eval; c.iter = st.iter; c.lam = st.lam;
    return c;
}

// ---- the fused evaluation pass ----------------------------------------------------------------
struct __align__(16) EvalSmem {
    ViewC sview[kBlockWarps][1];
    double sc[16];
    double sred[kBlockWarps][kNA];
    double sE[kBlockWarps];
};

// Block *vbp of the evaluation pass (warps vb * 8 .. vb * 8 + 7 of the partition, rec = this warp's
// record) at the point in parameter buffer `cur`: writes the piece rows of its warps and row vb of
// partA / partE.  vbp points to shared memory written before the call (not necessarily synchronised:
// it is read after the first barrier below) -- the index is needed after the sweeps only and must not
// cost a register inside them.  Called by all threads of a block; ends without a barrier after its
// last shared-memory reads.
template <int MODEL>
__device__ __forceinline__ void eval_block(const Dev& d, EvalSmem& sm, const volatile int* vbp, int cur, const WarpRec& rec) {
    using T = ModelTraits<MODEL>;
    constexpr int L = T::L;
    const int tid = tid_now();
    const int lane = tid & 31, warp = tid >> 5;
    STAMP(d, 0, 0);
    const double* P = d.Pbuf + (long long)cur * d.K;
    const ViewC* const views[1] = {d.views + (long long)cur * d.M};
    if (tid < L) sm.sc[tid] = P[tid];
    __syncthreads();
    // sweep 1: B_v, D_v, g_v -- one row of kNBD slots per piece
    {
        double acc[kNBD];
#pragma unroll
        for (int i = 0; i < kNBD; ++i) acc[i] = 0.0;
        sweep_pieces<1, 1>(d, d.ev_piece, rec, views, sm.sview[warp], lane,
            [&](const Piece&, int) {},
            [&](const Pt& p) {
                PointOut<MODEL> o;
                point_eval<MODEL, true, true>(sm.sc, sm.sview[warp][0], p.X, p.Y, p.Z, p.U, p.V, o);
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
    STAMP(d, 0, 1);
    // sweep 2: A, g_c, E -- accumulated over all pieces of the warp, one row per block.  The point
    // loop runs on a step count and a running index (COUNTED): the piece's start and count are not
    // live inside it, which removes its in-loop spill loads (profiles/README.md, static look at k_eval)
    {
        double acc[kNA];
#pragma unroll
        for (int i = 0; i < kNA; ++i) acc[i] = 0.0;
        double e = 0.0;
        sweep_pieces<1, 1, true>(d, d.ev_piece, rec, views, sm.sview[warp], lane,
            [&](const Piece&, int) {},
            [&](const Pt& p) {
                PointOut<MODEL> o;
                point_eval<MODEL, true, false>(sm.sc, sm.sview[warp][0], p.X, p.Y, p.Z, p.U, p.V, o);
                accum_A<MODEL>(o, acc, e);
            },
            [&](const Piece&, int) {});
        warp_reduce_scatter<kNA>(acc, lane);
        e = warp_sum(e);
        sm.sred[warp][2 * lane] = acc[0];
        sm.sred[warp][2 * lane + 1] = acc[1];
        if (lane == 0) sm.sE[warp] = e;
    }
    STAMP(d, 0, 2);
    __syncthreads();
    STAMP(d, 0, 3);
    const int vb = *vbp;
    if (tid < kNA) {
        double s = sm.sred[0][tid];
#pragma unroll
        for (int k = 1; k < kBlockWarps; ++k) s += sm.sred[k][tid];
        d.partA[(long long)vb * kNA + tid] = s;
    } else if (tid == kNA) {
        double s = sm.sE[0];
#pragma unroll
        for (int k = 1; k < kBlockWarps; ++k) s = __dadd_rn(s, sm.sE[k]);
        d.partE[vb] = s;
    }
    STAMP(d, 0, 4);
}

template <int MODEL>
__global__ void __launch_bounds__(kBlockThreads, 1) k_eval(Dev d) {
    const int warp = threadIdx.x >> 5;
    const int w = blockIdx.x * kBlockWarps + warp;
    WarpRec rec;                       // issued together with the state load below: one round trip
    rec.p0 = rec.p1 = 0;
    pdl_trigger();
    if (w < d.ev_warps) rec = d.ev_rec[w];        // static: may be read before the previous kernel is done
    pdl_wait();
    const LmState* st = d.st;
    if (st->done || !st->need_eval) return;
    __shared__ EvalSmem sm;
    __shared__ int s_vb;
    if (threadIdx.x == 0) s_vb = blockIdx.x;
    eval_block<MODEL>(d, sm, &s_vb, st->cur, rec);
}

// ---- accept / reject, lambda update, termination: src/calibrate.py:155-168 ------------------
// Replays the reference's decisions over the J candidates of this pass (calib_kernels.cuh, "damping
// ladder"): candidate j is the reference's iteration iter+j with damping lam * 10^j; the first
// accepted one ends the pass, the rest is dropped.  st: the state to advance (the device state, or a
// block's copy of it in the persistent kernel, where every block replays the same decisions from the
// same trial costs and only one writes the trace).  Returns done.
__device__ __forceinline__ int lm_accept_on(LmState* st, const double* etrial, double* trace) {
    const int J = st->J, cur = st->cur;
    double lam = st->lam;
    const double E = st->E_cur;
    int it = st->iter, done = 0;
    bool accepted = false;
    for (int j = 0; j < J && !done && !accepted; ++j) {
        const double E1 = etrial[j];
        const bool acc = E1 < E;   // strict; NaN rejects
        if (trace) {
            double* tr = trace + 4ll * it;
            tr[0] = E; tr[1] = E1; tr[2] = lam; tr[3] = acc ? 1.0 : 0.0;
        }
        if (acc) { st->cur = trial_buf(cur, j); lam = lam / 10.0; st->E_cur = E1; accepted = true; }
        else     { lam = lam * 10.0; }
        st->E_trial = E1;
        it += 1;
        done = (!(st->lam_min < lam && lam < st->lam_max) || E < st->err_min || it >= st->max_iters) ? 1 : 0;
    }
    st->need_eval = accepted ? 1 : 0;
    st->lam = lam;
    st->sse = E;
    st->iter = it;
    st->n_pass += 1;
    int nj = accepted ? 1 : 2 * J;
    if (nj > st->jmax) nj = st->jmax;
    if (nj > st->max_iters - it) nj = st->max_iters - it;
    st->J = nj < 1 ? 1 : nj;
    if (done) st->done = 1;
    return done;
}

// COND: the launch sits inside the WHILE node of the whole-loop graph and sets its condition.  (Kept
// out of the plain variants: a kernel that calls the device runtime cannot be profiled by ncu.)
template <bool COND>
__device__ __forceinline__ void lm_accept(const Dev& d) {
    const int done = lm_accept_on(d.st, d.etrial, d.trace);
    if (COND) cudaGraphSetConditional(d.cond, done ? 0u : 1u);
}

template <bool COND>
__global__ void k_accept(Dev d) {
    pdl_wait();
    if (d.st->done) return;
    lm_accept<COND>(d);
}

// ---- cost-only sweep at the J trial points ------------------------------------------------------
struct __align__(16) CostSmem {
    ViewC sview[kBlockWarps][kJMax];
    double sc[kJMax][16];
    double sE[kBlockWarps][kJMax];
    double s_tot[kJMax];
    unsigned long long bars[kBlockWarps][kCostStages];
    int s_last;
};

template <int MODEL, bool TMA, int NJ, int PF>
__device__ __forceinline__ void cost_sweep(const Dev& d, const WarpRec& rec, int cur, double (*sc)[16],
                                           ViewC (&sview)[kJMax], unsigned long long* bars, int warp, int lane,
                                           double (&e)[kJMax]) {
    const ViewC* views[NJ];
#pragma unroll
    for (int j = 0; j < NJ; ++j) views[j] = d.views + (long long)trial_buf(cur, j) * d.M;
    const ViewC* const (&cviews)[NJ] = views;
    ViewC (&sv)[NJ] = reinterpret_cast<ViewC (&)[NJ]>(sview);
    sweep<TMA, kCostStages, kWinSteps, NJ, PF>(d, d.co_piece, rec, cviews, sv, warp, bars, lane,
        [&](const Piece&, int) {},
        [&](const Pt& p) {
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
                PointOut<MODEL> o;
                point_eval<MODEL, false, false>(sc[j], sv[j], p.X, p.Y, p.Z, p.U, p.V, o);
                e[j] = __dadd_rn(e[j], fma(o.rv, o.rv, __dmul_rn(o.ru, o.ru)));
            }
        },
        [&](const Piece&, int) {});
}

// Block *vbp of the cost pass (vbp as in eval_block): the J trial costs of its warps' chunks -> row vb
// of partC.  Called by all threads of a block; starts with a barrier-protected load of the trial
// shared parameters.
template <int MODEL, bool TMA, int PF>
__device__ __forceinline__ void cost_block(const Dev& d, CostSmem& sm, const volatile int* vbp, int cur, int J, const WarpRec& rec) {
    constexpr int L = ModelTraits<MODEL>::L;
    const int tid = tid_now();
    const int lane = tid & 31, warp = tid >> 5;
    STAMP(d, 3, 0);
    if (tid < kJMax * 16) {
        const int j = tid >> 4, i = tid & 15;
        if (j < J && i < L) sm.sc[j][i] = d.Pbuf[(long long)trial_buf(cur, j) * d.K + i];
    }
    __syncthreads();
    double e[kJMax];
#pragma unroll
    for (int j = 0; j < kJMax; ++j) e[j] = 0.0;
    if (J == 1) cost_sweep<MODEL, TMA, 1, PF>(d, rec, cur, sm.sc, sm.sview[warp], sm.bars[warp], warp, lane, e);
    else if (J == 2) cost_sweep<MODEL, TMA, 2, PF>(d, rec, cur, sm.sc, sm.sview[warp], sm.bars[warp], warp, lane, e);
    else cost_sweep<MODEL, TMA, kJMax, PF>(d, rec, cur, sm.sc, sm.sview[warp], sm.bars[warp], warp, lane, e);
    STAMP(d, 3, 1);
#pragma unroll
    for (int j = 0; j < kJMax; ++j) {
        const double s = warp_sum(e[j]);
        if (lane == 0) sm.sE[warp][j] = s;
    }
    __syncthreads();
    if (tid < kJMax) {
        double s = sm.sE[0][tid];
#pragma unroll
        for (int k = 1; k < kBlockWarps; ++k) s = __dadd_rn(s, sm.sE[k][tid]);
        d.partC[(long long)*vbp * kJMax + tid] = s;
    }
}

// Sum of the block rows of partC per candidate, fixed order: warp j sums column j -> sm.s_tot[j].
// Called by all threads of a block (at least kJMax warps); ends with a barrier.
__device__ __forceinline__ void cost_total(const Dev& d, CostSmem& sm) {
    const int tid = tid_now();
    const int lane = tid & 31, warp = tid >> 5;
    if (warp < kJMax) {
        double acc = 0.0;
        for (int i = lane; i < d.co_blocks; i += 32) acc = __dadd_rn(acc, __ldcg(d.partC + (long long)i * kJMax + warp));
        acc = warp_sum(acc);
        if (lane == 0) sm.s_tot[warp] = acc;
    }
    __syncthreads();
}

template <int MODEL, bool COND, bool TMA>
__global__ void __launch_bounds__(kBlockThreads, 2) k_cost(Dev d, int fuse_accept) {
    const int warp = threadIdx.x >> 5;
    const int w = blockIdx.x * kBlockWarps + warp;
    WarpRec rec;
    rec.p0 = rec.p1 = 0;
    pdl_trigger();
    if (w < d.co_warps) rec = d.co_rec[w];
    pdl_wait();
    LmState* st = d.st;
    if (st->done) {   // also after a failed exchange: the WHILE node must not spin
        if (COND) { if (blockIdx.x == 0 && threadIdx.x == 0) cudaGraphSetConditional(d.cond, 0u); }
        return;
    }
    __shared__ CostSmem sm;
    __shared__ int s_vb;
    if (threadIdx.x == 0) s_vb = blockIdx.x;
    const int J = st->J;
    cost_block<MODEL, TMA, 1>(d, sm, &s_vb, st->cur, J, rec);
    if (threadIdx.x < kJMax) __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int prev = atomicAdd(&st->cnt_cost, 1u);
        sm.s_last = (prev == gridDim.x - 1);
    }
    __syncthreads();
    STAMP(d, 3, 2);
    if (!sm.s_last) return;
    __threadfence();
    cost_total(d, sm);
    bool ok = true;
    if (d.peer.nranks > 1) ok = peer_allreduce(d.peer, 1, sm.s_tot, J, &sm.s_last);   // views sharded: sum over the ranks
    if (threadIdx.x == 0) {
        for (int j = 0; j < kJMax; ++j) d.etrial[j] = sm.s_tot[j];
        st->cnt_cost = 0;
        if (!ok) {
            st->comm_fail = 1; st->done = 1;
            if (COND) cudaGraphSetConditional(d.cond, 0u);
        } else if (fuse_accept) {
            lm_accept<COND>(d);
        }
    }
    STAMP(d, 3, 3);
}

}  // namespace calib
