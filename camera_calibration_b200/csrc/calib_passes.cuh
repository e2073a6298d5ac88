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

template <bool TMA, int D, int WS, int NV, class Begin, class Body, class End>
__device__ __forceinline__ void sweep(const Dev& d, const Piece* __restrict__ pieces, const WarpRec& rec,
                                      const ViewC* const (&views)[NV], ViewC (&sview)[NV], int warp,
                                      unsigned long long* bars, int lane, Begin begin, Body body, End end) {
    if constexpr (TMA) {
        PointRing<D, WS>* ring = reinterpret_cast<PointRing<D, WS>*>(dyn_smem);
        sweep_windows<D, WS, NV>(d, pieces, rec, views, sview, ring[warp], bars, lane, begin, body, end);
    } else {
        sweep_pieces<NV>(d, pieces, rec, views, sview, lane, begin, body, end);
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
// A point's 2 x (L+6) Jacobian rows feed 143 running sums (56 of the intrinsic block A, g_c and the cost;
// 87 of B_v, D_v, g_v): more than one thread can hold.  Round 1 swept the points twice with one warp (the
// projection chain evaluated twice: 416 FP64 instructions per point, the detections read twice beyond L2);
// since round 2 a producer / consumer PAIR of warps shares ONE sweep (356 instructions, detections read once).
constexpr int kEvalPairs = kBlockWarps / 2;
constexpr int kEvalStages = 4;
template <int MODEL> struct EvalRing {
    static constexpr int NV2 = 8 + ModelTraits<MODEL>::ND;      // (xd,yd) (ru,rv) (ku,kv)[ND] (ju,jv)[6] -- or, chain cut: the 6 pairs the last stage starts from
    double2 v[kEvalStages][NV2][32];
};
constexpr size_t kEvalDynSmem = kEvalPairs * sizeof(EvalRing<0>);     // the radtan ring is the larger one

struct __align__(16) EvalSmem {
    ViewC sview[kEvalPairs], sviewA[kEvalPairs];    // the view constants of the chain warp / of the warp that finishes the extrinsic columns
    double sc[16];
    double sred[kEvalPairs][kNA];
    double sE[kEvalPairs];
    unsigned long long full[kEvalPairs][kEvalStages], empty[kEvalPairs][kEvalStages];
};

// try_wait with a way out: a hand-off partner that never arrives (a bug, not a run-time condition) must
// not hang the GPU.  After ~2 s the waiter raises the state's abort flag -- every other wait of the launch
// then falls through at once -- and the host reports the run as failed.
__device__ __forceinline__ bool mbar_try(unsigned long long* bar, unsigned parity) {
    unsigned ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait_guarded(unsigned long long* bar, unsigned parity, LmState* st, int who) {
    if (mbar_try(bar, parity)) return;
    const unsigned long long t0 = global_timer_ns();
    unsigned spins = 0;
    while (!mbar_try(bar, parity)) {
        if ((++spins & 63u) == 0) {
            if (*(volatile int*)&st->bar_abort) return;
            if (global_timer_ns() - t0 > 2000000000ull) {
                *(volatile int*)&st->bar_abort = 2 + who; return;
            }
        }
    }
}

__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// ---- producer / consumer warp PAIRS ---------------------------------------------------------------------
// A pair of warps (w, w + 4: same scheduler) shares one sweep over a chunk of points:
//   producer   loads the point, runs the dependent projection chain, keeps the 57 sums of A, g_c, E and hands what the
//              view blocks need to the consumer through a 4-stage ring in shared memory (one 16-byte slot per lane
//              and value pair, lane-private, mbarrier full / empty per stage);
//   consumer   keeps the 87 sums of B_v, D_v, g_v and writes one row per piece.
// Since point_eval's reciprocal and reciprocal square root lost the special-case branch of __drcp_rn / rsqrt
// (rcp_nr, rsqrt_nr in calib_device.cuh: the branch ended the basic block, nothing was scheduled into the ~70 cycles
// of the operation), k_eval<radtan> takes 53.9 us at cfg3 instead of 61.8 us, bit-identical.
// CUT (fisheye): the chain is cut after the distortion derivatives; the consumer finishes the extrinsic columns
// (extrinsic_columns, 43 FP64 instructions) from six value pairs that replace the six (ju,jv) pairs in the slot.  The
// fisheye chain is the longer one (rsqrt, atan): 309 -> 301 us at cfg5 x 0.2; radtan is faster uncut (53.9 against
// 56.9 us cut, 55.4 us cut with the producer's sums after the signal: r4r).  Per-lane accumulation order is the same in every variant: bit-identical results.
// What bounds it (profiles/README.md, r4n): an FP64 instruction holds a scheduler's issue port for two cycles and
// other instructions do not hide behind it (micro/fp64_issue_mix.cu), so a 32-point step cannot take less than
// 2 x 349 + ~0.9 x 187 = 866 cycles; it takes 1 100 (79 %).  Measured and rejected: warp TRIPLES with setmaxnreg
// (chain / A sums / view sums in three warps, experiments/calib_eval_triples.cuh: 57.9 us -- it was the faster kernel
// while the reciprocal still had its branch: 58.3 against 61.8 us), a three-stage pipeline (experiments/calib_eval_pipeline.cuh), producers and
// consumers on different schedulers (60.5 us), two steps per trip, an L2 prefetch, the chain without its `valid` branch.
template <int MODEL> struct EvalCut { static constexpr bool value = MODEL == 1; };

template <int MODEL>
__global__ void __launch_bounds__(kBlockThreads, 1) k_eval(Dev d) {
    constexpr bool CUT = EvalCut<MODEL>::value;
    using T = ModelTraits<MODEL>;
    constexpr int L = T::L, ND = T::ND, S = kEvalStages, JM = 2 + ND;
    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    const int pair = warp & (kEvalPairs - 1);
    const bool producer = warp < kEvalPairs;
    const int w = blockIdx.x * kEvalPairs + pair;
    WarpRec rec;
    rec.p0 = rec.p1 = 0;
    if (w < d.ev_warps) rec = d.ev_rec[w];
    const LmState* st = d.st;
    if (st->done || !st->need_eval) return;
    __shared__ EvalSmem sm;
    const int cur = st->cur;
    EvalRing<MODEL>& ring = reinterpret_cast<EvalRing<MODEL>*>(dyn_smem)[pair];
    unsigned long long* full = sm.full[pair];
    unsigned long long* empty = sm.empty[pair];
    if (tid < L) sm.sc[tid] = d.Pbuf[(long long)cur * d.K + tid];
    if (tid < kEvalPairs * S) { mbar_init(&sm.full[0][0] + tid, 1); mbar_init(&sm.empty[0][0] + tid, 1); }
    if (tid == 0) mbar_fence_init();
    __syncthreads();
    const int p0 = rec.p0, p1 = rec.p1;
    const ViewC* views = d.views + (long long)cur * d.M;
    if (producer) {
        double acc[kNA];
#pragma unroll
        for (int i = 0; i < kNA; ++i) acc[i] = 0.0;
        double e = 0.0;
        if (p0 < p1) {
            Piece pc = rec.first;
            Pt pn;
            if (lane < pc.count) load_point(d, (long long)pc.start + lane, pn);
            double vn = 0.0;
            if (lane < kViewDoubles) vn = reinterpret_cast<const double*>(views + pc.view)[lane];
            unsigned k = 0;                              // hand-offs so far
            for (int pi = p0; pi < p1; ++pi) {
                const bool has_next = pi + 1 < p1;
                Piece nxt = pc;
                if (has_next) nxt = d.ev_piece[pi + 1];
                __syncwarp();
                if (lane < kViewDoubles) reinterpret_cast<double*>(&sm.sview[pair])[lane] = vn;
                __syncwarp();
                if (has_next && lane < kViewDoubles) vn = reinterpret_cast<const double*>(views + nxt.view)[lane];
                const int nsteps = (pc.count + 31) >> 5;
                int idx = pc.start + lane;
                int left = pc.count - lane;              // > 0: this lane has a point in the current step
                for (int sidx = 0; sidx < nsteps; ++sidx) {
                    const Pt p = pn;
                    const bool valid = left > 0;
                    left -= 32;
                    idx += 32;
                    if (sidx + 1 < nsteps) { if (left > 0) load_point(d, (long long)idx, pn); }
                    else if (has_next && lane < nxt.count) load_point(d, (long long)nxt.start + lane, pn);
                    const unsigned s = k % S;
                    if (k >= S) mbar_wait_guarded(&empty[s], ((k / S) - 1u) & 1u, d.st, 0);
                    PointOut<MODEL> o;
                    if (valid) {
                        double2* slot = &ring.v[s][0][lane];
                        if constexpr (CUT) {
                            PointMid m;
                            point_eval<MODEL, true, 2>(sm.sc, sm.sview[pair], p.X, p.Y, p.Z, p.U, p.V, o, &m);
                            slot[(JM + 0) * 32] = make_double2(m.dxx, m.dxy);
                            slot[(JM + 1) * 32] = make_double2(m.dyy, m.iz);
                            slot[(JM + 2) * 32] = make_double2(m.xn, m.yn);
                            slot[(JM + 3) * 32] = make_double2(m.q0, m.q1);
                            slot[(JM + 4) * 32] = make_double2(m.q2, p.Y);
                            slot[(JM + 5) * 32] = make_double2(p.Z, 0.0);
                        } else {
                            point_eval<MODEL, true, 1>(sm.sc, sm.sview[pair], p.X, p.Y, p.Z, p.U, p.V, o);
#pragma unroll
                            for (int j = 0; j < 6; ++j) slot[(JM + j) * 32] = make_double2(o.ju[j], o.jv[j]);     // first: the extrinsic columns die here
                        }
                        slot[0 * 32] = make_double2(o.xd, o.yd);
                        slot[1 * 32] = make_double2(o.ru, o.rv);
#pragma unroll
                        for (int a = 0; a < ND; ++a) slot[(2 + a) * 32] = make_double2(o.ku[a], o.kv[a]);
                        if constexpr (CUT) accum_A<MODEL>(o, acc, e);      // measured: 301.4 against 303.7 us (cfg5 x 0.2)
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&full[s]);
                    // uncut chain: the consumer is released first, the producer's own sums follow (53.9 against 56.8 us, cfg3)
                    if constexpr (!CUT) { if (valid) accum_A<MODEL>(o, acc, e); }
                    ++k;
                }
                pc = nxt;
            }
        }
        warp_reduce_scatter<kNA>(acc, lane);
        e = warp_sum(e);
        sm.sred[pair][2 * lane] = acc[0];
        sm.sred[pair][2 * lane + 1] = acc[1];
        if (lane == 0) sm.sE[pair] = e;
    } else {
        double acc[kNBD];
#pragma unroll
        for (int i = 0; i < kNBD; ++i) acc[i] = 0.0;
        if (p0 < p1) {
            Piece pc = rec.first;
            double vn = 0.0;
            if (CUT && lane < kViewDoubles) vn = reinterpret_cast<const double*>(views + pc.view)[lane];
            unsigned k = 0;
            for (int pi = p0; pi < p1; ++pi) {
                const bool has_next = pi + 1 < p1;
                Piece nxt = pc;
                if (has_next) nxt = d.ev_piece[pi + 1];
                if constexpr (CUT) {                     // its own copy of the view constants: it lags the producer by up to S steps
                    __syncwarp();
                    if (lane < kViewDoubles) reinterpret_cast<double*>(&sm.sviewA[pair])[lane] = vn;
                    __syncwarp();
                    if (has_next && lane < kViewDoubles) vn = reinterpret_cast<const double*>(views + nxt.view)[lane];
                }
                const int nsteps = (pc.count + 31) >> 5;
                int left = pc.count - lane;
                for (int sidx = 0; sidx < nsteps; ++sidx) {
                    const bool valid = left > 0;
                    left -= 32;
                    const unsigned s = k % S;
                    mbar_wait_guarded(&full[s], (k / S) & 1u, d.st, 1);
                    if (valid) {
                        const double2* slot = &ring.v[s][0][lane];
                        if constexpr (CUT) {
                            PointMid m;
                            double Y, Z;
                            { const double2 t = slot[(JM + 0) * 32]; m.dxx = t.x; m.dxy = t.y; }
                            { const double2 t = slot[(JM + 1) * 32]; m.dyy = t.x; m.iz = t.y; }
                            { const double2 t = slot[(JM + 2) * 32]; m.xn = t.x; m.yn = t.y; }
                            { const double2 t = slot[(JM + 3) * 32]; m.q0 = t.x; m.q1 = t.y; }
                            { const double2 t = slot[(JM + 4) * 32]; m.q2 = t.x; Y = t.y; }
                            { const double2 t = slot[(JM + 5) * 32]; Z = t.x; }
                            double ju[6], jv[6];
                            extrinsic_columns(sm.sc, sm.sviewA[pair], m, Y, Z, ju, jv);
                            accum_BD_j<MODEL>(slot, ju, jv, acc);
                        } else {
                            accum_BD_ring<MODEL>(slot, acc);
                        }
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&empty[s]);
                    ++k;
                }
                warp_reduce_scatter<kNBD>(acc, lane);
                double* row = d.partBD + (long long)pi * kNBD + lane * (kNBD / 32);
#pragma unroll
                for (int j = 0; j < kNBD / 32; ++j) row[j] = acc[j];
#pragma unroll
                for (int i = 0; i < kNBD; ++i) acc[i] = 0.0;
                pc = nxt;
            }
        }
    }
    __syncthreads();
    if (tid < kNA) {
        double s = sm.sred[0][tid];
#pragma unroll
        for (int q = 1; q < kEvalPairs; ++q) s += sm.sred[q][tid];
        d.partA[(long long)blockIdx.x * kNA + tid] = s;
    } else if (tid == kNA) {
        double s = sm.sE[0];
#pragma unroll
        for (int q = 1; q < kEvalPairs; ++q) s = __dadd_rn(s, sm.sE[q]);
        d.partE[blockIdx.x] = s;
    }
}

// ---- accept / reject, lambda update, termination: src/calibrate.py:155-168 ------------------
// Replays the reference's decisions over the J candidates of this pass (calib_kernels.cuh, "damping
// ladder"): candidate j is the reference's iteration iter+j with damping lam * 10^j; the first
// accepted one ends the pass, the rest is dropped.  st: the state to advance (the device state, or a
// block's copy of it in the persistent kernel, where every block replays the same decisions from the
// same trial costs and only one writes the trace).  Returns done.
// ptrace / Pcur (with trace, when st->want_ptrace): the L shared parameters of the current iterate go beside
// every trace row -- what the reference prints per iteration (src/calibrate.py:158-159,269-274).
__device__ __forceinline__ int lm_accept_on(LmState* st, const double* etrial, double* trace, double* ptrace = nullptr,
                                            const double* Pcur = nullptr, int L = 0) {
    const int J = st->J, cur = st->cur;
    double lam = st->lam;
    const double E = st->E_cur;
    int it = st->iter, done = 0;
    bool accepted = false;
    const bool after_accept = st->need_eval != 0 && it > 0;     // this pass tries the step right after an accepted one
    for (int j = 0; j < J && !done && !accepted; ++j) {
        const double E1 = etrial[j];
        const bool acc = E1 < E;   // strict; NaN rejects
        if (j == 0 && after_accept) st->jdyn = acc ? 1 : 2;
        if (trace) {
            double* tr = trace + 4ll * it;
            tr[0] = E; tr[1] = E1; tr[2] = lam; tr[3] = acc ? 1.0 : 0.0;
            if (ptrace && st->want_ptrace)
                for (int i = 0; i < L; ++i) ptrace[16ll * it + i] = Pcur[i];
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
    // Candidates of the next pass: after rejections the ladder doubles; after an ACCEPTED step it starts at 1 or
    // 2.  LM runs bounce -- an accepted step divides lambda by 10 and the step tried next is often rejected and
    // retried at 10 lambda (itest: 35 of 44 accepted steps, cfg1: 9 of 16, cfg3: 3 of 8, cfg2: none) -- and with two
    // candidates the pass after an accept already carries the retry: one pass (schur + back-substitution + cost,
    // ~36 us) less per bounce for ~5 us of extra arithmetic in the cost sweep when there is none.  jstart = 0
    // (default) predicts from the last time: 2 while the step after an accepted one keeps being rejected, else 1.
    // Which candidates a pass carries never changes a result (test_damping_ladder_is_invisible).
    int nj = accepted ? (st->jstart ? st->jstart : st->jdyn) : 2 * J;
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
    const int done = lm_accept_on(d.st, d.etrial, d.trace, d.ptrace, d.Pbuf + (long long)d.st->cur * d.K, d.L);
    if (COND) cudaGraphSetConditional(d.cond, done ? 0u : 1u);
}

template <bool COND>
__global__ void k_accept(Dev d) {
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

template <int MODEL, bool TMA, int NJ>
__device__ __forceinline__ void cost_sweep(const Dev& d, const WarpRec& rec, int cur, double (*sc)[16],
                                           ViewC (&sview)[kJMax], unsigned long long* bars, int warp, int lane,
                                           double (&e)[kJMax]) {
    const ViewC* views[NJ];
#pragma unroll
    for (int j = 0; j < NJ; ++j) views[j] = d.views + (long long)trial_buf(cur, j) * d.M;
    const ViewC* const (&cviews)[NJ] = views;
    ViewC (&sv)[NJ] = reinterpret_cast<ViewC (&)[NJ]>(sview);
    sweep<TMA, kCostStages, kWinSteps, NJ>(d, d.co_piece, rec, cviews, sv, warp, bars, lane,
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
template <int MODEL, bool TMA>
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
    if (J == 1) cost_sweep<MODEL, TMA, 1>(d, rec, cur, sm.sc, sm.sview[warp], sm.bars[warp], warp, lane, e);
    else if (J == 2) cost_sweep<MODEL, TMA, 2>(d, rec, cur, sm.sc, sm.sview[warp], sm.bars[warp], warp, lane, e);
    else cost_sweep<MODEL, TMA, kJMax>(d, rec, cur, sm.sc, sm.sview[warp], sm.bars[warp], warp, lane, e);
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
    if (w < d.co_warps) rec = d.co_rec[w];
    LmState* st = d.st;
    if (st->done) {   // also after a failed exchange: the WHILE node must not spin
        if (COND) { if (blockIdx.x == 0 && threadIdx.x == 0) cudaGraphSetConditional(d.cond, 0u); }
        return;
    }
    __shared__ CostSmem sm;
    __shared__ int s_vb;
    if (threadIdx.x == 0) s_vb = blockIdx.x;
    const int J = st->J;
    cost_block<MODEL, TMA>(d, sm, &s_vb, st->cur, J, rec);
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
