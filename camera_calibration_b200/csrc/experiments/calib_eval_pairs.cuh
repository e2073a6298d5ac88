// EXPERIMENT / history, not in the product build: the evaluation pass as producer / consumer warp PAIRS
// (the product's kernel of round 2 until the warp triples of calib_passes.cuh replaced it; still what the
// persistent kernel of calib_persistent.cuh calls).  Bit-identical to the triples; cfg3 61.8 us against 58.3,
// cfg5 x 0.2 326 against 316 us (profiles/r4a_eval_roles.txt, r4b_eval_roles.txt).
#pragma once
#include "../calib_passes.cuh"

#ifndef CALIB_PAIRS_SPLIT
#define CALIB_PAIRS_SPLIT 0
#endif

namespace calib {

// Block *vbp of the evaluation pass (pairs vb * 4 .. vb * 4 + 3 of the partition, rec = the record of
// this warp's pair) at the point in parameter buffer `cur`: writes the piece rows of its pairs and row
// vb of partA / partE.  vbp points to shared memory written before the call (not necessarily
// synchronised: it is read after the barriers below) -- the index is needed after the sweep only and
// must not cost a register inside it.  Called by all threads of a block; ends without a barrier after
// its last shared-memory reads.  The ring lives in dynamic shared memory (kEvalDynSmem).
template <int MODEL>
__device__ __forceinline__ void eval_block(const Dev& d, EvalSmem& sm, const volatile int* vbp, int cur, const WarpRec& rec) {
    using T = ModelTraits<MODEL>;
    constexpr int L = T::L, ND = T::ND, S = kEvalStages;
    const int tid = tid_now();
    const int lane = tid & 31, warp = tid >> 5;
#if CALIB_PAIRS_SPLIT
    // producers on schedulers 0 and 1 (warps 0, 1, 4, 5), consumers on schedulers 2 and 3 (warps 2, 3, 6, 7)
    const int pair = (warp & 1) | ((warp >> 2) << 1);
    const bool producer = (warp & 2) == 0;
#else
    const int pair = warp & (kEvalPairs - 1);
    const bool producer = warp < kEvalPairs;
#endif
    EvalRing<MODEL>& ring = reinterpret_cast<EvalRing<MODEL>*>(dyn_smem)[pair];
    unsigned long long* full = sm.full[pair];
    unsigned long long* empty = sm.empty[pair];
    STAMP(d, 0, 0);
    if (tid < L) sm.sc[tid] = d.Pbuf[(long long)cur * d.K + tid];
    if (tid < kEvalPairs * S) { mbar_init(&sm.full[0][0] + tid, 1); mbar_init(&sm.empty[0][0] + tid, 1); }
    if (tid == 0) mbar_fence_init();
    __syncthreads();
    const int p0 = rec.p0, p1 = rec.p1;
    if (producer) {
        // ---- producer: projection chain, A / g_c / E, hand-off
        double acc[kNA];
#pragma unroll
        for (int i = 0; i < kNA; ++i) acc[i] = 0.0;
        double e = 0.0;
        if (p0 < p1) {
            const ViewC* views = d.views + (long long)cur * d.M;
            Piece pc = rec.first;
            Pt pn;
            if (lane < pc.count) load_point(d, (long long)pc.start + lane, pn);
            double vn = 0.0;
            if (lane < kViewDoubles) vn = reinterpret_cast<const double*>(views + pc.view)[lane];
            int k = 0;                                   // hand-offs so far
            for (int pi = p0; pi < p1; ++pi) {
                const bool has_next = pi + 1 < p1;
                Piece nxt = pc;
                if (has_next) nxt = d.ev_piece[pi + 1];
                __syncwarp();
                if (lane < kViewDoubles) reinterpret_cast<double*>(&sm.sview[pair])[lane] = vn;
                __syncwarp();
                if (has_next && lane < kViewDoubles) vn = reinterpret_cast<const double*>(views + nxt.view)[lane];
                const int nsteps = (pc.count + 31) >> 5;
                long long idx = (long long)pc.start + lane;
                int left = pc.count - lane;              // > 0: this lane has a point in the current step
                for (int sidx = 0; sidx < nsteps; ++sidx) {
                    const Pt p = pn;
                    const bool valid = left > 0;
                    left -= 32;
                    idx += 32;
                    if (sidx + 1 < nsteps) { if (left > 0) load_point(d, idx, pn); }
                    else if (has_next && lane < nxt.count) load_point(d, (long long)nxt.start + lane, pn);
                    // the stage is normally free long before (the consumer has less to do per point)
                    const int s = k % S;
                    if (k >= S) mbar_wait_guarded(&empty[s], (unsigned)((k / S) - 1) & 1u, d.st, 0);
                    PointOut<MODEL> o;
                    if (valid) {
                        point_eval<MODEL, true, true>(sm.sc, sm.sview[pair], p.X, p.Y, p.Z, p.U, p.V, o);
                        double2* slot = &ring.v[s][0][lane];     // hand-off first: the extrinsic columns die here
#pragma unroll
                        for (int j = 0; j < 6; ++j) slot[(2 + ND + j) * 32] = make_double2(o.ju[j], o.jv[j]);
                        slot[0 * 32] = make_double2(o.xd, o.yd);
                        slot[1 * 32] = make_double2(o.ru, o.rv);
#pragma unroll
                        for (int a = 0; a < ND; ++a) slot[(2 + a) * 32] = make_double2(o.ku[a], o.kv[a]);
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&full[s]);
                    if (valid) accum_A<MODEL>(o, acc, e);
                    ++k;
                }
                pc = nxt;
            }
        }
        STAMP(d, 0, 1);
        warp_reduce_scatter<kNA>(acc, lane);
        e = warp_sum(e);
        sm.sred[pair][2 * lane] = acc[0];
        sm.sred[pair][2 * lane + 1] = acc[1];
        if (lane == 0) sm.sE[pair] = e;
    } else {
        // ---- consumer: B_v, D_v, g_v -- one row of kNBD slots per piece
        double acc[kNBD];
#pragma unroll
        for (int i = 0; i < kNBD; ++i) acc[i] = 0.0;
        if (p0 < p1) {
            Piece pc = rec.first;
            int k = 0;
            for (int pi = p0; pi < p1; ++pi) {
                const bool has_next = pi + 1 < p1;
                Piece nxt = pc;
                if (has_next) nxt = d.ev_piece[pi + 1];
                const int nsteps = (pc.count + 31) >> 5;
                int left = pc.count - lane;
                for (int sidx = 0; sidx < nsteps; ++sidx) {
                    const bool valid = left > 0;
                    left -= 32;
                    const int s = k % S;
                    mbar_wait_guarded(&full[s], (unsigned)(k / S) & 1u, d.st, 1);
                    if (valid) {
                        accum_BD_ring<MODEL>(&ring.v[s][0][lane], acc);
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
        STAMP(d, 0, 2);
    }
    __syncthreads();
    STAMP(d, 0, 3);
    const int vb = *vbp;
    if (tid < kNA) {
        double s = sm.sred[0][tid];
#pragma unroll
        for (int k = 1; k < kEvalPairs; ++k) s += sm.sred[k][tid];
        d.partA[(long long)vb * kNA + tid] = s;
    } else if (tid == kNA) {
        double s = sm.sE[0];
#pragma unroll
        for (int k = 1; k < kEvalPairs; ++k) s = __dadd_rn(s, sm.sE[k]);
        d.partE[vb] = s;
    }
    STAMP(d, 0, 4);
}

template <int MODEL>
__global__ void __launch_bounds__(kBlockThreads, 1) k_eval_pairs(Dev d) {
#if CALIB_PAIRS_SPLIT
    const int pair = ((threadIdx.x >> 5) & 1) | ((threadIdx.x >> 7) << 1);
#else
    const int pair = (threadIdx.x >> 5) & (kEvalPairs - 1);
#endif
    const int w = blockIdx.x * kEvalPairs + pair;
    WarpRec rec;                       // issued together with the state load below: one round trip
    rec.p0 = rec.p1 = 0;
    if (w < d.ev_warps) rec = d.ev_rec[w];
    const LmState* st = d.st;
    if (st->done || !st->need_eval) return;
    __shared__ EvalSmem sm;
    __shared__ int s_vb;
    if (threadIdx.x == 0) s_vb = blockIdx.x;
    eval_block<MODEL>(d, sm, &s_vb, st->cur, rec);
}

}  // namespace calib
