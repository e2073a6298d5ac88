// EXPERIMENT, not in the product build (measured and rejected, profiles/r4b_eval_roles.txt, r4c_*):
// the evaluation pass as a three-stage PIPELINE of warps.  Bit-identical to k_eval, but slower: cfg3 67.5 us
// against 58.3 (k_eval) and 61.8 (warp pairs); cfg5 x 0.2 340 against 316 / 326 us.  ncu source sampling: with
// the chain cut in two ALL three warps of a scheduler are busy (P 208, CA 228, CB 205 instructions per 32-point
// step) and every one of them runs slower -- a step takes 1300 cycles, P waits on the FP64 pipe (math_pipe_throttle,
// not_selected 19 % of its samples against 3 %), and the second ring adds 12 shared-memory transactions per point
// to a pipe that already moves 528 B per point.  To build it: include this file after calib_passes.cuh, add the
// function attributes (kEval4DynSmem) and a launch of k_eval4<MODEL> with kEval3Threads threads.
#pragma once
#include "calib_eval_triples.cuh"

namespace calib {

// ---- warp triples as a PIPELINE: P -> CA -> CB -----------------------------------------------------------
// k_eval still leaves the whole projection chain to P.  Here the chain is cut: P stops at the distortion
// derivatives (point_eval<.., 2>) and hands over what the last stage starts from; CA finishes the extrinsic
// columns (extrinsic_columns: 43 FP64 instructions), passes them on through a second, smaller ring, and keeps
// the sums of the intrinsic block; CB reads (xd,yd) (ru,rv) (ku,kv) from P's ring and (ju,jv) from CA's.
// FP64 instructions per point: P 99, CA 106, CB 144 -- no warp carries more than the pipe can issue in the
// time the others need.  Ring 1: full 1 arrival (P), empty 2 (CA, CB); ring 2: full 1 (CA), empty 1 (CB).
template <int MODEL> struct Eval4Regs;
template <> struct Eval4Regs<0> { static constexpr int P = 104, A = 184, B = 216; };
template <> struct Eval4Regs<1> { static constexpr int P = 120, A = 160, B = 224; };
struct EvalRing2 { double2 v[kEvalStages][6][32]; };
constexpr size_t kEval4DynSmem = kEvalPairs * (sizeof(EvalRing<0>) + sizeof(EvalRing2));
struct __align__(16) Eval4Smem {
    ViewC sview[kEvalPairs], sviewA[kEvalPairs];
    double sc[16];
    double sred[kEvalPairs][kNA];
    double sE[kEvalPairs];
    unsigned long long full[kEvalPairs][kEvalStages], empty[kEvalPairs][kEvalStages];
    unsigned long long full2[kEvalPairs][kEvalStages], empty2[kEvalPairs][kEvalStages];
};

template <int MODEL>
__global__ void __launch_bounds__(kEval3Threads, 1) k_eval4(Dev d) {
    using T = ModelTraits<MODEL>;
    constexpr int L = T::L, ND = T::ND, S = kEvalStages, JM = 2 + ND;     // JM: first slot of the mid values
    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    const int trip = warp & (kEvalPairs - 1);
    const int role = warp >> 2;                    // 0: P, 1: CA, 2: CB
    const int w = blockIdx.x * kEvalPairs + trip;
    WarpRec rec;
    rec.p0 = rec.p1 = 0;
    if (w < d.ev_warps) rec = d.ev_rec[w];
    const LmState* st = d.st;
    if (st->done || !st->need_eval) return;
    __shared__ Eval4Smem sm;
    const int cur = st->cur;
    EvalRing<MODEL>& ring = reinterpret_cast<EvalRing<MODEL>*>(dyn_smem)[trip];
    EvalRing2& ring2 = reinterpret_cast<EvalRing2*>(dyn_smem + kEvalPairs * sizeof(EvalRing<0>))[trip];
    unsigned long long* full = sm.full[trip];
    unsigned long long* empty = sm.empty[trip];
    unsigned long long* full2 = sm.full2[trip];
    unsigned long long* empty2 = sm.empty2[trip];
    if (tid < L) sm.sc[tid] = d.Pbuf[(long long)cur * d.K + tid];
    if (tid < kEvalPairs * S) {
        mbar_init(&sm.full[0][0] + tid, 1); mbar_init(&sm.empty[0][0] + tid, 2);
        mbar_init(&sm.full2[0][0] + tid, 1); mbar_init(&sm.empty2[0][0] + tid, 1);
    }
    if (tid == 0) mbar_fence_init();
    __syncthreads();
    const int p0 = rec.p0, p1 = rec.p1;
    if (role == 0) {
        // ---- P: the projection chain up to the distortion derivatives, d/dk, and the hand-off
        reg_release<Eval4Regs<MODEL>::P>();
        if (p0 >= p1) return;
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
            if (lane < kViewDoubles) reinterpret_cast<double*>(&sm.sview[trip])[lane] = vn;
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
                const int s = k % S;
                if (k >= S) mbar_wait_guarded(&empty[s], (unsigned)((k / S) - 1) & 1u, d.st, 0);
                if (valid) {
                    PointOut<MODEL> o;
                    PointMid m;
                    point_eval<MODEL, true, 2>(sm.sc, sm.sview[trip], p.X, p.Y, p.Z, p.U, p.V, o, &m);
                    double2* slot = &ring.v[s][0][lane];
                    slot[(JM + 0) * 32] = make_double2(m.dxx, m.dxy);
                    slot[(JM + 1) * 32] = make_double2(m.dyy, m.iz);
                    slot[(JM + 2) * 32] = make_double2(m.xn, m.yn);
                    slot[(JM + 3) * 32] = make_double2(m.q0, m.q1);
                    slot[(JM + 4) * 32] = make_double2(m.q2, p.Y);
                    slot[(JM + 5) * 32] = make_double2(p.Z, 0.0);
                    slot[0 * 32] = make_double2(o.xd, o.yd);
                    slot[1 * 32] = make_double2(o.ru, o.rv);
#pragma unroll
                    for (int a = 0; a < ND; ++a) slot[(2 + a) * 32] = make_double2(o.ku[a], o.kv[a]);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&full[s]);
                ++k;
            }
            pc = nxt;
        }
    } else if (role == 1) {
        // ---- CA: last stage of the extrinsic columns -> ring 2; intrinsic block, its right-hand side, the cost
        reg_set<Eval4Regs<MODEL>::A>();
        double acc[kNA];
#pragma unroll
        for (int i = 0; i < kNA; ++i) acc[i] = 0.0;
        double e = 0.0;
        if (p0 < p1) {
            const ViewC* views = d.views + (long long)cur * d.M;
            Piece pc = rec.first;
            double vn = 0.0;
            if (lane < kViewDoubles) vn = reinterpret_cast<const double*>(views + pc.view)[lane];
            int k = 0;
            for (int pi = p0; pi < p1; ++pi) {
                const bool has_next = pi + 1 < p1;
                Piece nxt = pc;
                if (has_next) nxt = d.ev_piece[pi + 1];
                __syncwarp();
                if (lane < kViewDoubles) reinterpret_cast<double*>(&sm.sviewA[trip])[lane] = vn;
                __syncwarp();
                if (has_next && lane < kViewDoubles) vn = reinterpret_cast<const double*>(views + nxt.view)[lane];
                const int nsteps = (pc.count + 31) >> 5;
                int left = pc.count - lane;
                for (int sidx = 0; sidx < nsteps; ++sidx) {
                    const bool valid = left > 0;
                    left -= 32;
                    const int s = k % S;
                    mbar_wait_guarded(&full[s], (unsigned)(k / S) & 1u, d.st, 1);
                    if (k >= S) mbar_wait_guarded(&empty2[s], (unsigned)((k / S) - 1) & 1u, d.st, 1);
                    if (valid) {
                        const double2* slot = &ring.v[s][0][lane];
                        PointMid m;
                        double Y, Z;
                        { const double2 t = slot[(JM + 0) * 32]; m.dxx = t.x; m.dxy = t.y; }
                        { const double2 t = slot[(JM + 1) * 32]; m.dyy = t.x; m.iz = t.y; }
                        { const double2 t = slot[(JM + 2) * 32]; m.xn = t.x; m.yn = t.y; }
                        { const double2 t = slot[(JM + 3) * 32]; m.q0 = t.x; m.q1 = t.y; }
                        { const double2 t = slot[(JM + 4) * 32]; m.q2 = t.x; Y = t.y; }
                        { const double2 t = slot[(JM + 5) * 32]; Z = t.x; }
                        double ju[6], jv[6];
                        extrinsic_columns(sm.sc, sm.sviewA[trip], m, Y, Z, ju, jv);
                        double2* out = &ring2.v[s][0][lane];
#pragma unroll
                        for (int j = 0; j < 6; ++j) out[j * 32] = make_double2(ju[j], jv[j]);
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&full2[s]);
                    if (valid) accum_A_ring<MODEL>(&ring.v[s][0][lane], acc, e);
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&empty[s]);
                    ++k;
                }
                pc = nxt;
            }
        }
        warp_reduce_scatter<kNA>(acc, lane);
        e = warp_sum(e);
        sm.sred[trip][2 * lane] = acc[0];
        sm.sred[trip][2 * lane + 1] = acc[1];
        if (lane == 0) sm.sE[trip] = e;
        asm volatile("bar.sync 1, 128;" ::: "memory");       // the four CA warps
        const int t = tid - 128;
        if (t < kNA) {
            double s = sm.sred[0][t];
#pragma unroll
            for (int q = 1; q < kEvalPairs; ++q) s += sm.sred[q][t];
            d.partA[(long long)blockIdx.x * kNA + t] = s;
        } else if (t == kNA) {
            double s = sm.sE[0];
#pragma unroll
            for (int q = 1; q < kEvalPairs; ++q) s = __dadd_rn(s, sm.sE[q]);
            d.partE[blockIdx.x] = s;
        }
    } else {
        // ---- CB: B_v, D_v, g_v -- one row of kNBD slots per piece
        reg_acquire<Eval4Regs<MODEL>::B>();
        double acc[kNBD];
#pragma unroll
        for (int i = 0; i < kNBD; ++i) acc[i] = 0.0;
        if (p0 >= p1) return;
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
                mbar_wait_guarded(&full2[s], (unsigned)(k / S) & 1u, d.st, 2);      // ring 2 full implies ring 1 full
                if (valid) accum_BD_ring2<MODEL>(&ring.v[s][0][lane], &ring2.v[s][0][lane], acc);
                __syncwarp();
                if (lane == 0) { mbar_arrive(&empty[s]); mbar_arrive(&empty2[s]); }
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

}  // namespace calib
