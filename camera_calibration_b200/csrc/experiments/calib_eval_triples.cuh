// EXPERIMENT / history, not in the product build: the evaluation pass as warp TRIPLES with re-balanced registers.
// It was the product's kernel for a few hours of round 2: 57.9 us at cfg3 against 61.8 us for the warp pairs -- until
// the branch-free reciprocal (rcp_nr) let ptxas schedule the pair producer's independent sums into its chain's
// stalls: pairs 55.0 us, triples 57.9 us (profiles/r4m_pairs2.txt).  Bit-identical to the pairs.  ncu per role:
// profiles/r4c_eval3_ncu.md.  To build: include after calib_passes.cuh, set the dynamic-shared-memory attribute
// (kEvalDynSmem) on k_eval3<0/1> and launch with kEval3Threads threads.
#pragma once
#include "../calib_passes.cuh"

namespace calib {

// ---- warp triples with re-balanced registers ---------------------------------------------------------
// ncu's source sampling of the pair kernel (profiles/r2f): one 32-point step took 1 570 cycles, the producer --
// dependent projection chain AND the 57 sums of the intrinsic block, 352 instructions -- issued during 22 % of
// them while the consumer slept on the hand-off.  Here the chain is a warp's ONLY work:
//   P    (warps 0-3)   loads the point, runs point_eval, writes the hand-off slot -- no running sums;
//   CA   (warps 4-7)   A, g_c and the cost (56 sums + 1) from pairs 0 .. 1+ND of the slot;
//   CB   (warps 8-11)  B_v, D_v, g_v (87 sums) from all 8+ND pairs, one row per piece.
// Warps t, t+4, t+8 share a scheduler and one ring (full: 1 arrival, empty: 2).  Three warps per scheduler
// hold 168 registers each at launch; `setmaxnreg` moves them to where the sums live (radtan: P 112, CA 160,
// CB 232; ptxas allocates per region, USETMAXREG in SASS; no spill).  A step now takes ~1 200 cycles
// (profiles/r4c_eval3_ncu.md): P still bounds it -- 248 instructions at one issue per ~5 cycles (fixed-latency
// dependencies 34 % of its samples) -- with the FP64 pipe 59 % busy.  Measured and rejected on top of this
// (profiles/r4b, r4e): cutting the chain in two (P -> CA -> CB pipeline, experiments/calib_eval_pipeline.cuh:
// slower, every warp of the scheduler then waits on the pipe), two steps per loop trip with swapped point
// buffers (no gain), an L2 prefetch three steps ahead (-10 us on cold data, +1.4 us on L2-resident data), the
// chain without its `valid` branch (lanes without a point computing on stale registers: 57.9 against 57.9 us).
constexpr int kEval3Threads = 384;
// per model: radtan carries 57 sums in CA and a shorter chain in P, fisheye 46 sums and the atan chain (no spill in either)
template <int MODEL> struct Eval3Regs;
template <> struct Eval3Regs<0> { static constexpr int P = 112, A = 160, B = 232; };
template <> struct Eval3Regs<1> { static constexpr int P = 128, A = 144, B = 232; };
template <int N> __device__ __forceinline__ void reg_release() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N> __device__ __forceinline__ void reg_acquire() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
// from the 168 registers a thread of a 384-thread block starts with (dec above / inc below the current count is illegal)
template <int N> __device__ __forceinline__ void reg_set() { if constexpr (N > 168) reg_acquire<N>(); else reg_release<N>(); }

template <int MODEL>
__device__ __forceinline__ void accum_A_ring(const double2* slot, double (&acc)[kNA], double& e) {
    constexpr int ND = ModelTraits<MODEL>::ND;
    PointOut<MODEL> o;
    { const double2 t = slot[0 * 32]; o.xd = t.x; o.yd = t.y; }
    { const double2 t = slot[1 * 32]; o.ru = t.x; o.rv = t.y; }
#pragma unroll
    for (int a = 0; a < ND; ++a) { const double2 t = slot[(2 + a) * 32]; o.ku[a] = t.x; o.kv[a] = t.y; }
    accum_A<MODEL>(o, acc, e);
}

template <int MODEL>
__global__ void __launch_bounds__(kEval3Threads, 1) k_eval3(Dev d) {
    using T = ModelTraits<MODEL>;
    constexpr int L = T::L, ND = T::ND, S = kEvalStages;
    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    const int trip = warp & (kEvalPairs - 1);
    const int role = warp >> 2;                    // 0: P, 1: CA, 2: CB (warp-group uniform: setmaxnreg needs that)
    const int w = blockIdx.x * kEvalPairs + trip;
    WarpRec rec;
    rec.p0 = rec.p1 = 0;
    if (w < d.ev_warps) rec = d.ev_rec[w];
    const LmState* st = d.st;
    if (st->done || !st->need_eval) return;
    __shared__ EvalSmem sm;
    const int cur = st->cur;
    EvalRing<MODEL>& ring = reinterpret_cast<EvalRing<MODEL>*>(dyn_smem)[trip];
    unsigned long long* full = sm.full[trip];
    unsigned long long* empty = sm.empty[trip];
    if (tid < L) sm.sc[tid] = d.Pbuf[(long long)cur * d.K + tid];
    if (tid < kEvalPairs * S) { mbar_init(&sm.full[0][0] + tid, 1); mbar_init(&sm.empty[0][0] + tid, 2); }
    if (tid == 0) mbar_fence_init();
    __syncthreads();
    const int p0 = rec.p0, p1 = rec.p1;
    if (role == 0) {
        // ---- P: the projection chain and the hand-off, nothing else
        reg_release<Eval3Regs<MODEL>::P>();
        if (p0 >= p1) return;
        const ViewC* views = d.views + (long long)cur * d.M;
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
            // the consumers never read sview: only this warp's own steps of the previous piece must be done
            __syncwarp();
            if (lane < kViewDoubles) reinterpret_cast<double*>(&sm.sview[trip])[lane] = vn;
            __syncwarp();
            if (has_next && lane < kViewDoubles) vn = reinterpret_cast<const double*>(views + nxt.view)[lane];
            const int nsteps = (pc.count + 31) >> 5;
            int idx = pc.start + lane;       // point indices fit 32 bits (Piece::start is an int)
            int left = pc.count - lane;              // > 0: this lane has a point in the current step
            // one step: the point in `p` is evaluated and handed off, the next one is fetched into `nx`
            auto step = [&](const Pt& p, Pt& nx, bool last) {
                const bool valid = left > 0;
                left -= 32;
                idx += 32;
                if (!last) { if (left > 0) load_point(d, (long long)idx, nx); }
                else if (has_next && lane < nxt.count) load_point(d, (long long)nxt.start + lane, nx);
                const unsigned s = k % S;
                if (k >= S) mbar_wait_guarded(&empty[s], ((k / S) - 1u) & 1u, d.st, 0);
                if (valid) {
                    PointOut<MODEL> o;
                    point_eval<MODEL, true, true>(sm.sc, sm.sview[trip], p.X, p.Y, p.Z, p.U, p.V, o);
                    double2* slot = &ring.v[s][0][lane];
#pragma unroll
                    for (int j = 0; j < 6; ++j) slot[(2 + ND + j) * 32] = make_double2(o.ju[j], o.jv[j]);
                    slot[0 * 32] = make_double2(o.xd, o.yd);
                    slot[1 * 32] = make_double2(o.ru, o.rv);
#pragma unroll
                    for (int a = 0; a < ND; ++a) slot[(2 + a) * 32] = make_double2(o.ku[a], o.kv[a]);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&full[s]);
                ++k;
            };
            for (int sidx = 0; sidx < nsteps; ++sidx) {
                const Pt p = pn;
                step(p, pn, sidx + 1 == nsteps);
            }
            pc = nxt;
        }
    } else if (role == 1) {
        // ---- CA: intrinsic block, its right-hand side, the cost; then the block's row of partA / partE
        reg_release<Eval3Regs<MODEL>::A>();
        double acc[kNA];
#pragma unroll
        for (int i = 0; i < kNA; ++i) acc[i] = 0.0;
        double e = 0.0;
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
        reg_acquire<Eval3Regs<MODEL>::B>();
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
                mbar_wait_guarded(&full[s], (unsigned)(k / S) & 1u, d.st, 2);
                if (valid) accum_BD_ring<MODEL>(&ring.v[s][0][lane], acc);
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

}  // namespace calib
