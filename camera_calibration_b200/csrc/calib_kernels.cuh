// Kernels of the single-calibration ("large") path.  One LM iteration =
//   k_eval          fused projection + residual + analytic Jacobian + block normal equations.
//                   Persistent warps: warp w owns a fixed, contiguous range of points (a chunk,
//                   cut into pieces at view boundaries).  Sweep 1 accumulates the coupling block
//                   B_v, the view block D_v and the view right-hand side in registers and writes
//                   one 96-slot row per piece; sweep 2 accumulates the intrinsic block A, its
//                   right-hand side and the cost.  No J is ever stored.
//   k_schur         one warp per view: sums the view's piece rows, damps, 6x6 Cholesky,
//                   Z = L^-1 [B^T g_v], W = Z^T Z; cross-view sums; the last block writes the
//                   reduced system and (one GPU) solves it          (calib_solve.cuh)
//   k_backsub       one warp per view: back-substitution, trial parameters, trial view constants
//   k_cost          residual-only sweep at the trial point; the last block sums and (one GPU)
//                   takes the accept/reject decision
// Replaces the loop body of src/calibrate.py:143-168 (see include/calib_b200.h).
#pragma once
#include "calib_device.cuh"
#include "calib_peer.cuh"

namespace calib {

constexpr int kBlockThreads = 256;          // every kernel of this path: 8 warps per block
constexpr int kBlockWarps = kBlockThreads / 32;
constexpr int kNA = 64;                     // padded slot count of the intrinsic sweep
constexpr int kNBD = 96;                    // padded slot count of a piece row and of W
constexpr int kVtab = 88;                   // per view: Cholesky factor (21) + Z (6 (L+1)), padded

// Damping ladder.  After a rejected step the reference only multiplies lambda by 10 and tries again
// from the same point with the same normal equations (src/calibrate.py:161-166), so the next few
// trial steps can be taken AHEAD of the decisions: one pass of k_schur / k_backsub / k_cost handles
// J candidates lambda, 10 lambda, ... (J = 1 after an accepted step, doubling after rejected ones up
// to kJMax); the accept step then replays the reference's decisions one candidate at a time and
// drops the candidates after the first accepted one.  Iterates, costs, lambdas, trace rows and the
// iteration count are those of the one-at-a-time loop (the sums of candidate j do not depend on J);
// the latency-bound tail of the iteration and the sweep over the detections are paid once per J.
constexpr int kJMax = 4;
constexpr int kNBuf = kJMax + 1;            // parameter / view-constant buffers: current + kJMax trial
// red[]: A slots (and g_c), the cost, then W (and w) of every candidate
constexpr int kRedA = 0, kRedE = kNA, kRedW0 = kNA + 8;
__host__ __device__ constexpr int red_w(int j) { return kRedW0 + kNBD * j; }
__host__ __device__ constexpr int red_n(int J) { return kRedW0 + kNBD * J; }
constexpr int kRedN = red_n(kJMax);         // 456
__host__ __device__ inline int trial_buf(int cur, int j) { const int s = cur + 1 + j; return s >= kNBuf ? s - kNBuf : s; }

// <= count points of ONE view, contiguous in the point arrays
struct Piece { int view, start, count, pad; };
// what a persistent warp needs to start: its piece range and its first piece, in ONE 32-byte load
struct __align__(16) WarpRec { int p0, p1, end, pad1; Piece first; };   // end: one past the chunk's last point

struct LmState {
    double lam, E_cur, E_trial, sse;
    double lam_min, lam_max, err_min;
    int iter, max_iters, done, need_eval, cur, nranks;
    int comm_fail;         // a peer never arrived at an exchange (calib_peer.cuh)
    int J, jmax;           // candidates of this pass of the damping ladder; its cap (1 = ladder off)
    int jstart;            // candidates of the pass after an ACCEPTED step: 1, 2, or 0 = adaptive (see lm_accept_on)
    int jdyn;              // adaptive: 2 while the step tried right after an accepted one keeps being rejected, else 1
    int n_pass;            // passes (one k_schur/k_backsub/k_cost each) so far
    unsigned int cnt_schur, cnt_cost, cnt_eval, pad1;
    long long n_eval;   // Jacobian evaluations actually executed
    int want_ptrace;          // record the L shared parameters of the iterate beside every trace row
    unsigned int bar_count;   // persistent kernel: arrivals at its grid barriers so far (calib_persistent.cuh)
    int bar_abort;            // persistent kernel: a block gave up waiting at a barrier
};

struct Dev {  // everything the kernels need; passed by value
    int model, L, M;
    long long N, K;
    int ev_warps, ev_blocks, ev_pieces;   // partition of the points for k_eval
    int co_warps, co_blocks;              // partition for k_cost
    int sc_groups;                        // k_schur: view groups (grid = sc_groups x kJMax)
    int bs_groups;                        // k_backsub: view groups (grid = bs_groups x kJMax)
    const double *X, *Y, *Z, *U, *V;      // SoA detections
    const Piece* ev_piece;                // ev_pieces, in point (= view) order
    const WarpRec* ev_rec;                // ev_warps    : piece range (and first piece) of warp w
    const int* view_piece;                // M + 1       : piece range of view v
    const Piece* co_piece;
    const WarpRec* co_rec;                // co_warps
    double* Pbuf;                         // kNBuf x K
    ViewC* views;                         // kNBuf x M
    double* partA;                        // ev_blocks x kNA
    double* partE;                        // ev_blocks
    double* partBD;                       // ev_pieces x kNBD
    double* partW;                        // kJMax x sc_groups x kNBD
    double* red;                          // kRedN : the reduced system, all-reduced when sharded
    double* redA;                         // kRedW0: A slots, g_c and the cost of THIS rank at the current point
    double* vtab;                         // kJMax x M x kVtab
    double* partC;                        // co_blocks x kJMax (trial costs)
    double* etrial;                       // kJMax
    double* delta;                        // K
    double* dcs;                          // kJMax x 16: step of the shared parameters, per candidate
    double* trace;                        // max_iters x 4
    double* ptrace;                       // max_iters x 16: shared parameters of the iterate each row started from
    LmState* st;
    cudaGraphConditionalHandle cond;      // WHILE node of the whole-loop graph
    int use_cond;
    PeerDev peer;                         // nranks > 1: the kernels all-reduce over peer memory
#ifdef CALIB_STAMPS
    unsigned long long* stamps;           // [kernel 4][block kStampBlocks][slot 8] globaltimer (latency study builds only)
#endif
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

#ifdef CALIB_STAMPS
constexpr int kStampBlocks = 1280;
#define STAMP(d, kern, slot)                                                                           \
    do {                                                                                               \
        if (threadIdx.x == 0 && blockIdx.x < kStampBlocks)                                             \
            (d).stamps[((kern) * kStampBlocks + blockIdx.x) * 8 + (slot)] = global_timer_ns();        \
    } while (0)
// the persistent kernel: [first / last block][pass][slot] after the per-kernel stamps
constexpr int kPStampPasses = 128, kPStampSlots = 16;
constexpr size_t kStampWords = (size_t)4 * kStampBlocks * 8 + (size_t)2 * kPStampPasses * kPStampSlots;
#define PSTAMP(d, pass, slot)                                                                              \
    do {                                                                                                   \
        if (tid_now() == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1) && (pass) < kPStampPasses)  \
            (d).stamps[(size_t)4 * kStampBlocks * 8 +                                                      \
                       ((size_t)(blockIdx.x == 0 ? 0 : 1) * kPStampPasses + (pass)) * kPStampSlots + (slot)] = global_timer_ns(); \
    } while (0)
#else
#define STAMP(d, kern, slot) do {} while (0)
#define PSTAMP(d, pass, slot) do {} while (0)
#endif

// threadIdx.x through a volatile asm: the phase functions of the iteration take their thread indices
// this way so that, inlined one after the other into the persistent kernel, nothing derived from them
// is kept alive (hoisted, common-subexpression'd) from one phase across the sweeps of another -- the
// evaluation sweeps run at the 255-register limit and every such value became a spill inside them.
__device__ __forceinline__ int tid_now() {
    int t;
    asm volatile("mov.u32 %0, %%tid.x;" : "=r"(t));
    return t;
}

struct Pt { double X, Y, Z, U, V; };

__device__ __forceinline__ void load_point(const Dev& d, long long i, Pt& p) {
    p.X = __ldg(d.X + i); p.Y = __ldg(d.Y + i); p.Z = __ldg(d.Z + i);
    p.U = __ldg(d.U + i); p.V = __ldg(d.V + i);
}

// deterministic sum of n partials by one warp: strided per-lane chains, then the butterfly
__device__ __forceinline__ double warp_sum_array(const double* __restrict__ part, int n, int lane) {
    double acc = 0.0;
    for (int i = lane; i < n; i += 32) acc = __dadd_rn(acc, __ldcg(part + i));
    return warp_sum(acc);
}

// ---- per-point accumulation (shared with the batch kernel) -----------------------------------
// intrinsic block A (structure of d(u,v)/d(alpha,beta,gamma,uc,vc) = (xd,0,yd,1,0 ; 0,yd,0,0,1)
// exploited: 6 moments instead of 15 products), its right-hand side, and the cost
template <int MODEL>
__device__ __forceinline__ void accum_A(const PointOut<MODEL>& o, double (&acc)[kNA], double& e) {
    constexpr int ND = ModelTraits<MODEL>::ND;
    using S = SlotsA<ND>;
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
}

// coupling block B_v (L x 6), view block D_v (upper triangle) and view right-hand side
template <int MODEL>
__device__ __forceinline__ void accum_BD(const PointOut<MODEL>& o, double (&acc)[kNBD]) {
    constexpr int ND = ModelTraits<MODEL>::ND, L = ModelTraits<MODEL>::L;
    using S = SlotsBD<L>;
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
}

// The same sums from a hand-off slot of the evaluation pass (calib_passes.cuh): slot[k * 32] = pair k of
// this lane, pairs (xd,yd) (ru,rv) (ku,kv)[ND] (ju,jv)[6].  Every running sum gets the same single update per
// point as in accum_BD (bit-identical); only the order in which the inputs are touched differs: the
// six extrinsic column pairs are held, the left factors stream through two at a time, so that 14
// input values are live at once instead of 26 beside the 87 sums.
template <int MODEL>
__device__ __forceinline__ void accum_BD_j(const double2* slot, const double (&ju)[6], const double (&jv)[6], double (&acc)[kNBD]);

// (jslot[j * 32] = (ju,jv)[j]: the same slot, or the second ring of the pipelined pass.)
template <int MODEL>
__device__ __forceinline__ void accum_BD_ring2(const double2* slot, const double2* jslot, double (&acc)[kNBD]) {
    constexpr int ND = ModelTraits<MODEL>::ND, L = ModelTraits<MODEL>::L;
    using S = SlotsBD<L>;
    double ju[6], jv[6];
#pragma unroll
    for (int j = 0; j < 6; ++j) { const double2 t = jslot[j * 32]; ju[j] = t.x; jv[j] = t.y; }
    accum_BD_j<MODEL>(slot, ju, jv, acc);
}

// ... with the extrinsic columns already in registers
template <int MODEL>
__device__ __forceinline__ void accum_BD_j(const double2* slot, const double (&ju)[6], const double (&jv)[6], double (&acc)[kNBD]) {
    constexpr int ND = ModelTraits<MODEL>::ND, L = ModelTraits<MODEL>::L;
    using S = SlotsBD<L>;
    {
        const double2 r = slot[1 * 32];                  // (ru, rv)
#pragma unroll
        for (int j = 0; j < 6; ++j) {
            acc[S::B + 3 * 6 + j] += ju[j];                                   // uc
            acc[S::B + 4 * 6 + j] += jv[j];                                   // vc
#pragma unroll
            for (int b = j; b < 6; ++b)
                acc[S::dd(j, b)] = fma(ju[j], ju[b], fma(jv[j], jv[b], acc[S::dd(j, b)]));
            acc[S::GV + j] = fma(ju[j], r.x, fma(jv[j], r.y, acc[S::GV + j]));
        }
    }
    {
        const double2 x = slot[0 * 32];                  // (xd, yd)
#pragma unroll
        for (int j = 0; j < 6; ++j) {
            acc[S::B + 0 * 6 + j] = fma(x.x, ju[j], acc[S::B + 0 * 6 + j]);   // alpha
            acc[S::B + 1 * 6 + j] = fma(x.y, jv[j], acc[S::B + 1 * 6 + j]);   // beta
            acc[S::B + 2 * 6 + j] = fma(x.y, ju[j], acc[S::B + 2 * 6 + j]);   // gamma
        }
    }
#pragma unroll
    for (int a = 0; a < ND; ++a) {
        const double2 k = slot[(2 + a) * 32];            // (ku_a, kv_a)
#pragma unroll
        for (int j = 0; j < 6; ++j)
            acc[S::B + (5 + a) * 6 + j] = fma(k.x, ju[j], fma(k.y, jv[j], acc[S::B + (5 + a) * 6 + j]));
    }
}

template <int MODEL>
__device__ __forceinline__ void accum_BD_ring(const double2* slot, double (&acc)[kNBD]) {
    accum_BD_ring2<MODEL>(slot, slot + (2 + ModelTraits<MODEL>::ND) * 32, acc);
}

// ---- a warp's sweep over its pieces ---------------------------------------------------------
// The next point (and the next piece's view constants) are fetched one step ahead, across piece
// boundaries, so a load is never waited for inside the FP64 stream.  `begin(piece, index)` runs
// after the view constants are in sview (warp-synchronised), `body(point)` per point of the
// lane, `end(piece, index)` with all 32 lanes converged.
// (Measured and rejected, profiles/README.md: two steps of look-ahead, a counted loop, and a two-stage
// software-pipelined body.)
template <int NV, class Begin, class Body, class End>
__device__ __forceinline__ void sweep_pieces(const Dev& d, const Piece* __restrict__ pieces, const WarpRec& rec,
                                             const ViewC* const (&views)[NV], ViewC (&sview)[NV], int lane,
                                             Begin begin, Body body, End end) {
    const int p0 = rec.p0, p1 = rec.p1;
    if (p0 >= p1) return;
    Piece cur = rec.first;
    Pt pn;
    if (lane < cur.count) load_point(d, (long long)cur.start + lane, pn);
    double vn[NV];
#pragma unroll
    for (int j = 0; j < NV; ++j) {
        vn[j] = 0.0;
        if (lane < kViewDoubles) vn[j] = reinterpret_cast<const double*>(views[j] + cur.view)[lane];
    }
    for (int pi = p0; pi < p1; ++pi) {
        const bool has_next = pi + 1 < p1;
        Piece nxt = cur;
        if (has_next) nxt = pieces[pi + 1];
        __syncwarp();
        if (lane < kViewDoubles) {
#pragma unroll
            for (int j = 0; j < NV; ++j) reinterpret_cast<double*>(&sview[j])[lane] = vn[j];
        }
        __syncwarp();
        if (has_next && lane < kViewDoubles) {
#pragma unroll
            for (int j = 0; j < NV; ++j) vn[j] = reinterpret_cast<const double*>(views[j] + nxt.view)[lane];
        }
        begin(cur, pi);
        if (lane >= cur.count && has_next && lane < nxt.count) load_point(d, (long long)nxt.start + lane, pn);
        for (int i = lane; i < cur.count; i += 32) {
            const Pt p = pn;
            if (i + 32 < cur.count) load_point(d, (long long)cur.start + i + 32, pn);
            else if (has_next && lane < nxt.count) load_point(d, (long long)nxt.start + lane, pn);
            body(p);
        }
        __syncwarp();
        end(cur, pi);
        cur = nxt;
    }
}

// ---- view constants for parameter buffer `which` (start of a run) ---------------------------
__global__ void k_make_views(Dev d, int which_is_cur) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= d.M) return;
    const int b = which_is_cur ? d.st->cur : trial_buf(d.st->cur, 0);
    ViewC w;
    make_view(d.Pbuf + (long long)b * d.K + d.L + 6ll * v, w);
    d.views[(long long)b * d.M + v] = w;
}

}  // namespace calib
