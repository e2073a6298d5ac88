// The whole LM loop of src/calibrate.py:141-171 as ONE cooperative launch: a persistent grid (one
// block per SM at most) that runs the phases of every pass -- evaluation, per-view Schur stage, reduced
// solve, back-substitution, trial-cost sweep, accept / reject -- separated by grid barriers instead of
// kernel boundaries.  The per-launch pipeline (k_eval -> k_schur -> k_backsub -> k_cost, loop modes 1
// and 2) pays a launch, a prologue, a "last block" ticket and a tail per kernel: about 35 us per pass
// whatever the problem size (profiles/r1l_kernel_times_vs_size.txt); here a phase change is one
// barrier (an atomic arrive and a poll of one L2 line).
//
// The phases are the SAME device functions the per-launch kernels call (eval_block, schur_views_block,
// schur_tail, backsub_block, cost_block, cost_total, lm_accept_on), on the same virtual blocks and in
// the same summation orders, so iterates, costs, lambdas, trace rows and the iteration count are
// bit-identical to loop modes 1 and 2 (tests/test_gpu_parity.py::test_loop_modes_identical).
//
// Who does what:
//   every block  keeps its own copy of the LM state in shared memory and advances it by replaying the
//                accept rule on the same trial costs (no broadcast of the decision);
//   block 0      gathers the reduced system, exchanges it with the peer ranks when the views are
//                sharded (calib_peer.cuh), solves it, writes the trace and, at the end, the state.
// Barriers per pass: [evaluation] - view stage - reduced solve - back-substitution - cost sweep.  One GPU:
// every block sums the block costs itself (co_blocks x 4 doubles) and decides; sharded views: block 0
// sums, exchanges with the peers, publishes the totals, and one more barrier follows.
//
// EXPERIMENT, not in the default build (-DCALIB_EXPERIMENT_PERSISTENT adds it as loop mode 3).  Measured on
// B200 with the two-sweep evaluation pass of round 1 (profiles/r2b_persistent_first.txt,
// r2c_pstamps_first.txt): bit-identical to loop modes 1 / 2, but NOT faster -- cfg1 28.1 against 27.1 us
// per LM iteration, cfg3 1.61 against 1.21 ms per refinement.  A grid barrier (arrive + poll + two
// gpu-scope fences) costs 1.5-2 us, what a kernel boundary inside the WHILE graph costs; the phases
// themselves (view stage 3.6, reduced system 5.6, back-substitution 2.8 us at 540 points) are latency
// chains that no launch mechanism shortens; the cost sweep loses a third with 8 instead of 16 warps per
// SM, and ptxas spills inside the sweeps once they sit in one body with the other phases.  With the
// producer / consumer evaluation pass of round 2 the hand-off of a chunk whose first step is a partial
// warp never completes inside this kernel (the same code is fine as k_eval; not understood, the waits are
// bounded so it fails with CALIB_E_CUDA instead of hanging).
#pragma once
#include "../calib_solve.cuh"
#include "calib_eval_pairs.cuh"

namespace calib {

constexpr size_t kPersistDynSmem = kEvalDynSmem > kCostDynSmem ? kEvalDynSmem : kCostDynSmem;
constexpr unsigned long long kGridTimeoutNs = 40ull * 1000 * 1000 * 1000;   // > the peer exchange's own timeout

// Arrive-and-wait over all blocks of the (co-resident) grid.  `target` counts arrivals so far + one
// round, per block; the counter only grows within a run and is zeroed by the host before the launch.
// The fences on both sides are what make plain loads after the barrier see the other blocks' writes
// (fence.acq_rel.gpu invalidates the SM's L1; same scheme as cooperative groups' grid sync).
// Returns false when the grid was aborted (a block waited longer than kGridTimeoutNs: some block
// died or the launch was not co-resident) -- every block then leaves.
__device__ __forceinline__ bool grid_barrier(LmState* st, unsigned int& s_target, int* s_flag) {
    __syncthreads();
    if (tid_now() == 0) {
        const unsigned int target = s_target + gridDim.x;
        s_target = target;
        __threadfence();
        atomicAdd(&st->bar_count, 1u);
        volatile unsigned int* cnt = &st->bar_count;
        volatile int* abort_flag = &st->bar_abort;
        int ok = 1;
        if ((int)(*cnt - target) < 0) {
            const unsigned long long t0 = global_timer_ns();
            unsigned spins = 0;
            while ((int)(*cnt - target) < 0) {
                if ((++spins & 255u) == 0) {
                    if (*abort_flag) { ok = 0; break; }
                    if (global_timer_ns() - t0 > kGridTimeoutNs) { *abort_flag = 1; ok = 0; break; }
                }
            }
        }
        if (ok && *abort_flag) ok = 0;
        __threadfence();
        *s_flag = ok;
    }
    __syncthreads();
    return *s_flag != 0;
}

template <int MODEL>
union PersistSmem {
    EvalSmem e;
    SchurSmem<MODEL> s;
    BacksubSmem b;
    CostSmem c;
};

// The latency-bound phases as functions of their own (not inlined): inlined into one body with the
// sweeps, whatever ptxas kept alive of them cost the evaluation sweeps ~30 registers, i.e. spills
// inside their point loops.  d lives in the kernel's parameter space (__grid_constant__), so passing it
// by reference makes no copy; these phases can afford to read it through a pointer, the sweeps (which
// stay inlined and take their array bases as constant-bank operands) cannot.
template <int MODEL>
__device__ __noinline__ void schur_views_phase(const Dev& d, SchurSmem<MODEL>& sm, int J, double lam) {
    const int nvb = d.sc_groups;
    for (int vb = (int)blockIdx.x; vb < nvb * J; vb += (int)gridDim.x) {
        const int cand = vb / nvb, grp = vb - cand * nvb;
        schur_views_block<MODEL>(d, sm, cand, grp, lam);
        __syncthreads();
    }
}

template <int MODEL>
__device__ __noinline__ bool schur_tail_phase(const Dev& d, SchurSmem<MODEL>& sm, const LmState& st) {
    const PassCtx c = pass_ctx(st);
    return schur_tail<MODEL>(d, sm, c, 1);
}

template <int MODEL>
__device__ __noinline__ void backsub_phase(const Dev& d, BacksubSmem& sm, int J, int cur) {
    backsub_load_dc<MODEL>(d, sm, J);
    const int nb = d.bs_groups;
    for (int vb = (int)blockIdx.x; vb < nb * J; vb += (int)gridDim.x) {
        const int j = vb / nb, grp = vb - j * nb;
        backsub_block<MODEL>(d, sm, j, grp, cur);
    }
}

// view constants of the starting point (k_make_views of the per-launch modes)
__device__ __noinline__ void make_views_phase(const Dev& d, int cur) {
    for (int v = blockIdx.x * kBlockThreads + tid_now(); v < d.M; v += gridDim.x * kBlockThreads) {
        ViewC w;
        make_view(d.Pbuf + (long long)cur * d.K + d.L + 6ll * v, w);
        d.views[(long long)cur * d.M + v] = w;
    }
}

// one-GPU decision: every block sums the block costs itself and replays the accept rule on its copy
__device__ __noinline__ void accept_phase(const Dev& d, CostSmem& sm, LmState& st) {
    cost_total(d, sm);                          // every block: same rows, same order, same sums
    if (tid_now() == 0) {
        if (blockIdx.x == 0) for (int j = 0; j < kJMax; ++j) d.etrial[j] = sm.s_tot[j];
        lm_accept_on(&st, sm.s_tot, blockIdx.x == 0 ? d.trace : nullptr, d.ptrace, d.Pbuf + (long long)st.cur * d.K, d.L);
    }
}

// sharded views: block 0 sums the block costs and exchanges the totals with the peer ranks
__device__ __noinline__ void cost_exchange_phase(const Dev& d, CostSmem& sm, int J) {
    cost_total(d, sm);
    const bool ok = peer_allreduce(d.peer, 1, sm.s_tot, J, &sm.s_last);
    if (tid_now() == 0) {
        for (int j = 0; j < kJMax; ++j) d.etrial[j] = sm.s_tot[j];
        if (!ok) { d.st->comm_fail = 1; d.st->done = 1; }
    }
}

__device__ __noinline__ void accept_published_phase(const Dev& d, LmState& st) {
    if (tid_now() == 0) {
        double et[kJMax];
        for (int j = 0; j < kJMax; ++j) et[j] = __ldcg(d.etrial + j);
        lm_accept_on(&st, et, blockIdx.x == 0 ? d.trace : nullptr, d.ptrace, d.Pbuf + (long long)st.cur * d.K, d.L);
    }
}

// TMA: the cost sweep streams the detections through the bulk-copy ring (dynamic shared memory:
// kCostDynSmem).
template <int MODEL, bool TMA>
__global__ void __launch_bounds__(kBlockThreads, 1) k_lm_persistent(const __grid_constant__ Dev d) {
    // Nothing of the loop is kept in registers across a phase: the state copy, the barrier target and
    // the virtual-block counters live in shared memory and the grid geometry is re-read from the
    // special registers.  The evaluation sweeps sit exactly at the 255-register limit; every value
    // live across them turned into local-memory traffic inside their point loops (18 loads per
    // 32-point step with the loop state in registers).
    __shared__ PersistSmem<MODEL> sm;
    __shared__ LmState s_st;
    __shared__ int s_flag, s_vb;
    __shared__ unsigned int s_target;
    LmState* const gst = d.st;
    if (tid_now() == 0) { s_st = *gst; s_target = 0; }   // pushed by the host; nobody writes it before the first barrier
    __syncthreads();
    make_views_phase(d, s_st.cur);
    if (!grid_barrier(gst, s_target, &s_flag)) return;
    for (;;) {
        if (s_st.done) break;
        PSTAMP(d, s_st.n_pass, 0);
        // ---- evaluation at the current point (skipped after a rejected step: nothing changed)
        if (s_st.need_eval) {
            if (tid_now() == 0) s_vb = blockIdx.x;
            __syncthreads();
            while (s_vb < d.ev_blocks) {
                const int w = s_vb * kEvalPairs + ((tid_now() >> 5) & (kEvalPairs - 1));
                WarpRec rec;
                rec.p0 = rec.p1 = 0;
                if (w < d.ev_warps) rec = d.ev_rec[w];
                eval_block<MODEL>(d, sm.e, &s_vb, s_st.cur, rec);
                __syncthreads();
                if (tid_now() == 0) s_vb += gridDim.x;
                __syncthreads();
            }
            PSTAMP(d, s_st.n_pass, 1);
            if (!grid_barrier(gst, s_target, &s_flag)) return;
            PSTAMP(d, s_st.n_pass, 2);
        }
        // ---- per-view Schur stage, one virtual block per (candidate, view group)
        schur_views_phase<MODEL>(d, sm.s, s_st.J, s_st.lam);
        PSTAMP(d, s_st.n_pass, 3);
        if (!grid_barrier(gst, s_target, &s_flag)) return;
        PSTAMP(d, s_st.n_pass, 4);
        // ---- reduced system: gather, exchange, solve (block 0)
        if (blockIdx.x == 0) {
            const bool ok = schur_tail_phase<MODEL>(d, sm.s, s_st);
            if (tid_now() == 0) {
                if (!ok) { gst->comm_fail = 1; gst->done = 1; }
                else if (s_st.iter == 0) gst->E_cur = sm.s.sred[kRedE];
            }
        }
        PSTAMP(d, s_st.n_pass, 5);
        if (!grid_barrier(gst, s_target, &s_flag)) return;
        PSTAMP(d, s_st.n_pass, 6);
        if (d.peer.nranks > 1 && gst->comm_fail) return;            // uniform: written before the barrier
        if (s_st.iter == 0 && tid_now() == 0) s_st.E_cur = gst->E_cur;
        // ---- back-substitution, trial parameters and trial view constants
        backsub_phase<MODEL>(d, sm.b, s_st.J, s_st.cur);
        PSTAMP(d, s_st.n_pass, 7);
        if (!grid_barrier(gst, s_target, &s_flag)) return;
        PSTAMP(d, s_st.n_pass, 8);
        // ---- trial costs
        if (tid_now() == 0) s_vb = blockIdx.x;
        __syncthreads();
        while (s_vb < d.co_blocks) {
            const int w = s_vb * kBlockWarps + (tid_now() >> 5);
            WarpRec rec;
            rec.p0 = rec.p1 = 0;
            if (w < d.co_warps) rec = d.co_rec[w];
            cost_block<MODEL, TMA>(d, sm.c, &s_vb, s_st.cur, s_st.J, rec);
            __syncthreads();
            if (tid_now() == 0) s_vb += gridDim.x;
            __syncthreads();
        }
        PSTAMP(d, s_st.n_pass, 9);
        if (!grid_barrier(gst, s_target, &s_flag)) return;
        PSTAMP(d, s_st.n_pass, 10);
        // ---- totals and the accept / reject decision
        if (d.peer.nranks <= 1) {
            accept_phase(d, sm.c, s_st);
        } else {
            if (blockIdx.x == 0) cost_exchange_phase(d, sm.c, s_st.J);
            if (!grid_barrier(gst, s_target, &s_flag)) return;
            if (gst->comm_fail) return;
            accept_published_phase(d, s_st);
        }
        __syncthreads();
    }
    if (blockIdx.x == 0 && tid_now() == 0) {
        gst->lam = s_st.lam; gst->E_cur = s_st.E_cur; gst->E_trial = s_st.E_trial; gst->sse = s_st.sse;
        gst->iter = s_st.iter; gst->done = s_st.done; gst->need_eval = s_st.need_eval; gst->cur = s_st.cur;
        gst->J = s_st.J; gst->n_pass = s_st.n_pass;
    }
}

}  // namespace calib
