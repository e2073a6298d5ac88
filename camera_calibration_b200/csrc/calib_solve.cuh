// Schur complement over the views, reduced solve and back-substitution (device side of
// `np.linalg.inv(JTJ + lam*diagJTJ) @ J.T @ r`, src/calibrate.py:147,152, in block form):
//   S = A' - sum_v B_v D_v'^-1 B_v^T,  s = g_c - sum_v B_v D_v'^-1 g_v,   S dc = s   (L x L)
//   dv = D_v'^-1 (g_v - B_v^T dc)                                         (6 x 6 per view)
// with A' = A + lam diag(A), D_v' = D_v + lam diag(D_v).
#pragma once
#include "calib_passes.cuh"

namespace calib {

// slot of A(i,j), i <= j, in the intrinsic-sweep layout; -1 = structurally zero
template <int ND>
__host__ __device__ inline int a_slot(int i, int j) {
    using S = SlotsA<ND>;
    if (j < 5) {
        // moments of (xd, yd, 1): rows/columns alpha, beta, gamma, uc, vc; 4-bit entries (slot + 1)
        //   {0,-1,1,2,-1}, {-1,3,-1,-1,4}, {1,-1,3,4,-1}, {2,-1,4,5,-1}, {-1,4,-1,-1,5}
        const int c = i * 5 + j;
        const unsigned long long t = c < 16 ? 0x3054025004003201ull : 0x600500650ull;
        return (int)((t >> (4 * (c & 15))) & 15) - 1;
    }
    if (i < 5) {
        const int a = j - 5;
        switch (i) {
            case 0: return S::UK + 3 * a;
            case 2: return S::UK + 3 * a + 1;
            case 3: return S::UK + 3 * a + 2;
            case 1: return S::VK + 2 * a;
            default: return S::VK + 2 * a + 1;
        }
    }
    return S::kk(i - 5, j - 5);
}

// ---- reduced L x L system: Cholesky solve by one warp ------------------------------------------
// Lane i keeps row i of S in registers; the factorisation is right-looking (after column j every
// lane updates the rest of its row), so the serial chain per column is one reciprocal square root,
// one multiply and two shuffles.  The factor goes back to shared memory so that the back-substitution
// reads L[j][i] instead of shuffling whole rows.  S: L x L row-major, then the right-hand side.
// Returns dc[lane] for lane < L.
struct LdCg { __device__ __forceinline__ double operator()(const double* p) const { return __ldcg(p); } };
struct LdPlain { __device__ __forceinline__ double operator()(const double* p) const { return *p; } };

template <int L>
__device__ __forceinline__ double solve_dense_core(double* S, int lane) {
    const int i = lane < L ? lane : L - 1;       // idle lanes mirror the last row (no divergence)
    double row[L];
#pragma unroll
    for (int j = 0; j < L; ++j) row[j] = S[i * L + j];
    double rhs = S[L * L + i];
    double inv[L];
#pragma unroll
    for (int j = 0; j < L; ++j) {
        const double dj = __shfl_sync(0xffffffffu, row[j], j);
        const double id = rsqrt_nr(dj);          // dj <= 0 or NaN poisons the step (rejected)
        inv[j] = id;
        const double lij = row[j] * id;
        row[j] = lij;
#pragma unroll
        for (int k = 0; k < L; ++k) {
            if (k > j) {
                const double lkj = __shfl_sync(0xffffffffu, lij, k);
                row[k] = fma(-lij, lkj, row[k]);
            }
        }
    }
    __syncwarp();
    if (lane < L) {
#pragma unroll
        for (int j = 0; j < L; ++j) S[i * L + j] = row[j];
    }
    __syncwarp();
#pragma unroll
    for (int j = 0; j < L; ++j) {
        const double yj = __shfl_sync(0xffffffffu, rhs * inv[j], j);
        if (i == j) rhs = yj;
        else if (i > j) rhs = fma(-row[j], yj, rhs);
    }
#pragma unroll
    for (int j = L - 1; j >= 0; --j) {
        const double xj = __shfl_sync(0xffffffffu, rhs * inv[j], j);
        const double mine = S[j * L + i];        // L[j][i]
        if (i == j) rhs = xj;
        else if (i < j) rhs = fma(-mine, xj, rhs);
    }
    return rhs;
}

// Called by a whole block after red (shared or global, layout of red[]) is final: builds the dense
// systems of the J candidates with all threads, then warp j solves candidate j and writes the step
// and the trial shared parameters.  E seeds the loop (see lm_accept): from the second iteration on
// the current cost IS the trial cost of the last accepted step, as in the reference where both
// numbers come from one function (calibrate.py:155-156).
template <int MODEL, class Ld>
__device__ __forceinline__ void solve_candidates(const Dev& d, const double* red, double* dense, const PassCtx& c, Ld ld) {
    using T = ModelTraits<MODEL>;
    constexpr int L = T::L, ND = T::ND, NE = L * L + L;
    using SA = SlotsA<ND>;
    const int tid = tid_now(), lane = tid & 31, warp = tid >> 5;
    const double lam0 = c.lam;
    const int cur = c.cur, J = c.J;
    double pcur = 0.0;
    if (warp < J && lane < L) pcur = d.Pbuf[(long long)cur * d.K + lane];
    // dense S = A + lam diag(A) - W (L x L, row-major), then s = g_c - w (L), per candidate; red may be
    // global memory (NCCL path), hence ld
    for (int e = tid; e < J * NE; e += (int)blockDim.x) {
        const int j = e / NE, el = e - j * NE;
        double lam = lam0;
        for (int k = 0; k < j; ++k) lam = lam * 10.0;      // the reference's own update chain
        const double* redA = red + kRedA;
        const double* redW = red + red_w(j);
        double out;
        if (el < L * L) {
            const int i = el / L, cc = el - i * L;
            const int lo = i < cc ? i : cc, hi = i < cc ? cc : i;
            const int sl = a_slot<ND>(lo, hi);
            double v = sl >= 0 ? ld(redA + sl) : 0.0;
            if (i == cc) v = fma(lam, v, v);
            out = v - ld(redW + tri(L, lo, hi));
        } else {
            const int i = el - L * L;
            out = ld(redA + SA::G + i) - ld(redW + L * (L + 1) / 2 + i);
        }
        dense[e] = out;
    }
    __syncthreads();
    if (warp < J) {
        const double dc = solve_dense_core<L>(dense + warp * NE, lane);
        if (lane < L) {
            if (warp == 0) d.delta[lane] = dc;
            d.dcs[warp * 16 + lane] = dc;
            d.Pbuf[(long long)trial_buf(cur, warp) * d.K + lane] = pcur + dc;
        }
    }
}

// ---- per-view Schur stage -------------------------------------------------------------------------
template <int MODEL> struct SchurSmem {
    static constexpr int L = ModelTraits<MODEL>::L;
    double sv[kBlockWarps][kNBD];          // the view's summed piece rows; reused for sums
    double sz[kBlockWarps][(L + 1) * 6];   // Z = L^-1 [B^T | g_v], column-major by c
    double sl[kBlockWarps][24];            // packed lower factor of the view block
    double sga[kBlockWarps][kNA], sge[kBlockWarps];
    double sred[kRedN];                    // the reduced system in the layout of red[]
    double sdense[kJMax * (L * L + L)];    // its dense form, per candidate
    int s_last;
};

// Block (grp, cand) of the per-view stage: candidate `cand` of the damping ladder on views
// 8 grp + warp + k * 8 sc_groups, so the views and the summation order of a candidate do not depend on
// how many candidates the pass has.  lam0: damping of candidate 0.  Writes vtab and row (cand, grp) of
// partW.  Called by all threads of a block; no barrier at entry, none after its last shared read.
template <int MODEL>
__device__ __forceinline__ void schur_views_block(const Dev& d, SchurSmem<MODEL>& sm, int cand, int grp, double lam0) {
    using T = ModelTraits<MODEL>;
    constexpr int L = T::L, NZ = L + 1, NT = L * (L + 1) / 2, NW = NT + L;
    using S = SlotsBD<L>;
    const int tid = tid_now(), lane = tid & 31, warp = tid >> 5;
    const int nvb = d.sc_groups;
    // which entries of W this lane accumulates: e = lane + 32 k  ->  (p, c), c in [p, L]
    int ep[3], ec[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        int e = lane + 32 * k;
        ep[k] = -1; ec[k] = 0;
        if (e < NT) {
            int p = 0;
            while (e >= L - p) { e -= L - p; ++p; }
            ep[k] = p; ec[k] = p + e;
        } else if (e < NW) {
            ep[k] = e - NT; ec[k] = L;
        }
    }
    double wacc[3] = {0.0, 0.0, 0.0};
    double lam = lam0;
    for (int k = 0; k < cand; ++k) lam = lam * 10.0;
    double* vtab = d.vtab + (long long)cand * d.M * kVtab;
    for (int v = grp * kBlockWarps + warp; v < d.M; v += nvb * kBlockWarps) {
        const int r0 = d.view_piece[v], r1 = d.view_piece[v + 1];
        double s0 = 0.0, s1 = 0.0, s2 = 0.0;
        for (int r = r0; r < r1; ++r) {
            const double* row = d.partBD + (long long)r * kNBD;
            s0 += __ldcg(row + lane); s1 += __ldcg(row + lane + 32); s2 += __ldcg(row + lane + 64);
        }
        __syncwarp();
        STAMP(d, 1, 1);
        sm.sv[warp][lane] = s0; sm.sv[warp][lane + 32] = s1; sm.sv[warp][lane + 64] = s2;
        __syncwarp();
        // D' = D + lam diag(D), lower Cholesky, every lane redundantly (registers)
        double Lc[36];
#pragma unroll
        for (int a = 0; a < 6; ++a)
#pragma unroll
            for (int b = a; b < 6; ++b) {
                const double s = sm.sv[warp][S::dd(a, b)];
                Lc[b * 6 + a] = (a == b) ? fma(lam, s, s) : s;
            }
        double id[6];
        chol_inplace<6>(Lc, id);
        double* vt = vtab + (long long)v * kVtab;
        if (lane == 0) {
#pragma unroll
            for (int a = 0; a < 6; ++a)
#pragma unroll
                for (int b = 0; b <= a; ++b) sm.sl[warp][a * (a + 1) / 2 + b] = (a == b) ? id[a] : Lc[a * 6 + b];
        }
        // forward substitution, lane c <= L solves L z = column c of [B^T | g_v]
        if (lane < NZ) {
            double z[6];
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                double s = sm.sv[warp][(lane < L) ? (S::B + lane * 6 + i) : (S::GV + i)];
#pragma unroll
                for (int k = 0; k < i; ++k) s = fma(-Lc[i * 6 + k], z[k], s);
                z[i] = s * id[i];
                sm.sz[warp][lane * 6 + i] = z[i];
            }
        }
        __syncwarp();
        if (lane < 21) vt[lane] = sm.sl[warp][lane];     // strictly-lower factor, 1/diagonal on the diagonal
        for (int i = lane; i < NZ * 6; i += 32) vt[21 + i] = sm.sz[warp][i];
        // W = Z^T Z (upper triangle) and w = Z^T z_g
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            if (ep[k] >= 0) {
                const double* zp = &sm.sz[warp][ep[k] * 6];
                const double* zc = &sm.sz[warp][ec[k] * 6];
                double s = 0.0;
#pragma unroll
                for (int i = 0; i < 6; ++i) s = fma(zp[i], zc[i], s);
                wacc[k] += s;
            }
        }
    }
    __syncwarp();
    STAMP(d, 1, 2);
    sm.sv[warp][lane] = wacc[0]; sm.sv[warp][lane + 32] = wacc[1]; sm.sv[warp][lane + 64] = wacc[2];
    __syncthreads();
    if (tid < kNBD) {
        double s = sm.sv[0][tid];
#pragma unroll
        for (int k = 1; k < kBlockWarps; ++k) s += sm.sv[k][tid];
        d.partW[((long long)cand * nvb + grp) * kNBD + tid] = s;
    }
}

// By ONE block once every row of partW (and, after an evaluation, of partA / partE) is visible:
// gathers them in a fixed order (all loads of a warp in flight at once), publishes the reduced system
// (summed over the ranks when the views are sharded) and, with fuse_solve, solves it from shared
// memory, one warp per candidate.  Returns false when a peer never arrived at the exchange.
template <int MODEL>
__device__ __forceinline__ bool schur_tail(const Dev& d, SchurSmem<MODEL>& sm, const PassCtx& c, int fuse_solve) {
    const int tid = tid_now(), lane = tid & 31, warp = tid >> 5;
    const int nvb = d.sc_groups, J = c.J;
    // The intrinsic block, g_c and the cost at the current point: after an evaluation the block rows
    // the evaluation pass wrote are summed here and kept in redA; while steps from this point are
    // rejected they do not change and 65 numbers are read back.
    if (c.need_eval) {
        constexpr int R = 19;                       // rows per warp held in flight (148 blocks / 8 warps)
        double a0 = 0.0, a1 = 0.0, e = 0.0;
        for (int r0 = warp; r0 < d.ev_blocks; r0 += kBlockWarps * R) {
            double t0[R], t1[R];
#pragma unroll
            for (int k = 0; k < R; ++k) {
                const int r = r0 + k * kBlockWarps;
                const double* row = d.partA + (long long)r * kNA;
                t0[k] = r < d.ev_blocks ? __ldcg(row + lane) : 0.0;
                t1[k] = r < d.ev_blocks ? __ldcg(row + lane + 32) : 0.0;
            }
#pragma unroll
            for (int k = 0; k < R; ++k) { a0 += t0[k]; a1 += t1[k]; }
        }
        for (int r = tid; r < d.ev_blocks; r += kBlockThreads) e = __dadd_rn(e, __ldcg(d.partE + r));
        e = warp_sum(e);
        sm.sga[warp][lane] = a0; sm.sga[warp][lane + 32] = a1;
        if (lane == 0) sm.sge[warp] = e;
        __syncthreads();
        if (tid < kNA) {
            double s = sm.sga[0][tid];
#pragma unroll
            for (int k = 1; k < kBlockWarps; ++k) s += sm.sga[k][tid];
            sm.sred[kRedA + tid] = s;
            d.redA[kRedA + tid] = s;
        } else if (tid == kNA) {
            double s = sm.sge[0];
#pragma unroll
            for (int k = 1; k < kBlockWarps; ++k) s = __dadd_rn(s, sm.sge[k]);
            sm.sred[kRedE] = s;
            d.redA[kRedE] = s;
            d.st->n_eval += 1;
        } else if (tid < kRedW0) {
            sm.sred[tid] = 0.0;
        }
    } else if (tid < kRedW0) {
        sm.sred[tid] = tid <= kRedE ? __ldcg(d.redA + tid) : 0.0;
    }
    STAMP(d, 1, 4);
    for (int j = 0; j < J; ++j) {
        constexpr int R = 8;                        // rows per warp held in flight
        double s0 = 0.0, s1 = 0.0, s2 = 0.0;
        for (int r0 = warp; r0 < nvb; r0 += kBlockWarps * R) {
            double t0[R], t1[R], t2[R];
#pragma unroll
            for (int k = 0; k < R; ++k) {
                const int r = r0 + k * kBlockWarps;
                const double* row = d.partW + ((long long)j * nvb + r) * kNBD;
                const bool in = r < nvb;
                t0[k] = in ? __ldcg(row + lane) : 0.0;
                t1[k] = in ? __ldcg(row + lane + 32) : 0.0;
                t2[k] = in ? __ldcg(row + lane + 64) : 0.0;
            }
#pragma unroll
            for (int k = 0; k < R; ++k) { s0 += t0[k]; s1 += t1[k]; s2 += t2[k]; }
        }
        __syncthreads();          // sv of the previous candidate has been consumed
        sm.sv[warp][lane] = s0; sm.sv[warp][lane + 32] = s1; sm.sv[warp][lane + 64] = s2;
        __syncthreads();
        if (tid < kNBD) {
            double s = sm.sv[0][tid];
#pragma unroll
            for (int k = 1; k < kBlockWarps; ++k) s += sm.sv[k][tid];
            sm.sred[red_w(j) + tid] = s;
        }
    }
    __syncthreads();
    // views sharded over the GPUs of the node: sum the reduced systems of all ranks, right here
    if (d.peer.nranks > 1 && !peer_allreduce(d.peer, 0, sm.sred, red_n(J), &sm.s_last)) return false;
    for (int e = tid; e < red_n(J); e += kBlockThreads) d.red[e] = sm.sred[e];
    STAMP(d, 1, 5);
    if (fuse_solve) solve_candidates<MODEL>(d, sm.sred, sm.sdense, c, LdPlain());
    STAMP(d, 1, 6);
    return true;
}

// Grid = sc_groups x kJMax blocks: block (g, j) works for candidate j of the damping ladder (and
// leaves the view stage at once when the pass has fewer candidates); the last block to finish (atomic
// ticket) runs the tail.
template <int MODEL>
__global__ void __launch_bounds__(kBlockThreads) k_schur(Dev d, int fuse_solve) {
    LmState* st = d.st;
    if (st->done) return;
    __shared__ SchurSmem<MODEL> sm;
    const int tid = threadIdx.x;
    const int nvb = d.sc_groups;
    const PassCtx c = pass_ctx(*st);
    const int cand = blockIdx.x / nvb, grp = blockIdx.x - cand * nvb;
    STAMP(d, 1, 0);
    if (cand < c.J) schur_views_block<MODEL>(d, sm, cand, grp, c.lam);
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const unsigned int prev = atomicAdd(&st->cnt_schur, 1u);
        sm.s_last = (prev == gridDim.x - 1);
    }
    __syncthreads();
    STAMP(d, 1, 3);
    if (!sm.s_last) return;
    __threadfence();
    if (tid == 0) st->cnt_schur = 0;
    if (!schur_tail<MODEL>(d, sm, c, fuse_solve)) {
        if (tid == 0) { st->comm_fail = 1; st->done = 1; }
        return;
    }
    if (fuse_solve && tid == 0 && c.iter == 0) st->E_cur = sm.sred[kRedE];
}

// reduced solve as its own launch (after the all-reduce when NCCL sums the reduced systems):
// one warp per candidate
template <int MODEL>
__global__ void k_solve(Dev d) {
    constexpr int L = ModelTraits<MODEL>::L;
    if (d.st->done) return;
    __shared__ double sdense[kJMax * (L * L + L)];
    const PassCtx c = pass_ctx(*d.st);
    solve_candidates<MODEL>(d, d.red, sdense, c, LdCg());
    if (threadIdx.x == 0 && c.iter == 0) d.st->E_cur = __ldcg(d.red + kRedE);
}

// ---- back-substitution, trial parameters and trial view constants: one warp per view; block (g, j)
// works for candidate j on views 8 g + warp + k * 8 bs_groups -----------------------------------------
struct __align__(16) BacksubSmem {
    ViewC sview[kBlockWarps];
    double srow[kBlockWarps][kVtab];
    double sx[kBlockWarps][8];
    double sdc[kJMax][16];
};

// All threads of a block; sm.sdc must hold the shared-parameter steps of the pass (backsub_load_dc).
template <int MODEL>
__device__ __forceinline__ void backsub_block(const Dev& d, BacksubSmem& sm, int j, int grp, int cur) {
    constexpr int L = ModelTraits<MODEL>::L;
    const int tid = tid_now(), lane = tid & 31, warp = tid >> 5;
    const double* P = d.Pbuf + (long long)cur * d.K;
    const int nb = d.bs_groups;
    const int tb = trial_buf(cur, j);
    STAMP(d, 2, 0);
    for (int v = grp * kBlockWarps + warp; v < d.M; v += nb * kBlockWarps) {
        double* Pt = d.Pbuf + (long long)tb * d.K;
        const double* vt = d.vtab + ((long long)j * d.M + v) * kVtab;
        __syncwarp();
        sm.srow[warp][lane] = vt[lane];
        sm.srow[warp][lane + 32] = vt[lane + 32];
        if (lane + 64 < kVtab) sm.srow[warp][lane + 64] = vt[lane + 64];
        __syncwarp();
        const double* Lc = &sm.srow[warp][0];          // packed lower factor
        const double* Zm = &sm.srow[warp][21];         // Z columns, 6 each
        if (lane < 6) {
            double s = Zm[L * 6 + lane];
#pragma unroll
            for (int p = 0; p < L; ++p) s = fma(-Zm[p * 6 + lane], sm.sdc[j][p], s);
            sm.sx[warp][lane] = s;
        }
        __syncwarp();
        if (lane == 0) {
            double x[6];
#pragma unroll
            for (int i = 0; i < 6; ++i) x[i] = sm.sx[warp][i];
#pragma unroll
            for (int i = 5; i >= 0; --i) {
                double s = x[i];
#pragma unroll
                for (int k = i + 1; k < 6; ++k) s = fma(-Lc[k * (k + 1) / 2 + i], x[k], s);
                x[i] = __dmul_rn(s, Lc[i * (i + 1) / 2 + i]);      // the diagonal slot holds 1 / L_ii (explicit roundings,
                                                                   // as in backsub_thread: the two must agree bit for bit)
            }
            double e[6];
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                const long long idx = L + 6ll * v + i;
                e[i] = __dadd_rn(P[idx], x[i]);
                Pt[idx] = e[i];
                if (j == 0) d.delta[idx] = x[i];
            }
            make_view(e, sm.sview[warp]);
        }
        __syncwarp();
        if (lane < kViewDoubles)
            reinterpret_cast<double*>(d.views + (long long)tb * d.M + v)[lane] =
                reinterpret_cast<const double*>(&sm.sview[warp])[lane];
    }
    STAMP(d, 2, 1);
}

template <int MODEL>
__device__ __forceinline__ void backsub_load_dc(const Dev& d, BacksubSmem& sm, int J) {
    constexpr int L = ModelTraits<MODEL>::L;
    const int tid = tid_now();
    if (tid < kJMax * 16) {
        const int j = tid >> 4, i = tid & 15;
        if (j < J && i < L) sm.sdc[j][i] = d.dcs[j * 16 + i];
    }
    __syncthreads();
}

template <int MODEL>
__global__ void __launch_bounds__(kBlockThreads) k_backsub(Dev d) {
    LmState* st = d.st;
    if (st->done) return;
    __shared__ BacksubSmem sm;
    const int cur = st->cur, J = st->J;
    backsub_load_dc<MODEL>(d, sm, J);
    const int nb = d.bs_groups;
    const int j = blockIdx.x / nb, grp = blockIdx.x - j * nb;
    if (j >= J) return;
    backsub_block<MODEL>(d, sm, j, grp, cur);
}

// ---- parity / debug kernels (not on the timed path) ------------------------------------------
// Per point: y (N x 2), r (2N), Jc (2N x L), Jv (2N x 6); any pointer may be null.
template <int MODEL>
__global__ void k_debug_point(Dev d, double* y, double* r, double* Jc, double* Jv) {
    using T = ModelTraits<MODEL>;
    constexpr int L = T::L, ND = T::ND;
    const int lane = threadIdx.x & 31;
    const int t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (t >= d.ev_pieces) return;
    const Piece task = d.ev_piece[t];
    const int cur = d.st->cur;
    const double* P = d.Pbuf + (long long)cur * d.K;
    double c[L];
#pragma unroll
    for (int i = 0; i < L; ++i) c[i] = P[i];
    const ViewC w = d.views[(long long)cur * d.M + task.view];
    for (int i = lane; i < task.count; i += 32) {
        const long long n = task.start + i;
        Pt p;
        load_point(d, n, p);
        PointOut<MODEL> o;
        point_eval<MODEL, true, true>(c, w, p.X, p.Y, p.Z, p.U, p.V, o);
        if (y) { y[2 * n] = o.u; y[2 * n + 1] = o.v; }
        if (r) { r[2 * n] = o.ru; r[2 * n + 1] = o.rv; }
        if (Jc) {
            double* ju = Jc + 2 * n * L;
            double* jv = ju + L;
            ju[0] = o.xd; ju[1] = 0; ju[2] = o.yd; ju[3] = 1; ju[4] = 0;
            jv[0] = 0; jv[1] = o.yd; jv[2] = 0; jv[3] = 0; jv[4] = 1;
#pragma unroll
            for (int a = 0; a < ND; ++a) { ju[5 + a] = o.ku[a]; jv[5 + a] = o.kv[a]; }
        }
        if (Jv) {
#pragma unroll
            for (int j = 0; j < 6; ++j) { Jv[12 * n + j] = o.ju[j]; Jv[12 * n + 6 + j] = o.jv[j]; }
        }
    }
}

// Unpacks the partials into the reference's block layout: A L x L, B M x L x 6,
// D M x 6 x 6, g K.  red[] must already hold the column sums (k_schur ran).
template <int MODEL>
__global__ void k_debug_unpack(Dev d, double* A, double* B, double* D, double* g) {
    using T = ModelTraits<MODEL>;
    constexpr int L = T::L, ND = T::ND;
    using SA = SlotsA<ND>;
    using S = SlotsBD<L>;
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v == 0) {
        for (int i = 0; i < L; ++i) {
            for (int j = 0; j < L; ++j) {
                const int sl = a_slot<ND>(i < j ? i : j, i < j ? j : i);
                A[i * L + j] = sl >= 0 ? d.red[kRedA + sl] : 0.0;
            }
            g[i] = d.red[kRedA + SA::G + i];
        }
    }
    if (v >= d.M) return;
    const int r0 = d.view_piece[v], r1 = d.view_piece[v + 1];
    auto slot_sum = [&](int slot) {
        double s = 0.0;
        for (int r = r0; r < r1; ++r) s += d.partBD[(long long)r * kNBD + slot];
        return s;
    };
    for (int p = 0; p < L; ++p)
        for (int j = 0; j < 6; ++j) B[((long long)v * L + p) * 6 + j] = slot_sum(S::B + p * 6 + j);
    for (int a = 0; a < 6; ++a)
        for (int b = a; b < 6; ++b) {
            const double s = slot_sum(S::dd(a, b));
            D[(long long)v * 36 + a * 6 + b] = s;
            D[(long long)v * 36 + b * 6 + a] = s;
        }
    for (int j = 0; j < 6; ++j) g[L + 6ll * v + j] = slot_sum(S::GV + j);
}

// FP64 FMA pipe micro-benchmark: 8 independent chains per thread
__global__ void k_fp64_peak(double* out, int iters, double a, double b) {
    double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; ++i) {
        x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
        x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}

__global__ void k_fill(double* p, long long n, double v) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) p[i] = v;
}

// AoS (model N x 3, sensor N x 2, as numpy holds them) -> the SoA arrays the sweeps stream
__global__ void k_repack(const double* __restrict__ model, const double* __restrict__ sensor, long long n,
                         double* X, double* Y, double* Z, double* U, double* V) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        X[i] = model[3 * i]; Y[i] = model[3 * i + 1]; Z[i] = model[3 * i + 2];
        U[i] = sensor[2 * i]; V[i] = sensor[2 * i + 1];
    }
}

}  // namespace calib
