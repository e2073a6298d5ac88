// Streaming of the detections into shared memory with the bulk-copy engine (TMA, 1-D
// cp.async.bulk; UBLKCP in SASS) and mbarrier completion -- one private ring of D stages per warp.
//
// A persistent warp owns a contiguous chunk [cs, ce) of the point arrays, cs a multiple of 32, cut
// into WINDOWS of 32*WS points.  For window k one elected lane arms the stage's mbarrier with the
// byte count and issues five bulk copies (X, Y, Z, U, V: 256*WS bytes each, 256-B aligned) D windows
// ahead of the lanes that consume them, so no load is waited for inside the FP64 stream and no
// register is spent on prefetch.  (Copies of one 32-point step, 256 B, were tried first: the copy
// engine's request rate, not bandwidth, then limits the sweep; WS = 4 gives 1 KB requests.)
// Lane l reads slots l, l+32, ... of a stage only: the data are lane-private, a __syncwarp() is all
// that orders "every lane has read the last step of stage s" before the stage is re-armed.  View
// boundaries (pieces) are independent of windows: a 32-point step that straddles two pieces is
// consumed twice with complementary lane masks and read from shared memory once.
#pragma once
#include "calib_kernels.cuh"

namespace calib {

template <int D, int WS> struct __align__(128) PointRing { double v[D][5][32 * WS]; };

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// One warp's sweep over its chunk.  body(point) runs for every point of the current piece owned by
// this lane; begin / end run once per piece with the warp converged (end after a __syncwarp()).
// bars: D mbarriers of this warp, used by this sweep only.
template <int D, int WS, int NV, class Begin, class Body, class End>
__device__ __forceinline__ void sweep_windows(const Dev& d, const Piece* __restrict__ pieces, const WarpRec& rec,
                                              const ViewC* const (&views)[NV], ViewC (&sview)[NV], PointRing<D, WS>& ring,
                                              unsigned long long* bars, int lane, Begin begin, Body body, End end) {
    constexpr int WP = 32 * WS;           // points per window
    const int p0 = rec.p0, p1 = rec.p1;
    if (p0 >= p1) return;
    const int cs = rec.first.start, ce = rec.end;
    const int nwin = (ce - cs + WP - 1) / WP;
    auto produce = [&](int k) {           // lane 0 only
        const int s = k % D;
        const long long i0 = (long long)cs + (long long)WP * k;
        int cnt = ce - (int)i0;
        cnt = cnt > WP ? WP : ((cnt + 1) & ~1);           // 16-byte granules; the arrays are padded
        const unsigned bytes = (unsigned)cnt * 8u;
        mbar_expect_tx(&bars[s], 5u * bytes);
        bulk_g2s(&ring.v[s][0][0], d.X + i0, bytes, &bars[s]);
        bulk_g2s(&ring.v[s][1][0], d.Y + i0, bytes, &bars[s]);
        bulk_g2s(&ring.v[s][2][0], d.Z + i0, bytes, &bars[s]);
        bulk_g2s(&ring.v[s][3][0], d.U + i0, bytes, &bars[s]);
        bulk_g2s(&ring.v[s][4][0], d.V + i0, bytes, &bars[s]);
    };
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < D; ++s) mbar_init(&bars[s], 1);
        mbar_fence_init();
#pragma unroll
        for (int k = 0; k < D; ++k)
            if (k < nwin) produce(k);
    }
    __syncwarp();
    Piece cur = rec.first;
    double vn[NV];
#pragma unroll
    for (int j = 0; j < NV; ++j) {
        vn[j] = 0.0;
        if (lane < kViewDoubles) vn[j] = reinterpret_cast<const double*>(views[j] + cur.view)[lane];
    }
    int loaded = -1, ready = -1;          // last 32-point step held in registers; last window waited for
    Pt pt;
    pt.X = pt.Y = pt.Z = pt.U = pt.V = 0.0;
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
        const int t0 = (cur.start - cs) >> 5, t1 = (cur.start + cur.count - 1 - cs) >> 5;
        for (int t = t0; t <= t1; ++t) {
            if (t != loaded) {
                const int k = t / WS, j = t - k * WS, s = k % D;
                if (k != ready) { mbar_wait(&bars[s], (unsigned)(k / D) & 1u); ready = k; }
                const int o = 32 * j + lane;
                pt.X = ring.v[s][0][o]; pt.Y = ring.v[s][1][o]; pt.Z = ring.v[s][2][o];
                pt.U = ring.v[s][3][o]; pt.V = ring.v[s][4][o];
                loaded = t;
                if (j == WS - 1) {                      // last step of the window: the stage is free
                    __syncwarp();
                    if (lane == 0 && k + D < nwin) produce(k + D);
                }
            }
            const int idx = cs + 32 * t + lane;
            if (idx >= cur.start && idx < cur.start + cur.count) body(pt);
        }
        __syncwarp();
        end(cur, pi);
        cur = nxt;
    }
}

}  // namespace calib
