// All-reduce of the reduced system over NVLink peer memory, inside the kernels that produce it.
//
// When the views of one calibration are sharded over the GPUs of a node the
// only exchange per LM iteration is the reduced system (72 + 96 doubles per candidate of the damping
// ladder) and the trial costs.  A library collective costs a launch (plus a 1-thread kernel to consume it) per exchange and
// keeps the WHILE-graph loop off the table; here the last thread block of k_schur / k_cost does it:
//   push   thread t stores value t of its rank's contribution into a slot of EVERY peer's inbox,
//          through the NVLink mapping of the peer's memory, as ONE 16-byte store of two 64-bit
//          elements {seq:low word, seq:high word}: the data carry their own flag.  A vector access
//          is a set of scalar accesses of its element type under the PTX memory model, and an
//          aligned 64-bit scalar is single-copy atomic -- so each half arrives with ITS sequence
//          number or not at all, and there is no fence and no separate flag round trip (a .v4.u32
//          store, the first version, only promised that per 32-bit word);
//   wait   thread t polls element t of every peer's slot in its LOCAL inbox until both sequence
//          words match (a first version with fence.sys + a release flag cost 5-6 us per exchange);
//   sum    in RANK ORDER (its own contribution from shared memory), so all ranks hold
//          bit-identical sums and take identical LM decisions.
// One NVLink store latency per exchange, no extra launch, no host involvement: the body of the
// whole-loop graph stays free of library nodes.  Slots are double-buffered by sequence parity;
// channel 0 carries the reduced system, channel 1 the trial cost.  A rank that never shows up is
// detected by a timeout (the run ends with CALIB_E_COMM instead of hanging the GPU).
#pragma once
#include <cuda_runtime.h>

namespace calib {

constexpr int kPeerMaxRanks = 8;
constexpr int kPeerSlot = 464;            // elements per slot (>= kRedN, the reduced systems of a full damping ladder)
constexpr int kPeerChannels = 2;
constexpr size_t kPeerInboxElems = (size_t)kPeerChannels * 2 * kPeerMaxRanks * kPeerSlot;   // uint4 each
constexpr size_t kPeerInboxBytes = kPeerInboxElems * 16;
constexpr unsigned long long kPeerTimeoutNs = 20ull * 1000 * 1000 * 1000;

struct PeerDev {
    int nranks, rank;                         // nranks <= 1: no exchange
    uint4* box[kPeerMaxRanks];                // box[p]: inbox of rank p as mapped into this process
    unsigned int* seq;                        // local, kPeerChannels counters (exchanges done so far)
};

// one inbox element: two 64-bit halves, each {sequence number : 32 bits of the double}
__device__ __forceinline__ void st_volatile_v2(uint4* p, unsigned long long lo, unsigned long long hi) {
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(lo), "l"(hi) : "memory");
}
__device__ __forceinline__ void ld_volatile_v2(const uint4* p, unsigned long long& lo, unsigned long long& hi) {
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(lo), "=l"(hi) : "l"(p) : "memory");
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

__device__ __forceinline__ size_t peer_slot_off(int ch, int par, int src) {
    return ((size_t)(ch * 2 + par) * kPeerMaxRanks + src) * kPeerSlot;
}

// Called by ALL threads of one block.  val[0..n) in shared memory is this rank's
// contribution and is replaced by the sum over ranks.  Returns false on timeout.
__device__ __forceinline__ bool peer_allreduce(const PeerDev& pc, int ch, double* val, int n, int* s_flag) {
    const int tid = threadIdx.x;
    unsigned int seq = *reinterpret_cast<volatile unsigned int*>(pc.seq + ch) + 1u;
    if (seq == 0u) seq = 1u;                  // 0 is what a fresh inbox holds
    const int par = (int)(seq & 1u);
    if (tid == 0) *s_flag = 1;
    __syncthreads();
    for (int e = tid; e < n; e += (int)blockDim.x) {
        const double mine = val[e];
        const unsigned long long bits = (unsigned long long)__double_as_longlong(mine);
        const unsigned long long tag = (unsigned long long)seq << 32;
        const unsigned long long lo = tag | (bits & 0xffffffffull), hi = tag | (bits >> 32);
        const size_t off = peer_slot_off(ch, par, pc.rank) + e;
        for (int p = 0; p < pc.nranks; ++p)
            if (p != pc.rank) st_volatile_v2(pc.box[p] + off, lo, hi);
    }
    for (int e = tid; e < n; e += (int)blockDim.x) {
        const double mine = val[e];
        double s = 0.0;
        for (int r = 0; r < pc.nranks; ++r) {
            double v = mine;
            if (r != pc.rank) {
                const uint4* src = pc.box[pc.rank] + peer_slot_off(ch, par, r) + e;
                unsigned long long lo, hi;
                ld_volatile_v2(src, lo, hi);
                if ((unsigned)(lo >> 32) != seq || (unsigned)(hi >> 32) != seq) {
                    const unsigned long long t0 = global_timer_ns();
                    unsigned spins = 0;
                    do {
                        ld_volatile_v2(src, lo, hi);
                        if ((++spins & 1023u) == 0 && global_timer_ns() - t0 > kPeerTimeoutNs) { *s_flag = 0; break; }
                    } while ((unsigned)(lo >> 32) != seq || (unsigned)(hi >> 32) != seq);
                }
                v = __longlong_as_double((long long)((hi << 32) | (lo & 0xffffffffull)));
            }
            s = __dadd_rn(s, v);
        }
        val[e] = s;
    }
    __syncthreads();
    if (tid == 0) pc.seq[ch] = seq;
    return *s_flag != 0;
}

}  // namespace calib
