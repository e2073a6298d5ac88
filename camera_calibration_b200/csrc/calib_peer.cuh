// All-reduce of the reduced system over NVLink peer memory, inside the kernels that produce it.
//
// When the views of one calibration are sharded over the GPUs of a node the
// only exchange per LM iteration is the reduced system (72 + 96 doubles per candidate of the damping
// ladder) and the trial costs.  A library collective costs a launch (plus a 1-thread kernel to consume it) per exchange and
// keeps the WHILE-graph loop off the table; here the last thread block of k_schur / k_cost does it:
//   push   thread t stores value t of its rank's contribution into a slot of EVERY peer's inbox,
//          through the NVLink mapping of the peer's memory, as ONE 16-byte store
//          {low word, seq, high word, seq}: the data carry their own flag (each 8-byte half is
//          written atomically), so there is no fence and no separate flag round trip;
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

__device__ __forceinline__ void st_volatile_v4(uint4* p, const uint4& v) {
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ uint4 ld_volatile_v4(const uint4* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p)
                 : "memory");
    return v;
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
        const uint4 pk = make_uint4((unsigned)bits, seq, (unsigned)(bits >> 32), seq);
        const size_t off = peer_slot_off(ch, par, pc.rank) + e;
        for (int p = 0; p < pc.nranks; ++p)
            if (p != pc.rank) st_volatile_v4(pc.box[p] + off, pk);
    }
    for (int e = tid; e < n; e += (int)blockDim.x) {
        const double mine = val[e];
        double s = 0.0;
        for (int r = 0; r < pc.nranks; ++r) {
            double v = mine;
            if (r != pc.rank) {
                const uint4* src = pc.box[pc.rank] + peer_slot_off(ch, par, r) + e;
                uint4 g = ld_volatile_v4(src);
                if (g.y != seq || g.w != seq) {
                    const unsigned long long t0 = global_timer_ns();
                    unsigned spins = 0;
                    do {
                        g = ld_volatile_v4(src);
                        if ((++spins & 1023u) == 0 && global_timer_ns() - t0 > kPeerTimeoutNs) { *s_flag = 0; break; }
                    } while (g.y != seq || g.w != seq);
                }
                v = __longlong_as_double((long long)(((unsigned long long)g.z << 32) | g.x));
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
