// C-ABI of libcalib_b200.so (see include/calib_b200.h).  Host side: device memory, the task
// list, the per-iteration CUDA graph, the LM driver loop, NCCL plumbing for sharded views.
#include <cuda_runtime.h>
#include <dlfcn.h>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif
#include <nccl.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <mutex>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "../../include/calib_b200.h"
#include "calib_solve.cuh"
#ifdef CALIB_EXPERIMENT_PERSISTENT
#include "experiments/calib_persistent.cuh"     // loop mode 3: measured, not faster, not in the default build
#endif
#include "calib_batch.cuh"
#include "calib_homography.cuh"
#include "calib_linear.cuh"

using namespace calib;

static thread_local std::string g_last_error;

// All device memory of the library comes from ONE stream-ordered pool per device that the library owns
// (not the device's default pool: its settings belong to the process, e.g. to a co-resident torch).  The pool
// keeps what is freed -- creating the next problem costs no cudaMalloc -- until calib_trim(device) hands the
// unused part back to the driver.
static std::mutex g_pool_mu;
static cudaMemPool_t g_pool[64] = {};
static cudaError_t lib_pool(int device, cudaMemPool_t* out) {
    std::lock_guard<std::mutex> lk(g_pool_mu);
    if (device < 0 || device >= 64) return cudaErrorInvalidDevice;
    if (!g_pool[device]) {
        cudaMemPoolProps props;
        memset(&props, 0, sizeof props);
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = device;
        cudaMemPool_t pool = nullptr;
        cudaError_t e = cudaMemPoolCreate(&pool, &props);
        if (e != cudaSuccess) return e;
        uint64_t thr = UINT64_MAX;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
        g_pool[device] = pool;
    }
    *out = g_pool[device];
    return cudaSuccess;
}
// stream-ordered allocation on `device` (the device of `stream`)
static cudaError_t lib_malloc_async(void** p, size_t bytes, cudaStream_t stream, int device) {
    cudaMemPool_t pool;
    cudaError_t e = lib_pool(device, &pool);
    if (e != cudaSuccess) return e;
    return cudaMallocFromPoolAsync(p, bytes, pool, stream);
}

extern "C" int calib_trim(int device) {
    cudaMemPool_t pool = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_pool_mu);
        if (device < 0 || device >= 64) return CALIB_E_ARG;
        pool = g_pool[device];
    }
    if (!pool) return CALIB_OK;                  // nothing was ever allocated on that device
    return cudaMemPoolTrimTo(pool, 0) == cudaSuccess ? CALIB_OK : CALIB_E_CUDA;
}

// Scratch buffers of one call: stream-ordered allocations from the library's pool, so a call costs no
// cudaMalloc / cudaFree.
template <typename T>
static cudaError_t dalloc(std::vector<void*>& pool, T** p, size_t count, cudaStream_t stream, int device) {
    void* q = nullptr;
    cudaError_t e = lib_malloc_async(&q, std::max<size_t>(count, 1) * sizeof(T), stream, device);
    if (e == cudaSuccess) { pool.push_back(q); *p = (T*)q; }
    return e;
}

static void free_pool(std::vector<void*>& pool, cudaStream_t stream) {
    for (void* p : pool) cudaFreeAsync(p, stream);
    pool.clear();
}

// Inbox of this rank for the in-kernel all-reduce (calib_peer.cuh) and the peers' inboxes as mapped
// into this process (CUDA IPC).  Owned by the handle that exported it, borrowed by calib_comm_share.
struct PeerComm {
    int nranks = 1, rank = 0, device = 0;
    uint4* inbox = nullptr;                  // cudaMalloc: kPeerInboxBytes, then the counters
    void* mapped[kPeerMaxRanks] = {};        // peers' inboxes (cudaIpcOpenMemHandle; plain pointers in one process)
    bool ipc = true;                         // false: single-process handle, peers reached through peer access
    ~PeerComm() {
        cudaSetDevice(device);
        if (ipc) for (void* m : mapped) if (m) cudaIpcCloseMemHandle(m);
        if (inbox) cudaFree(inbox);
    }
    bool mapped_ok() const {
        for (int r = 0; r < nranks; ++r) if (r != rank && !mapped[r]) return false;
        return true;
    }
};

struct calib_handle {
    int model = 0, L = 0, device = 0, n_sm = 148;
    int64_t M = 0, N = 0, K = 0;
    std::vector<int64_t> view_offsets;
    std::vector<char> meta_host;           // staging of the partition tables (build_tasks)
    int seg = 0;
    Dev dev{};
    cudaStream_t stream = nullptr;
    cudaGraphExec_t graph = nullptr;       // one iteration (host-driven loop)
    cudaGraphExec_t loop_graph = nullptr;  // WHILE node around the iteration (device-driven loop)
    struct CachedLoop* loop = nullptr;     // its source graph, nodes and key: parked for the next handle of that shape
    int loop_mode = 0;                     // 0 auto, 1 host-driven, 2 device WHILE graph, 3 persistent kernel
    int persist_blocks = 0;                // co-resident blocks of the persistent kernel on this device (0: not asked yet)
    int cost_tma = -1;                     // -1 auto (by size), 0 plain loads, 1 bulk-copy ring
    LmState* h_state = nullptr;   // pinned
    double* h_pinned = nullptr;   // pinned scratch (kRedN + 8 doubles)
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    void* arena_fixed = nullptr;  // detections, parameters, per-view tables
    void* arena_part = nullptr;   // everything that depends on the partition of the points
    int trace_cap = 0;
    double* trace_own = nullptr;  // only when a run asks for more iterations than the arena holds
    int cur = 0;
    std::string err;
    double last_ms = 0;
    int64_t last_launches = 0, last_evals = 0;
    bool profiling = false;
    double ktime[5] = {0, 0, 0, 0, 0};
    int64_t kcount[5] = {0, 0, 0, 0, 0};
    std::vector<std::pair<int, std::pair<cudaEvent_t, cudaEvent_t>>> prof_events;
    ncclComm_t comm = nullptr;
    bool comm_owned = false;
    int nranks = 1, rank = 0;
    std::shared_ptr<PeerComm> peer;        // in-kernel exchange over NVLink peer memory
    int ladder = kJMax;                    // cap of the damping ladder (1 = one candidate per pass)
    int ladder_start = 0;                  // candidates of the pass that follows an accepted step: 1, 2, 0 = adaptive
    int transport = 0;                     // 0 auto (peer memory when attached, else NCCL), 1 NCCL, 2 peer memory
    double* l2buf = nullptr;
    size_t l2bytes = 0;
    bool want_ptrace = false;              // calib_set_param_trace
    bool run_pending = false;              // lm_start enqueued a whole-loop launch that lm_finish has not collected yet
    int64_t pending_count = 0;
};

#define CUDA_TRY(h, call)                                                                     \
    do {                                                                                      \
        cudaError_t e__ = (call);                                                             \
        if (e__ != cudaSuccess) {                                                             \
            char b__[512];                                                                    \
            snprintf(b__, sizeof b__, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), \
                     __FILE__, __LINE__);                                                     \
            if (h) (h)->err = b__;                                                            \
            g_last_error = b__;                                                               \
            return CALIB_E_CUDA;                                                              \
        }                                                                                     \
    } while (0)

static int fail(calib_handle* h, int code, const char* msg) {
    if (h) h->err = msg;
    g_last_error = msg;
    return code;
}

// ---- NCCL through dlopen (no link-time dependency; torch's copy is reused when loaded) ------
namespace {
struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
} g_nccl;

bool nccl_load() {
    if (g_nccl.ok) return true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (g_nccl.lib) break;
    }
    if (!g_nccl.lib) return false;
    g_nccl.GetUniqueId = (decltype(g_nccl.GetUniqueId))dlsym(g_nccl.lib, "ncclGetUniqueId");
    g_nccl.CommInitRank = (decltype(g_nccl.CommInitRank))dlsym(g_nccl.lib, "ncclCommInitRank");
    g_nccl.AllReduce = (decltype(g_nccl.AllReduce))dlsym(g_nccl.lib, "ncclAllReduce");
    g_nccl.CommDestroy = (decltype(g_nccl.CommDestroy))dlsym(g_nccl.lib, "ncclCommDestroy");
    g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString))dlsym(g_nccl.lib, "ncclGetErrorString");
    g_nccl.ok = g_nccl.GetUniqueId && g_nccl.CommInitRank && g_nccl.AllReduce && g_nccl.CommDestroy;
    return g_nccl.ok;
}

// Bump allocator over one stream-ordered device allocation (lib_malloc_async: the library's pool keeps the
// memory of a destroyed handle, so creating the next problem does not pay for cudaMalloc again).
struct Arena {
    size_t bytes = 0;
    char* base = nullptr;
    template <typename T> size_t reserve(size_t count) {
        const size_t off = bytes;
        bytes += (std::max<size_t>(count, 1) * sizeof(T) + 255) & ~(size_t)255;
        return off;
    }
    template <typename T> T* at(size_t off) const { return reinterpret_cast<T*>(base + off); }
};

}  // namespace

// Streams, events and the page-locked state block are expensive to create (cudaMallocHost takes
// about a millisecond); a destroyed handle parks them here for the next calib_create on that GPU.
namespace {
struct DevCtx { int device; cudaStream_t stream; cudaEvent_t ev0, ev1; LmState* h_state; };
std::mutex g_ctx_mu;
std::vector<DevCtx> g_ctx_free;

bool debug_timing() {
    static const bool on = getenv("CALIB_DEBUG_TIMING") != nullptr;
    return on;
}
struct ScopeTimer {
    const char* what;
    std::chrono::steady_clock::time_point t0;
    explicit ScopeTimer(const char* w) : what(w), t0(std::chrono::steady_clock::now()) {}
    ~ScopeTimer() {
        if (!debug_timing()) return;
        const double us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count();
        fprintf(stderr, "[calib] %-28s %9.1f us\n", what, us);
    }
};
}  // namespace

static int ctx_acquire(calib_handle* h) {
    {
        std::lock_guard<std::mutex> lk(g_ctx_mu);
        for (size_t i = 0; i < g_ctx_free.size(); ++i)
            if (g_ctx_free[i].device == h->device) {
                const DevCtx c = g_ctx_free[i];
                g_ctx_free.erase(g_ctx_free.begin() + i);
                h->stream = c.stream; h->ev0 = c.ev0; h->ev1 = c.ev1; h->h_state = c.h_state;
                h->h_pinned = reinterpret_cast<double*>(h->h_state + 1);
                return CALIB_OK;
            }
    }
    CUDA_TRY(h, cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    CUDA_TRY(h, cudaEventCreate(&h->ev0));
    CUDA_TRY(h, cudaEventCreate(&h->ev1));
    CUDA_TRY(h, cudaMallocHost((void**)&h->h_state, sizeof(LmState) + (kRedN + 8) * sizeof(double)));
    h->h_pinned = reinterpret_cast<double*>(h->h_state + 1);
    return CALIB_OK;
}

static void ctx_release(calib_handle* h) {
    if (h->stream && h->ev0 && h->ev1 && h->h_state) {
        std::lock_guard<std::mutex> lk(g_ctx_mu);
        g_ctx_free.push_back(DevCtx{h->device, h->stream, h->ev0, h->ev1, h->h_state});
    } else {
        if (h->h_state) cudaFreeHost(h->h_state);
        if (h->ev0) cudaEventDestroy(h->ev0);
        if (h->ev1) cudaEventDestroy(h->ev1);
        if (h->stream) cudaStreamDestroy(h->stream);
    }
    h->stream = nullptr; h->ev0 = h->ev1 = nullptr; h->h_state = nullptr; h->h_pinned = nullptr;
}

// An instantiated WHILE-loop graph is expensive to make (capture + instantiate: 0.3-0.5 ms of host time, and
// the driver serialises it across the host threads of a process) but depends on the problem only through the
// kernels' by-value argument (Dev: pointers, sizes) and the grid sizes.  So a handle that lets go of its loop
// graph parks it here -- source graph, executable graph, the kernel nodes of the WHILE body -- and the next
// handle on that device whose iteration has the same shape (model, grid sizes, cost-pass variant) takes it
// and re-points the nodes at its own Dev with cudaGraphExecKernelNodeSetParams: a few microseconds
// (profiles/r2q_graph_patch.txt: 4 us against 540 us).  At most kLoopCacheMax graphs wait per process; the
// oldest beyond that are destroyed when the next handle is created (under its upload).
struct LoopKey {
    int device, model, ev_blocks, co_blocks, sc_groups, bs_groups, tma;
    bool operator==(const LoopKey& o) const {
        return device == o.device && model == o.model && ev_blocks == o.ev_blocks && co_blocks == o.co_blocks &&
               sc_groups == o.sc_groups && bs_groups == o.bs_groups && tma == o.tma;
    }
};
struct CachedLoop {
    LoopKey key;
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    cudaGraphConditionalHandle cond{};
    struct Node { cudaGraphNode_t node; cudaKernelNodeParams kp; int n_int; int iarg[2]; };
    std::vector<Node> nodes;               // kernel nodes of the WHILE body
};
static std::mutex g_loop_mu;
static std::vector<CachedLoop*> g_loop_cache;
constexpr size_t kLoopCacheMax = 16;

static void destroy_loop(CachedLoop* c) {
    if (!c) return;
    int cur = 0;
    cudaGetDevice(&cur);
    if (cur != c->key.device) cudaSetDevice(c->key.device);
    if (c->exec) cudaGraphExecDestroy(c->exec);
    if (c->graph) cudaGraphDestroy(c->graph);
    if (cur != c->key.device) cudaSetDevice(cur);
    delete c;
}

// hands the handle's loop graph to the cache (destroy = true: a handle that holds an NCCL communicator must
// tear its graphs down at once; such graphs are never cached anyway)
static void park_loop(calib_handle* h, bool destroy = false) {
    CachedLoop* c = h->loop;
    h->loop = nullptr; h->loop_graph = nullptr;
    if (!c) return;
    if (destroy) { destroy_loop(c); return; }
    std::lock_guard<std::mutex> lk(g_loop_mu);
    g_loop_cache.push_back(c);
}

// destroys the oldest parked graphs beyond `keep` (called where host time is free: under an upload)
static void trim_loop_cache(size_t keep) {
    std::vector<CachedLoop*> doomed;
    {
        std::lock_guard<std::mutex> lk(g_loop_mu);
        while (g_loop_cache.size() > keep) { doomed.push_back(g_loop_cache.front()); g_loop_cache.erase(g_loop_cache.begin()); }
    }
    for (CachedLoop* c : doomed) destroy_loop(c);
}

static void drop_graph(calib_handle* h) {
    if (h->graph) { cudaGraphExecDestroy(h->graph); h->graph = nullptr; }
    park_loop(h);
}

static void empty_trash(int, size_t) { trim_loop_cache(kLoopCacheMax); }

static void park_graphs(calib_handle* h) {
    if (h->graph) { cudaGraphExecDestroy(h->graph); h->graph = nullptr; }
    park_loop(h);
    trim_loop_cache(2 * kLoopCacheMax);
    cudaSetDevice(h->device);
}

// Cuts the N points into n_warps contiguous, near-equal chunks (one per persistent warp) and each
// chunk into pieces at the view boundaries.
static void cut_points(const std::vector<int64_t>& offs, int64_t M, int64_t N, int n_warps,
                       std::vector<Piece>& pieces, std::vector<int>& chunk, std::vector<int>& chunk_end) {
    pieces.clear();
    chunk.assign((size_t)n_warps + 1, 0);
    chunk_end.assign((size_t)n_warps, 0);
    // boundaries on multiples of 32 points: the windows the bulk copies fetch are 256-byte aligned
    auto bound = [&](int k) -> int64_t {
        if (k >= n_warps) return N;
        const int64_t b = (int64_t)((__int128)N * k / n_warps);
        return std::min<int64_t>(N, (b + 16) / 32 * 32);
    };
    int64_t v = 0;
    for (int k = 0; k < n_warps; ++k) {
        int64_t s = bound(k), e = bound(k + 1);
        chunk[k] = (int)pieces.size();
        chunk_end[k] = (int)e;
        while (s < e) {
            while (v < M && offs[v + 1] <= s) ++v;
            const int64_t cnt = std::min(e, offs[v + 1]) - s;
            pieces.push_back(Piece{(int)v, (int)s, (int)cnt, 0});
            s += cnt;
        }
    }
    chunk[n_warps] = (int)pieces.size();
}

// (Re)builds the partition of the points and everything sized by it.  seg = points per warp
// (0 = automatic: one chunk per resident warp of the persistent grids).
static int build_tasks(calib_handle* h, int seg, bool sync_now = true) {
    drop_graph(h);
    if (h->arena_part) { cudaFreeAsync(h->arena_part, h->stream); h->arena_part = nullptr; }
    h->seg = seg;
    const int64_t N = h->N, M = h->M;
    int ev_warps, co_warps;
    if (seg > 0) {
        ev_warps = co_warps = (int)std::max<int64_t>(1, (N + seg - 1) / seg);
    } else {
        const int64_t by_size = std::max<int64_t>(1, (N + 63) / 64);
        ev_warps = (int)std::min<int64_t>((int64_t)h->n_sm * kEvalPairs, by_size);      // one chunk per producer / consumer pair
        co_warps = (int)std::min<int64_t>((int64_t)h->n_sm * kBlockWarps * 2, by_size);
    }
    std::vector<Piece> evp, cop;
    std::vector<int> evc, coc, eve, coe, vp((size_t)M + 1, 0);
    cut_points(h->view_offsets, M, N, ev_warps, evp, evc, eve);
    cut_points(h->view_offsets, M, N, co_warps, cop, coc, coe);
    {   // piece range of every view (pieces are in view order)
        size_t p = 0;
        for (int64_t v = 0; v < M; ++v) {
            vp[v] = (int)p;
            while (p < evp.size() && evp[p].view == v) ++p;
        }
        vp[M] = (int)evp.size();
    }
    Dev& d = h->dev;
    d.ev_warps = ev_warps; d.ev_blocks = (ev_warps + kEvalPairs - 1) / kEvalPairs; d.ev_pieces = (int)evp.size();
    d.co_warps = co_warps; d.co_blocks = (co_warps + kBlockWarps - 1) / kBlockWarps;
    // k_schur / k_backsub: one warp per view; the grid repeats the view groups once per candidate
    // of the damping ladder (blocks of candidates a pass does not have leave at once)
    d.sc_groups = (int)std::max<int64_t>(1, std::min<int64_t>((M + kBlockWarps - 1) / kBlockWarps, 2 * h->n_sm));
    d.bs_groups = d.sc_groups;
    Arena a;
    auto records = [](const std::vector<Piece>& pc, const std::vector<int>& ch, const std::vector<int>& ce) {
        std::vector<WarpRec> r(ch.size() - 1);
        for (size_t k = 0; k + 1 < ch.size(); ++k) {
            r[k].p0 = ch[k]; r[k].p1 = ch[k + 1]; r[k].end = ce[k]; r[k].pad1 = 0;
            r[k].first = ch[k] < ch[k + 1] ? pc[ch[k]] : Piece{0, 0, 0, 0};
        }
        return r;
    };
    const std::vector<WarpRec> evr = records(evp, evc, eve), cor = records(cop, coc, coe);
    const size_t o_evp = a.reserve<Piece>(evp.size()), o_evc = a.reserve<WarpRec>(evr.size());
    const size_t o_vp = a.reserve<int>(vp.size());
    const size_t o_cop = a.reserve<Piece>(cop.size()), o_coc = a.reserve<WarpRec>(cor.size());
    const size_t meta_bytes = a.bytes;
    const size_t o_pA = a.reserve<double>((size_t)d.ev_blocks * kNA), o_pE = a.reserve<double>(d.ev_blocks);
    const size_t o_pBD = a.reserve<double>((size_t)std::max(d.ev_pieces, 1) * kNBD);
    const size_t o_pW = a.reserve<double>((size_t)d.sc_groups * kJMax * kNBD), o_pC = a.reserve<double>((size_t)d.co_blocks * kJMax);
    CUDA_TRY(h, lib_malloc_async((void**)&h->arena_part, a.bytes, h->stream, h->device));
    a.base = (char*)h->arena_part;
    // one upload for the five index tables
    std::vector<char>& meta = h->meta_host;      // stays alive until the stream has been synchronised
    meta.assign(meta_bytes, 0);
    memcpy(meta.data() + o_evp, evp.data(), evp.size() * sizeof(Piece));
    memcpy(meta.data() + o_evc, evr.data(), evr.size() * sizeof(WarpRec));
    memcpy(meta.data() + o_vp, vp.data(), vp.size() * sizeof(int));
    memcpy(meta.data() + o_cop, cop.data(), cop.size() * sizeof(Piece));
    memcpy(meta.data() + o_coc, cor.data(), cor.size() * sizeof(WarpRec));
    CUDA_TRY(h, cudaMemcpyAsync(a.base, meta.data(), meta_bytes, cudaMemcpyHostToDevice, h->stream));
    if (sync_now) CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    d.ev_piece = a.at<Piece>(o_evp); d.ev_rec = a.at<WarpRec>(o_evc); d.view_piece = a.at<int>(o_vp);
    d.co_piece = a.at<Piece>(o_cop); d.co_rec = a.at<WarpRec>(o_coc);
    d.partA = a.at<double>(o_pA); d.partE = a.at<double>(o_pE); d.partBD = a.at<double>(o_pBD);
    d.partW = a.at<double>(o_pW); d.partC = a.at<double>(o_pC);
    return CALIB_OK;
}

static int ensure_trace(calib_handle* h, int iters) {
    if (iters <= h->trace_cap) return CALIB_OK;
    drop_graph(h);
    double* t = nullptr;
    CUDA_TRY(h, cudaMalloc((void**)&t, (size_t)iters * 20 * sizeof(double)));
    if (h->trace_own) cudaFree(h->trace_own);
    h->trace_own = t;
    h->dev.trace = t;
    h->dev.ptrace = t + 4 * (size_t)iters;
    h->trace_cap = iters;
    return CALIB_OK;
}

namespace {
int default_loop_mode(const calib_handle* h);
bool cost_uses_tma(const calib_handle* h);
int build_loop_graph(calib_handle* h);
void apply_transport(calib_handle* h);
std::mutex g_default_mu;
const calib_handle* g_default_comm[64] = {nullptr};   // per device: handle whose communicator new handles borrow
}

extern "C" const char* calib_last_error(const calib_t* h) {
    return h ? h->err.c_str() : g_last_error.c_str();
}

// ---- detections held as per-view arrays: gather and upload, pipelined ---------------------------
namespace {

constexpr size_t kGatherChunkMax = (size_t)4 << 20; // bytes per staged piece, at most
constexpr int kGatherThreads = 8;                   // at most
constexpr int kGatherSlots = 4;                     // staging slots per thread, at most
constexpr size_t kGatherChunkDefault = (size_t)1 << 20;
constexpr int kGatherSlotsDefault = 2;
constexpr int kGatherDefault = 6;                   // with streaming stores 4 threads are gather-bound, 8 wait for PCIe; both ~1.6 ms at 62 MB

struct GatherPool {                                 // per device; created on first use, kept
    std::mutex mu;
    char* pinned = nullptr;
    cudaStream_t stream[kGatherThreads] = {};
    cudaEvent_t slot_free[kGatherThreads][kGatherSlots] = {};
    cudaEvent_t done[kGatherThreads] = {};
    bool ok = false;
};
GatherPool g_gather[64];

cudaError_t gather_pool_init(GatherPool& g) {
    if (g.ok) return cudaSuccess;
    cudaError_t e = cudaHostAlloc((void**)&g.pinned, kGatherThreads * kGatherSlots * kGatherChunkMax, cudaHostAllocPortable);
    if (e != cudaSuccess) return e;
    for (int t = 0; t < kGatherThreads && e == cudaSuccess; ++t) {
        e = cudaStreamCreateWithFlags(&g.stream[t], cudaStreamNonBlocking);
        for (int k = 0; k < kGatherSlots && e == cudaSuccess; ++k) e = cudaEventCreateWithFlags(&g.slot_free[t][k], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&g.done[t], cudaEventDisableTiming);
    }
    g.ok = e == cudaSuccess;
    return e;
}

// memcpy with streaming (non-temporal) stores where the target has SSE2.  The staging slot is read next by
// the GPU's DMA engine, not by this core: ordinary stores leave the lines dirty in this core's caches and every
// DMA read then has to snoop them out, streaming stores put them in memory (measured, profiles/README.md).
// The caller fences (_mm_sfence) before handing the slot to the copy.
inline void copy_streaming(char* d, const char* s, size_t n) {
#if defined(__SSE2__)
    size_t head = (16 - ((uintptr_t)d & 15)) & 15;
    if (head > n) head = n;
    if (head) { memcpy(d, s, head); d += head; s += head; n -= head; }
    for (; n >= 64; d += 64, s += 64, n -= 64) {
        const __m128i a = _mm_loadu_si128((const __m128i*)s), b = _mm_loadu_si128((const __m128i*)(s + 16));
        const __m128i c = _mm_loadu_si128((const __m128i*)(s + 32)), e = _mm_loadu_si128((const __m128i*)(s + 48));
        _mm_stream_si128((__m128i*)d, a); _mm_stream_si128((__m128i*)(d + 16), b);
        _mm_stream_si128((__m128i*)(d + 32), c); _mm_stream_si128((__m128i*)(d + 48), e);
    }
    for (; n >= 16; d += 16, s += 16, n -= 16) _mm_stream_si128((__m128i*)d, _mm_loadu_si128((const __m128i*)s));
    if (n) memcpy(d, s, n);
#else
    memcpy(d, s, n);
#endif
}

// Copies `bytes` bytes starting at byte `lo` of the concatenation of the views' arrays (stride bytes per
// point) into dst.
void gather_range(char* dst, const double* const* ptrs, const int64_t* offs, int64_t M, size_t stride, size_t lo, size_t bytes,
                  bool streaming) {
    const size_t hi = lo + bytes;
    int64_t i = (std::upper_bound(offs, offs + M + 1, (int64_t)(lo / stride)) - offs) - 1;
    for (; i < M && (size_t)offs[i] * stride < hi; ++i) {
        const size_t a = std::max((size_t)offs[i] * stride, lo), b = std::min((size_t)offs[i + 1] * stride, hi);
        if (b <= a) continue;
        const char* src = (const char*)ptrs[i] + (a - (size_t)offs[i] * stride);
        if (streaming) copy_streaming(dst + (a - lo), src, b - a);
        else memcpy(dst + (a - lo), src, b - a);
    }
#if defined(__SSE2__)
    if (streaming) _mm_sfence();
#endif
}

// model points -> aos_model (N x 3), sensor points -> aos_sensor (N x 2), both device and ready to be
// written.  Touches none of the caller's streams (it runs beside a stream capture of the creating thread):
// the copies go out on the pool's streams, *n_streams of them; upload_views_join orders a stream after them.
cudaError_t upload_views(int device, const int64_t* offs, int64_t M, const double* const* vm,
                         const double* const* vs, double* aos_model, double* aos_sensor, int* n_streams) {
    GatherPool& g = g_gather[device < 64 ? device : 63];
    std::lock_guard<std::mutex> lk(g.mu);
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = gather_pool_init(g);
    if (e != cudaSuccess) return e;
    const size_t N = (size_t)offs[M];
    const size_t bytes_m = N * 24, bytes_s = N * 16;
    static const size_t kGatherChunk = [] {
        const char* ev = getenv("CALIB_GATHER_CHUNK_KB");
        return ev ? std::max<size_t>(64 << 10, std::min(kGatherChunkMax, (size_t)atoi(ev) << 10)) : kGatherChunkDefault;
    }();
    static const int slots = [] {
        const char* ev = getenv("CALIB_GATHER_SLOTS");
        return ev ? std::max(1, std::min(kGatherSlots, atoi(ev))) : kGatherSlotsDefault;
    }();
    static const bool one_stream = getenv("CALIB_GATHER_ONE_STREAM") != nullptr;
    static const bool streaming = [] { const char* ev = getenv("CALIB_GATHER_NT"); return ev ? atoi(ev) != 0 : true; }();
    const size_t cm = (bytes_m + kGatherChunk - 1) / kGatherChunk, cs = (bytes_s + kGatherChunk - 1) / kGatherChunk;
    const size_t nchunks = cm + cs;
    int T = kGatherDefault;
    if (const char* ev = getenv("CALIB_GATHER_THREADS")) T = std::max(1, std::min(kGatherThreads, atoi(ev)));
    T = (int)std::min<size_t>({(size_t)T, nchunks, (size_t)std::max(1u, std::thread::hardware_concurrency())});
    *n_streams = std::max(T, 1);
    std::atomic<size_t> next{0};
    std::atomic<int> err{(int)cudaSuccess};
    ScopeTimer tm_all("upload_views (helper thread)");
    const bool timing = debug_timing();
    double t_gather[kGatherThreads] = {}, t_wait[kGatherThreads] = {}, t_issue[kGatherThreads] = {}, t_start[kGatherThreads] = {};
    const auto t_zero = std::chrono::steady_clock::now();
    auto us_since = [](std::chrono::steady_clock::time_point a) {
        return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - a).count();
    };
    auto work = [&](int t) {
        if (timing) t_start[t] = us_since(t_zero);
        cudaError_t r = cudaSetDevice(device);
        for (int k = 0; r == cudaSuccess; ++k) {
            const size_t c = next.fetch_add(1);
            if (c >= nchunks) break;
            const bool is_m = c < cm;
            const size_t lo = (is_m ? c : c - cm) * kGatherChunk;
            const size_t n = std::min(kGatherChunk, (is_m ? bytes_m : bytes_s) - lo);
            const int sl = k % slots;
            char* slot = g.pinned + ((size_t)t * kGatherSlots + sl) * kGatherChunkMax;
            auto ta = std::chrono::steady_clock::now();
            r = cudaEventSynchronize(g.slot_free[t][sl]);   // the copy that last used this slot (any call) has left it
            if (r != cudaSuccess) break;
            if (timing) { t_wait[t] += us_since(ta); ta = std::chrono::steady_clock::now(); }
            gather_range(slot, is_m ? vm : vs, offs, M, is_m ? 24 : 16, lo, n, streaming);
            if (timing) { t_gather[t] += us_since(ta); ta = std::chrono::steady_clock::now(); }
            cudaStream_t cs_ = g.stream[one_stream ? 0 : t];
            r = cudaMemcpyAsync((char*)(is_m ? aos_model : aos_sensor) + lo, slot, n, cudaMemcpyHostToDevice, cs_);
            if (r == cudaSuccess) r = cudaEventRecord(g.slot_free[t][sl], cs_);
            if (timing) t_issue[t] += us_since(ta);
        }
        if (r == cudaSuccess) r = cudaEventRecord(g.done[t], g.stream[one_stream ? 0 : t]);
        if (r != cudaSuccess) err.store((int)r);
    };
    if (T <= 1) work(0);
    else {
        std::vector<std::thread> th;
        std::vector<int> inline_t;                 // a thread that could not be started: its share runs here
        for (int t = 1; t < T; ++t) {
            try { th.emplace_back(work, t); } catch (...) { inline_t.push_back(t); }
        }
        work(0);
        for (int t : inline_t) work(t);
        for (auto& x : th) x.join();
    }
    if (timing)
        for (int t = 0; t < std::max(T, 1); ++t)
            fprintf(stderr, "[calib]   gather thread %2d: start +%.0f us, wait %.0f, gather %.0f, issue %.0f us\n", t, t_start[t],
                    t_wait[t], t_gather[t], t_issue[t]);
    if (err.load() != (int)cudaSuccess) {
        for (int t = 0; t < std::max(T, 1); ++t) cudaStreamSynchronize(g.stream[t]);
        return (cudaError_t)err.load();
    }
    return cudaSuccess;   // copies still in flight: the slot events guard the staging against the next call
}

cudaError_t upload_views_join(int device, cudaStream_t stream, int n_streams) {
    GatherPool& g = g_gather[device < 64 ? device : 63];
    for (int t = 0; t < n_streams; ++t) {
        cudaError_t e = cudaStreamWaitEvent(stream, g.done[t], 0);
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

}  // namespace

// Per-device function attributes (cudaFuncSetAttribute applies to the CURRENT device only): the
// per-warp rings of the cost pass exceed the 48 KB default.  Done once per device, under a lock.
static cudaError_t device_func_attributes(int device) {
    static std::mutex mu;
    static bool done[64] = {false};
    std::lock_guard<std::mutex> lk(mu);
    if (device >= 0 && device < 64 && done[device]) return cudaSuccess;
    cudaError_t e = cudaSuccess;
    auto set = [&](auto kern, size_t bytes) {
        if (e == cudaSuccess) e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    };
    set(k_cost<0, false, true>, kCostDynSmem); set(k_cost<0, true, true>, kCostDynSmem);
    set(k_cost<1, false, true>, kCostDynSmem); set(k_cost<1, true, true>, kCostDynSmem);
    set(k_eval<0>, kEvalDynSmem); set(k_eval<1>, kEvalDynSmem);
#ifdef CALIB_EXPERIMENT_PERSISTENT
    set(k_lm_persistent<0, true>, kPersistDynSmem); set(k_lm_persistent<1, true>, kPersistDynSmem);
    set(k_lm_persistent<0, false>, kPersistDynSmem); set(k_lm_persistent<1, false>, kPersistDynSmem);
#endif
    if (e == cudaSuccess && device >= 0 && device < 64) done[device] = true;
    return e;
}

static int create_impl(calib_t** out, int model, int64_t M, const int64_t* view_offsets, const double* model_pts,
                       const double* sensor_pts, const double* const* vmodel, const double* const* vsensor, int device_id) {
    if (!out) return fail(nullptr, CALIB_E_ARG, "calib_create: out is NULL");
    *out = nullptr;
    if (model != CALIB_MODEL_RADTAN && model != CALIB_MODEL_FISHEYE)
        return fail(nullptr, CALIB_E_ARG, "calib_create: unknown distortion model");
    if (M <= 0 || !view_offsets) return fail(nullptr, CALIB_E_ARG, "calib_create: need M >= 1 views");
    if (view_offsets[0] != 0) return fail(nullptr, CALIB_E_ARG, "calib_create: view_offsets[0] != 0");
    for (int64_t v = 0; v < M; ++v)
        if (view_offsets[v + 1] < view_offsets[v])
            return fail(nullptr, CALIB_E_ARG, "calib_create: view_offsets not non-decreasing");
    const int64_t N = view_offsets[M];
    if (N > 0 && !vmodel && (!model_pts || !sensor_pts)) return fail(nullptr, CALIB_E_ARG, "calib_create: NULL points");
    if (N >= (1ll << 31) - 64 || M >= (1ll << 27)) return fail(nullptr, CALIB_E_ARG, "calib_create: too large");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0)
        return fail(nullptr, CALIB_E_CUDA, "calib_create: no CUDA device (this library has no CPU fallback)");
    if (device_id < 0 || device_id >= ndev) return fail(nullptr, CALIB_E_ARG, "calib_create: bad device_id");
    ScopeTimer tm_("calib_create");
    calib_handle* h = new calib_handle();
    h->model = model; h->L = model == 0 ? 10 : 9; h->device = device_id;
    h->M = M; h->N = N; h->K = h->L + 6 * M;
    h->view_offsets.assign(view_offsets, view_offsets + M + 1);
    auto bail = [&](int rc) { calib_destroy(h); return rc; };
    std::thread uploader;
    cudaError_t upload_rc = cudaSuccess;
    int upload_streams = 0;
    bool upload_inline = false;
    std::function<cudaError_t()> repack;
#define TRY_(call) do { int rc__ = [&]() -> int { CUDA_TRY(h, call); return 0; }(); if (rc__) { g_last_error = h->err; return bail(rc__); } } while (0)
    TRY_(cudaSetDevice(device_id));
    TRY_(device_func_attributes(device_id));
    TRY_(cudaDeviceGetAttribute(&h->n_sm, cudaDevAttrMultiProcessorCount, device_id));
    { int rc = ctx_acquire(h); if (rc) { g_last_error = h->err; return bail(rc); } }
    Dev& d = h->dev;
    d.model = model; d.L = h->L; d.M = (int)M; d.N = N; d.K = h->K;
    {
        const size_t n = ((size_t)std::max<int64_t>(N, 1) + 31) / 32 * 32;   // array stride: 256-byte aligned
        Arena a;
        const size_t o_soa = a.reserve<double>(5 * n + 64);
        const size_t o_P = a.reserve<double>((size_t)kNBuf * h->K), o_views = a.reserve<ViewC>((size_t)kNBuf * M);
        const size_t o_red = a.reserve<double>(kRedN), o_vtab = a.reserve<double>((size_t)kJMax * M * kVtab);
        const size_t o_et = a.reserve<double>(8), o_delta = a.reserve<double>((size_t)h->K);
        const size_t o_dcs = a.reserve<double>(kJMax * 16), o_redA = a.reserve<double>(kRedW0);
#ifdef CALIB_STAMPS
        const size_t o_stamps = a.reserve<unsigned long long>(kStampWords);
#endif
        const size_t o_st = a.reserve<LmState>(1);
        constexpr int kTraceIters = 256;
        const size_t o_trace = a.reserve<double>(20 * kTraceIters);     // trace rows, then the parameter trace
        const size_t zero_from = o_P, zero_bytes = a.bytes - o_P;
        const size_t o_aos = a.reserve<double>(5 * n);     // upload staging, AoS as the caller holds it
        TRY_(lib_malloc_async((void**)&h->arena_fixed, a.bytes, h->stream, h->device));
        a.base = (char*)h->arena_fixed;
        double* soa = a.at<double>(o_soa);
        d.X = soa; d.Y = soa + n; d.Z = soa + 2 * n; d.U = soa + 3 * n; d.V = soa + 4 * n;
        d.Pbuf = a.at<double>(o_P); d.views = a.at<ViewC>(o_views); d.red = a.at<double>(o_red);
        d.vtab = a.at<double>(o_vtab); d.etrial = a.at<double>(o_et); d.delta = a.at<double>(o_delta);
        d.dcs = a.at<double>(o_dcs); d.redA = a.at<double>(o_redA);
#ifdef CALIB_STAMPS
        d.stamps = a.at<unsigned long long>(o_stamps);
#endif
        d.st = a.at<LmState>(o_st);
        d.trace = a.at<double>(o_trace);
        d.ptrace = d.trace + 4 * kTraceIters;
        h->trace_cap = kTraceIters;
        TRY_(cudaMemsetAsync(a.base + zero_from, 0, zero_bytes, h->stream));
        if (N > 0) {
            double* aos = a.at<double>(o_aos);
            repack = [=]() {
                const int nb = (int)std::min<int64_t>((N + 255) / 256, (int64_t)h->n_sm * 8);
                k_repack<<<nb, 256, 0, h->stream>>>(aos, aos + 3 * n, N, soa, soa + n, soa + 2 * n, soa + 3 * n, soa + 4 * n);
                return cudaGetLastError();
            };
            if (vmodel) {
                // gather + copies run on helper threads (the copies on their own streams, after everything
                // enqueued so far) while this thread builds the partition and the loop graph below
                TRY_(cudaStreamSynchronize(h->stream));   // allocation + memset done: the helper streams need no edge
                auto run_upload = [=, &upload_rc, &upload_streams]() {
                    upload_rc = upload_views(device_id, h->view_offsets.data(), M, vmodel, vsensor, aos, aos + 3 * n, &upload_streams);
                };
                try { uploader = std::thread(run_upload); } catch (...) { run_upload(); upload_inline = true; }
            } else {
                TRY_(cudaMemcpyAsync(aos, model_pts, (size_t)N * 3 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
                TRY_(cudaMemcpyAsync(aos + 3 * n, sensor_pts, (size_t)N * 2 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
                TRY_(repack());
            }
        }
    }
    int rc = build_tasks(h, 0, false);
    if (rc) { if (uploader.joinable()) uploader.join(); g_last_error = h->err; return bail(rc); }
    {   // borrow the device's default communicator (calib_comm_set_default) before the graph is built
        std::lock_guard<std::mutex> lk(g_default_mu);
        const calib_handle* o = device_id < 64 ? g_default_comm[device_id] : nullptr;
        if (o) {
            h->comm = o->comm; h->comm_owned = false; h->nranks = o->nranks; h->rank = o->rank;
            h->peer = o->peer; h->transport = o->transport;
            apply_transport(h);
        }
    }
    // the default loop graph is built here, on the host, while the detections are still crossing PCIe
    // (a setting that changes the iteration drops it and lm_run rebuilds it)
    if (default_loop_mode(h) == 2 && build_loop_graph(h) != CALIB_OK) { cudaGetLastError(); drop_graph(h); }
    empty_trash(device_id, 0);      // graphs of destroyed handles: torn down while the detections cross PCIe
    if (uploader.joinable() || upload_inline) {
        if (uploader.joinable()) uploader.join();
        TRY_(upload_rc);
        TRY_(upload_views_join(device_id, h->stream, upload_streams));
        TRY_(repack());
    }
#undef TRY_
    {   // the caller's buffers are free again once the stream is idle
        cudaError_t e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) { h->err = cudaGetErrorString(e); g_last_error = h->err; return bail(CALIB_E_CUDA); }
    }
    *out = h;
    return CALIB_OK;
}

extern "C" int calib_create(calib_t** out, int model, int64_t M, const int64_t* view_offsets,
                            const double* model_pts, const double* sensor_pts, int device_id) {
    return create_impl(out, model, M, view_offsets, model_pts, sensor_pts, nullptr, nullptr, device_id);
}

extern "C" int calib_create_from_views(calib_t** out, int model, int64_t M, const int64_t* counts,
                                       const double* const* model_ptrs, const double* const* sensor_ptrs, int device_id) {
    if (out) *out = nullptr;
    if (M <= 0 || !counts || !model_ptrs || !sensor_ptrs)
        return fail(nullptr, CALIB_E_ARG, "calib_create_from_views: need M >= 1 views, counts and pointers");
    std::vector<int64_t> offs((size_t)M + 1, 0);
    for (int64_t v = 0; v < M; ++v) {
        if (counts[v] < 0) return fail(nullptr, CALIB_E_ARG, "calib_create_from_views: negative count");
        if (counts[v] > 0 && (!model_ptrs[v] || !sensor_ptrs[v]))
            return fail(nullptr, CALIB_E_ARG, "calib_create_from_views: NULL points in a non-empty view");
        offs[v + 1] = offs[v] + counts[v];
    }
    return create_impl(out, model, M, offs.data(), nullptr, nullptr, model_ptrs, sensor_ptrs, device_id);
}

extern "C" void calib_destroy(calib_t* h) {
    if (!h) return;
    ScopeTimer tm_("calib_destroy");
    {
        std::lock_guard<std::mutex> lk(g_default_mu);
        if (h->device < 64 && g_default_comm[h->device] == h) g_default_comm[h->device] = nullptr;
    }
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    // The graphs are destroyed by the next calib_create, under its upload -- unless they hold NCCL nodes:
    // ncclCommDestroy waits for every graph that captured its collectives, so those go first, now.
    if (h->comm) { if (h->graph) { cudaGraphExecDestroy(h->graph); h->graph = nullptr; } park_loop(h, true); }
    else park_graphs(h);
    if (h->comm && h->comm_owned && g_nccl.ok) g_nccl.CommDestroy(h->comm);
    for (auto& pe : h->prof_events) { cudaEventDestroy(pe.second.first); cudaEventDestroy(pe.second.second); }
    if (h->arena_part && h->stream) cudaFreeAsync(h->arena_part, h->stream);
    if (h->arena_fixed && h->stream) cudaFreeAsync(h->arena_fixed, h->stream);
    if (h->trace_own) cudaFree(h->trace_own);
    if (h->l2buf) cudaFree(h->l2buf);
    if (h->stream) cudaStreamSynchronize(h->stream);
    ctx_release(h);
    delete h;
}

extern "C" int64_t calib_num_params(const calib_t* h) { return h ? h->K : -1; }
extern "C" int64_t calib_num_points(const calib_t* h) { return h ? h->N : -1; }

extern "C" int calib_set_segment(calib_t* h, int points_per_task) {
    if (!h) return fail(nullptr, CALIB_E_ARG, "NULL handle");
    if (points_per_task < 0) return fail(h, CALIB_E_ARG, "calib_set_segment: negative");
    CUDA_TRY(h, cudaSetDevice(h->device));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return build_tasks(h, points_per_task);
}

extern "C" int calib_set_loop_mode(calib_t* h, int mode) {
    if (!h) return fail(nullptr, CALIB_E_ARG, "NULL handle");
#ifdef CALIB_EXPERIMENT_PERSISTENT
    if (mode < 0 || mode > 3) return fail(h, CALIB_E_ARG, "calib_set_loop_mode: 0 auto, 1 host, 2 device graph, 3 persistent kernel");
#else
    if (mode < 0 || mode > 2) return fail(h, CALIB_E_ARG, "calib_set_loop_mode: 0 auto, 1 host, 2 device");
#endif
    h->loop_mode = mode;
    return CALIB_OK;
}

extern "C" int calib_set_ladder(calib_t* h, int max_candidates) {
    if (!h) return fail(nullptr, CALIB_E_ARG, "NULL handle");
    if (max_candidates != 1 && max_candidates != 2 && max_candidates != kJMax)
        return fail(h, CALIB_E_ARG, "calib_set_ladder: 1, 2 or 4 candidates");
    h->ladder = max_candidates;
    return CALIB_OK;
}

extern "C" int calib_set_ladder_start(calib_t* h, int candidates) {
    if (!h) return fail(nullptr, CALIB_E_ARG, "NULL handle");
    if (candidates < 0 || candidates > 2) return fail(h, CALIB_E_ARG, "calib_set_ladder_start: 0 (adaptive), 1 or 2 candidates");
    h->ladder_start = candidates;
    return CALIB_OK;
}

extern "C" int calib_set_cost_staging(calib_t* h, int mode) {
    if (!h) return fail(nullptr, CALIB_E_ARG, "NULL handle");
    if (mode < -1 || mode > 1) return fail(h, CALIB_E_ARG, "calib_set_cost_staging: -1 auto, 0 loads, 1 bulk copies");
    if (h->cost_tma != mode) drop_graph(h);
    h->cost_tma = mode;
    return CALIB_OK;
}

// Host-side helper of the Python mirror: concatenates the per-view arrays of allDetections (a list of
// separately allocated numpy arrays) into one buffer with a few threads -- numpy's concatenate is one
// thread and, at 60 MB, costs more than the whole refinement.
extern "C" int calib_host_gather(void* dst, const void* const* srcs, const int64_t* bytes, int64_t n, int nthreads) {
    if (!dst || !srcs || !bytes || n < 0) return fail(nullptr, CALIB_E_ARG, "calib_host_gather: bad args");
    std::vector<size_t> off((size_t)n + 1, 0);
    for (int64_t i = 0; i < n; ++i) {
        if (bytes[i] < 0 || (bytes[i] > 0 && !srcs[i])) return fail(nullptr, CALIB_E_ARG, "calib_host_gather: bad piece");
        off[i + 1] = off[i] + (size_t)bytes[i];
    }
    const size_t total = off[(size_t)n];
    int T = nthreads > 0 ? nthreads : (int)std::min<unsigned>(8u, std::max(1u, std::thread::hardware_concurrency()));
    if (total < ((size_t)4 << 20)) T = 1;
    auto work = [&](int t) {
        const size_t lo = total * (size_t)t / (size_t)T, hi = total * (size_t)(t + 1) / (size_t)T;   // byte range of thread t
        size_t i = (size_t)(std::upper_bound(off.begin(), off.end(), lo) - off.begin()) - 1;
        for (; i < (size_t)n && off[i] < hi; ++i) {
            const size_t a = std::max(off[i], lo), b = std::min(off[i + 1], hi);
            if (b > a) copy_streaming((char*)dst + a, (const char*)srcs[i] + (a - off[i]), b - a);   // dst is staging for a DMA
        }
#if defined(__SSE2__)
        _mm_sfence();
#endif
    };
    if (T == 1) { work(0); return CALIB_OK; }
    std::vector<std::thread> th;
    std::vector<int> inline_t;
    for (int t = 1; t < T; ++t) {
        try { th.emplace_back(work, t); } catch (...) { inline_t.push_back(t); }
    }
    work(0);
    for (int t : inline_t) work(t);
    for (auto& x : th) x.join();
    return CALIB_OK;
}

extern "C" void* calib_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaMallocHost(&p, std::max<size_t>(bytes, 1)) != cudaSuccess) {
        g_last_error = "calib_host_alloc: cudaMallocHost failed";
        return nullptr;
    }
    return p;
}
extern "C" void calib_host_free(void* p) { if (p) cudaFreeHost(p); }

// ---- launches ---------------------------------------------------------------------------------
namespace {

enum { KEVAL = 0, KSCHUR = 1, KBACK = 2, KCOST = 3, KSOLVE = 4 };

// Kernel launch of the iteration on the handle's stream.
template <typename... KArgs, typename... Args>
cudaError_t launch_iter(calib_handle* h, void (*kern)(KArgs...), int grid, int block, size_t smem, Args... args) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3((unsigned)block);
    cfg.dynamicSmemBytes = smem; cfg.stream = h->stream;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

struct Launcher {
    calib_handle* h;
    bool prof;
    int64_t count = 0;
    void begin(cudaEvent_t* a, cudaEvent_t* b) {
        if (!prof) return;
        cudaEventCreate(a); cudaEventCreate(b);
        cudaEventRecord(*a, h->stream);
    }
    void end(int which, cudaEvent_t a, cudaEvent_t b) {
        if (!prof) return;
        cudaEventRecord(b, h->stream);
        h->prof_events.push_back({which, {a, b}});
    }
};

template <int MODEL>
void enqueue_eval(calib_handle* h, Launcher& ln) {
    const Dev& d = h->dev;
    cudaEvent_t a, b;
    ln.begin(&a, &b);
    launch_iter(h, k_eval<MODEL>, d.ev_blocks, kBlockThreads, kEvalDynSmem, d);
    ln.end(KEVAL, a, b);
    ln.count += 1;
}

template <int MODEL>
void enqueue_schur(calib_handle* h, Launcher& ln, int fuse_solve) {
    const Dev& d = h->dev;
    cudaEvent_t a, b;
    ln.begin(&a, &b);
    launch_iter(h, k_schur<MODEL>, d.sc_groups * kJMax, kBlockThreads, 0, d, fuse_solve);
    ln.end(KSCHUR, a, b);
    ln.count += 1;
}

template <int MODEL>
void enqueue_solve(calib_handle* h, Launcher& ln) {
    cudaEvent_t a, b;
    ln.begin(&a, &b);
    k_solve<MODEL><<<1, 32 * kJMax, 0, h->stream>>>(h->dev);
    ln.end(KSOLVE, a, b);
    ln.count += 1;
}

template <int MODEL>
void enqueue_backsub(calib_handle* h, Launcher& ln) {
    const Dev& d = h->dev;
    cudaEvent_t a, b;
    ln.begin(&a, &b);
    launch_iter(h, k_backsub<MODEL>, d.bs_groups * kJMax, kBlockThreads, 0, d);
    ln.end(KBACK, a, b);
    ln.count += 1;
}

template <int MODEL>
void enqueue_cost(calib_handle* h, Launcher& ln, int fuse_accept) {
    const Dev& d = h->dev;
    cudaEvent_t a, b;
    ln.begin(&a, &b);
    const bool tma = cost_uses_tma(h);
    if (tma) {
        if (d.use_cond) launch_iter(h, k_cost<MODEL, true, true>, d.co_blocks, kBlockThreads, kCostDynSmem, d, fuse_accept);
        else launch_iter(h, k_cost<MODEL, false, true>, d.co_blocks, kBlockThreads, kCostDynSmem, d, fuse_accept);
    } else {
        if (d.use_cond) launch_iter(h, k_cost<MODEL, true, false>, d.co_blocks, kBlockThreads, 0, d, fuse_accept);
        else launch_iter(h, k_cost<MODEL, false, false>, d.co_blocks, kBlockThreads, 0, d, fuse_accept);
    }
    ln.end(KCOST, a, b);
    ln.count += 1;
}

bool peer_active(const calib_handle* h);
bool nccl_active(const calib_handle* h);
// Loop mode 0 (automatic): the WHILE graph; NCCL collectives cannot sit inside a WHILE body, that transport is
// driven by the host.  (A persistent cooperative kernel with grid barriers between the phases was built and
// measured, profiles/r2b_*, r2c_*: a grid barrier costs what a kernel boundary inside the WHILE graph costs,
// 1.5-2 us, and the phases run no faster, so it is not in the default build: csrc/experiments/.)
// CALIB_LOOP_MODE overrides.
int default_loop_mode(const calib_handle* h) {
    static const int env = [] { const char* e = getenv("CALIB_LOOP_MODE"); return e ? atoi(e) : 0; }();
    if (nccl_active(h)) return 1;
    return env >= 1 && env <= 2 ? env : 2;
}

bool peer_active(const calib_handle* h) {
    return h->peer && h->peer->nranks > 1 && h->peer->mapped_ok() && h->transport != 1;
}
bool nccl_active(const calib_handle* h) { return h->comm && !peer_active(h); }

// one LM iteration: src/calibrate.py:144-168.  Sharded views: the reduced system and the trial cost
// are summed over the ranks either inside k_schur / k_cost through peer memory (4 launches, as on one
// GPU) or by two NCCL all-reduces, each followed by a small kernel that consumes the sum.
template <int MODEL>
int enqueue_iteration(calib_handle* h, Launcher& ln) {
    const Dev& d = h->dev;
    const bool nccl = nccl_active(h);
    enqueue_eval<MODEL>(h, ln);
    enqueue_schur<MODEL>(h, ln, nccl ? 0 : 1);
    if (nccl) {
        if (g_nccl.AllReduce(d.red, d.red, kRedN, ncclDouble, ncclSum, h->comm, h->stream) != ncclSuccess)
            return fail(h, CALIB_E_COMM, "ncclAllReduce(reduced system) failed");
        enqueue_solve<MODEL>(h, ln);
    }
    enqueue_backsub<MODEL>(h, ln);
    enqueue_cost<MODEL>(h, ln, nccl ? 0 : 1);
    if (nccl) {
        if (g_nccl.AllReduce(d.etrial, d.etrial, kJMax, ncclDouble, ncclSum, h->comm, h->stream) != ncclSuccess)
            return fail(h, CALIB_E_COMM, "ncclAllReduce(trial cost) failed");
        if (d.use_cond) k_accept<true><<<1, 1, 0, h->stream>>>(d);
        else k_accept<false><<<1, 1, 0, h->stream>>>(d);
        ln.count += 1;
    }
    return CALIB_OK;
}

int enqueue_iteration_any(calib_handle* h, Launcher& ln) {
    return h->model == 0 ? enqueue_iteration<0>(h, ln) : enqueue_iteration<1>(h, ln);
}

int kernels_per_iteration(const calib_handle* h) { return nccl_active(h) ? 6 : 4; }

// fills / clears the device view of the peer communicator according to the chosen transport
void apply_transport(calib_handle* h) {
    drop_graph(h);
    PeerDev& p = h->dev.peer;
    memset(&p, 0, sizeof p);
    if (!peer_active(h)) return;
    const PeerComm& c = *h->peer;
    p.nranks = c.nranks; p.rank = c.rank;
    for (int r = 0; r < c.nranks; ++r) p.box[r] = r == c.rank ? c.inbox : (uint4*)c.mapped[r];
    p.seq = reinterpret_cast<unsigned int*>(c.inbox + kPeerInboxElems);
}

// host-driven loop: one iteration as a graph
int build_graph(calib_handle* h) {
    if (h->graph) return CALIB_OK;
    cudaGraph_t g = nullptr;
    h->dev.use_cond = 0;
    CUDA_TRY(h, cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
    Launcher ln{h, false};
    int rc = enqueue_iteration_any(h, ln);
    cudaError_t e = cudaStreamEndCapture(h->stream, &g);
    if (rc) { if (g) cudaGraphDestroy(g); return rc; }
    CUDA_TRY(h, e);
    e = cudaGraphInstantiate(&h->graph, g, 0);
    cudaGraphDestroy(g);
    CUDA_TRY(h, e);
    return CALIB_OK;
}

// device-driven loop: a WHILE conditional node whose body is the iteration; the accept step sets
// the condition (cudaGraphSetConditional), so a whole refinement is ONE graph launch and the host
// reads the state back once.
int build_loop_graph(calib_handle* h) {
    if (h->loop_graph) return CALIB_OK;
    ScopeTimer tm_("build_loop_graph");
    const Dev& d0 = h->dev;
    const LoopKey key{h->device, h->model, d0.ev_blocks, d0.co_blocks, d0.sc_groups, d0.bs_groups,
                      cost_uses_tma(h) ? 1 : 0};
    const bool cacheable = !nccl_active(h);
    if (cacheable) {
        CachedLoop* c = nullptr;
        {
            std::lock_guard<std::mutex> lk(g_loop_mu);
            for (size_t i = g_loop_cache.size(); i-- > 0;)
                if (g_loop_cache[i]->key == key) { c = g_loop_cache[i]; g_loop_cache.erase(g_loop_cache.begin() + i); break; }
        }
        if (c) {
            // same shape: re-point the kernel nodes of the WHILE body at this handle's Dev
            h->dev.cond = c->cond;
            Dev dcopy = h->dev;
            dcopy.use_cond = 1;
            cudaError_t e = cudaSuccess;
            for (CachedLoop::Node& n : c->nodes) {
                void* args[3] = {&dcopy, &n.iarg[0], &n.iarg[1]};
                cudaKernelNodeParams kp = n.kp;
                kp.kernelParams = args;
                kp.extra = nullptr;
                e = cudaGraphExecKernelNodeSetParams(c->exec, n.node, &kp);
                if (e != cudaSuccess) break;
            }
            if (e == cudaSuccess) {
                h->loop = c; h->loop_graph = c->exec;
                return CALIB_OK;
            }
            cudaGetLastError();
            destroy_loop(c);             // could not be re-pointed: build a fresh one below
        }
    }
    cudaGraph_t g = nullptr;
    CUDA_TRY(h, cudaGraphCreate(&g, 0));
    cudaGraphConditionalHandle handle;
    cudaError_t e = cudaGraphConditionalHandleCreate(&handle, g, 1, cudaGraphCondAssignDefault);
    if (e != cudaSuccess) { cudaGraphDestroy(g); CUDA_TRY(h, e); }
    cudaGraphNodeParams np = {cudaGraphNodeTypeConditional};
    np.type = cudaGraphNodeTypeConditional;
    np.conditional.handle = handle;
    np.conditional.type = cudaGraphCondTypeWhile;
    np.conditional.size = 1;
    cudaGraphNode_t node;
    e = cudaGraphAddNode(&node, g, nullptr, 0, &np);
    if (e != cudaSuccess) { cudaGraphDestroy(g); CUDA_TRY(h, e); }
    cudaGraph_t body = np.conditional.phGraph_out[0];
    h->dev.cond = handle;
    h->dev.use_cond = 1;
    e = cudaStreamBeginCaptureToGraph(h->stream, body, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal);
    if (e != cudaSuccess) { h->dev.use_cond = 0; cudaGraphDestroy(g); CUDA_TRY(h, e); }
    Launcher ln{h, false};
    int rc = enqueue_iteration_any(h, ln);
    cudaGraph_t got = nullptr;
    e = cudaStreamEndCapture(h->stream, &got);
    h->dev.use_cond = 0;
    if (rc) { cudaGraphDestroy(g); return rc; }
    if (e != cudaSuccess) { cudaGraphDestroy(g); CUDA_TRY(h, e); }
    cudaGraphExec_t exec = nullptr;
    e = cudaGraphInstantiate(&exec, g, 0);
    if (e != cudaSuccess) { cudaGraphDestroy(g); CUDA_TRY(h, e); }
    CachedLoop* c = new CachedLoop();
    c->key = key; c->graph = g; c->exec = exec; c->cond = handle;
    if (cacheable) {
        // the kernel nodes of the body with their launch geometry and trailing int arguments (k_schur: fuse_solve,
        // k_cost: fuse_accept); a graph whose nodes cannot be read is simply not reusable
        size_t n = 0;
        bool ok = cudaGraphGetNodes(body, nullptr, &n) == cudaSuccess && n > 0;
        std::vector<cudaGraphNode_t> nodes(n);
        ok = ok && cudaGraphGetNodes(body, nodes.data(), &n) == cudaSuccess;
        for (size_t i = 0; ok && i < n; ++i) {
            cudaGraphNodeType ty;
            if (cudaGraphNodeGetType(nodes[i], &ty) != cudaSuccess) { ok = false; break; }
            if (ty != cudaGraphNodeTypeKernel) continue;
            CachedLoop::Node cn{};
            cn.node = nodes[i];
            if (cudaGraphKernelNodeGetParams(nodes[i], &cn.kp) != cudaSuccess) { ok = false; break; }
            const void* f = cn.kp.func;
            const bool is_schur = f == (const void*)k_schur<0> || f == (const void*)k_schur<1>;
            const bool is_cost = f == (const void*)k_cost<0, true, false> || f == (const void*)k_cost<0, true, true> ||
                                 f == (const void*)k_cost<1, true, false> || f == (const void*)k_cost<1, true, true>;
            cn.n_int = (is_schur || is_cost) ? 1 : 0;
            cn.iarg[0] = cn.n_int ? *(const int*)cn.kp.kernelParams[1] : 0;
            cn.iarg[1] = 0;
            cn.kp.kernelParams = nullptr;
            c->nodes.push_back(cn);
        }
        if (!ok) { cudaGetLastError(); c->nodes.clear(); c->key.device = -1; }    // never matches a lookup
    } else {
        c->key.device = -1;
    }
    h->loop = c; h->loop_graph = exec;
    return CALIB_OK;
}

int upload_to(calib_handle* h, const double* P, int buf) {
    CUDA_TRY(h, cudaMemcpyAsync(h->dev.Pbuf + (size_t)buf * h->K, P, h->K * sizeof(double),
                                cudaMemcpyHostToDevice, h->stream));
    return CALIB_OK;
}

int push_state(calib_handle* h, double lam, double lam_min, double lam_max, double err_min, int max_iters, int jmax = 1) {
    LmState s;
    memset(&s, 0, sizeof s);
    s.lam = lam; s.lam_min = lam_min; s.lam_max = lam_max; s.err_min = err_min;
    s.max_iters = max_iters; s.need_eval = 1; s.cur = h->cur; s.nranks = h->nranks;
    s.J = 1; s.jmax = std::max(1, std::min(jmax, (int)kJMax));
    s.jstart = std::min(h->ladder_start, s.jmax);      // 0: adaptive
    s.jdyn = 1;
    s.want_ptrace = h->want_ptrace ? 1 : 0;
    *h->h_state = s;
    CUDA_TRY(h, cudaMemcpyAsync(h->dev.st, h->h_state, sizeof(LmState), cudaMemcpyHostToDevice, h->stream));
    return CALIB_OK;
}

int make_views(calib_handle* h, int which_is_cur) {
    const int nb = (int)((h->M + 127) / 128);
    k_make_views<<<nb, 128, 0, h->stream>>>(h->dev, which_is_cur);
    CUDA_TRY(h, cudaGetLastError());
    return CALIB_OK;
}

int sync(calib_handle* h) {
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    CUDA_TRY(h, cudaGetLastError());
    return CALIB_OK;
}

#define RC(x) do { int rc__ = (x); if (rc__) return rc__; } while (0)

int check_handle(calib_handle* h) {
    if (!h) return fail(nullptr, CALIB_E_ARG, "NULL handle");
    CUDA_TRY(h, cudaSetDevice(h->device));
    return CALIB_OK;
}

}  // namespace

// ---- the persistent kernel (experiments/calib_persistent.cuh; -DCALIB_EXPERIMENT_PERSISTENT) ---------------------------------------------------
namespace {
bool cost_uses_tma(const calib_handle* h) {
    // bulk-copy ring when the detections cannot stay in L2 between the sweeps, plain loads otherwise
    return h->cost_tma >= 0 ? h->cost_tma != 0 : (size_t)h->N * 40 > ((size_t)96 << 20);
}

#ifdef CALIB_EXPERIMENT_PERSISTENT
#include "experiments/calib_persistent_host.inc"
#else
int launch_persistent(calib_handle* h) { return fail(h, CALIB_E_STATE, "loop mode 3 is not in this build"); }
#endif
}  // namespace

// ---- LM loop ------------------------------------------------------------------------------------
extern "C" int calib_upload_params(calib_t* h, const double* P) {
    RC(check_handle(h));
    if (!P) return fail(h, CALIB_E_ARG, "calib_upload_params: P is NULL");
    h->cur = 0;
    RC(upload_to(h, P, 0));
    return sync(h);
}

extern "C" int calib_download_params(calib_t* h, double* P) {
    RC(check_handle(h));
    if (!P) return fail(h, CALIB_E_ARG, "calib_download_params: P is NULL");
    CUDA_TRY(h, cudaMemcpyAsync(P, h->dev.Pbuf + (size_t)h->cur * h->K, h->K * sizeof(double),
                                cudaMemcpyDeviceToHost, h->stream));
    return sync(h);
}

extern "C" int calib_download_trace(calib_t* h, int iters, double* trace) {
    RC(check_handle(h));
    if (!trace || iters < 0 || iters > h->trace_cap) return fail(h, CALIB_E_ARG, "calib_download_trace: bad args");
    CUDA_TRY(h, cudaMemcpyAsync(trace, h->dev.trace, (size_t)iters * 4 * sizeof(double),
                                cudaMemcpyDeviceToHost, h->stream));
    return sync(h);
}

// The loop in two halves so that one host thread can start it on several devices before waiting for any
// (calib_multi.inc).  lm_start: everything up to and including the launch(es) of loop modes 2 / 3 and the
// read-back of the state; returns with the work in flight (run_pending).  Loop mode 1 and profiling runs
// are driven by the host and complete inside lm_start.
namespace {
int lm_finish(calib_handle* h);

int lm_start(calib_handle* h, int maxIters, double lam0, double lam_min, double lam_max, double err_min) {
    ScopeTimer tm_("lm_start");
    RC(ensure_trace(h, maxIters));
    const bool prof = h->profiling;
    // the device-driven loop needs every node of the body on this GPU; with a communicator the
    // loop is host-driven unless asked for explicitly
    if (h->loop_mode == 2 && nccl_active(h))
        return fail(h, CALIB_E_STATE, "loop mode 2: a WHILE-node body cannot hold NCCL nodes (cudaGraphInstantiate rejects "
                                      "them); use the peer-memory transport (calib_peer_attach) or loop mode 1");
    if (h->loop_mode == 3 && nccl_active(h))
        return fail(h, CALIB_E_STATE, "loop mode 3: the persistent kernel sums over the ranks through peer memory "
                                      "(calib_peer_attach); NCCL all-reduces need loop mode 1");
    const int mode = h->loop_mode ? h->loop_mode : default_loop_mode(h);
    const bool persistent = !prof && mode == 3;
    const bool device_loop = !prof && mode == 2;
    if (!prof && !persistent) {
        int rc = device_loop ? build_loop_graph(h) : build_graph(h);
        RC(rc);
    }
    for (auto& pe : h->prof_events) { cudaEventDestroy(pe.second.first); cudaEventDestroy(pe.second.second); }
    h->prof_events.clear();
    for (int i = 0; i < 5; ++i) { h->ktime[i] = 0; h->kcount[i] = 0; }
    RC(push_state(h, lam0, lam_min, lam_max, err_min, maxIters, h->ladder));
    CUDA_TRY(h, cudaEventRecord(h->ev0, h->stream));
    Launcher ln{h, prof};
    const int kpi = kernels_per_iteration(h);
    if (persistent) {
        RC(launch_persistent(h));
        ln.count = 1;
        CUDA_TRY(h, cudaMemcpyAsync(h->h_state, h->dev.st, sizeof(LmState), cudaMemcpyDeviceToHost, h->stream));
        CUDA_TRY(h, cudaEventRecord(h->ev1, h->stream));
    } else if (device_loop) {
        RC(make_views(h, 1));
        ln.count = 1;
        CUDA_TRY(h, cudaGraphLaunch(h->loop_graph, h->stream));
        CUDA_TRY(h, cudaMemcpyAsync(h->h_state, h->dev.st, sizeof(LmState), cudaMemcpyDeviceToHost, h->stream));
        CUDA_TRY(h, cudaEventRecord(h->ev1, h->stream));
    } else {
        RC(make_views(h, 1));
        ln.count = 1;
        int enq = 0;
        int batch = 2;   // iterations enqueued between two looks at the device state
        while (enq < maxIters) {
            const int n = std::min(batch, maxIters - enq);
            for (int i = 0; i < n; ++i) {
                if (prof) RC(enqueue_iteration_any(h, ln));
                else { CUDA_TRY(h, cudaGraphLaunch(h->graph, h->stream)); ln.count += kpi; }
            }
            enq += n;
            CUDA_TRY(h, cudaMemcpyAsync(h->h_state, h->dev.st, sizeof(LmState), cudaMemcpyDeviceToHost, h->stream));
            RC(sync(h));
            if (h->h_state->done) break;
            batch = std::min(batch * 2, 16);
        }
        CUDA_TRY(h, cudaEventRecord(h->ev1, h->stream));
    }
    h->pending_count = ln.count;
    h->run_pending = true;
    return CALIB_OK;
}

int lm_finish(calib_handle* h) {
    if (!h->run_pending) return fail(h, CALIB_E_STATE, "lm_finish: no run in flight");
    h->run_pending = false;
    RC(sync(h));
    const int mode = h->loop_mode ? h->loop_mode : default_loop_mode(h);
    int64_t count = h->pending_count;
    if (!h->profiling && mode == 2)      // skipped evaluations (after a rejected step) still launch and return at once
        count += (int64_t)kernels_per_iteration(h) * h->h_state->n_pass;
    if (h->h_state->bar_abort) {
        char b[160];
        snprintf(b, sizeof b, "a wait inside the kernels timed out (code %d: 1 grid barrier of the persistent kernel, 2 / 3 / 4 "
                              "hand-off of the evaluation pass: producer / consumer warp)", h->h_state->bar_abort);
        return fail(h, CALIB_E_CUDA, b);
    }
    float ms = 0;
    CUDA_TRY(h, cudaEventElapsedTime(&ms, h->ev0, h->ev1));
    h->last_ms = ms;
    h->last_launches = count;
    h->last_evals = h->h_state->n_eval;
    h->cur = h->h_state->cur;
    for (auto& pe : h->prof_events) {
        float t = 0;
        cudaEventElapsedTime(&t, pe.second.first, pe.second.second);
        h->ktime[pe.first] += t; h->kcount[pe.first] += 1;
    }
    if (h->h_state->comm_fail)
        return fail(h, CALIB_E_COMM, "a peer rank never arrived at the in-kernel all-reduce (timeout)");
    return CALIB_OK;
}
}  // namespace

extern "C" int calib_set_param_trace(calib_t* h, int on) {
    if (!h) return fail(nullptr, CALIB_E_ARG, "NULL handle");
    h->want_ptrace = on != 0;
    return CALIB_OK;
}

extern "C" int calib_download_param_trace(calib_t* h, int iters, double* out) {
    RC(check_handle(h));
    if (!out || iters < 0 || iters > h->trace_cap) return fail(h, CALIB_E_ARG, "calib_download_param_trace: bad args");
    if (!h->want_ptrace) return fail(h, CALIB_E_STATE, "calib_download_param_trace: calib_set_param_trace(h, 1) before the run");
    std::vector<double> tmp((size_t)iters * 16);
    CUDA_TRY(h, cudaMemcpyAsync(tmp.data(), h->dev.ptrace, tmp.size() * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    RC(sync(h));
    for (int i = 0; i < iters; ++i) memcpy(out + (size_t)i * h->L, tmp.data() + (size_t)i * 16, (size_t)h->L * sizeof(double));
    return CALIB_OK;
}

extern "C" int calib_lm_run_resident(calib_t* h, int maxIters, double lam0, double lam_min,
                                     double lam_max, double err_min, double* sse_out, int* iters_out) {
    RC(check_handle(h));
    if (maxIters <= 0)
        return fail(h, CALIB_E_STATE, "calib_lm_run: maxIters <= 0 (the reference raises UnboundLocalError)");
    ScopeTimer tm_("calib_lm_run_resident");
    RC(lm_start(h, maxIters, lam0, lam_min, lam_max, err_min));
    const int rc = lm_finish(h);
    if (sse_out) *sse_out = h->h_state->sse;
    if (iters_out) *iters_out = h->h_state->iter;
    return rc;
}

extern "C" int calib_lm_run(calib_t* h, double* P_inout, int maxIters, double lam0, double lam_min,
                            double lam_max, double err_min, double* sse_out, int* iters_out, double* trace) {
    RC(check_handle(h));
    if (!P_inout) return fail(h, CALIB_E_ARG, "calib_lm_run: P is NULL");
    if (maxIters <= 0)
        return fail(h, CALIB_E_STATE, "calib_lm_run: maxIters <= 0 (the reference raises UnboundLocalError)");
    h->cur = 0;
    RC(upload_to(h, P_inout, 0));
    int iters = 0;
    RC(calib_lm_run_resident(h, maxIters, lam0, lam_min, lam_max, err_min, sse_out, &iters));
    if (iters_out) *iters_out = iters;
    if (trace) {      // both downloads behind one synchronisation
        if (iters > h->trace_cap) return fail(h, CALIB_E_ARG, "calib_lm_run: trace longer than its buffer");
        CUDA_TRY(h, cudaMemcpyAsync(trace, h->dev.trace, (size_t)iters * 4 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    }
    return calib_download_params(h, P_inout);
}

extern "C" int calib_last_run_stats(const calib_t* h, double* device_ms, int64_t* kernel_launches,
                                    int64_t* jacobian_evals) {
    if (!h) return fail(nullptr, CALIB_E_ARG, "NULL handle");
    if (device_ms) *device_ms = h->last_ms;
    if (kernel_launches) *kernel_launches = h->last_launches;
    if (jacobian_evals) *jacobian_evals = h->last_evals;
    return CALIB_OK;
}

extern "C" int calib_set_profiling(calib_t* h, int on) {
    if (!h) return fail(nullptr, CALIB_E_ARG, "NULL handle");
    h->profiling = on != 0;
    return CALIB_OK;
}

extern "C" int calib_kernel_time(const calib_t* h, int which, double* total_ms, int64_t* launches) {
    if (!h || which < 0 || which > 4) return fail(nullptr, CALIB_E_ARG, "calib_kernel_time: bad args");
    if (total_ms) *total_ms = h->ktime[which];
    if (launches) *launches = h->kcount[which];
    return CALIB_OK;
}

// ---- parity / debug entry points -------------------------------------------------------------------
namespace {
template <int MODEL>
void launch_debug_point(calib_handle* h, double* y, double* r, double* Jc, double* Jv) {
    const Dev& d = h->dev;
    k_debug_point<MODEL><<<(d.ev_pieces + kBlockWarps - 1) / kBlockWarps + 1, kBlockThreads, 0, h->stream>>>(d, y, r, Jc, Jv);
}

int debug_point(calib_handle* h, const double* P, double* y, double* r, double* Jc, double* Jv) {
    RC(check_handle(h));
    if (!P) return fail(h, CALIB_E_ARG, "P is NULL");
    h->cur = 0;
    RC(upload_to(h, P, 0));
    RC(push_state(h, 0.0, 0, 0, 0, 1));
    RC(make_views(h, 1));
    const size_t N = (size_t)std::max<int64_t>(h->N, 1);
    double *dy = nullptr, *dr = nullptr, *dJc = nullptr, *dJv = nullptr;
    std::vector<void*> tmp;
    if (y) CUDA_TRY(h, dalloc(tmp, &dy, 2 * N, h->stream, h->device));
    if (r) CUDA_TRY(h, dalloc(tmp, &dr, 2 * N, h->stream, h->device));
    if (Jc) CUDA_TRY(h, dalloc(tmp, &dJc, 2 * N * h->L, h->stream, h->device));
    if (Jv) CUDA_TRY(h, dalloc(tmp, &dJv, 12 * N, h->stream, h->device));
    if (h->model == 0) launch_debug_point<0>(h, dy, dr, dJc, dJv);
    else launch_debug_point<1>(h, dy, dr, dJc, dJv);
    int rc = sync(h);
    const size_t n = (size_t)h->N;
    if (!rc && y) cudaMemcpy(y, dy, 2 * n * sizeof(double), cudaMemcpyDeviceToHost);
    if (!rc && r) cudaMemcpy(r, dr, 2 * n * sizeof(double), cudaMemcpyDeviceToHost);
    if (!rc && Jc) cudaMemcpy(Jc, dJc, 2 * n * h->L * sizeof(double), cudaMemcpyDeviceToHost);
    if (!rc && Jv) cudaMemcpy(Jv, dJv, 12 * n * sizeof(double), cudaMemcpyDeviceToHost);
    free_pool(tmp, h->stream);
    return rc;
}

// uploads P, evaluates the normal equations and the Schur stage with damping lam (red[] valid
// afterwards); solve: also the reduced solve and the back-substitution (delta valid afterwards)
int eval_system(calib_handle* h, const double* P, double lam, bool solve) {
    RC(check_handle(h));
    if (!P) return fail(h, CALIB_E_ARG, "P is NULL");
    h->cur = 0;
    h->dev.use_cond = 0;
    RC(upload_to(h, P, 0));
    RC(push_state(h, lam, 0, 0, 0, 1));
    RC(make_views(h, 1));
    Launcher ln{h, false};
    if (h->model == 0) { enqueue_eval<0>(h, ln); enqueue_schur<0>(h, ln, 0); }
    else { enqueue_eval<1>(h, ln); enqueue_schur<1>(h, ln, 0); }
    if (nccl_active(h) && g_nccl.AllReduce(h->dev.red, h->dev.red, kRedN, ncclDouble, ncclSum, h->comm, h->stream) != ncclSuccess)
        return fail(h, CALIB_E_COMM, "ncclAllReduce failed");
    if (solve) {
        if (h->model == 0) { enqueue_solve<0>(h, ln); enqueue_backsub<0>(h, ln); }
        else { enqueue_solve<1>(h, ln); enqueue_backsub<1>(h, ln); }
    }
    return sync(h);
}
}  // namespace

extern "C" int calib_project(calib_t* h, const double* P, double* y) {
    if (!y) return fail(h, CALIB_E_ARG, "calib_project: y is NULL");
    return debug_point(h, P, y, nullptr, nullptr, nullptr);
}

extern "C" int calib_eval_residual_jacobian(calib_t* h, const double* P, double* r, double* Jc, double* Jv) {
    return debug_point(h, P, nullptr, r, Jc, Jv);
}

extern "C" int calib_cost(calib_t* h, const double* P, double* E) {
    RC(check_handle(h));
    if (!P || !E) return fail(h, CALIB_E_ARG, "calib_cost: NULL argument");
    h->cur = 0;
    h->dev.use_cond = 0;
    RC(upload_to(h, P, 1));
    RC(push_state(h, 0.0, 0, 0, 0, 1));
    RC(make_views(h, 0));
    Launcher ln{h, false};
    if (h->model == 0) enqueue_cost<0>(h, ln, 0); else enqueue_cost<1>(h, ln, 0);
    if (nccl_active(h) && g_nccl.AllReduce(h->dev.etrial, h->dev.etrial, 1, ncclDouble, ncclSum, h->comm, h->stream) != ncclSuccess)
        return fail(h, CALIB_E_COMM, "ncclAllReduce failed");
    CUDA_TRY(h, cudaMemcpyAsync(h->h_pinned, h->dev.etrial, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    RC(sync(h));
    *E = h->h_pinned[0];
    return CALIB_OK;
}

extern "C" int calib_eval_normal_equations(calib_t* h, const double* P, double* A, double* B, double* D,
                                           double* g, double* E) {
    if (!h) return fail(nullptr, CALIB_E_ARG, "NULL handle");
    if (!A || !B || !D || !g || !E) return fail(h, CALIB_E_ARG, "calib_eval_normal_equations: NULL output");
    RC(eval_system(h, P, 0.0, false));
    const int L = h->L;
    const size_t M = (size_t)h->M;
    double *dA = nullptr, *dB = nullptr, *dD = nullptr, *dg = nullptr;
    std::vector<void*> tmp;
    CUDA_TRY(h, dalloc(tmp, &dA, (size_t)L * L, h->stream, h->device));
    CUDA_TRY(h, dalloc(tmp, &dB, M * L * 6, h->stream, h->device));
    CUDA_TRY(h, dalloc(tmp, &dD, M * 36, h->stream, h->device));
    CUDA_TRY(h, dalloc(tmp, &dg, (size_t)h->K, h->stream, h->device));
    const int nb = (int)((h->M + 127) / 128);
    if (h->model == 0) k_debug_unpack<0><<<nb, 128, 0, h->stream>>>(h->dev, dA, dB, dD, dg);
    else k_debug_unpack<1><<<nb, 128, 0, h->stream>>>(h->dev, dA, dB, dD, dg);
    int rc = sync(h);
    if (!rc) {
        cudaMemcpy(A, dA, (size_t)L * L * sizeof(double), cudaMemcpyDeviceToHost);
        cudaMemcpy(B, dB, M * L * 6 * sizeof(double), cudaMemcpyDeviceToHost);
        cudaMemcpy(D, dD, M * 36 * sizeof(double), cudaMemcpyDeviceToHost);
        cudaMemcpy(g, dg, (size_t)h->K * sizeof(double), cudaMemcpyDeviceToHost);
        cudaMemcpy(E, h->dev.red + kRedE, sizeof(double), cudaMemcpyDeviceToHost);
    }
    free_pool(tmp, h->stream);
    return rc;
}

extern "C" int calib_lm_step(calib_t* h, const double* P, double lam, double* delta) {
    if (!h) return fail(nullptr, CALIB_E_ARG, "NULL handle");
    if (!delta) return fail(h, CALIB_E_ARG, "calib_lm_step: delta is NULL");
    RC(eval_system(h, P, lam, true));
    CUDA_TRY(h, cudaMemcpy(delta, h->dev.delta, (size_t)h->K * sizeof(double), cudaMemcpyDeviceToHost));
    return CALIB_OK;
}

// Average device time of ONE kernel of the iteration, launched `reps` times back to back at P with
// damping lam (events around the whole batch, so launch gaps between identical independent
// launches are included but no host time).  flush != 0: the 256 MB L2 flush before every launch
// (then the time of the flush kernel alone is measured the same way and subtracted).
extern "C" int calib_time_kernel(calib_t* h, const double* P, double lam, int which, int reps, int flush,
                                 double* avg_us) {
    RC(check_handle(h));
    if (!P || !avg_us || reps <= 0 || which < 0 || which > 3) return fail(h, CALIB_E_ARG, "calib_time_kernel: bad args");
    h->cur = 0;
    h->dev.use_cond = 0;
    RC(upload_to(h, P, 0));
    RC(push_state(h, lam, 0, 0, 0, 1 << 30));
    RC(make_views(h, 1));
    Launcher ln{h, false};
    auto one = [&](int k) {
        if (h->model == 0) {
            if (k == 0) enqueue_eval<0>(h, ln); else if (k == 1) enqueue_schur<0>(h, ln, 1);
            else if (k == 2) enqueue_backsub<0>(h, ln); else enqueue_cost<0>(h, ln, 0);
        } else {
            if (k == 0) enqueue_eval<1>(h, ln); else if (k == 1) enqueue_schur<1>(h, ln, 1);
            else if (k == 2) enqueue_backsub<1>(h, ln); else enqueue_cost<1>(h, ln, 0);
        }
    };
    for (int k = 0; k < 4; ++k) one(k);       // a full iteration first: every input of `which` is valid
    RC(sync(h));
    if (flush && !h->l2buf) RC(calib_flush_l2(h));
    const long long nfl = (long long)(h->l2bytes / sizeof(double));
    float ms_flush = 0.f, ms = 0.f;
    if (flush) {
        CUDA_TRY(h, cudaEventRecord(h->ev0, h->stream));
        for (int r = 0; r < reps; ++r) k_fill<<<1184, 256, 0, h->stream>>>(h->l2buf, nfl, 1.0);
        CUDA_TRY(h, cudaEventRecord(h->ev1, h->stream));
        RC(sync(h));
        CUDA_TRY(h, cudaEventElapsedTime(&ms_flush, h->ev0, h->ev1));
    }
    for (int r = 0; r < 3; ++r) one(which);
    CUDA_TRY(h, cudaEventRecord(h->ev0, h->stream));
    for (int r = 0; r < reps; ++r) {
        if (flush) k_fill<<<1184, 256, 0, h->stream>>>(h->l2buf, nfl, 1.0);
        one(which);
    }
    CUDA_TRY(h, cudaEventRecord(h->ev1, h->stream));
    RC(sync(h));
    CUDA_TRY(h, cudaEventElapsedTime(&ms, h->ev0, h->ev1));
    *avg_us = 1e3 * (double)(ms - ms_flush) / reps;
    return CALIB_OK;
}

// ---- roofline helpers ----------------------------------------------------------------------------
extern "C" int calib_measure_fp64_peak(int device_id, double* tflops) {
    if (!tflops) return fail(nullptr, CALIB_E_ARG, "tflops is NULL");
    calib_handle* h = nullptr;
    CUDA_TRY(h, cudaSetDevice(device_id));
    cudaDeviceProp prop;
    CUDA_TRY(h, cudaGetDeviceProperties(&prop, device_id));
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 4096;
    double* out = nullptr;
    CUDA_TRY(h, cudaMalloc((void**)&out, (size_t)blocks * threads * sizeof(double)));
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    double best = 0;
    for (int rep = 0; rep < 6; ++rep) {
        cudaEventRecord(a);
        k_fp64_peak<<<blocks, threads>>>(out, iters, 1.0000001, 1e-9);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms = 0;
        cudaEventElapsedTime(&ms, a, b);
        const double fl = 2.0 * 8.0 * iters * (double)blocks * threads;
        if (rep > 0) best = std::max(best, fl / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(a); cudaEventDestroy(b);
    cudaFree(out);
    CUDA_TRY(h, cudaGetLastError());
    *tflops = best;
    return CALIB_OK;
}

extern "C" int calib_flush_l2(calib_t* h) {
    RC(check_handle(h));
    if (!h->l2buf) {
        h->l2bytes = (size_t)256 << 20;   // > 126 MB L2
        CUDA_TRY(h, cudaMalloc((void**)&h->l2buf, h->l2bytes));
    }
    k_fill<<<1184, 256, 0, h->stream>>>(h->l2buf, (long long)(h->l2bytes / sizeof(double)), 1.0);
    return sync(h);
}

// ---- multi-GPU ------------------------------------------------------------------------------------
extern "C" int calib_comm_unique_id(void* id128) {
    if (!id128) return fail(nullptr, CALIB_E_ARG, "id128 is NULL");
    if (!nccl_load()) return fail(nullptr, CALIB_E_COMM, "libnccl.so.2 not found");
    ncclUniqueId id;
    if (g_nccl.GetUniqueId(&id) != ncclSuccess) return fail(nullptr, CALIB_E_COMM, "ncclGetUniqueId failed");
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    memcpy(id128, &id, 128);
    return CALIB_OK;
}

extern "C" int calib_comm_init(calib_t* h, int nranks, int rank, const void* id128) {
    RC(check_handle(h));
    if (!id128 || nranks < 1 || rank < 0 || rank >= nranks) return fail(h, CALIB_E_ARG, "calib_comm_init: bad args");
    if (!nccl_load()) return fail(h, CALIB_E_COMM, "libnccl.so.2 not found");
    drop_graph(h);
    if (h->comm && h->comm_owned) g_nccl.CommDestroy(h->comm);
    h->comm = nullptr; h->comm_owned = false;
    h->nranks = nranks; h->rank = rank;
    if (nranks == 1) return CALIB_OK;
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    ncclResult_t r = g_nccl.CommInitRank(&h->comm, nranks, id, rank);
    if (r != ncclSuccess) {
        h->comm = nullptr;
        return fail(h, CALIB_E_COMM, g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "ncclCommInitRank failed");
    }
    h->comm_owned = true;
    apply_transport(h);
    return CALIB_OK;
}

extern "C" int calib_comm_share(calib_t* h, const calib_t* owner) {
    RC(check_handle(h));
    if (!owner) return fail(h, CALIB_E_ARG, "calib_comm_share: owner is NULL");
    if (owner->device != h->device) return fail(h, CALIB_E_ARG, "calib_comm_share: handles on different devices");
    drop_graph(h);
    if (h->comm && h->comm_owned) g_nccl.CommDestroy(h->comm);
    h->comm = owner->comm; h->comm_owned = false;
    h->nranks = owner->nranks; h->rank = owner->rank;
    h->peer = owner->peer; h->transport = owner->transport;
    apply_transport(h);
    return CALIB_OK;
}

extern "C" int calib_comm_set_default(const calib_t* owner, int device_id) {
    if (device_id < 0 || device_id >= 64) return fail(nullptr, CALIB_E_ARG, "calib_comm_set_default: bad device_id");
    if (owner && owner->device != device_id) return fail(nullptr, CALIB_E_ARG, "calib_comm_set_default: owner is on another device");
    std::lock_guard<std::mutex> lk(g_default_mu);
    g_default_comm[device_id] = owner;
    return CALIB_OK;
}

// ---- in-kernel all-reduce over NVLink peer memory (calib_peer.cuh) -------------------------------------
extern "C" int calib_peer_export(calib_t* h, void* handle64) {
    RC(check_handle(h));
    if (!handle64) return fail(h, CALIB_E_ARG, "calib_peer_export: handle64 is NULL");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    if (!h->peer || h->peer->nranks > 1) {   // (re-)export: a fresh inbox, sequence numbers start over
        auto pc = std::make_shared<PeerComm>();
        pc->device = h->device;
        const size_t bytes = kPeerInboxBytes + 64;
        CUDA_TRY(h, cudaMalloc((void**)&pc->inbox, bytes));
        CUDA_TRY(h, cudaMemset(pc->inbox, 0, bytes));
        CUDA_TRY(h, cudaDeviceSynchronize());
        h->peer = pc;
        apply_transport(h);
    }
    cudaIpcMemHandle_t mh;
    CUDA_TRY(h, cudaIpcGetMemHandle(&mh, h->peer->inbox));
    memcpy(handle64, &mh, 64);
    return CALIB_OK;
}

extern "C" int calib_peer_attach(calib_t* h, int nranks, int rank, const void* handles) {
    RC(check_handle(h));
    if (!handles || nranks < 1 || nranks > kPeerMaxRanks || rank < 0 || rank >= nranks)
        return fail(h, CALIB_E_ARG, "calib_peer_attach: bad args (at most 8 ranks)");
    if (!h->peer || !h->peer->inbox) return fail(h, CALIB_E_STATE, "calib_peer_attach: call calib_peer_export first");
    PeerComm& pc = *h->peer;
    for (void*& m : pc.mapped) if (m) { cudaIpcCloseMemHandle(m); m = nullptr; }
    pc.nranks = nranks; pc.rank = rank;
    for (int r = 0; r < nranks; ++r) {
        if (r == rank) continue;
        cudaIpcMemHandle_t mh;
        memcpy(&mh, (const char*)handles + 64 * (size_t)r, 64);
        cudaError_t e = cudaIpcOpenMemHandle(&pc.mapped[r], mh, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            pc.mapped[r] = nullptr;
            pc.nranks = 1;
            cudaGetLastError();
            apply_transport(h);
            char b[256];
            snprintf(b, sizeof b, "calib_peer_attach: cudaIpcOpenMemHandle(rank %d) failed: %s", r, cudaGetErrorString(e));
            return fail(h, CALIB_E_COMM, b);
        }
    }
    h->nranks = nranks; h->rank = rank;
    apply_transport(h);
    return CALIB_OK;
}

extern "C" int calib_set_transport(calib_t* h, int mode) {
    if (!h) return fail(nullptr, CALIB_E_ARG, "NULL handle");
    if (mode < 0 || mode > 2) return fail(h, CALIB_E_ARG, "calib_set_transport: 0 auto, 1 NCCL, 2 peer memory");
    if (mode == 2 && !(h->peer && h->peer->nranks > 1 && h->peer->mapped_ok()))
        return fail(h, CALIB_E_STATE, "calib_set_transport: no peer communicator attached");
    h->transport = mode;
    apply_transport(h);
    return CALIB_OK;
}

extern "C" int calib_transport(const calib_t* h) {
    if (!h) return -1;
    return peer_active(h) ? 2 : (h->comm ? 1 : 0);
}

// ---- per-view homography refinement (calib_homography.cuh) -------------------------------------------
extern "C" int calib_refine_homographies(calib_t* h, double* H_inout, int maxIters, double lam0, double lam_min,
                                         double lam_max, double err_min, int* iters_out, double* trace) {
    RC(check_handle(h));
    if (!H_inout) return fail(h, CALIB_E_ARG, "calib_refine_homographies: H is NULL");
    if (maxIters <= 0) return fail(h, CALIB_E_ARG, "calib_refine_homographies: maxIters <= 0");
    const size_t M = (size_t)h->M;
    std::vector<void*> tmp;
    long long* d_offs = nullptr;
    double *d_H = nullptr, *d_tr = nullptr;
    int* d_it = nullptr;
    int rc = [&]() -> int {
        CUDA_TRY(h, dalloc(tmp, &d_offs, M + 1, h->stream, h->device));
        CUDA_TRY(h, dalloc(tmp, &d_H, M * 9, h->stream, h->device));
        CUDA_TRY(h, dalloc(tmp, &d_it, M, h->stream, h->device));
        if (trace) CUDA_TRY(h, dalloc(tmp, &d_tr, M * (size_t)maxIters * 4, h->stream, h->device));
        static_assert(sizeof(long long) == sizeof(int64_t), "int64 layout");
        CUDA_TRY(h, cudaMemcpyAsync(d_offs, h->view_offsets.data(), (M + 1) * 8, cudaMemcpyHostToDevice, h->stream));
        CUDA_TRY(h, cudaMemcpyAsync(d_H, H_inout, M * 72, cudaMemcpyHostToDevice, h->stream));
        if (trace) CUDA_TRY(h, cudaMemsetAsync(d_tr, 0, M * (size_t)maxIters * 32, h->stream));
        HomDev d;
        d.X = h->dev.X; d.Y = h->dev.Y; d.U = h->dev.U; d.V = h->dev.V;
        d.offs = d_offs; d.M = (int)M; d.max_iters = maxIters; d.want_trace = trace ? 1 : 0;
        d.lam0 = lam0; d.lam_min = lam_min; d.lam_max = lam_max; d.err_min = err_min;
        d.H = d_H; d.iters = d_it; d.trace = d_tr;
        const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>(((int64_t)M + kBlockWarps - 1) / kBlockWarps, 8 * h->n_sm));
        k_homography_lm<<<blocks, kBlockThreads, 0, h->stream>>>(d);
        CUDA_TRY(h, cudaGetLastError());
        CUDA_TRY(h, cudaMemcpyAsync(H_inout, d_H, M * 72, cudaMemcpyDeviceToHost, h->stream));
        if (iters_out) CUDA_TRY(h, cudaMemcpyAsync(iters_out, d_it, M * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
        if (trace) CUDA_TRY(h, cudaMemcpyAsync(trace, d_tr, M * (size_t)maxIters * 32, cudaMemcpyDeviceToHost, h->stream));
        return sync(h);
    }();
    free_pool(tmp, h->stream);
    return rc;
}

// ---- per-point parts of the closed-form initial estimate (calib_linear.cuh) ---------------------------
namespace {
int linear_call(calib_handle* h, int which, const double* A9, const double* W12, double* out0, size_t n0,
                double* out1, size_t n1) {
    const size_t M = (size_t)h->M;
    std::vector<void*> tmp;
    int rc = [&]() -> int {
        long long* d_offs = nullptr;
        double *d0 = nullptr, *d1 = nullptr, *dA = nullptr, *dW = nullptr;
        CUDA_TRY(h, dalloc(tmp, &d_offs, M + 1, h->stream, h->device));
        CUDA_TRY(h, dalloc(tmp, &d0, M * n0, h->stream, h->device));
        if (n1) CUDA_TRY(h, dalloc(tmp, &d1, M * n1, h->stream, h->device));
        CUDA_TRY(h, cudaMemcpyAsync(d_offs, h->view_offsets.data(), (M + 1) * 8, cudaMemcpyHostToDevice, h->stream));
        LinDev d;
        memset(&d, 0, sizeof d);
        d.X = h->dev.X; d.Y = h->dev.Y; d.Z = h->dev.Z; d.U = h->dev.U; d.V = h->dev.V;
        d.offs = d_offs; d.M = (int)M; d.model = h->model;
        const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>(((int64_t)M + kBlockWarps - 1) / kBlockWarps, 8 * h->n_sm));
        if (which == 0 || which == 2) {
            if (which == 0) { d.gram = d0; d.norm = d1; }
            else {          // the homographies are what leaves the device; Gram sums and normalisations stay
                CUDA_TRY(h, dalloc(tmp, &d.gram, M * kGramN, h->stream, h->device));
                CUDA_TRY(h, dalloc(tmp, &d.norm, M * kNormN, h->stream, h->device));
                d.Hdlt = d0;
            }
            k_dlt_gram<<<blocks, kBlockThreads, 0, h->stream>>>(d);
        } else {
            CUDA_TRY(h, dalloc(tmp, &dA, 9, h->stream, h->device));
            CUDA_TRY(h, dalloc(tmp, &dW, M * 12, h->stream, h->device));
            CUDA_TRY(h, cudaMemcpyAsync(dA, A9, 72, cudaMemcpyHostToDevice, h->stream));
            CUDA_TRY(h, cudaMemcpyAsync(dW, W12, M * 96, cudaMemcpyHostToDevice, h->stream));
            d.A = dA; d.W = dW; d.drows = d0;
            if (h->model == 0) k_distortion_rows<0><<<blocks, kBlockThreads, 0, h->stream>>>(d);
            else k_distortion_rows<1><<<blocks, kBlockThreads, 0, h->stream>>>(d);
        }
        CUDA_TRY(h, cudaGetLastError());
        CUDA_TRY(h, cudaMemcpyAsync(out0, d0, M * n0 * 8, cudaMemcpyDeviceToHost, h->stream));
        if (n1) CUDA_TRY(h, cudaMemcpyAsync(out1, d1, M * n1 * 8, cudaMemcpyDeviceToHost, h->stream));
        return sync(h);
    }();
    free_pool(tmp, h->stream);
    return rc;
}
}  // namespace

extern "C" int calib_dlt_gram(calib_t* h, double* gram, double* norm) {
    RC(check_handle(h));
    if (!gram || !norm) return fail(h, CALIB_E_ARG, "calib_dlt_gram: NULL output");
    return linear_call(h, 0, nullptr, nullptr, gram, kGramN, norm, kNormN);
}

extern "C" int calib_dlt_homographies(calib_t* h, double* H) {
    RC(check_handle(h));
    if (!H) return fail(h, CALIB_E_ARG, "calib_dlt_homographies: NULL output");
    return linear_call(h, 2, nullptr, nullptr, H, 9, nullptr, 0);
}

extern "C" int calib_distortion_rows(calib_t* h, const double* A, const double* W, double* rows) {
    RC(check_handle(h));
    if (!A || !W || !rows) return fail(h, CALIB_E_ARG, "calib_distortion_rows: NULL argument");
    return linear_call(h, 1, A, W, rows, kDistN, nullptr, 0);
}

// ---- batch of independent calibrations (calib_batch.cuh) -------------------------------------------
struct calib_batch_handle {
    int model = 0, L = 0, device = 0;
    int64_t n_problems = 0, total_views = 0, N = 0, total_params = 0;
    BatchDev dev{};
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    void* arena = nullptr;     // one stream-ordered allocation (the device pool keeps freed memory: a create /
                               // destroy per batch does not go through cudaMalloc / cudaFree, 10-40 ms at cfg4)
    double* trace_dev = nullptr;
    size_t trace_cap = 0;      // doubles
    size_t smem_bytes = 0;
    std::string err;
};

#define BATCH_TRY(h, call)                                                                    \
    do {                                                                                      \
        cudaError_t e__ = (call);                                                             \
        if (e__ != cudaSuccess) {                                                             \
            char b__[512];                                                                    \
            snprintf(b__, sizeof b__, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), \
                     __FILE__, __LINE__);                                                     \
            if (h) (h)->err = b__;                                                            \
            g_last_error = b__;                                                               \
            return CALIB_E_CUDA;                                                              \
        }                                                                                     \
    } while (0)

static int bfail(calib_batch_handle* h, int code, const char* msg) {
    if (h) h->err = msg;
    g_last_error = msg;
    return code;
}

extern "C" void calib_batch_destroy(calib_batch_t* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->arena && h->stream) cudaFreeAsync(h->arena, h->stream);
    if (h->trace_dev && h->stream) cudaFreeAsync(h->trace_dev, h->stream);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

extern "C" const char* calib_batch_last_error(const calib_batch_t* h) {
    return h ? h->err.c_str() : g_last_error.c_str();
}

extern "C" int calib_batch_create(calib_batch_t** out, int model, int64_t n_problems,
                                  const int64_t* problem_view_offsets, const int64_t* view_offsets,
                                  const double* model_pts, const double* sensor_pts, int device_id) {
    if (!out) return bfail(nullptr, CALIB_E_ARG, "calib_batch_create: out is NULL");
    *out = nullptr;
    if (model != CALIB_MODEL_RADTAN && model != CALIB_MODEL_FISHEYE)
        return bfail(nullptr, CALIB_E_ARG, "calib_batch_create: unknown distortion model");
    if (n_problems <= 0 || !problem_view_offsets || !view_offsets)
        return bfail(nullptr, CALIB_E_ARG, "calib_batch_create: need n_problems >= 1 and both offset arrays");
    if (problem_view_offsets[0] != 0 || view_offsets[0] != 0)
        return bfail(nullptr, CALIB_E_ARG, "calib_batch_create: offsets must start at 0");
    const int L = model == 0 ? 10 : 9;
    int64_t max_views = 0, max_points = 0;
    for (int64_t p = 0; p < n_problems; ++p) {
        const int64_t a = problem_view_offsets[p], b = problem_view_offsets[p + 1];
        if (b <= a) return bfail(nullptr, CALIB_E_ARG, "calib_batch_create: every problem needs >= 1 view");
        max_views = std::max(max_views, b - a);
    }
    const int64_t total_views = problem_view_offsets[n_problems];
    for (int64_t v = 0; v < total_views; ++v)
        if (view_offsets[v + 1] < view_offsets[v])
            return bfail(nullptr, CALIB_E_ARG, "calib_batch_create: view_offsets not non-decreasing");
    for (int64_t p = 0; p < n_problems; ++p)
        max_points = std::max(max_points, view_offsets[problem_view_offsets[p + 1]] - view_offsets[problem_view_offsets[p]]);
    const int64_t N = view_offsets[total_views];
    if (N > 0 && (!model_pts || !sensor_pts)) return bfail(nullptr, CALIB_E_ARG, "calib_batch_create: NULL points");
    if (N >= (1ll << 31) - 64) return bfail(nullptr, CALIB_E_ARG, "calib_batch_create: too many points");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0)
        return bfail(nullptr, CALIB_E_CUDA, "calib_batch_create: no CUDA device (this library has no CPU fallback)");
    if (device_id < 0 || device_id >= ndev) return bfail(nullptr, CALIB_E_ARG, "calib_batch_create: bad device_id");
    // shared memory: stage the detections when two blocks per SM still fit, else stream them from L2
    int stage = 1;
    BatchSmem lay = batch_smem_layout(L, (int)max_views, (int)max_points, 1);
    if ((size_t)lay.total * 8 > (size_t)(220 / kBatchBlocks) * 1024) { stage = 0; lay = batch_smem_layout(L, (int)max_views, (int)max_points, 0); }
    if ((size_t)lay.total * 8 > 220 * 1024)
        return bfail(nullptr, CALIB_E_ARG, "calib_batch_create: a problem has too many views for the one-block-per-problem "
                                           "path (about 100); use calib_create for it");
    calib_batch_handle* h = new calib_batch_handle();
    h->model = model; h->L = L; h->device = device_id; h->n_problems = n_problems;
    h->total_views = total_views; h->N = N; h->total_params = (int64_t)L * n_problems + 6 * total_views;
    h->smem_bytes = (size_t)lay.total * 8;
    auto bail = [&](int rc) { calib_batch_destroy(h); return rc; };
#define TRYB_(call) do { int rc__ = [&]() -> int { BATCH_TRY(h, call); return 0; }(); if (rc__) return bail(rc__); } while (0)
    TRYB_(cudaSetDevice(device_id));
    TRYB_(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    TRYB_(cudaEventCreate(&h->ev0));
    TRYB_(cudaEventCreate(&h->ev1));
    BatchDev& d = h->dev;
    d.model = model; d.L = L; d.n_problems = (int)n_problems; d.max_views = (int)max_views;
    d.max_points = (int)max_points; d.stage_points = stage;
    const size_t n = (size_t)std::max<int64_t>(N, 1);
    double *soa = nullptr, *aos = nullptr;
    long long *pv = nullptr, *vo = nullptr;
    {
            Arena a;
        const size_t o_soa = a.reserve<double>(5 * n + 64), o_aos = a.reserve<double>(5 * n);
        const size_t o_pv = a.reserve<long long>((size_t)n_problems + 1), o_vo = a.reserve<long long>((size_t)total_views + 1);
        const size_t o_P = a.reserve<double>((size_t)h->total_params), o_sse = a.reserve<double>((size_t)n_problems);
        const size_t o_it = a.reserve<int>((size_t)n_problems);
        TRYB_(lib_malloc_async((void**)&h->arena, a.bytes, h->stream, h->device));
        a.base = (char*)h->arena;
        soa = a.at<double>(o_soa); aos = a.at<double>(o_aos);
        pv = a.at<long long>(o_pv); vo = a.at<long long>(o_vo);
        d.P = a.at<double>(o_P); d.sse = a.at<double>(o_sse); d.iters = a.at<int>(o_it);
    }
    static_assert(sizeof(long long) == sizeof(int64_t), "int64 layout");
    TRYB_(cudaMemcpyAsync(pv, problem_view_offsets, ((size_t)n_problems + 1) * 8, cudaMemcpyHostToDevice, h->stream));
    TRYB_(cudaMemcpyAsync(vo, view_offsets, ((size_t)total_views + 1) * 8, cudaMemcpyHostToDevice, h->stream));
    if (N > 0) {
        TRYB_(cudaMemcpyAsync(aos, model_pts, (size_t)N * 3 * 8, cudaMemcpyHostToDevice, h->stream));
        TRYB_(cudaMemcpyAsync(aos + 3 * n, sensor_pts, (size_t)N * 2 * 8, cudaMemcpyHostToDevice, h->stream));
        k_repack<<<(int)std::min<int64_t>((N + 255) / 256, 1184), 256, 0, h->stream>>>(aos, aos + 3 * n, N, soa, soa + n,
                                                                                  soa + 2 * n, soa + 3 * n, soa + 4 * n);
        TRYB_(cudaGetLastError());
    }
    d.prob_view = pv; d.view_off = vo;
    d.X = soa; d.Y = soa + n; d.Z = soa + 2 * n; d.U = soa + 3 * n; d.V = soa + 4 * n;
    TRYB_(cudaStreamSynchronize(h->stream));
#undef TRYB_
    *out = h;
    return CALIB_OK;
}

extern "C" int calib_batch_lm_run(calib_batch_t* h, double* P_inout, int maxIters, double lam0,
                                  double lam_min, double lam_max, double err_min, double* sse_out,
                                  int* iters_out, double* trace, double* device_ms) {
    if (!h) return bfail(nullptr, CALIB_E_ARG, "NULL handle");
    if (!P_inout) return bfail(h, CALIB_E_ARG, "calib_batch_lm_run: P is NULL");
    if (maxIters <= 0)
        return bfail(h, CALIB_E_STATE, "calib_batch_lm_run: maxIters <= 0 (the reference raises UnboundLocalError)");
    BATCH_TRY(h, cudaSetDevice(h->device));
    BatchDev& d = h->dev;
    d.max_iters = maxIters; d.lam0 = lam0; d.lam_min = lam_min; d.lam_max = lam_max; d.err_min = err_min;
    d.want_trace = trace ? 1 : 0;
    if (trace) {
        const size_t need = (size_t)h->n_problems * maxIters * 4;
        if (need > h->trace_cap) {
            if (h->trace_dev) cudaFreeAsync(h->trace_dev, h->stream);
            h->trace_dev = nullptr; h->trace_cap = 0;
            BATCH_TRY(h, lib_malloc_async((void**)&h->trace_dev, need * 8, h->stream, h->device));
            h->trace_cap = need;
        }
        BATCH_TRY(h, cudaMemsetAsync(h->trace_dev, 0, need * 8, h->stream));
        d.trace = h->trace_dev;
    }
    BATCH_TRY(h, cudaMemcpyAsync(d.P, P_inout, (size_t)h->total_params * 8, cudaMemcpyHostToDevice, h->stream));
    // per launch and device: another live batch handle may need a different size
    if (h->model == 0) BATCH_TRY(h, cudaFuncSetAttribute(k_batch_lm<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_bytes));
    else BATCH_TRY(h, cudaFuncSetAttribute(k_batch_lm<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_bytes));
    BATCH_TRY(h, cudaEventRecord(h->ev0, h->stream));
    if (h->model == 0) k_batch_lm<0><<<(int)h->n_problems, kBatchThreads, h->smem_bytes, h->stream>>>(d);
    else k_batch_lm<1><<<(int)h->n_problems, kBatchThreads, h->smem_bytes, h->stream>>>(d);
    BATCH_TRY(h, cudaGetLastError());
    BATCH_TRY(h, cudaEventRecord(h->ev1, h->stream));
    BATCH_TRY(h, cudaMemcpyAsync(P_inout, d.P, (size_t)h->total_params * 8, cudaMemcpyDeviceToHost, h->stream));
    if (sse_out) BATCH_TRY(h, cudaMemcpyAsync(sse_out, d.sse, (size_t)h->n_problems * 8, cudaMemcpyDeviceToHost, h->stream));
    if (iters_out) BATCH_TRY(h, cudaMemcpyAsync(iters_out, d.iters, (size_t)h->n_problems * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    if (trace) BATCH_TRY(h, cudaMemcpyAsync(trace, h->trace_dev, (size_t)h->n_problems * maxIters * 32, cudaMemcpyDeviceToHost, h->stream));
    BATCH_TRY(h, cudaStreamSynchronize(h->stream));
    BATCH_TRY(h, cudaGetLastError());
    if (device_ms) {
        float ms = 0;
        BATCH_TRY(h, cudaEventElapsedTime(&ms, h->ev0, h->ev1));
        *device_ms = ms;
    }
    return CALIB_OK;
}

#ifdef CALIB_STAMPS
// latency-study builds only: cycle counters per phase of k_batch_lm, summed over the blocks of the runs since the
// last call (the first call allocates them; out = kBatchPhases values)
extern "C" int calib_batch_debug_phases(calib_batch_t* h, unsigned long long* out) {
    if (!h) return bfail(nullptr, CALIB_E_ARG, "NULL handle");
    BATCH_TRY(h, cudaSetDevice(h->device));
    if (!h->dev.phase) {
        BATCH_TRY(h, cudaMalloc((void**)&h->dev.phase, kBatchPhases * 8));
        BATCH_TRY(h, cudaMemset(h->dev.phase, 0, kBatchPhases * 8));
        for (int i = 0; i < kBatchPhases; ++i) out[i] = 0;
        return CALIB_OK;
    }
    BATCH_TRY(h, cudaMemcpy(out, h->dev.phase, kBatchPhases * 8, cudaMemcpyDeviceToHost));
    BATCH_TRY(h, cudaMemset(h->dev.phase, 0, kBatchPhases * 8));
    return CALIB_OK;
}

// latency-study builds only: the persistent kernel's stamps of the last run, [2][kPStampPasses][kPStampSlots]
extern "C" int calib_debug_pstamps(calib_t* h, unsigned long long* out) {
    RC(check_handle(h));
    CUDA_TRY(h, cudaMemcpy(out, h->dev.stamps + (size_t)4 * kStampBlocks * 8, (size_t)2 * kPStampPasses * kPStampSlots * 8,
                           cudaMemcpyDeviceToHost));
    CUDA_TRY(h, cudaMemset(h->dev.stamps + (size_t)4 * kStampBlocks * 8, 0, (size_t)2 * kPStampPasses * kPStampSlots * 8));
    return CALIB_OK;
}
// latency-study builds only (not in the header): one pass at P, then the globaltimer stamps
extern "C" int calib_debug_stamps(calib_t* h, const double* P, double lam, unsigned long long* out) {
    RC(check_handle(h));
    if (!P) {      // just read what the last run left (the stamps of its last pass)
        CUDA_TRY(h, cudaMemcpy(out, h->dev.stamps, (size_t)4 * kStampBlocks * 8 * 8, cudaMemcpyDeviceToHost));
        return CALIB_OK;
    }
    h->cur = 0;
    h->dev.use_cond = 0;
    RC(upload_to(h, P, 0));
    RC(push_state(h, lam, 0, 0, 0, 1 << 30));
    RC(make_views(h, 1));
    Launcher ln{h, false};
    for (int rep = 0; rep < 3; ++rep) {
        if (rep == 2) CUDA_TRY(h, cudaMemsetAsync(h->dev.stamps, 0, (size_t)4 * kStampBlocks * 8 * 8, h->stream));
        if (h->model == 0) { enqueue_eval<0>(h, ln); enqueue_schur<0>(h, ln, 1); enqueue_backsub<0>(h, ln); enqueue_cost<0>(h, ln, 0); }
        else { enqueue_eval<1>(h, ln); enqueue_schur<1>(h, ln, 1); enqueue_backsub<1>(h, ln); enqueue_cost<1>(h, ln, 0); }
    }
    RC(sync(h));
    CUDA_TRY(h, cudaMemcpy(out, h->dev.stamps, (size_t)4 * kStampBlocks * 8 * 8, cudaMemcpyDeviceToHost));
    return CALIB_OK;
}
#endif

#include "calib_multi.inc"
#include "calib_noalloc.inc"
