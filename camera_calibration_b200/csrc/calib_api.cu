// C-ABI of libcalib_b200.so (see include/calib_b200.h).  Host side: device memory, the task
// list, the per-iteration CUDA graph, the LM driver loop, NCCL plumbing for sharded views.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/calib_b200.h"
#include "calib_solve.cuh"
#include "calib_batch.cuh"

using namespace calib;

static thread_local std::string g_last_error;

struct calib_handle {
    int model = 0, L = 0, device = 0;
    int64_t M = 0, N = 0, K = 0;
    std::vector<int64_t> view_offsets;
    int seg = 0;
    Dev dev{};
    cudaStream_t stream = nullptr;
    cudaGraphExec_t graph = nullptr;
    LmState* h_state = nullptr;   // pinned
    double* h_pinned = nullptr;   // pinned scratch (kRedN + 8 doubles)
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::vector<void*> allocs;    // device allocations tied to the task list
    std::vector<void*> fixed;     // device allocations for the lifetime of the handle
    int trace_cap = 0;
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
    double* l2buf = nullptr;
    size_t l2bytes = 0;
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
}  // namespace

template <typename T>
static cudaError_t dalloc(std::vector<void*>& pool, T** p, size_t count) {
    void* q = nullptr;
    cudaError_t e = cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T));
    if (e == cudaSuccess) { pool.push_back(q); *p = (T*)q; }
    return e;
}

static void free_pool(std::vector<void*>& pool) {
    for (void* p : pool) cudaFree(p);
    pool.clear();
}

static void drop_graph(calib_handle* h) {
    if (h->graph) { cudaGraphExecDestroy(h->graph); h->graph = nullptr; }
}

// Splits every view into near-equal tasks of <= seg points and (re)allocates what depends on it.
static int build_tasks(calib_handle* h, int seg) {
    drop_graph(h);
    free_pool(h->allocs);
    if (seg <= 0) {
        // automatic: about 6 waves of warp tasks over 148 SMs x 8 resident warps
        const int64_t target = 148 * 8 * 6;
        int64_t s = (h->N + target - 1) / std::max<int64_t>(target, 1);
        s = ((s + 31) / 32) * 32;
        seg = (int)std::min<int64_t>(512, std::max<int64_t>(64, s));
    }
    h->seg = seg;
    std::vector<Task> tasks;
    std::vector<int> off(h->M + 1, 0);
    for (int64_t v = 0; v < h->M; ++v) {
        const int64_t s0 = h->view_offsets[v], n = h->view_offsets[v + 1] - s0;
        const int64_t parts = (n + seg - 1) / seg;
        int64_t pos = s0;
        for (int64_t k = 0; k < parts; ++k) {
            const int64_t cnt = n / parts + (k < n % parts ? 1 : 0);
            tasks.push_back(Task{(int)v, (int)pos, (int)cnt, 0});
            pos += cnt;
        }
        off[v + 1] = (int)tasks.size();
    }
    Dev& d = h->dev;
    d.ntasks = (int)tasks.size();
    d.nblocks = std::max(1, (d.ntasks + kWarpsPerBlock - 1) / kWarpsPerBlock);
    d.Tpad = ((d.ntasks + 31) / 32) * 32 + 32;
    Task* dt = nullptr; int* doff = nullptr;
    CUDA_TRY(h, dalloc(h->allocs, &dt, tasks.size()));
    CUDA_TRY(h, dalloc(h->allocs, &doff, off.size()));
    CUDA_TRY(h, cudaMemcpy(dt, tasks.data(), tasks.size() * sizeof(Task), cudaMemcpyHostToDevice));
    CUDA_TRY(h, cudaMemcpy(doff, off.data(), off.size() * sizeof(int), cudaMemcpyHostToDevice));
    d.tasks = dt; d.view_task_off = doff;
    CUDA_TRY(h, dalloc(h->allocs, &d.partA, (size_t)d.nblocks * kNA));
    CUDA_TRY(h, dalloc(h->allocs, &d.partE, (size_t)d.nblocks));
    CUDA_TRY(h, dalloc(h->allocs, &d.partC, (size_t)d.nblocks));
    CUDA_TRY(h, dalloc(h->allocs, &d.partBD, (size_t)kNBD * d.Tpad));
    CUDA_TRY(h, cudaMemset(d.partBD, 0, (size_t)kNBD * d.Tpad * sizeof(double)));
    return CALIB_OK;
}

static int ensure_trace(calib_handle* h, int iters) {
    if (iters <= h->trace_cap) return CALIB_OK;
    drop_graph(h);
    double* t = nullptr;
    CUDA_TRY(h, cudaMalloc((void**)&t, (size_t)iters * 4 * sizeof(double)));
    if (h->dev.trace) cudaFree(h->dev.trace);
    h->dev.trace = t;
    h->trace_cap = iters;
    return CALIB_OK;
}

extern "C" const char* calib_last_error(const calib_t* h) {
    return h ? h->err.c_str() : g_last_error.c_str();
}

extern "C" int calib_create(calib_t** out, int model, int64_t M, const int64_t* view_offsets,
                            const double* model_pts, const double* sensor_pts, int device_id) {
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
    if (N > 0 && (!model_pts || !sensor_pts)) return fail(nullptr, CALIB_E_ARG, "calib_create: NULL points");
    if (N >= (1ll << 31) - 64 || M >= (1ll << 27)) return fail(nullptr, CALIB_E_ARG, "calib_create: too large");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0)
        return fail(nullptr, CALIB_E_CUDA, "calib_create: no CUDA device (this library has no CPU fallback)");
    if (device_id < 0 || device_id >= ndev) return fail(nullptr, CALIB_E_ARG, "calib_create: bad device_id");
    calib_handle* h = new calib_handle();
    h->model = model; h->L = model == 0 ? 10 : 9; h->device = device_id;
    h->M = M; h->N = N; h->K = h->L + 6 * M;
    h->view_offsets.assign(view_offsets, view_offsets + M + 1);
    auto bail = [&](int rc) { calib_destroy(h); return rc; };
#define TRY_(call) do { int rc__ = [&]() -> int { CUDA_TRY(h, call); return 0; }(); if (rc__) { g_last_error = h->err; return bail(rc__); } } while (0)
    TRY_(cudaSetDevice(device_id));
    TRY_(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    TRY_(cudaEventCreate(&h->ev0));
    TRY_(cudaEventCreate(&h->ev1));
    TRY_(cudaMallocHost((void**)&h->h_state, sizeof(LmState)));
    TRY_(cudaMallocHost((void**)&h->h_pinned, (kRedN + 8) * sizeof(double)));
    Dev& d = h->dev;
    d.model = model; d.L = h->L; d.M = (int)M; d.N = N; d.K = h->K;
    d.Mpad = (int)(((M + 31) / 32) * 32);
    // SoA repack on the host, one upload
    {
        std::vector<double> soa((size_t)5 * std::max<int64_t>(N, 1));
        for (int64_t i = 0; i < N; ++i) {
            soa[i] = model_pts[3 * i]; soa[N + i] = model_pts[3 * i + 1]; soa[2 * N + i] = model_pts[3 * i + 2];
            soa[3 * N + i] = sensor_pts[2 * i]; soa[4 * N + i] = sensor_pts[2 * i + 1];
        }
        double* pts = nullptr;
        TRY_(dalloc(h->fixed, &pts, soa.size() + 64));
        TRY_(cudaMemcpy(pts, soa.data(), soa.size() * sizeof(double), cudaMemcpyHostToDevice));
        d.X = pts; d.Y = pts + N; d.Z = pts + 2 * N; d.U = pts + 3 * N; d.V = pts + 4 * N;
    }
    TRY_(dalloc(h->fixed, &d.Pbuf, (size_t)2 * h->K));
    TRY_(cudaMemset(d.Pbuf, 0, (size_t)2 * h->K * sizeof(double)));
    TRY_(dalloc(h->fixed, &d.views, (size_t)2 * M));
    const int nvb = (int)((M + kSchurThreads - 1) / kSchurThreads);
    TRY_(dalloc(h->fixed, &d.partW, (size_t)nvb * (kSchurThreads / 32) * kNBD));
    TRY_(dalloc(h->fixed, &d.asum, (size_t)kNA));
    TRY_(dalloc(h->fixed, &d.red, (size_t)kRedN));
    TRY_(dalloc(h->fixed, &d.vtab, (size_t)vtab_items(h->L) * d.Mpad));
    TRY_(dalloc(h->fixed, &d.etrial, (size_t)8));
    TRY_(dalloc(h->fixed, &d.delta, (size_t)h->K));
    TRY_(dalloc(h->fixed, &d.st, (size_t)1));
    TRY_(cudaMemset(d.st, 0, sizeof(LmState)));
#undef TRY_
    int rc = build_tasks(h, 0);
    if (rc) { g_last_error = h->err; return bail(rc); }
    rc = ensure_trace(h, 128);
    if (rc) { g_last_error = h->err; return bail(rc); }
    *out = h;
    return CALIB_OK;
}

extern "C" void calib_destroy(calib_t* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    drop_graph(h);
    if (h->comm && h->comm_owned && g_nccl.ok) g_nccl.CommDestroy(h->comm);
    for (auto& pe : h->prof_events) { cudaEventDestroy(pe.second.first); cudaEventDestroy(pe.second.second); }
    free_pool(h->allocs);
    free_pool(h->fixed);
    if (h->dev.trace) cudaFree(h->dev.trace);
    if (h->l2buf) cudaFree(h->l2buf);
    if (h->h_state) cudaFreeHost(h->h_state);
    if (h->h_pinned) cudaFreeHost(h->h_pinned);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->stream) cudaStreamDestroy(h->stream);
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

// ---- launches ---------------------------------------------------------------------------------
namespace {

enum { KA = 0, KBD = 1, KCOST = 2, KSCHUR = 3, KSOLVE = 4 };

struct Launcher {
    calib_handle* h;
    bool prof;
    int64_t count = 0;
    void begin(int which, cudaEvent_t* a, cudaEvent_t* b) {
        if (!prof) return;
        cudaEventCreate(a); cudaEventCreate(b);
        cudaEventRecord(*a, h->stream);
        (void)which;
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
    ln.begin(KA, &a, &b);
    k_eval_A<MODEL><<<d.nblocks, kEvalThreads, 0, h->stream>>>(d);
    ln.end(KA, a, b);
    ln.begin(KBD, &a, &b);
    k_eval_BD<MODEL><<<d.nblocks, kEvalThreads, 0, h->stream>>>(d);
    ln.end(KBD, a, b);
    ln.count += 2;
}

template <int MODEL>
void enqueue_schur(calib_handle* h, Launcher& ln) {
    const Dev& d = h->dev;
    const int nvb = (d.M + kSchurThreads - 1) / kSchurThreads;
    cudaEvent_t a, b;
    ln.begin(KSCHUR, &a, &b);
    k_schur<MODEL><<<nvb + 1, kSchurThreads, 0, h->stream>>>(d);
    ln.end(KSCHUR, a, b);
    ln.count += 1;
}

template <int MODEL>
void enqueue_solve(calib_handle* h, Launcher& ln) {
    const Dev& d = h->dev;
    const int nvb = (d.M + kSchurThreads - 1) / kSchurThreads;
    cudaEvent_t a, b;
    ln.begin(KSOLVE, &a, &b);
    k_solve_backsub<MODEL><<<nvb, kSchurThreads, 0, h->stream>>>(d);
    ln.end(KSOLVE, a, b);
    ln.count += 1;
}

template <int MODEL>
void enqueue_cost(calib_handle* h, Launcher& ln, int fuse_accept) {
    const Dev& d = h->dev;
    cudaEvent_t a, b;
    ln.begin(KCOST, &a, &b);
    k_cost<MODEL><<<d.nblocks, kEvalThreads, 0, h->stream>>>(d, fuse_accept);
    ln.end(KCOST, a, b);
    ln.count += 1;
}

// one LM iteration: src/calibrate.py:144-168
template <int MODEL>
int enqueue_iteration(calib_handle* h, Launcher& ln) {
    const Dev& d = h->dev;
    enqueue_eval<MODEL>(h, ln);
    enqueue_schur<MODEL>(h, ln);
    if (h->comm) {
        if (g_nccl.AllReduce(d.red, d.red, kRedN, ncclDouble, ncclSum, h->comm, h->stream) != ncclSuccess)
            return fail(h, CALIB_E_COMM, "ncclAllReduce(reduced system) failed");
    }
    enqueue_solve<MODEL>(h, ln);
    enqueue_cost<MODEL>(h, ln, h->comm ? 0 : 1);
    if (h->comm) {
        if (g_nccl.AllReduce(d.etrial, d.etrial, 1, ncclDouble, ncclSum, h->comm, h->stream) != ncclSuccess)
            return fail(h, CALIB_E_COMM, "ncclAllReduce(trial cost) failed");
        k_accept<<<1, 1, 0, h->stream>>>(d);
        ln.count += 1;
    }
    return CALIB_OK;
}

int enqueue_iteration_any(calib_handle* h, Launcher& ln) {
    return h->model == 0 ? enqueue_iteration<0>(h, ln) : enqueue_iteration<1>(h, ln);
}

int kernels_per_iteration(const calib_handle* h) { return h->comm ? 6 : 5; }

int build_graph(calib_handle* h) {
    if (h->graph) return CALIB_OK;
    cudaGraph_t g = nullptr;
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

int upload_to(calib_handle* h, const double* P, int buf) {
    CUDA_TRY(h, cudaMemcpyAsync(h->dev.Pbuf + (size_t)buf * h->K, P, h->K * sizeof(double),
                                cudaMemcpyHostToDevice, h->stream));
    return CALIB_OK;
}

int push_state(calib_handle* h, double lam, double lam_min, double lam_max, double err_min, int max_iters) {
    LmState s;
    memset(&s, 0, sizeof s);
    s.lam = lam; s.lam_min = lam_min; s.lam_max = lam_max; s.err_min = err_min;
    s.max_iters = max_iters; s.need_eval = 1; s.cur = h->cur; s.nranks = h->nranks;
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

extern "C" int calib_lm_run_resident(calib_t* h, int maxIters, double lam0, double lam_min,
                                     double lam_max, double err_min, double* sse_out, int* iters_out) {
    RC(check_handle(h));
    if (maxIters <= 0)
        return fail(h, CALIB_E_STATE, "calib_lm_run: maxIters <= 0 (the reference raises UnboundLocalError)");
    RC(ensure_trace(h, maxIters));
    const bool prof = h->profiling;
    if (!prof) RC(build_graph(h));
    for (auto& pe : h->prof_events) { cudaEventDestroy(pe.second.first); cudaEventDestroy(pe.second.second); }
    h->prof_events.clear();
    for (int i = 0; i < 5; ++i) { h->ktime[i] = 0; h->kcount[i] = 0; }
    RC(push_state(h, lam0, lam_min, lam_max, err_min, maxIters));
    CUDA_TRY(h, cudaEventRecord(h->ev0, h->stream));
    RC(make_views(h, 1));
    Launcher ln{h, prof};
    ln.count = 1;
    const int kpi = kernels_per_iteration(h);
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
        batch = std::min(batch * 2, 8);
    }
    CUDA_TRY(h, cudaEventRecord(h->ev1, h->stream));
    RC(sync(h));
    float ms = 0;
    CUDA_TRY(h, cudaEventElapsedTime(&ms, h->ev0, h->ev1));
    h->last_ms = ms;
    h->last_launches = ln.count;
    h->last_evals = h->h_state->n_eval;
    h->cur = h->h_state->cur;
    for (auto& pe : h->prof_events) {
        float t = 0;
        cudaEventElapsedTime(&t, pe.second.first, pe.second.second);
        h->ktime[pe.first] += t; h->kcount[pe.first] += 1;
    }
    if (sse_out) *sse_out = h->h_state->sse;
    if (iters_out) *iters_out = h->h_state->iter;
    return CALIB_OK;
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
    if (trace) RC(calib_download_trace(h, iters, trace));
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
    k_debug_point<MODEL><<<d.nblocks, kEvalThreads, 0, h->stream>>>(d, y, r, Jc, Jv);
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
    if (y) CUDA_TRY(h, dalloc(tmp, &dy, 2 * N));
    if (r) CUDA_TRY(h, dalloc(tmp, &dr, 2 * N));
    if (Jc) CUDA_TRY(h, dalloc(tmp, &dJc, 2 * N * h->L));
    if (Jv) CUDA_TRY(h, dalloc(tmp, &dJv, 12 * N));
    if (h->model == 0) launch_debug_point<0>(h, dy, dr, dJc, dJv);
    else launch_debug_point<1>(h, dy, dr, dJc, dJv);
    int rc = sync(h);
    const size_t n = (size_t)h->N;
    if (!rc && y) cudaMemcpy(y, dy, 2 * n * sizeof(double), cudaMemcpyDeviceToHost);
    if (!rc && r) cudaMemcpy(r, dr, 2 * n * sizeof(double), cudaMemcpyDeviceToHost);
    if (!rc && Jc) cudaMemcpy(Jc, dJc, 2 * n * h->L * sizeof(double), cudaMemcpyDeviceToHost);
    if (!rc && Jv) cudaMemcpy(Jv, dJv, 12 * n * sizeof(double), cudaMemcpyDeviceToHost);
    free_pool(tmp);
    return rc;
}

// uploads P, evaluates both passes and the Schur stage with damping lam (red[] valid afterwards)
int eval_system(calib_handle* h, const double* P, double lam, bool solve) {
    RC(check_handle(h));
    if (!P) return fail(h, CALIB_E_ARG, "P is NULL");
    h->cur = 0;
    RC(upload_to(h, P, 0));
    RC(push_state(h, lam, 0, 0, 0, 1));
    RC(make_views(h, 1));
    Launcher ln{h, false};
    if (h->model == 0) { enqueue_eval<0>(h, ln); enqueue_schur<0>(h, ln); }
    else { enqueue_eval<1>(h, ln); enqueue_schur<1>(h, ln); }
    if (h->comm && g_nccl.AllReduce(h->dev.red, h->dev.red, kRedN, ncclDouble, ncclSum, h->comm, h->stream) != ncclSuccess)
        return fail(h, CALIB_E_COMM, "ncclAllReduce failed");
    if (solve) { if (h->model == 0) enqueue_solve<0>(h, ln); else enqueue_solve<1>(h, ln); }
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
    RC(upload_to(h, P, 1));
    RC(push_state(h, 0.0, 0, 0, 0, 1));
    RC(make_views(h, 0));
    Launcher ln{h, false};
    if (h->model == 0) enqueue_cost<0>(h, ln, 0); else enqueue_cost<1>(h, ln, 0);
    if (h->comm && g_nccl.AllReduce(h->dev.etrial, h->dev.etrial, 1, ncclDouble, ncclSum, h->comm, h->stream) != ncclSuccess)
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
    CUDA_TRY(h, dalloc(tmp, &dA, (size_t)L * L));
    CUDA_TRY(h, dalloc(tmp, &dB, M * L * 6));
    CUDA_TRY(h, dalloc(tmp, &dD, M * 36));
    CUDA_TRY(h, dalloc(tmp, &dg, (size_t)h->K));
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
    free_pool(tmp);
    return rc;
}

extern "C" int calib_lm_step(calib_t* h, const double* P, double lam, double* delta) {
    if (!h) return fail(nullptr, CALIB_E_ARG, "NULL handle");
    if (!delta) return fail(h, CALIB_E_ARG, "calib_lm_step: delta is NULL");
    RC(eval_system(h, P, lam, true));
    CUDA_TRY(h, cudaMemcpy(delta, h->dev.delta, (size_t)h->K * sizeof(double), cudaMemcpyDeviceToHost));
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
    return CALIB_OK;
}

// ---- batch of independent calibrations (calib_batch.cuh) -------------------------------------------
struct calib_batch_handle { std::string err; };
extern "C" int calib_batch_create(calib_batch_t** out, int, int64_t, const int64_t*, const int64_t*,
                                  const double*, const double*, int) {
    if (out) *out = nullptr;
    return fail(nullptr, CALIB_E_STATE, "calib_batch_create: not implemented in this build");
}
extern "C" void calib_batch_destroy(calib_batch_t* h) { delete h; }
extern "C" const char* calib_batch_last_error(const calib_batch_t* h) { return h ? h->err.c_str() : g_last_error.c_str(); }
extern "C" int calib_batch_lm_run(calib_batch_t*, double*, int, double, double, double, double, double*, int*, double*) {
    return fail(nullptr, CALIB_E_STATE, "calib_batch_lm_run: not implemented in this build");
}
