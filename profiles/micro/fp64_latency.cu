// Micro-benchmarks behind the evaluation pass's design (run: profiles/micro/fp64_latency on a B200):
//   dependent-issue latency of DFMA / DMUL / DADD, of the double reciprocal and rsqrt sequences,
//   FP64 issue rate of one warp alone and of two warps on one scheduler (ILP 1..8 chains per thread).
#include <cstdio>
#include <cuda_runtime.h>

__global__ void lat_dfma(double* out, long long* cyc, int n, double a, double b) {
    double x = threadIdx.x;
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) { x = fma(x, a, b); x = fma(x, a, b); x = fma(x, a, b); x = fma(x, a, b); }
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void lat_dadd(double* out, long long* cyc, int n, double b) {
    double x = threadIdx.x;
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) { x = x + b; x = x + b; x = x + b; x = x + b; }
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void lat_rcp(double* out, long long* cyc, int n) {
    double x = 1.5 + threadIdx.x;
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) { x = __drcp_rn(x); x = __drcp_rn(x); x = __drcp_rn(x); x = __drcp_rn(x); }
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void lat_rsqrt(double* out, long long* cyc, int n) {
    double x = 1.5 + threadIdx.x;
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) { x = rsqrt(x); x = rsqrt(x); x = rsqrt(x); x = rsqrt(x); }
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
// ILP independent chains per thread; blockDim warps all on one SM
template <int ILP>
__global__ void thr_dfma(double* out, long long* cyc, int n, double a, double b) {
    double x[ILP];
#pragma unroll
    for (int k = 0; k < ILP; ++k) x[k] = threadIdx.x + k;
    __syncthreads();
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) {
#pragma unroll
        for (int k = 0; k < ILP; ++k) x[k] = fma(x[k], a, b);
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int k = 0; k < ILP; ++k) s += x[k];
    out[threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
int main() {
    double* out; long long* cyc;
    cudaMalloc(&out, 1024 * 8); cudaMallocManaged(&cyc, 8);
    const int n = 4096;
    for (int rep = 0; rep < 2; ++rep) {
        lat_dfma<<<1, 32>>>(out, cyc, n, 1.0000001, 1e-9); cudaDeviceSynchronize();
        if (rep) printf("DFMA dependent latency      %6.2f cycles\n", (double)cyc[0] / (4.0 * n));
        lat_dadd<<<1, 32>>>(out, cyc, n, 1e-9); cudaDeviceSynchronize();
        if (rep) printf("DADD dependent latency      %6.2f cycles\n", (double)cyc[0] / (4.0 * n));
        lat_rcp<<<1, 32>>>(out, cyc, n); cudaDeviceSynchronize();
        if (rep) printf("__drcp_rn dependent latency %6.2f cycles\n", (double)cyc[0] / (4.0 * n));
        lat_rsqrt<<<1, 32>>>(out, cyc, n); cudaDeviceSynchronize();
        if (rep) printf("rsqrt(double) dep. latency  %6.2f cycles\n", (double)cyc[0] / (4.0 * n));
    }
#define THR(ILP, W) thr_dfma<ILP><<<1, 32 * W>>>(out, cyc, n, 1.0000001, 1e-9); cudaDeviceSynchronize(); \
    thr_dfma<ILP><<<1, 32 * W>>>(out, cyc, n, 1.0000001, 1e-9); cudaDeviceSynchronize(); \
    printf("warps/SM %2d (per scheduler %d)  ILP %d: %6.2f cycles per DFMA per warp, scheduler issues one DFMA per %5.2f cycles\n", W, (W + 3) / 4, ILP, \
           (double)cyc[0] / ((double)n * ILP), (double)cyc[0] / ((double)n * ILP * ((W + 3) / 4)));
    THR(1, 1) THR(2, 1) THR(4, 1) THR(8, 1)
    THR(1, 8) THR(2, 8) THR(4, 8) THR(8, 8)
    THR(1, 16) THR(2, 16) THR(4, 16) THR(8, 16)
    return 0;
}
