// Does an FP64 instruction occupy a scheduler's issue port for one cycle or for two?
// Per scheduler: W warps, each with 8 independent DFMA chains and, per DFMA, K independent 32-bit integer
// instructions (LOP3/IADD on private registers).  If the FP64 pipe's half rate only limits DFMA-to-DFMA issue, K = 1
// costs nothing (2 cycles per DFMA + 1 ALU slot in the shadow); if every DFMA blocks the port for 2 cycles, K = 1
// gives 3 cycles per pair.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_issue_mix fp64_issue_mix.cu && ./fp64_issue_mix
#include <cstdio>
#include <cuda_runtime.h>

template <int K>
__global__ void k(double* out, unsigned* iout, int iters, double a, double b, long long* cycles) {
    double x[8];
    unsigned u[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { x[i] = threadIdx.x + i; u[i] = threadIdx.x * 7 + i; }
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            x[i] = fma(x[i], a, b);
#pragma unroll
            for (int j = 0; j < K; ++j) u[(i + j) & 7] = (u[(i + j) & 7] ^ (unsigned)it) + 0x9e3779b9u;
        }
    }
    const long long t1 = clock64();
    double s = 0; unsigned v = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) { s += x[i]; v ^= u[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    iout[blockIdx.x * blockDim.x + threadIdx.x] = v;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int K> void run(int warps_per_sched, double* out, unsigned* iout, long long* dc) {
    const int iters = 20000, threads = 128 * warps_per_sched;
    k<K><<<148, threads>>>(out, iout, 100, 1.0000001, 1e-9, dc);
    k<K><<<148, threads>>>(out, iout, iters, 1.0000001, 1e-9, dc);
    long long c;
    cudaMemcpy(&c, dc, 8, cudaMemcpyDeviceToHost);
    const double per = (double)c / ((double)iters * 8 * warps_per_sched);
    printf("warps/scheduler %d, %d integer instruction(s) per DFMA: %.2f cycles per DFMA (+%d ALU) per scheduler\n", warps_per_sched, K, per, K);
}

int main() {
    double* out; unsigned* iout; long long* dc;
    cudaMalloc(&out, 148 * 1024 * 8); cudaMalloc(&iout, 148 * 1024 * 4); cudaMalloc(&dc, 8);
    for (int w = 1; w <= 4; w *= 2) { run<0>(w, out, iout, dc); run<1>(w, out, iout, dc); run<2>(w, out, iout, dc); run<4>(w, out, iout, dc); }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
