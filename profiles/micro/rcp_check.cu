// Is the branch-free reciprocal of calib_device.cuh (MUFU seed + two Newton steps) the correctly rounded one?
// Counts the inputs on which it differs from __drcp_rn (IEEE round-to-nearest) and from 1.0 / x.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o rcp_check rcp_check.cu && ./rcp_check
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../camera_calibration_b200/csrc/calib_device.cuh"


__device__ uint64_t splitmix(uint64_t& s) {
    uint64_t z = (s += 0x9e3779b97f4a7c15ull);
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}

__global__ void k(unsigned long long* diff, unsigned long long* diff_rsq, int per_thread, int mode) {
    uint64_t s = 0x1234567ull + 7919ull * (blockIdx.x * blockDim.x + threadIdx.x);
    unsigned long long d = 0, dr = 0;
    for (int i = 0; i < per_thread; ++i) {
        const uint64_t r = splitmix(s);
        double x;
        if (mode == 0) x = 0.05 + 100.0 * (double)(r >> 11) * (1.0 / 9007199254740992.0);      // depths
        else x = __longlong_as_double((long long)((r & 0x800fffffffffffffull) | ((uint64_t)(1023 - 300 + (r >> 52) % 600) << 52)));
        if (calib::rcp_nr(x) != __drcp_rn(x)) ++d;
        const double ax = fabs(x);
        if (calib::rsqrt_nr(ax) != rsqrt(ax)) ++dr;
    }
    atomicAdd(diff, d);
    atomicAdd(diff_rsq, dr);
}

int main() {
    unsigned long long *d, h[2];
    cudaMalloc(&d, 16);
    for (int mode = 0; mode < 2; ++mode) {
        cudaMemset(d, 0, 16);
        k<<<1024, 256>>>(d, d + 1, 4096, mode);
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        printf("mode %d (%s): %llu of %llu reciprocals differ from __drcp_rn, %llu rsqrt differ from libdevice rsqrt\n", mode,
               mode ? "random sign, exponents 2^-300 .. 2^300" : "uniform in [0.05, 100]", h[0], 1024ull * 256 * 4096, h[1]);
    }
    return 0;
}
