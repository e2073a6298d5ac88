// Can the kernel nodes INSIDE the body of a WHILE conditional node be re-pointed in an instantiated graph
// (cudaGraphExecKernelNodeSetParams), and what does it cost against instantiating again?
#include <chrono>
#include <cstdio>
#include <cuda_runtime.h>
#include <vector>
struct Big { double* out; int n; double pad[40]; cudaGraphConditionalHandle cond; };
__global__ void k_body(Big b, int add) {
    int* cnt = (int*)(b.out + 1);
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        b.out[0] += add;
        *cnt += 1;
        cudaGraphSetConditional(b.cond, *cnt < b.n ? 1u : 0u);
    }
}
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s -> %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
int main() {
    cudaStream_t s; CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    double *a, *b; CK(cudaMalloc(&a, 64)); CK(cudaMalloc(&b, 64)); CK(cudaMemset(a, 0, 64)); CK(cudaMemset(b, 0, 64));
    auto t0 = std::chrono::steady_clock::now();
    cudaGraph_t g; CK(cudaGraphCreate(&g, 0));
    cudaGraphConditionalHandle h; CK(cudaGraphConditionalHandleCreate(&h, g, 1, cudaGraphCondAssignDefault));
    cudaGraphNodeParams np = {cudaGraphNodeTypeConditional};
    np.conditional.handle = h; np.conditional.type = cudaGraphCondTypeWhile; np.conditional.size = 1;
    cudaGraphNode_t node; CK(cudaGraphAddNode(&node, g, nullptr, 0, &np));
    cudaGraph_t body = np.conditional.phGraph_out[0];
    Big pa{a, 5, {}, h};
    CK(cudaStreamBeginCaptureToGraph(s, body, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
    k_body<<<1, 32, 0, s>>>(pa, 1);
    k_body<<<2, 32, 0, s>>>(pa, 0);
    cudaGraph_t got; CK(cudaStreamEndCapture(s, &got));
    cudaGraphExec_t ex; CK(cudaGraphInstantiate(&ex, g, 0));
    auto t1 = std::chrono::steady_clock::now();
    CK(cudaGraphLaunch(ex, s)); CK(cudaStreamSynchronize(s));
    double ha[2]; CK(cudaMemcpy(ha, a, 16, cudaMemcpyDeviceToHost));
    printf("first graph: out=%g (expect 5), build+instantiate %.1f us\n", ha[0], std::chrono::duration<double, std::micro>(t1 - t0).count());
    // patch: point both kernels at b, 7 iterations
    size_t n = 0; CK(cudaGraphGetNodes(body, nullptr, &n));
    std::vector<cudaGraphNode_t> nodes(n); CK(cudaGraphGetNodes(body, nodes.data(), &n));
    printf("body has %zu nodes\n", n);
    Big pb{b, 14, {}, h};     // two kernels per iteration increment the counter twice
    auto t2 = std::chrono::steady_clock::now();
    for (size_t i = 0; i < n; ++i) {
        cudaKernelNodeParams kp; CK(cudaGraphKernelNodeGetParams(nodes[i], &kp));
        int add = *(int*)kp.kernelParams[1];
        void* args[2] = {&pb, &add};
        kp.kernelParams = args;
        cudaError_t e = cudaGraphExecKernelNodeSetParams(ex, nodes[i], &kp);
        if (e != cudaSuccess) { printf("cudaGraphExecKernelNodeSetParams on a WHILE-body node: %s\n", cudaGetErrorString(e)); return 2; }
    }
    auto t3 = std::chrono::steady_clock::now();
    CK(cudaGraphLaunch(ex, s)); CK(cudaStreamSynchronize(s));
    double hb[2]; CK(cudaMemcpy(hb, b, 16, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(ha, a, 16, cudaMemcpyDeviceToHost));
    printf("patched graph: b.out=%g (expect 7), a.out=%g (expect 5 still), patch %.1f us\n", hb[0], ha[0],
           std::chrono::duration<double, std::micro>(t3 - t2).count());
    return 0;
}
