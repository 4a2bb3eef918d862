// Probe: do conditional WHILE graph nodes work on this driver, and what does one loop trip cost against the same
// kernels replayed from a flat graph?   nvcc -gencode arch=compute_100a,code=sm_100a -o /tmp/cond cond_graph_probe.cu
#include <cuda_runtime.h>
#include <cstdio>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s -> %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)
__global__ void work(int* cnt) { if (threadIdx.x == 0 && blockIdx.x == 0) cnt[1] += 1; }
__global__ void body(int* cnt, cudaGraphConditionalHandle h, int n) {
    int c = ++(*cnt);
    cudaGraphSetConditional(h, c < n ? 1 : 0);
}
__global__ void body_flat(int* cnt) { ++(*cnt); }
int main() {
    const int N = 2000;
    cudaStream_t s; CK(cudaStreamCreate(&s));
    int* d; CK(cudaMalloc(&d, 8)); CK(cudaMemset(d, 0, 8));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    // while graph
    cudaGraph_t g; CK(cudaGraphCreate(&g, 0));
    cudaGraphConditionalHandle h;
    CK(cudaGraphConditionalHandleCreate(&h, g, 1, cudaGraphCondAssignDefault));
    cudaGraphNodeParams p = {};
    p.type = cudaGraphNodeTypeConditional;
    p.conditional.handle = h; p.conditional.type = cudaGraphCondTypeWhile; p.conditional.size = 1;
    cudaGraphNode_t node;
    CK(cudaGraphAddNode(&node, g, nullptr, 0, &p));
    cudaGraph_t bodyg = p.conditional.phGraph_out[0];
    CK(cudaStreamBeginCaptureToGraph(s, bodyg, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
    work<<<148, 256, 0, s>>>(d);
    body<<<1, 32, 0, s>>>(d, h, N);
    CK(cudaStreamEndCapture(s, nullptr));
    cudaGraphExec_t x; CK(cudaGraphInstantiate(&x, g, 0));
    for (int rep = 0; rep < 3; ++rep) {
        CK(cudaMemsetAsync(d, 0, 8, s));
        CK(cudaEventRecord(e0, s));
        CK(cudaGraphLaunch(x, s));
        CK(cudaEventRecord(e1, s));
        CK(cudaStreamSynchronize(s));
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        int hc[2]; CK(cudaMemcpy(hc, d, 8, cudaMemcpyDeviceToHost));
        printf("while graph: %d trips (%d), %.3f ms, %.2f us per trip (2 kernels)\n", hc[0], hc[1], ms, 1e3 * ms / hc[0]);
    }
    // flat graph of the same 2N kernels
    cudaGraph_t gf; cudaGraphExec_t xf;
    CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
    for (int i = 0; i < N; ++i) { work<<<148, 256, 0, s>>>(d); body_flat<<<1, 32, 0, s>>>(d); }
    CK(cudaStreamEndCapture(s, &gf));
    CK(cudaGraphInstantiate(&xf, gf, 0));
    for (int rep = 0; rep < 3; ++rep) {
        CK(cudaMemsetAsync(d, 0, 8, s));
        CK(cudaEventRecord(e0, s));
        CK(cudaGraphLaunch(xf, s));
        CK(cudaEventRecord(e1, s));
        CK(cudaStreamSynchronize(s));
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        int hc[2]; CK(cudaMemcpy(hc, d, 8, cudaMemcpyDeviceToHost));
        printf("flat graph : %d trips (%d), %.3f ms, %.2f us per trip (2 kernels)\n", hc[0], hc[1], ms, 1e3 * ms / hc[0]);
    }
    // plain stream launches
    for (int rep = 0; rep < 2; ++rep) {
        CK(cudaMemsetAsync(d, 0, 8, s));
        CK(cudaEventRecord(e0, s));
        for (int i = 0; i < N; ++i) { work<<<148, 256, 0, s>>>(d); body_flat<<<1, 32, 0, s>>>(d); }
        CK(cudaEventRecord(e1, s));
        CK(cudaStreamSynchronize(s));
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("stream     : %.3f ms, %.2f us per trip (2 kernels)\n", ms, 1e3 * ms / N);
    }
    return 0;
}
