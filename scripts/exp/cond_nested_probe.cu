// Probe: nested conditional WHILE nodes built with capture-to-graph in pieces (outer body part 1 -> inner WHILE -> outer
// body part 2), a cluster kernel inside the inner body, handles created on the ROOT graph or on the body graph.
#include <cuda_runtime.h>
#include <cooperative_groups.h>
#include <cstdio>
namespace cg = cooperative_groups;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s -> %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)
__global__ void a_kernel(int* c, cudaGraphConditionalHandle hin) { c[0] += 1; c[3] = 0; cudaGraphSetConditional(hin, 1); }
__global__ void __cluster_dims__(4, 1, 1) inner_kernel(int* c, cudaGraphConditionalHandle hin) {
    cg::this_cluster().sync();
    if (blockIdx.x == 0 && threadIdx.x == 0) { c[1] += 1; c[3] += 1; cudaGraphSetConditional(hin, c[3] < 3 ? 1 : 0); }
}
__global__ void b_kernel(int* c, cudaGraphConditionalHandle hout) { c[2] += 1; cudaGraphSetConditional(hout, c[2] < 5 ? 1 : 0); }
static int build(int handle_on_root, cudaStream_t s, int* d) {
    cudaGraph_t g; CK(cudaGraphCreate(&g, 0));
    cudaGraphConditionalHandle hout, hin;
    CK(cudaGraphConditionalHandleCreate(&hout, g, 1, cudaGraphCondAssignDefault));
    cudaGraphNodeParams p = {};
    p.type = cudaGraphNodeTypeConditional;
    p.conditional.handle = hout; p.conditional.type = cudaGraphCondTypeWhile; p.conditional.size = 1;
    cudaGraphNode_t outer; CK(cudaGraphAddNode(&outer, g, nullptr, 0, &p));
    cudaGraph_t bl = p.conditional.phGraph_out[0];
    CK(cudaGraphConditionalHandleCreate(&hin, handle_on_root ? g : bl, 0, 0));
    // part 1
    CK(cudaStreamBeginCaptureToGraph(s, bl, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
    a_kernel<<<1, 1, 0, s>>>(d, hin);
    cudaStreamCaptureStatus st; const cudaGraphNode_t* deps = nullptr; size_t nd = 0; cudaGraph_t cg_ = nullptr;
    CK(cudaStreamGetCaptureInfo(s, &st, nullptr, &cg_, &deps, &nd));
    cudaGraphNode_t last = nd ? deps[0] : nullptr;
    CK(cudaStreamEndCapture(s, nullptr));
    cudaGraphNodeParams q = {};
    q.type = cudaGraphNodeTypeConditional;
    q.conditional.handle = hin; q.conditional.type = cudaGraphCondTypeWhile; q.conditional.size = 1;
    cudaGraphNode_t inner; CK(cudaGraphAddNode(&inner, bl, &last, 1, &q));
    cudaGraph_t bp = q.conditional.phGraph_out[0];
    CK(cudaStreamBeginCaptureToGraph(s, bp, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
    inner_kernel<<<4, 64, 0, s>>>(d, hin);
    CK(cudaStreamEndCapture(s, nullptr));
    CK(cudaStreamBeginCaptureToGraph(s, bl, &inner, nullptr, 1, cudaStreamCaptureModeThreadLocal));
    b_kernel<<<1, 1, 0, s>>>(d, hout);
    CK(cudaStreamEndCapture(s, nullptr));
    cudaGraphExec_t x; CK(cudaGraphInstantiate(&x, g, 0));
    for (int rep = 0; rep < 2; ++rep) {
        CK(cudaMemsetAsync(d, 0, 16, s));
        CK(cudaGraphLaunch(x, s));
        CK(cudaStreamSynchronize(s));
        int hc[4]; CK(cudaMemcpy(hc, d, 16, cudaMemcpyDeviceToHost));
        printf("handle on %s: outer part1 %d, inner trips %d, outer part2 %d (want 5, 15, 5)\n", handle_on_root ? "root" : "body", hc[0], hc[1], hc[2]);
    }
    return 0;
}
int main() {
    cudaStream_t s; CK(cudaStreamCreate(&s));
    int* d; CK(cudaMalloc(&d, 16));
    int r0 = build(1, s, d);
    cudaGetLastError();
    int r1 = build(0, s, d);
    printf("root %d body %d\n", r0, r1);
    return 0;
}
