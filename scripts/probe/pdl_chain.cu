// Probe: cost of one programmatic-dependent-launch boundary on B200.  A chain of N kernels on one stream, each waiting for its
// predecessor; variants: (0) trigger at top, wait, exit; (1) + one dependent 16-byte load/store per thread (L2 round trip);
// (2) like 1 but trigger AFTER the wait (the r01d policy); (3) plain launches without the PDL attribute.
#include <cuda_runtime.h>
#include <cstdio>
__global__ void k(const uint4* in, uint4* out, int mode) {
    if (mode != 2) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (mode == 2) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (mode >= 1) {
        uint4 v = __ldcg(in + blockIdx.x * blockDim.x + threadIdx.x);
        v.x += 1;
        out[blockIdx.x * blockDim.x + threadIdx.x] = v;
    }
}
int main() {
    uint4 *a, *b;
    cudaMalloc(&a, 1 << 20); cudaMalloc(&b, 1 << 20);
    cudaMemset(a, 0, 1 << 20); cudaMemset(b, 0, 1 << 20);
    cudaStream_t s; cudaStreamCreate(&s);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int N = 200;
    for (int blocks : {64, 1024}) for (int mode = 0; mode < 4; ++mode) {
        cudaGraph_t g; cudaGraphExec_t ge;
        cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
        for (int i = 0; i < N; ++i) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(blocks); cfg.blockDim = dim3(128); cfg.stream = s;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = at; cfg.numAttrs = mode == 3 ? 0 : 1;
            cudaLaunchKernelEx(&cfg, k, (const uint4*)((i & 1) ? b : a), (i & 1) ? a : b, mode == 3 ? 1 : mode);
        }
        cudaStreamEndCapture(s, &g);
        cudaGraphInstantiate(&ge, g, 0);
        for (int w = 0; w < 3; ++w) cudaGraphLaunch(ge, s);
        cudaStreamSynchronize(s);
        cudaEventRecord(e0, s);
        for (int w = 0; w < 5; ++w) cudaGraphLaunch(ge, s);
        cudaEventRecord(e1, s);
        cudaStreamSynchronize(s);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("blocks %4d mode %d: %.2f us per kernel (%s)\n", blocks, mode, ms * 1e3 / (5 * N), cudaGetErrorString(cudaGetLastError()));
        cudaGraphExecDestroy(ge); cudaGraphDestroy(g);
    }
    return 0;
}
