// probe: time of a chain of small-N tcgen05.mma (bf16, SS) under different accumulator / shape choices
#include <cstdio>
#include "../../chatterbox-vllm2_b200/csrc/t3b_common.cuh"
void t3b_set_error(const char*, ...) {}
int t3b_num_sms() { return 148; }
__device__ __forceinline__ bool elect_one() {
    uint32_t p;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(p));
    return p != 0;
}
__global__ void __launch_bounds__(128, 1) k(int M, int N, int nacc, int nmma, int a_stride, long long* out) {
    extern __shared__ unsigned char sm_[];
    unsigned char* base = (unsigned char*)(((uintptr_t)sm_ + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar; __shared__ uint32_t tptr;
    for (int i = threadIdx.x; i < 160 * 1024 / 4; i += 128) ((uint32_t*)base)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    if (threadIdx.x < 32) tmem_alloc(&tptr, 512);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before(); __syncthreads(); tc_fence_after();
    if (threadIdx.x < 32) {          // whole warp, uniform control flow; one elected lane issues
        const uint32_t idesc = umma_idesc_bf16(M, N);
        const uint32_t tp = tptr;
        for (int rep = 0; rep < 3; ++rep) {
            long long t0 = clock64();
            for (int i = 0; i < nmma; ++i) {
                const uint32_t a = smem_u32(base + (size_t)(i / 4) * a_stride), b = smem_u32(base + 140 * 1024);
                const uint64_t ad = umma_smem_desc_sw128(a) + 2 * (i & 3), bd = umma_smem_desc_sw128(b) + 2 * (i & 3);
                if (elect_one()) umma_bf16(tp + (i % nacc) * 32, ad, bd, idesc, i >= nacc);
            }
            long long t1 = clock64();
            if (elect_one()) umma_commit(&bar);
            mbar_wait(&bar, rep & 1);
            long long t2 = clock64();
            if (rep == 2 && threadIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
        }
    }
    tc_fence_before(); __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc(tptr, 512);
}
int main() {
    long long* d; cudaMalloc(&d, 16); long long h[2];
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    struct { int M, N, nacc, astride; const char* name; } cs[] = {
        {128, 16, 1, 16384, "M128 N16 1 acc"}, {128, 16, 4, 16384, "M128 N16 4 acc"}, {64, 16, 1, 8192, "M64  N16 1 acc"},
        {64, 16, 4, 8192, "M64  N16 4 acc"}, {128, 64, 1, 16384, "M128 N64 1 acc"}, {128, 64, 4, 16384, "M128 N64 4 acc"},
        {128, 128, 1, 16384, "M128 N128 1 acc"}, {64, 64, 1, 8192, "M64  N64 1 acc"}, {128, 16, 1, 0, "M128 N16 same A"}};
    for (auto& c : cs) {
        k<<<1, 128, 200 * 1024>>>(c.M, c.N, c.nacc, 32, c.astride, d);
        cudaError_t e = cudaDeviceSynchronize();
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        printf("%-18s 32 MMAs: issue %lld cyc, complete %lld cyc (%.0f cyc/MMA) %s\n", c.name, h[0], h[1], h[1] / 32.0, cudaGetErrorString(e));
    }
    return 0;
}
