// Probe: how fast can 148 SMs pull scattered 2..16 KB pieces out of HBM with cp.async.bulk (1-D bulk copies into shared memory)
// versus plain 128-bit loads, as a function of piece size and of the bytes each SM keeps in flight?  (Sizing the K/V ring of
// csrc/attention.cu and the weight streams of the GEMMs.)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/probe/bin/bulk_bw scripts/probe/bulk_bw.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(b)), "r"(c)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t n) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(b)), "r"(n) : "memory"); }
__device__ __forceinline__ bool mbar_try(uint64_t* b, uint32_t ph) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(b)), "r"(ph) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void bulk(void* dst, const void* src, uint32_t n, uint64_t* b) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" :: "r"(smem_u32(dst)), "l"(src), "r"(n), "r"(smem_u32(b)) : "memory");
}

// every warp owns `slots` ring slots of `piece` bytes; pieces are taken from `src` at a scattered (hashed) piece index
__global__ void __launch_bounds__(1024, 1) bulk_kernel(const unsigned char* src, size_t n_pieces, int piece, int slots, int per_warp, unsigned* sink) {
    extern __shared__ __align__(128) unsigned char sm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    unsigned char* ring = sm + (size_t)warp * slots * piece;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm + (size_t)nw * slots * piece) + warp * slots;
    if (lane == 0) for (int i = 0; i < slots; ++i) mbar_init(&bars[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncwarp();
    const size_t gw = (size_t)blockIdx.x * nw + warp, tw = (size_t)gridDim.x * nw;
    auto idx = [&](int i) { return ((gw + (size_t)i * tw) * 2654435761ull) % n_pieces; };
    unsigned acc = 0;
    for (int i = 0; i < slots && i < per_warp; ++i)
        if (lane == 0) { mbar_expect(&bars[i], piece); bulk(ring + (size_t)i * piece, src + idx(i) * piece, piece, &bars[i]); }
    for (int i = 0; i < per_warp; ++i) {
        const int s = i % slots;
        while (!mbar_try(&bars[s], (i / slots) & 1)) {}
        acc += reinterpret_cast<const unsigned*>(ring + (size_t)s * piece)[lane];
        __syncwarp();
        if (i + slots < per_warp && lane == 0) { mbar_expect(&bars[s], piece); bulk(ring + (size_t)s * piece, src + idx(i + slots) * piece, piece, &bars[s]); }
    }
    if (acc == 0x12345678u) *sink = acc;
}

// attention-like: a slot = G pieces of `piece` bytes on one mbarrier; issued by lane 0 alone (par = 0) or by lanes 0..G-1 (par = 1)
__global__ void __launch_bounds__(1024, 1) group_kernel(const unsigned char* src, size_t n_pieces, int piece, int G, int slots, int per_warp, int par,
                                                        unsigned* sink, long long* cyc) {
    extern __shared__ __align__(128) unsigned char sm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    const int slot_b = G * piece;
    unsigned char* ring = sm + (size_t)warp * slots * slot_b;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm + (size_t)nw * slots * slot_b) + warp * slots;
    if (lane == 0) for (int i = 0; i < slots; ++i) mbar_init(&bars[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncwarp();
    const size_t gw = (size_t)blockIdx.x * nw + warp, tw = (size_t)gridDim.x * nw;
    auto idx = [&](int i, int g) { return (((gw + (size_t)i * tw) * G + g) * 2654435761ull) % n_pieces; };
    long long issue_cyc = 0;
    auto issue = [&](int i, int s) {
        const long long t0 = clock64();
        if (par) {
            if (lane == 0) mbar_expect(&bars[s], slot_b);
            __syncwarp();
            if (lane < G) bulk(ring + (size_t)s * slot_b + lane * piece, src + idx(i, lane) * piece, piece, &bars[s]);
        } else if (lane == 0) {
            mbar_expect(&bars[s], slot_b);
            for (int g = 0; g < G; ++g) bulk(ring + (size_t)s * slot_b + g * piece, src + idx(i, g) * piece, piece, &bars[s]);
        }
        issue_cyc += clock64() - t0;
    };
    unsigned acc = 0;
    for (int i = 0; i < slots && i < per_warp; ++i) issue(i, i);
    for (int i = 0; i < per_warp; ++i) {
        const int s = i % slots;
        while (!mbar_try(&bars[s], (i / slots) & 1)) {}
        acc += reinterpret_cast<const unsigned*>(ring + (size_t)s * slot_b)[lane];
        __syncwarp();
        if (i + slots < per_warp) issue(i + slots, s);
    }
    if (acc == 0x12345678u) *sink = acc;
    if (blockIdx.x == 0 && threadIdx.x == 0) *cyc = issue_cyc / per_warp;
}

// the same pieces with 128-bit loads: a warp reads a piece with `piece / 512` loads per lane, `unroll` pieces in flight
template <int U>
__global__ void __launch_bounds__(1024, 1) ldg_kernel(const unsigned char* src, size_t n_pieces, int piece, int per_warp, unsigned* sink) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    const size_t gw = (size_t)blockIdx.x * nw + warp, tw = (size_t)gridDim.x * nw;
    unsigned acc = 0;
    const int lp = piece / 512;
    for (int i = 0; i < per_warp; i += U) {
        uint4 v[U][4];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t p = ((gw + (size_t)(i + u) * tw) * 2654435761ull) % n_pieces;
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (j < lp) asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u][j].x), "=r"(v[u][j].y), "=r"(v[u][j].z), "=r"(v[u][j].w)
                                         : "l"(src + p * piece + j * 512 + lane * 16));
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int j = 0; j < 4; ++j) if (j < lp) acc += v[u][j].x ^ v[u][j].w;
    }
    if (acc == 0x12345678u) *sink = acc;
}

int main() {
    const size_t bytes = 4ull << 30;
    unsigned char* src; unsigned* sink;
    cudaMalloc(&src, bytes); cudaMalloc(&sink, 4);
    cudaMemset(src, 1, bytes);
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const size_t total = 1ull << 30;            // bytes moved per measurement
    printf("mode piece warps slots inflight_KB_per_SM GB/s\n");
    cudaFuncSetAttribute(bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    for (int piece : {2048, 4096, 8192, 16384}) {
        for (int warps : {4, 8, 16, 32}) {
            for (int slots : {1, 2, 3, 4, 6, 8}) {
                const size_t smem = (size_t)warps * slots * piece + warps * slots * 8;
                if (smem > 216 * 1024) continue;
                const int per_warp = (int)(total / piece / ((size_t)sms * warps));
                float best = 1e9f;
                for (int rep = 0; rep < 3; ++rep) {
                    cudaEventRecord(e0);
                    bulk_kernel<<<sms, warps * 32, smem>>>(src, bytes / piece, piece, slots, per_warp, sink);
                    cudaEventRecord(e1); cudaEventSynchronize(e1);
                    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
                }
                const double moved = (double)per_warp * piece * sms * warps;
                printf("bulk %5d %2d %d %6.1f %7.1f\n", piece, warps, slots, warps * slots * piece / 1024.0, moved / best / 1e6);
            }
        }
    }
    if (getenv("BULK_ONLY_GROUP")) {
        long long* cyc; cudaMalloc(&cyc, 8);
        cudaFuncSetAttribute(group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        printf("group piece G warps slots par inflight_KB GB/s issue_cycles_per_slot\n");
        for (int piece : {1024, 2048, 4096})
            for (int G : {2, 4})
                for (int warps : {7, 8, 14})
                    for (int slots : {2, 3})
                        for (int par : {0, 1}) {
                            const size_t smem = (size_t)warps * slots * G * piece + warps * slots * 8;
                            if (smem > 216 * 1024) continue;
                            const int per_warp = (int)(total / (G * piece) / ((size_t)sms * warps));
                            float best = 1e9f;
                            for (int rep = 0; rep < 3; ++rep) {
                                cudaEventRecord(e0);
                                group_kernel<<<sms, warps * 32, smem>>>(src, bytes / piece, piece, G, slots, per_warp, par, sink, cyc);
                                cudaEventRecord(e1); cudaEventSynchronize(e1);
                                float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
                            }
                            long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
                            printf("group %5d %d %2d %d %d %6.1f %7.1f %lld\n", piece, G, warps, slots, par, warps * slots * G * piece / 1024.0,
                                   (double)per_warp * G * piece * sms * warps / best / 1e6, h);
                        }
        printf("status %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
        return 0;
    }
    for (int piece : {2048}) {
        for (int warps : {8, 16, 32}) {
            const int per_warp = (int)(total / piece / ((size_t)sms * warps)) & ~7;
            for (int U : {1, 2, 4, 8}) {
                float best = 1e9f;
                for (int rep = 0; rep < 3; ++rep) {
                    cudaEventRecord(e0);
                    if (U == 1) ldg_kernel<1><<<sms, warps * 32>>>(src, bytes / piece, piece, per_warp, sink);
                    if (U == 2) ldg_kernel<2><<<sms, warps * 32>>>(src, bytes / piece, piece, per_warp, sink);
                    if (U == 4) ldg_kernel<4><<<sms, warps * 32>>>(src, bytes / piece, piece, per_warp, sink);
                    if (U == 8) ldg_kernel<8><<<sms, warps * 32>>>(src, bytes / piece, piece, per_warp, sink);
                    cudaEventRecord(e1); cudaEventSynchronize(e1);
                    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
                }
                const double moved = (double)per_warp * piece * sms * warps;
                printf("ldg  %5d %2d %d %6.1f %7.1f\n", piece, warps, U, warps * U * piece / 1024.0, moved / best / 1e6);
            }
        }
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status %s\n", cudaGetErrorString(e));
    return 0;
}
