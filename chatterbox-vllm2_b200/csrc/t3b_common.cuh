// Shared device/host helpers for libt3b200 (sm_100a only).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/t3b200.h"

// ------------------------------------------------------------------ host side
void t3b_set_error(const char* fmt, ...);
#define T3B_FAIL(code, ...) do { t3b_set_error(__VA_ARGS__); return (code); } while (0)
#define T3B_CHECK_ARG(cond, ...) do { if (!(cond)) T3B_FAIL(T3B_ERR_ARG, __VA_ARGS__); } while (0)
#define T3B_CHECK_SHAPE(cond, ...) do { if (!(cond)) T3B_FAIL(T3B_ERR_SHAPE, __VA_ARGS__); } while (0)
#define T3B_CHECK_ALIGN(p) do { if ((reinterpret_cast<uintptr_t>(p) & 15u) != 0) \
        T3B_FAIL(T3B_ERR_ALIGN, "%s: pointer %s not 16-byte aligned", __func__, #p); } while (0)
#define T3B_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
        T3B_FAIL(T3B_ERR_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); } while (0)
#define T3B_LAUNCH_CHECK() T3B_CUDA(cudaPeekAtLastError())

static inline cudaStream_t t3b_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

// A region of weights the NEXT kernels of the decode chain will stream (t3b_decode_step hands every launch the weights of the
// projection two launches ahead): the kernel's blocks pull their slices into L2 with bulk prefetches before they wait for their
// own inputs, so HBM keeps streaming while the dependent chain (wait -> load x -> MMA -> exchange -> store) runs from L2.
// bytes == 0: nothing.  A hint only - never affects results.
// F2.1 folded into the tcgen05 projections of the decode chain (linear_umma.cu; rows <= 128, hidden = 8 tiles of 128).
// resid != nullptr: producer (o / down) - the epilogue adds the residual, rewrites it, and leaves the fp32 sum + per-tile sums of squares.
// xs32 != nullptr: consumer (qkv / gate_up) - the activation operand is normalised from those on the fly.
struct UmmaFuse {
    void* resid;             // bf16 [rows, N]
    float* s32;              // fp32 [rows, N]
    float* ssq;              // fp32 [N / 128, rows]
    const float* xs32;       // fp32 [rows, K]
    const float* xssq;       // fp32 [ssq_tiles, rows]
    const void* ln_w;        // bf16 [K]
    float eps;
    int ssq_tiles;
};
struct PfArgs {
    const unsigned char* p;
    unsigned long long bytes;      // multiple of 16
};
static inline PfArgs pf_none() { PfArgs a; a.p = nullptr; a.bytes = 0; return a; }
static inline PfArgs pf_linear(const t3b_linear_t* w) {
    PfArgs a;
    a.p = static_cast<const unsigned char*>(w->w);
    a.bytes = w->kind == T3B_W_AWQ ? (unsigned long long)w->n * w->k / 2 : (unsigned long long)w->n * w->k * 2;
    a.bytes &= ~15ull;
    return a;
}
int t3b_num_sms();

// launch with the programmatic-dependent-launch attribute: the kernel may start while its
// predecessor drains; every kernel calls pdl_wait() before touching activations.
template <typename... KArgs, typename... Args>
static inline cudaError_t t3b_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                     cudaStream_t stream, Args... args) {
    // Every kernel of the chain asks for the same L1 / shared-memory split (all shared): kernels with little shared memory would
    // otherwise prefer L1, and an SM cannot hold blocks of two kernels that want different splits - which is what the programmatic
    // launches are for (64 rows: 1.351 -> 1.339 ms/step; AWQ 1.527 -> 1.466; profiles/r2_notes.md r4g)
    cudaFuncSetAttribute((const void*)kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ------------------------------------------------------------------ device side
#ifdef __CUDACC__
typedef __nv_bfloat16 bf16;

__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ float bf16_round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
__device__ __forceinline__ float bf16lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ void unpack8(const uint4& v, float* f) {
    f[0] = bf16lo(v.x); f[1] = bf16hi(v.x); f[2] = bf16lo(v.y); f[3] = bf16hi(v.y);
    f[4] = bf16lo(v.z); f[5] = bf16hi(v.z); f[6] = bf16lo(v.w); f[7] = bf16hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float* f) {
    return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}
// streaming 128-bit load that does not allocate in L1 (weights / KV are read once per step)
__device__ __forceinline__ uint4 ldg_stream(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
// Loads of data that other kernels of the step write (activations, decode state): L2-coherent, never L1.  Kernels trigger
// their dependents early, so several kernels of the chain can be resident on one SM; an L1 line filled by an older
// kernel's read of a reused buffer would otherwise be served stale to a younger one (L1 is only invalidated at launch).
__device__ __forceinline__ uint4 ld_act(const uint4* p) { return __ldcg(p); }
__device__ __forceinline__ uint2 ld_act(const uint2* p) { return __ldcg(p); }
__device__ __forceinline__ float4 ld_act(const float4* p) { return __ldcg(p); }
__device__ __forceinline__ float ld_act(const float* p) { return __ldcg(p); }
__device__ __forceinline__ int ld_act(const int* p) { return __ldcg(p); }
__device__ __forceinline__ unsigned ld_act(const unsigned* p) { return __ldcg(p); }
__device__ __forceinline__ unsigned long long ld_act(const unsigned long long* p) { return __ldcg(p); }
__device__ __forceinline__ float ld_act(const bf16* p) { return bf16lo((uint32_t)__ldcg(reinterpret_cast<const unsigned short*>(p))); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

// ---- mbarrier -------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                 "selp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
// Bounded wait: a protocol bug traps (-> cudaErrorLaunchFailure) instead of hanging the GPU box.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try_wait(bar, parity)) {
        if (clock64() - t0 > 4000000000LL) { printf("t3b: mbarrier timeout block %d thread %d\n", blockIdx.x, threadIdx.x); __trap(); }
    }
}

// ---- TMA ------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
    asm volatile("prefetch.tensormap [%0];" :: "l"(m) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int32_t c0, int32_t c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 :: "r"(smem_u32(smem_dst)), "l"(m), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
// 1-D bulk copy global -> shared (16-byte aligned, size multiple of 16)
__device__ __forceinline__ void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// L2 prefetch of [p, p + bytes) (16-byte aligned, multiple of 16): fire and forget, nothing to wait for
__device__ __forceinline__ void l2_prefetch_bulk(const void* p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(p), "r"(bytes) : "memory");
}
// block `cta` of `ncta` prefetches its 4 KB-aligned slice of the region in 16 KB pieces; call from ONE thread
__device__ __forceinline__ void pf_slice(const PfArgs pf, int cta, int ncta) {
    if (pf.bytes == 0) return;
    const unsigned long long per = ((pf.bytes + ncta - 1) / ncta + 4095ull) & ~4095ull;
    unsigned long long a = per * (unsigned long long)cta;
    const unsigned long long e = a + per < pf.bytes ? a + per : pf.bytes;
    for (; a < e; a += 16384ull) l2_prefetch_bulk(pf.p + a, (uint32_t)(e - a < 16384ull ? e - a : 16384ull));
}
// the same slice, issued by a whole WARP: lane i takes pieces i, i + 32, ... - a bulk instruction costs the issuing thread several
// hundred cycles (scripts/probe/bulk_bw.cu), 32 lanes issue theirs in one warp instruction; call with all 32 lanes
__device__ __forceinline__ void pf_slice_warp(const PfArgs pf, int cta, int ncta, int lane) {
    if (pf.bytes == 0) return;
    const unsigned long long per = ((pf.bytes + ncta - 1) / ncta + 4095ull) & ~4095ull;
    const unsigned long long a0 = per * (unsigned long long)cta;
    const unsigned long long e = a0 + per < pf.bytes ? a0 + per : pf.bytes;
    for (unsigned long long a = a0 + 16384ull * lane; a < e; a += 16384ull * 32)
        l2_prefetch_bulk(pf.p + a, (uint32_t)(e - a < 16384ull ? e - a : 16384ull));
}
// 1-D bulk copy with an L2 evict-first hint: data read once per step (K/V) must not push the prefetched weights out of L2
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void bulk_load_1d_hint(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar, uint64_t policy) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 :: "r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(policy) : "memory");
}

// ---- tcgen05 / TMEM -------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(smem_dst)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// D[tmem] (+)= A[smem desc] * B[smem desc], kind::f16 (bf16 in, fp32 accumulate)
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 :: "r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrives on `bar` when all previously issued tcgen05.mma of this thread have completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(bar)) : "memory");
}
// 32 lanes x 16 consecutive fp32 columns -> 16 registers per thread (thread t <-> lane base+t)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major, 128-byte-swizzled shared-memory operand descriptor (rows of 64 bf16 = 128 B; 8-row
// groups 1024 B apart).  Field layout: cute/arch/mma_sm100_desc.hpp SmemDescriptor.
__device__ __forceinline__ uint64_t umma_smem_desc_sw128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);       // start address  [0,14)
    d |= (uint64_t)1 << 16;                             // leading byte offset (unused for swizzled K-major) [16,30)
    d |= (uint64_t)(1024 >> 4) << 32;                   // stride byte offset = 8 rows * 128 B [32,46)
    d |= (uint64_t)1 << 46;                             // descriptor version 1 (sm_100)
    d |= (uint64_t)2 << 61;                             // SWIZZLE_128B
    return d;
}
// kind::f16 instruction descriptor: D=f32, A=B=bf16, both K-major, M x N.
__device__ __forceinline__ uint32_t umma_idesc_bf16(uint32_t M, uint32_t N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((N >> 3) << 17) | ((M >> 4) << 24);
}
#endif  // __CUDACC__
