// F2.2 / F2.5 / F2.6 / C1 for rows > 4: TMA-fed tcgen05.mma with the accumulator in TMEM.
//
// Swap-AB skinny GEMM: the weight tile is the UMMA "A" operand (M = 128 output features, K-major
// straight from the nn.Linear layout), the activations are the "B" operand (N = CFG rows padded to
// 16, <= 256, also K-major), so D^T[128 features x rows] lives in 128 TMEM lanes x `rows` fp32 columns
// and a 2..256-row decode batch uses full-width MMAs without padding the batch to 128.
//
//   warp 0  TMA producer   W tile {64 k, 128 n} + X tile {64 k, Rp rows}, SWIZZLE_128B, `stages`-deep ring.
//                          The first ring of W tiles is issued BEFORE griddepcontrol.wait (weights never
//                          depend on the previous kernel) so HBM keeps streaming across kernel boundaries.
//   warp 1  MMA issuer     one elected lane, 4 x tcgen05.mma (K = 16 each) per stage, tcgen05.commit
//                          frees the stage / publishes the accumulator.
//   warp 2  TMEM allocate / free.
//   warps 4-7 epilogue     tcgen05.ld 32x32b (lane = feature, columns = rows) -> bf16 y[r][n] when the CTA
//                          owns the whole K range, else fp32 split-K partials [split][r][n] that the
//                          consumer kernel (add_rmsnorm / rope / silu) sums in fixed order.
//
// grid = (ceil(N/128), n_split, ceil(rows/256)).  Memory-bound for rows < ~216.
#include <cudaTypedefs.h>

#include "t3b_common.cuh"

#define UM_M 128
#define UM_K 64
#define UM_A_BYTES (UM_M * UM_K * 2)
#define UM_MAX_STAGES 8

__global__ void __launch_bounds__(256, 1) linear_umma_kernel(const __grid_constant__ CUtensorMap tmap_w,
                                                             const __grid_constant__ CUtensorMap tmap_x, bf16* __restrict__ y,
                                                             float* __restrict__ partials, int rows, int N, int kb_total, int n_split,
                                                             int Rp, int row_chunk, int stages, uint32_t tmem_cols) {
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~uintptr_t(1023));
    const uint32_t stage_bytes = UM_A_BYTES + (uint32_t)Rp * 128u;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)stages * stage_bytes);
    uint64_t* empty_bar = full_bar + UM_MAX_STAGES;
    uint64_t* tmem_full_bar = empty_bar + UM_MAX_STAGES;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x, split = blockIdx.y, rchunk = blockIdx.z;
    const int kb_begin = (int)(((long)split * kb_total) / n_split);
    const int kb_end = (int)(((long)(split + 1) * kb_total) / n_split);
    const int nkb = kb_end - kb_begin;

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmap_w);
        tma_prefetch_desc(&tmap_x);
        for (int i = 0; i < stages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
        mbar_init(tmem_full_bar, 1);
        mbar_fence_init();
    }
    if (warp == 2) tmem_alloc(tmem_ptr, tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    pdl_launch_dependents();

    if (warp == 0) {
        if (lane == 0) {
            const int pre = nkb < stages ? nkb : stages;
            for (int i = 0; i < pre; ++i) {
                mbar_expect_tx(&full_bar[i], stage_bytes);
                tma_load_2d(smem + (size_t)i * stage_bytes, &tmap_w, &full_bar[i], (kb_begin + i) * UM_K, tile * UM_M);
            }
            pdl_wait();
            for (int i = 0; i < pre; ++i)
                tma_load_2d(smem + (size_t)i * stage_bytes + UM_A_BYTES, &tmap_x, &full_bar[i], (kb_begin + i) * UM_K, rchunk * row_chunk);
            for (int i = pre; i < nkb; ++i) {
                const int s = i % stages, round = i / stages;
                mbar_wait(&empty_bar[s], (round & 1) ^ 1);
                mbar_expect_tx(&full_bar[s], stage_bytes);
                tma_load_2d(smem + (size_t)s * stage_bytes, &tmap_w, &full_bar[s], (kb_begin + i) * UM_K, tile * UM_M);
                tma_load_2d(smem + (size_t)s * stage_bytes + UM_A_BYTES, &tmap_x, &full_bar[s], (kb_begin + i) * UM_K, rchunk * row_chunk);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_bf16(UM_M, (uint32_t)Rp);
            for (int i = 0; i < nkb; ++i) {
                const int s = i % stages, round = i / stages;
                mbar_wait(&full_bar[s], round & 1);
                tc_fence_after();
                const uint32_t a_addr = smem_u32(smem + (size_t)s * stage_bytes);
                const uint64_t a_desc = umma_smem_desc_sw128(a_addr);
                const uint64_t b_desc = umma_smem_desc_sw128(a_addr + UM_A_BYTES);
#pragma unroll
                for (int k = 0; k < UM_K / 16; ++k)   // +32 B per K=16 step inside the 128 B swizzle atom
                    umma_bf16(tmem_base, a_desc + (uint64_t)(2 * k), b_desc + (uint64_t)(2 * k), idesc, (uint32_t)((i | k) != 0));
                umma_commit(&empty_bar[s]);
            }
            umma_commit(tmem_full_bar);
        }
    } else if (warp >= 4) {
        pdl_wait();
        const int quad = warp & 3;
        const int n = tile * UM_M + quad * 32 + lane;
        mbar_wait(tmem_full_bar, 0);
        tc_fence_after();
        const int r0 = rchunk * row_chunk;
        for (int c0 = 0; c0 < Rp; c0 += 16) {
            uint32_t v[16];
            tmem_ld16(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
            tmem_ld_wait();
            if (n < N) {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const int r = r0 + c0 + j;
                    if (r < rows) {
                        const float f = __uint_as_float(v[j]);
                        if (partials != nullptr) partials[((size_t)split * rows + r) * N + n] = f;
                        else y[(size_t)r * N + n] = __float2bfloat16_rn(f);
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, tmem_cols);
}

// ------------------------------------------------------------------ host
static PFN_cuTensorMapEncodeTiled_v12000 get_encode_fn() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
    if (fn == nullptr) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
    }
    return fn;
}

// bf16 [outer, inner] row-major, box {64, box_outer}, 128-byte swizzle, OOB reads return 0
static int make_tmap(CUtensorMap* m, const void* ptr, uint64_t inner, uint64_t outer, uint32_t box_outer, bool weights) {
    auto enc = get_encode_fn();
    if (!enc) return -1;
    cuuint64_t gdim[2] = {inner, outer};
    cuuint64_t gstride[1] = {inner * 2};
    cuuint32_t box[2] = {UM_K, box_outer};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     weights ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : (int)r;
}

int t3b_umma_available() { return get_encode_fn() != nullptr; }

// Split-K factor: fill the SMs, keep >= 1 k-block per CTA, bound the fp32 partial traffic.
int t3b_umma_pick_split(int rows, int N, int K) {
    const int tiles = (N + UM_M - 1) / UM_M * ((rows + 255) / 256);
    const int kb = K / UM_K;
    const int sms = t3b_num_sms();
    int s = (sms + tiles / 2) / tiles;
    if (s < 1) s = 1;
    if (s > kb) s = kb;
    const long budget = 8l << 20;                       // bytes of partials per GEMM
    while (s > 1 && (long)s * rows * N * 4 > budget) --s;
    return s;
}

int32_t t3b_linear_umma_launch(const void* x, const void* w, int rows, int N, int K, void* y, float* partials, int n_split,
                               cudaStream_t stream) {
    T3B_CHECK_SHAPE(K % UM_K == 0 && K >= UM_K, "umma: K %d must be a multiple of 64", K);
    T3B_CHECK_SHAPE(rows >= 1, "umma: rows %d", rows);
    const int kb_total = K / UM_K;
    T3B_CHECK_ARG(n_split >= 1 && n_split <= kb_total, "umma: n_split %d not in [1, %d]", n_split, kb_total);
    T3B_CHECK_ARG((n_split == 1) ? (y != nullptr) : (partials != nullptr), "umma: output pointer for n_split %d is null", n_split);
    const int row_chunk = rows > 256 ? 256 : rows;
    const int chunks = (rows + row_chunk - 1) / row_chunk;
    const int Rp = (row_chunk + 15) & ~15;
    uint32_t tmem_cols = 32;
    while ((int)tmem_cols < Rp) tmem_cols <<= 1;
    const uint32_t stage_bytes = UM_A_BYTES + (uint32_t)Rp * 128u;
    const uint32_t budget = Rp >= 128 ? 200u * 1024u : 96u * 1024u;   // small batch: two CTAs per SM
    int stages = (int)(budget / stage_bytes);
    if (stages > UM_MAX_STAGES) stages = UM_MAX_STAGES;
    if (stages < 2) stages = 2;
    const size_t smem = (size_t)stages * stage_bytes + 1024 + (2 * UM_MAX_STAGES + 2) * 8;
    CUtensorMap tw, tx;
    int e1 = make_tmap(&tw, w, (uint64_t)K, (uint64_t)N, UM_M, true);
    int e2 = make_tmap(&tx, x, (uint64_t)K, (uint64_t)rows, (uint32_t)Rp, false);
    if (e1 || e2) T3B_FAIL(T3B_ERR_CUDA, "umma: cuTensorMapEncodeTiled failed (%d, %d)", e1, e2);
    T3B_CUDA(cudaFuncSetAttribute(linear_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((N + UM_M - 1) / UM_M, n_split, chunks);
    T3B_CUDA(t3b_launch(linear_umma_kernel, grid, dim3(256), smem, stream, tw, tx, (bf16*)(n_split == 1 ? y : nullptr),
                        n_split == 1 ? (float*)nullptr : partials, rows, N, kb_total, n_split, Rp, row_chunk, stages, tmem_cols));
    return T3B_OK;
}
