// F2.2 / F2.5 / F2.6 for passes of thousands of rows (prefill, t3.py:542-561: 6,336 rows for 32 prompts; AWQ layers arrive here as
// their dequantised bf16 copy): a PERSISTENT tcgen05 GEMM with a double-buffered TMEM accumulator.
//
// linear_umma_kernel runs such passes as one block per (128 features x 256 rows) tile: two blocks per SM with two 48 KB stages
// each, so the TMA latency of a stage is exposed in every k-block and the 256-column epilogue of a block only overlaps the OTHER
// block's main loop - 0.47-0.64 of the sustained bf16 peak (profiles/r3d_gemm_ncu_summary.csv).  Here ONE block per SM walks its
// tiles: a 4-deep ring (192 KB) keeps the tensor pipe fed, and the accumulator of tile j (256 TMEM columns) is drained by the
// epilogue warps while the MMA warp already fills the other 256 columns with tile j + 1.
//
//   warp 0    TMA producer: W tile 2 x {64 k, 64 n} + X tile {64 k, 256 rows}, SWIZZLE_128B, ring shared by all tiles of the block
//   warp 1    MMA issuer: 4 x tcgen05.mma (M = 128, N = 256, K = 16) per k-block into accumulator (tile & 1); tcgen05.commit frees
//             the stage; after the last k-block it publishes the accumulator (tmem_full) - and waits for tmem_empty before reuse
//   warp 2    TMEM allocate (all 512 columns) / free
//   warps 4-11 epilogue, two groups of four (a warp reads the TMEM lanes of subpartition warp % 4): group g takes the 16-row column
//             groups 2i + g - tcgen05.ld 32x32b, through its own 8 KB exchange buffer -> 16-byte bf16 stores (or SiLU * mul).  With
//             four warps the SiLU epilogue of a tile (15 us) outlasted its main loop (7.4 us) and the MMA warp waited for a free
//             accumulator; eight warps and ex2 / rcp approximations (the decode kernels' formula) halve it.
//
// Tiles are numbered row-chunk-major, so the blocks running at any moment share their activation tile in L2.  No split-K, one
// accumulation order per output element: a prompt's result does not depend on the prompts it is batched with.
#include <cudaTypedefs.h>

#include <cstring>

#include "t3b_common.cuh"

#define BG_M 128
#define BG_R 256                           // rows per tile = MMA N
#define BG_STAGES 4
#define BG_STAGE_BYTES (BG_M * 64 * 2 + BG_R * 128)        // 16 KB W + 32 KB X

struct BigParams {
    bf16* y;
    int rows, N, kb_total, n_ft, n_tiles, epi;            // n_ft feature tiles, n_tiles = n_ft * row chunks
};

int t3b_make_tmap_bf16(CUtensorMap* m, const void* ptr, uint64_t inner, uint64_t outer, uint32_t box_outer, bool weights);   // linear_umma.cu

__global__ void __launch_bounds__(384, 1) linear_big_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x,
                                                            const BigParams p) {
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
    float* ex_all = reinterpret_cast<float*>(smem + BG_STAGES * BG_STAGE_BYTES);        // [2 groups][16][128] fp32
    uint64_t* full = reinterpret_cast<uint64_t*>(ex_all + 2 * 16 * BG_M);
    uint64_t* empty = full + BG_STAGES;
    uint64_t* tfull = empty + BG_STAGES;        // [2] accumulator ready
    uint64_t* tempty = tfull + 2;               // [2] accumulator drained
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tempty + 2);

    pdl_launch_dependents();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmap_w);
        tma_prefetch_desc(&tmap_x);
        for (int i = 0; i < BG_STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 8); }
        mbar_fence_init();
    }
    if (warp == 2) tmem_alloc(tmem_ptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    const int nkb = p.kb_total;
    const int my_tiles = ((int)blockIdx.x < p.n_tiles) ? (p.n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

    if (warp == 0) {
        if (lane == 0) {
            pdl_wait();
            int it = 0;                                            // k-blocks issued so far (ring position)
            for (int j = 0; j < my_tiles; ++j) {
                const int tile = (int)blockIdx.x + j * (int)gridDim.x, rc = tile / p.n_ft, ft = tile - rc * p.n_ft;
                const int n_a = p.epi ? ft * 64 : ft * BG_M, n_b = p.epi ? p.N / 2 + ft * 64 : ft * BG_M + 64;
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int s = it % BG_STAGES, round = it / BG_STAGES;
                    if (it >= BG_STAGES) mbar_wait(&empty[s], (round & 1) ^ 1);
                    unsigned char* st = smem + (size_t)s * BG_STAGE_BYTES;
                    mbar_expect_tx(&full[s], BG_STAGE_BYTES);
                    tma_load_2d(st, &tmap_w, &full[s], kb * 64, n_a);
                    tma_load_2d(st + BG_M * 64, &tmap_w, &full[s], kb * 64, n_b);
                    tma_load_2d(st + BG_M * 128, &tmap_x, &full[s], kb * 64, rc * BG_R);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_bf16(BG_M, BG_R);
            int it = 0;
            for (int j = 0; j < my_tiles; ++j) {
                const int buf = j & 1;
                if (j >= 2) mbar_wait(&tempty[buf], ((j >> 1) & 1) ^ 1);        // the epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t acc = tmem_base + (uint32_t)buf * BG_R;
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int s = it % BG_STAGES, round = it / BG_STAGES;
                    mbar_wait(&full[s], round & 1);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smem + (size_t)s * BG_STAGE_BYTES);
                    const uint64_t a_desc = umma_smem_desc_sw128(a_addr), b_desc = umma_smem_desc_sw128(a_addr + BG_M * 128);
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_bf16(acc, a_desc + (uint64_t)(2 * k), b_desc + (uint64_t)(2 * k), idesc, (uint32_t)((kb | k) != 0));
                    umma_commit(&empty[s]);
                }
                umma_commit(&tfull[buf]);
            }
        }
    } else if (warp >= 4) {
        pdl_wait();
        const int quad = warp & 3, grp = (warp - 4) >> 2, f = quad * 32 + lane, t = (threadIdx.x - 128) & 127;
        float* ex = ex_all + grp * 16 * BG_M;
        for (int j = 0; j < my_tiles; ++j) {
            const int tile = (int)blockIdx.x + j * (int)gridDim.x, rc = tile / p.n_ft, ft = tile - rc * p.n_ft;
            const int buf = j & 1, r0 = rc * BG_R;
            const int valid = min(BG_R, p.rows - r0);
            mbar_wait(&tfull[buf], (j >> 1) & 1);
            tc_fence_after();
            for (int c0 = 16 * grp; c0 < valid; c0 += 32) {
                uint32_t v[16];
                tmem_ld16(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(buf * BG_R + c0), v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 16; ++i) ex[i * BG_M + f] = __uint_as_float(v[i]);
                asm volatile("bar.sync %0, 128;" :: "r"(1 + grp) : "memory");
                if (p.epi == 0 && (p.N & 7) == 0) {
#pragma unroll
                    for (int it = t; it < 16 * 16; it += 128) {
                        const int i = it >> 4, g = it & 15, n = ft * BG_M + g * 8;
                        if (c0 + i < valid && n < p.N) {
                            const float4* src = reinterpret_cast<const float4*>(ex + i * BG_M + g * 8);
                            const float4 a = src[0], b = src[1];
                            const float o[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
                            *reinterpret_cast<uint4*>(p.y + (size_t)(r0 + c0 + i) * p.N + n) = pack8(o);
                        }
                    }
                } else if (p.epi == 0) {
                    const int n = ft * BG_M + t;
                    if (n < p.N)
                        for (int i = 0; i < 16; ++i)
                            if (c0 + i < valid) p.y[(size_t)(r0 + c0 + i) * p.N + n] = __float2bfloat16_rn(ex[i * BG_M + t]);
                } else {
                    const int i = t >> 3, g8 = (t & 7) * 8, ffn = p.N / 2;
                    if (c0 + i < valid) {
                        const float4* gs = reinterpret_cast<const float4*>(ex + i * BG_M + g8);
                        const float4* us = reinterpret_cast<const float4*>(ex + i * BG_M + 64 + g8);
                        const float4 ga = gs[0], gb = gs[1], ua = us[0], ub = us[1];
                        const float gv[8] = {ga.x, ga.y, ga.z, ga.w, gb.x, gb.y, gb.z, gb.w}, uv[8] = {ua.x, ua.y, ua.z, ua.w, ub.x, ub.y, ub.z, ub.w};
                        float o[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) {
                            const float g = bf16_round(gv[e]);
                            // ex2.approx / rcp.approx: relative error ~1e-6, three orders below the bf16 rounding that follows
                            o[e] = bf16_round(__fdividef(g, 1.0f + __expf(-g))) * bf16_round(uv[e]);
                        }
                        *reinterpret_cast<uint4*>(p.y + (size_t)(r0 + c0 + i) * ffn + ft * 64 + g8) = pack8(o);
                    }
                }
                asm volatile("bar.sync %0, 128;" :: "r"(1 + grp) : "memory");
            }
            // this accumulator may be overwritten: every epilogue warp says so once its TMEM reads are complete
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[buf]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, 512);
}

// rows > 256 recommended; bf16 weights [N, K], K % 64 == 0; epi 0 = y[rows, N], 1 = gate_up -> y[rows, N / 2] (N % 128 == 0)
int32_t t3b_linear_big_launch(const void* x, const void* w, int rows, int N, int K, void* y, int epi, cudaStream_t stream) {
    T3B_CHECK_ARG(x && w && y && rows >= 1, "linear_big: bad arguments");
    T3B_CHECK_SHAPE(K % 64 == 0 && K >= 64, "linear_big: K %d must be a multiple of 64", K);
    if (epi == 1) T3B_CHECK_SHAPE(N % 128 == 0, "linear_big: gate_up N %d must be a multiple of 128", N);
    BigParams p = {};
    p.y = (bf16*)y; p.rows = rows; p.N = N; p.kb_total = K / 64; p.epi = epi;
    p.n_ft = epi == 1 ? N / 128 : (N + BG_M - 1) / BG_M;
    p.n_tiles = p.n_ft * ((rows + BG_R - 1) / BG_R);
    CUtensorMap tw, tx;
    memset(&tw, 0, sizeof(tw)); memset(&tx, 0, sizeof(tx));
    const int e1 = t3b_make_tmap_bf16(&tw, w, (uint64_t)K, (uint64_t)N, 64, true);
    const int e2 = t3b_make_tmap_bf16(&tx, x, (uint64_t)K, (uint64_t)rows, BG_R, false);
    if (e1 || e2) T3B_FAIL(T3B_ERR_CUDA, "linear_big: cuTensorMapEncodeTiled failed (%d, %d)", e1, e2);
    const size_t smem = 1024 + (size_t)BG_STAGES * BG_STAGE_BYTES + 2 * 16 * BG_M * 4 + (2 * BG_STAGES + 4) * 8 + 16;
    T3B_CUDA(cudaFuncSetAttribute(linear_big_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int sms = t3b_num_sms();
    const int grid = p.n_tiles < sms ? p.n_tiles : sms;
    T3B_CUDA(t3b_launch(linear_big_kernel, dim3((unsigned)grid), dim3(384), smem, stream, tw, tx, p));
    return T3B_OK;
}
