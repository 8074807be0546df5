// F2.2 / F2.5 / F2.6 / C1 / Q for rows > 4: TMA-fed tcgen05.mma with the accumulator in TMEM.
//
// Swap-AB skinny GEMM: the weight tile is the UMMA "A" operand (M = 128 output features, K-major
// straight from the nn.Linear layout), the activations are the "B" operand (N = CFG rows padded to
// 16, <= 256, also K-major), so D^T[128 features x rows] lives in 128 TMEM lanes x `rows` fp32 columns
// and a 2..256-row decode batch uses full-width MMAs without padding the batch to 128.
//
// Parallelism for the short-and-wide decode shapes (N/128 is only 8..64 tiles) comes from split-K over a
// THREAD-BLOCK CLUSTER: the n_split CTAs of a cluster each own a K slice of the same 128-feature tile,
// park their fp32 partial tile in shared memory, and after one barrier.cluster each CTA reduces a
// disjoint set of rows across all peers through distributed shared memory (fixed peer order 0..n-1 ->
// deterministic) and runs the epilogue for those rows.  No fp32 partials ever reach HBM/L2.
//
//   warp 0  TMA producer   W tile 2 x {64 k, 64 n} + X tile {64 k, Rp rows}, SWIZZLE_128B, `stages`-deep ring.
//                          The first ring of W tiles is issued BEFORE griddepcontrol.wait (weights never
//                          depend on the previous kernel) so HBM keeps streaming across kernel boundaries.
//                          AWQ: the W tile is a 4 KB block of packed nibbles fetched with a 1-D bulk copy.
//   warp 1  MMA issuer     one elected lane, 4 x tcgen05.mma (K = 16 each) per stage, tcgen05.commit
//                          frees the stage / publishes the accumulator.
//   warp 2  TMEM allocate / free.
//   warps 4-7 epilogue     tcgen05.ld 32x32b (lane = feature, columns = rows) -> fp32 tile in smem ->
//                          cluster reduce -> epilogue -> bf16 global stores (256 B per row segment).
//   warps 8-15 (AWQ only)  dequantise: nibble -> bf16 via the 0x4300 magic, (q - z) * s in bf16x2
//                          (exactly bf16_rn((q - z) * s), the oracle's definition), written as the
//                          128B-swizzled K-major A tile the MMA reads.
//
// Epilogues: STORE  y[r, n] = bf16(acc)
//            SILU   tile t = gate features [64t, 64t+64) in lanes 0..63 and up features [ffn + 64t, ..) in
//                   lanes 64..127;  act[r, 64t + p] = bf16(bf16(silu(bf16 g)) * bf16 u)
//                   (llama.py:117-121, activation.py:138-141).
//            RESID  (o / down of a small decode step) s = bf16(acc) + resid in fp32; resid := bf16(s), s32 := s,
//                   ssq[tile][r] := sum of s^2 over the tile's 128 features: F2.1's first half, done where the sum is born.
// Normalising operand (xsrc, qkv / gate_up of the same steps): tmap_x describes the fp32 sums; the producer stages a
//   [Rp][64] fp32 slice per k-block next to the operand stage and warps 4-11 (4-7 with AWQ) write
//   bf16(bf16(s * rstd) * ln_w) into the swizzled X stage - F2.1's second half, so a layer is 5 launches instead of 7.
// Legacy mode (partials != NULL): each split writes its fp32 tile to global memory [split][rows][N].
//
// grid = (ceil(N/128), n_split, ceil(rows/256)), cluster = (1, n_split, 1).
#include <cudaTypedefs.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "t3b_common.cuh"

#define UM_M 128
#define UM_K 64
#define UM_A_BYTES (UM_M * UM_K * 2)
#define UM_RAW_BYTES 4096                  // AWQ: 128 x 64 nibbles
#define UM_MAX_STAGES 8

enum { EPI_STORE = 0, EPI_SILU = 1, EPI_RESID = 2 };

struct UmmaParams {
    bf16* y;                 // bf16 [rows, N] (STORE) or [rows, N/2] (SILU); NULL in legacy mode
    float* partials;         // legacy: fp32 [n_split, rows, N]
    int rows, N, kb_total, n_split, Rp, row_chunk, stages;
    uint32_t tmem_cols;
    int epi;
    // AWQ (WSRC == 1)
    const unsigned char* qw; // repacked nibbles [N/128][K/64][2][128][16 B]
    const bf16* scales;      // bf16 [K/128, N]
    const unsigned char* zeros;   // u8 [K/128, N]
    PfArgs pf;               // weights of a later launch to pull into L2 (hint)
    int w_wait;              // 1 = the weights were written by an earlier kernel of the chain (AWQ scratch): no loads before the wait
    // ---- F2.1 folded into the projections on either side of it (decode chain, rows <= 128; t3b_decode_step):
    // producer (o / down, EPI_RESID): s = bf16(acc) + resid; resid := bf16(s); s32 := s (fp32); ssq[tile][row] := sum of s^2 over the tile
    bf16* resid;             // bf16 [rows, N], updated in place
    float* s32;              // fp32 [rows, N]
    float* ssq;              // fp32 [N / 128, rows]
    // consumer (qkv / gate_up, xsrc = 1): the activation operand is not fetched by TMA but built by the epilogue warps:
    // x[r, k] = bf16(bf16(s32[r, k] * rstd[r]) * ln_w[k]),  rstd[r] = rsqrt(sum_t ssq[t][r] / K + eps)   (layernorm.py:40-59)
    int xsrc, ssq_tiles;
    const float* xs32;       // fp32 [rows, K]
    const float* xssq;       // fp32 [ssq_tiles, rows]
    const bf16* ln_w;        // bf16 [K]
    float eps;
};

__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
__device__ __forceinline__ float ld_dsmem_f32(uint32_t addr) {
    float v;
    asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) { asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(nthreads) : "memory"); }

// actual output feature held by row `r` of tile `tile`
__device__ __forceinline__ int tile_feature(int tile, int r, int N, int epi) {
    if (epi == EPI_SILU) return r < 64 ? tile * 64 + r : N / 2 + tile * 64 + (r - 64);
    return tile * UM_M + r;
}

// acc[e] = sum over the splits c = 0..NS-1 (in that order) of red[c][j][col + e], e < 8; red = [NS][per][128] fp32 in shared memory
template <int NS>
__device__ __forceinline__ void red_sum8_n(const float* red, int per, int j, int col, float* acc) {
    float4 a[NS], b[NS];
#pragma unroll
    for (int c = 0; c < NS; ++c) {
        const float4* src = reinterpret_cast<const float4*>(red + ((size_t)(c * per + j) * UM_M + col));
        a[c] = src[0]; b[c] = src[1];
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll
    for (int c = 0; c < NS; ++c) {
        acc[0] += a[c].x; acc[1] += a[c].y; acc[2] += a[c].z; acc[3] += a[c].w;
        acc[4] += b[c].x; acc[5] += b[c].y; acc[6] += b[c].z; acc[7] += b[c].w;
    }
}
__device__ __forceinline__ void red_sum8(const float* red, int n_split, int per, int j, int col, float* acc) {
    if (n_split == 8) red_sum8_n<8>(red, per, j, col, acc);
    else if (n_split == 4) red_sum8_n<4>(red, per, j, col, acc);
    else red_sum8_n<2>(red, per, j, col, acc);
}

// Split-K push: row rl of the block's partial tile belongs to CTA (rl % NS) of the cluster; thread `f` (one accumulator lane = one
// output feature) stores its 4-byte partial of it into the owner's buffer red[src split][rl / NS][128] through distributed shared
// memory.  The owner addresses are mapped once (mapa), the TMEM loads of 32 columns are issued back to back before the stores.
template <int NS>
__device__ __forceinline__ void push_partials(uint32_t tmem_base, int quad, int f, int valid, int Rp, int split, uint32_t red_s) {
    const int per = Rp / NS;                                   // rows each CTA owns
    uint32_t owner[NS];
#pragma unroll
    for (int o = 0; o < NS; ++o) owner[o] = map_to_cta(red_s + (uint32_t)(split * per * UM_M + f) * 4u, (uint32_t)o);
    for (int c0 = 0; c0 < valid; c0 += 32) {
        uint32_t v[32];
        tmem_ld16(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
        if (c0 + 16 < valid) tmem_ld16(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(c0 + 16), v + 16);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            if (c0 + j < valid)                                // c0 is a multiple of 32: (c0 + j) % NS == j % NS, a compile-time index
                asm volatile("st.shared::cluster.u32 [%0], %1;"
                             :: "r"(owner[j % NS] + (uint32_t)(((c0 + j) / NS) * UM_M) * 4u), "r"(v[j]) : "memory");
        }
    }
}

template <int WSRC>
__global__ void __launch_bounds__(WSRC ? 512 : 384, 1) linear_umma_kernel(const __grid_constant__ CUtensorMap tmap_w,
                                                                          const __grid_constant__ CUtensorMap tmap_x,
                                                                          const UmmaParams p) {
    extern __shared__ unsigned char smem_dyn[];
    // 1024-byte alignment as an OFFSET into the __shared__ array: pointer arithmetic keeps the shared address space, so every
    // access below compiles to LDS / STS (a round trip through uintptr_t made them generic LD.E / ST.E)
    unsigned char* smem = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
    const int stages = p.stages, Rp = p.Rp;
    const uint32_t x_bytes = (uint32_t)Rp * 128u;
    const uint32_t sraw_off = UM_A_BYTES + x_bytes + (WSRC ? UM_RAW_BYTES : 0);
    const uint32_t sraw_bytes = p.xsrc ? (uint32_t)Rp * 256u : 0u;                   // fp32 [Rp][64] slice of the residual sum
    const uint32_t stage_bytes = sraw_off + sraw_bytes;                              // [A | X | raw nibbles | raw fp32 rows]
    // fp32 reduction buffer behind the ring (peers push into it while this CTA's pipeline may still be running):
    // n_split > 1: [n_split][Rp / n_split][128]; n_split == 1: a [16][128] exchange buffer; legacy: none
    const size_t ring_bytes = (size_t)stages * stage_bytes;
    const size_t red_bytes = p.partials ? 0 : (p.n_split > 1 ? (size_t)Rp * UM_M * 4 : (size_t)16 * UM_M * 4);
    const size_t body = ring_bytes + red_bytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + body);
    uint64_t* empty_bar = full_bar + UM_MAX_STAGES;
    uint64_t* raw_bar = empty_bar + UM_MAX_STAGES;
    uint64_t* sraw_bar = raw_bar + UM_MAX_STAGES;
    uint64_t* tmem_full_bar = sraw_bar + UM_MAX_STAGES;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);
    float* red = reinterpret_cast<float*>(smem + ring_bytes);
    float* rstd_s = reinterpret_cast<float*>(smem + body + 512);      // [128] (xsrc)
    uint4* lnw_s = reinterpret_cast<uint4*>(smem + body + 1024);      // this split's slice of the norm weight, <= 16 k-blocks x 128 B (xsrc)

    pdl_launch_dependents();                     // the next kernel's launch latency overlaps this whole kernel (it waits before reading)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x, split = blockIdx.y, rchunk = blockIdx.z;
    const int kb_begin = (int)(((long)split * p.kb_total) / p.n_split);
    const int kb_end = (int)(((long)(split + 1) * p.kb_total) / p.n_split);
    const int nkb = kb_end - kb_begin;
    const int n_a = p.epi == EPI_SILU ? tile * 64 : tile * UM_M;                 // first 64 weight rows of the tile
    const int n_b = p.epi == EPI_SILU ? p.N / 2 + tile * 64 : tile * UM_M + 64;   // second 64

    if (threadIdx.x == 0) {
        if (!WSRC) tma_prefetch_desc(&tmap_w);
        tma_prefetch_desc(&tmap_x);
        for (int i = 0; i < stages; ++i) {
            mbar_init(&full_bar[i], (WSRC ? 1 + 8 : 1) + (p.xsrc ? (WSRC ? 4 : 8) : 0));      // producer (tx) [+ 8 dequant warps] [+ operand-build warps]
            mbar_init(&empty_bar[i], 1);
            mbar_init(&raw_bar[i], 1);
            mbar_init(&sraw_bar[i], 1);
        }
        mbar_init(tmem_full_bar, 1);
        mbar_fence_init();
    }
    if (warp == 2) tmem_alloc(tmem_ptr, p.tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    auto load_w = [&](int i, int s) {
        unsigned char* st = smem + (size_t)s * stage_bytes;
        if (WSRC) {
            mbar_expect_tx(&raw_bar[s], UM_RAW_BYTES);
            bulk_load_1d(st + UM_A_BYTES + x_bytes, p.qw + ((size_t)tile * p.kb_total + (kb_begin + i)) * UM_RAW_BYTES, UM_RAW_BYTES, &raw_bar[s]);
        } else {
            tma_load_2d(st, &tmap_w, &full_bar[s], (kb_begin + i) * UM_K, n_a);
            tma_load_2d(st + UM_A_BYTES / 2, &tmap_w, &full_bar[s], (kb_begin + i) * UM_K, n_b);
        }
    };
    // xsrc: builder thread bt of n_build turns the staged fp32 rows of every k-block into the swizzled bf16 operand
    // (bf16: warps 4..11, AWQ: warps 4..7 - the dequant warps sit behind them)
    const int n_build = WSRC ? 128 : 256;
    auto build_x = [&](int bt) {
        named_bar_sync(2, n_build);                          // rstd_s / lnw_s written by warps 4..7
        const int ch = bt & 7, first = (bt >> 2) & 1;        // lanes 4..7 of each quarter-warp read their second half first: no bank conflicts
        for (int i = 0; i < nkb; ++i) {
            const int s = i % stages, round = i / stages;
            unsigned char* st = smem + (size_t)s * stage_bytes;
            unsigned char* xdst = st + UM_A_BYTES;
            const uint4 wraw = lnw_s[i * 8 + ch];
            mbar_wait(&sraw_bar[s], round & 1);              // (on a wrapped stage this also implies the MMAs that read it last round are done)
            for (int row = bt >> 3; row < Rp; row += n_build / 8) {
                const float4* sp = reinterpret_cast<const float4*>(st + sraw_off + row * 256 + ch * 32);   // rows past the end arrive as zeros
                const float4 u = sp[first], v = sp[first ^ 1];
                const float4 a = first ? v : u, b = first ? u : v;
                const float sv[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
                float w8[8], o[8];
                unpack8(wraw, w8);
                const float rs = rstd_s[row];
#pragma unroll
                for (int e = 0; e < 8; ++e) o[e] = bf16_round(sv[e] * rs) * w8[e];
                *reinterpret_cast<uint4*>(xdst + (row >> 3) * 1024 + (row & 7) * 128 + ((ch ^ (row & 7)) << 4)) = pack8(o);
            }
            fence_proxy_async_smem();
            __syncwarp();
            if ((threadIdx.x & 31) == 0) mbar_arrive(&full_bar[s]);
        }
    };
    const uint32_t tx = (WSRC ? 0u : (uint32_t)UM_A_BYTES) + (p.xsrc ? 0u : x_bytes);
    const int pre = nkb < stages ? nkb : stages;
    // The first ring of weight tiles goes out before anything else: weights never depend on the previous kernel.  A TMA / bulk
    // copy instruction costs the issuing THREAD several hundred cycles (scripts/probe/bulk_bw.cu), so lane i of warp 0 issues the
    // loads of stage i: one warp instruction for the whole ring instead of a serial loop.
    if (warp == 0 && lane < pre) {
        if (p.w_wait) pdl_wait();
        mbar_expect_tx(&full_bar[lane], tx);
        load_w(lane, lane);
    }
    // an otherwise idle warp pulls this block's slice of a later launch's weights into L2
    if (warp == 3 && lane == 0)
        pf_slice(p.pf, (int)(blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z)), (int)(gridDim.x * gridDim.y * gridDim.z));
    // Every CTA of the cluster must be running before a peer stores into its shared memory (hidden under the
    // first-byte latency of the loads above).
    if (p.partials == nullptr && p.n_split > 1) cluster_sync_all();

    if (warp == 0) {
        pdl_wait();
        // xsrc: tmap_x describes the fp32 residual sums; the slice lands in the stage's raw area and the build warps turn it into the operand
        if (lane < pre) {
            if (p.xsrc) {
                mbar_expect_tx(&sraw_bar[lane], sraw_bytes);
                tma_load_2d(smem + (size_t)lane * stage_bytes + sraw_off, &tmap_x, &sraw_bar[lane], (kb_begin + lane) * UM_K, rchunk * p.row_chunk);
            } else {
                tma_load_2d(smem + (size_t)lane * stage_bytes + UM_A_BYTES, &tmap_x, &full_bar[lane], (kb_begin + lane) * UM_K, rchunk * p.row_chunk);
            }
        }
        if (lane == 0) {
            for (int i = pre; i < nkb; ++i) {
                const int s = i % stages, round = i / stages;
                mbar_wait(&empty_bar[s], (round & 1) ^ 1);
                mbar_expect_tx(&full_bar[s], tx);
                load_w(i, s);
                if (p.xsrc) {
                    mbar_expect_tx(&sraw_bar[s], sraw_bytes);
                    tma_load_2d(smem + (size_t)s * stage_bytes + sraw_off, &tmap_x, &sraw_bar[s], (kb_begin + i) * UM_K, rchunk * p.row_chunk);
                } else {
                    tma_load_2d(smem + (size_t)s * stage_bytes + UM_A_BYTES, &tmap_x, &full_bar[s], (kb_begin + i) * UM_K, rchunk * p.row_chunk);
                }
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
    } else if (WSRC && warp >= 8) {
        // ---- AWQ dequant: thread d owns weight row (d & 127) and k half (d >> 7) of every k-block
        const int d = threadIdx.x - 256, row = d & 127, half = d >> 7;
        const int n = tile_feature(tile, row, p.N, p.epi);
        const uint32_t a_off = (uint32_t)(row >> 3) * 1024u + (uint32_t)(row & 7) * 128u;
        int g_cur = -1;
        __nv_bfloat162 s2 = __float2bfloat162_rn(0.f), zb2 = s2;
        for (int i = 0; i < nkb; ++i) {
            const int s = i % stages, round = i / stages;
            const int g = (kb_begin + i) >> 1;                       // 128-wide quantisation group
            if (g != g_cur) {
                g_cur = g;
                s2 = __bfloat162bfloat162(p.scales[(size_t)g * p.N + n]);
                zb2 = __float2bfloat162_rn(128.f + (float)p.zeros[(size_t)g * p.N + n]);
            }
            unsigned char* st = smem + (size_t)s * stage_bytes;
            mbar_wait(&raw_bar[s], round & 1);
            const uint4 raw = *reinterpret_cast<const uint4*>(st + UM_A_BYTES + x_bytes + half * 2048 + row * 16);
            const uint32_t words[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
            for (int w = 0; w < 4; ++w) {
                uint32_t o[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    uint32_t t = ((words[w] >> (4 * j)) & 0x000F000Fu) | 0x43004300u;   // bf16x2 (128+q_lo, 128+q_hi)
                    __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&t);
                    v = __hmul2(__hsub2(v, zb2), s2);                                    // bf16_rn((q - z) * s)
                    o[j] = *reinterpret_cast<uint32_t*>(&v);
                }
                const uint32_t chunk = (uint32_t)(half * 4 + w) ^ (uint32_t)(row & 7);
                *reinterpret_cast<uint4*>(st + a_off + chunk * 16u) = make_uint4(o[0], o[1], o[2], o[3]);
            }
            fence_proxy_async_smem();                                 // generic-proxy writes -> visible to the MMA (async proxy)
            __syncwarp();
            if (lane == 0) mbar_arrive(&full_bar[s]);
        }
    } else if (warp >= 4 && warp < 8) {
        if (p.xsrc)          // norm weights are constants: fetched before the dependency wait
            for (int c = threadIdx.x - 128; c < nkb * 8; c += 128) lnw_s[c] = __ldg(reinterpret_cast<const uint4*>(p.ln_w + (size_t)kb_begin * UM_K) + c);
        pdl_wait();
        const int quad = warp & 3;
        const int f = quad * 32 + lane;          // feature row of the tile this thread reads from TMEM
        const int r0 = rchunk * p.row_chunk;
        const int valid = min(Rp, p.rows - r0);  // rows of this chunk that exist
        if (p.xsrc) {
            // ---- F2.1 on the way in: these four warps write the normalised activation tile of every k-block straight into the
            // swizzled operand stage (the layout TMA would have produced: row r at (r >> 3) * 1024 + (r & 7) * 128, 16-byte chunk
            // c at c ^ (r & 7)).  The row's sum of squares arrives as ssq_tiles partials, added in tile order.
            const int tt = threadIdx.x - 128;
            if (tt < Rp) {
                float ss = 0.f;
                if (r0 + tt < p.rows)
                    for (int t8 = 0; t8 < p.ssq_tiles; ++t8) ss += ld_act(p.xssq + (size_t)t8 * p.rows + r0 + tt);
                rstd_s[tt] = rsqrtf(ss / (float)(p.kb_total * UM_K) + p.eps);
            }
            build_x(tt);
        }
        mbar_wait(tmem_full_bar, 0);
        tc_fence_after();
        if (p.partials != nullptr) {
            // legacy: fp32 partial tile straight to global memory
            const int n = tile * UM_M + f;
            for (int c0 = 0; c0 < valid; c0 += 16) {
                uint32_t v[16];
                tmem_ld16(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
                tmem_ld_wait();
                if (n < p.N) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const int r = r0 + c0 + j;
                        if (r < p.rows) p.partials[((size_t)split * p.rows + r) * p.N + n] = __uint_as_float(v[j]);
                    }
                }
            }
        } else if (p.n_split > 1) {
            if (p.n_split == 8) push_partials<8>(tmem_base, quad, f, valid, Rp, split, smem_u32(red));
            else if (p.n_split == 4) push_partials<4>(tmem_base, quad, f, valid, Rp, split, smem_u32(red));
            else push_partials<2>(tmem_base, quad, f, valid, Rp, split, smem_u32(red));
        } else {
            // n_split == 1: 16 rows at a time through a small exchange buffer so that the stores of a row are
            // 256 B contiguous and gate / up (lanes f and f + 64) meet
            const int t = threadIdx.x - 128;
            for (int c0 = 0; c0 < valid; c0 += 16) {
                uint32_t v[16];
                tmem_ld16(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 16; ++j) red[j * UM_M + f] = __uint_as_float(v[j]);
                named_bar_sync(1, 128);
                if (p.epi == EPI_STORE && (p.N & 7) == 0) {
                    // (row, 8 adjacent features) items: two LDS.128, one 16-byte store (the prefill shapes live here)
#pragma unroll
                    for (int it = t; it < 16 * 16; it += 128) {
                        const int j = it >> 4, g = it & 15, n = tile * UM_M + g * 8;
                        if (c0 + j < valid && n < p.N) {
                            const float4* src = reinterpret_cast<const float4*>(red + j * UM_M + g * 8);
                            const float4 a = src[0], b = src[1];
                            const float o[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
                            *reinterpret_cast<uint4*>(p.y + (size_t)(r0 + c0 + j) * p.N + n) = pack8(o);
                        }
                    }
                } else if (p.epi == EPI_STORE) {
                    const int n = tile * UM_M + t;
                    if (n < p.N) {
#pragma unroll 4
                        for (int j = 0; j < 16; ++j)
                            if (c0 + j < valid) p.y[(size_t)(r0 + c0 + j) * p.N + n] = __float2bfloat16_rn(red[j * UM_M + t]);
                    }
                } else {
                    // one (row, 8 gate/up pairs) item per thread: four LDS.128, one 16-byte store
                    const int j = t >> 3, g8 = (t & 7) * 8, ffn = p.N / 2;
                    if (c0 + j < valid) {
                        const float4* gs = reinterpret_cast<const float4*>(red + j * UM_M + g8);
                        const float4* us = reinterpret_cast<const float4*>(red + j * UM_M + 64 + g8);
                        const float4 ga = gs[0], gb = gs[1], ua = us[0], ub = us[1];
                        const float gv[8] = {ga.x, ga.y, ga.z, ga.w, gb.x, gb.y, gb.z, gb.w}, uv[8] = {ua.x, ua.y, ua.z, ua.w, ub.x, ub.y, ub.z, ub.w};
                        float o[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) {
                            const float g = bf16_round(gv[e]);
                            o[e] = bf16_round(__fdividef(g, 1.0f + __expf(-g))) * bf16_round(uv[e]);       // the split-K epilogue's formula
                        }
                        *reinterpret_cast<uint4*>(p.y + (size_t)(r0 + c0 + j) * ffn + tile * 64 + g8) = pack8(o);
                    }
                }
                named_bar_sync(1, 128);
            }
        }
    }
    if (!WSRC && warp >= 8 && p.xsrc) build_x((int)threadIdx.x - 128);      // bf16 + xsrc blocks carry four extra builder warps
    tc_fence_before();
    if (p.partials == nullptr && p.n_split > 1) {
        // EPI_RESID: the residual values of this thread's first item are fetched while the cluster barrier collects the peers
        uint4 r_pre = make_uint4(0u, 0u, 0u, 0u);
        if (p.epi == EPI_RESID) {
            pdl_wait();
            const int it = (int)threadIdx.x, per = Rp / p.n_split;
            const int r = rchunk * p.row_chunk + (it >> 4) * p.n_split + split;
            if (it < per * 16 && r < p.rows) r_pre = ld_act(reinterpret_cast<const uint4*>(p.resid + (size_t)r * p.N + tile * UM_M + (it & 15) * 8));
        }
        cluster_sync_all();                      // release / acquire: every peer's pushes into this CTA's buffer have landed
        // (per-owner mbarriers with remote arrives + an acquire.cluster wait instead of this barrier measured SLOWER: qkv 5.6 -> 6.0 us,
        //  o 4.0 -> 4.4, down 6.8 -> 7.5 - barrier.cluster is the cheapest cluster-wide rendezvous there is; profiles/r2_notes.md r3b)
        // after the barrier the producer / MMA warps are idle too: the pure store epilogues run on every thread of the block
        const bool wide = p.epi == EPI_SILU || p.epi == EPI_RESID || (p.epi == EPI_STORE && (p.N & 7) == 0);
        const int nthr = wide ? (int)blockDim.x : 128;
        if (wide || (warp >= 4 && warp < 8)) {
            // The CTA owns rows {r0 + j * n_split + split}.  Item = (row j, 8 adjacent features): two LDS.128 per split, one
            // 16-byte global store - 8x fewer memory instructions than a thread per feature.  Splits are summed in peer order
            // 0..n-1 (deterministic, the same order for every batch).
            const int t = wide ? (int)threadIdx.x : (int)threadIdx.x - 128;
            const int r0 = rchunk * p.row_chunk;
            const int per = Rp / p.n_split;
            if (p.epi == EPI_STORE && (p.N & 7) == 0) {
                for (int it = t; it < per * 16; it += nthr) {
                    const int j = it >> 4, g = it & 15, r = r0 + j * p.n_split + split, n = tile * UM_M + g * 8;
                    if (r >= p.rows || n >= p.N) continue;
                    float acc[8];
                    red_sum8(red, p.n_split, per, j, g * 8, acc);
                    *reinterpret_cast<uint4*>(p.y + (size_t)r * p.N + n) = pack8(acc);
                }
            } else if (p.epi == EPI_RESID) {
                // ---- F2.1 on the way out (o / down): residual add, the fp32 sum for the consumer's normalisation and this tile's
                // share of the row's sum of squares.  16 adjacent lanes hold one row of the tile.
                for (int it0 = 0; it0 < per * 16; it0 += nthr) {
                    const int it = it0 + t;
                    const int j = it >> 4, g = it & 15, r = r0 + j * p.n_split + split, n = tile * UM_M + g * 8;
                    const bool ok = it < per * 16 && r < p.rows;
                    float ss = 0.f;
                    if (ok) {
                        float acc[8], rv[8];
                        red_sum8(red, p.n_split, per, j, g * 8, acc);
                        uint4* rp = reinterpret_cast<uint4*>(p.resid + (size_t)r * p.N + n);
                        unpack8(it0 == 0 ? r_pre : ld_act(rp), rv);
#pragma unroll
                        for (int e = 0; e < 8; ++e) {
                            acc[e] = bf16_round(acc[e]) + rv[e];
                            ss += acc[e] * acc[e];
                        }
                        *rp = pack8(acc);
                        float4* sp = reinterpret_cast<float4*>(p.s32 + (size_t)r * p.N + n);
                        sp[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
                        sp[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
                    }
#pragma unroll
                    for (int m = 8; m >= 1; m >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, m);
                    if (ok && g == 0) p.ssq[(size_t)tile * p.rows + r] = ss;
                }
            } else if (p.epi == EPI_STORE) {     // N not a multiple of 8 (the 8194-wide speech head): rows are not 16-byte aligned
                const int n = tile * UM_M + t;
#pragma unroll 4
                for (int j = 0; j < per; ++j) {
                    const int r = r0 + j * p.n_split + split;
                    if (r >= p.rows) break;
                    float acc = 0.f;
                    for (int c = 0; c < p.n_split; ++c) acc += red[(c * per + j) * UM_M + t];      // fixed order -> deterministic
                    if (n < p.N) p.y[(size_t)r * p.N + n] = __float2bfloat16_rn(acc);
                }
            } else {
                const int ffn = p.N / 2;
                for (int it = t; it < per * 8; it += nthr) {
                    const int j = it >> 3, g = it & 7, r = r0 + j * p.n_split + split;
                    if (r >= p.rows) continue;
                    float gt[8], up[8], o[8];
                    red_sum8(red, p.n_split, per, j, g * 8, gt);
                    red_sum8(red, p.n_split, per, j, 64 + g * 8, up);
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const float gg = bf16_round(gt[e]);
                        // ex2.approx / rcp.approx: relative error ~1e-6, three orders below the bf16 rounding that follows
                        o[e] = bf16_round(__fdividef(gg, 1.0f + __expf(-gg))) * bf16_round(up[e]);
                    }
                    *reinterpret_cast<uint4*>(p.y + (size_t)r * ffn + tile * 64 + g * 8) = pack8(o);
                }
            }
        }
    }
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, p.tmem_cols);
}

// ------------------------------------------------------------------ AWQ repack (one-off, at load)
// AutoAWQ GEMM tensors -> kernel layout.  Word (tile, kb, half, row, w) holds the 8 nibbles of weight
// row n(tile,row), k = kb*64 + half*32 + w*8 + {0,2,4,6,1,3,5,7}[nibble position] so that
// (word >> 4j) & 0x000F000F yields the k-adjacent pair (2j, 2j+1) as a bf16x2.
__global__ void awq_repack_kernel(const int32_t* __restrict__ qweight, const int32_t* __restrict__ qzeros, const __half* __restrict__ scales,
                                  int N, int K, int gate_up, uint32_t* __restrict__ w_out, unsigned char* __restrict__ zeros_out,
                                  bf16* __restrict__ scales_out) {
    const int rev[8] = {0, 4, 1, 5, 2, 6, 3, 7};          // column n % 8 sits at bits 4 * rev[n % 8] (awq_triton.py:56-62)
    const int koff[8] = {0, 2, 4, 6, 1, 3, 5, 7};
    const size_t n_words = (size_t)N * K / 8;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_words; i += (size_t)gridDim.x * blockDim.x) {
        const int w = (int)(i & 3), row = (int)((i >> 2) & 127), half = (int)((i >> 9) & 1);
        const size_t blk = i >> 10;
        const int kb = (int)(blk % (size_t)(K / 64)), tile = (int)(blk / (size_t)(K / 64));
        const int n = gate_up ? (row < 64 ? tile * 64 + row : N / 2 + tile * 64 + (row - 64)) : tile * 128 + row;
        uint32_t word = 0;
#pragma unroll
        for (int pos = 0; pos < 8; ++pos) {
            const int k = kb * 64 + half * 32 + w * 8 + koff[pos];
            const uint32_t q = ((uint32_t)qweight[(size_t)k * (N / 8) + n / 8] >> (4 * rev[n & 7])) & 15u;
            word |= q << (4 * pos);
        }
        w_out[i] = word;
    }
    const size_t n_sz = (size_t)(K / 128) * N;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_sz; i += (size_t)gridDim.x * blockDim.x) {
        const int n = (int)(i % (size_t)N);
        const size_t g = i / (size_t)N;
        zeros_out[i] = (unsigned char)(((uint32_t)qzeros[g * (N / 8) + n / 8] >> (4 * rev[n & 7])) & 15u);
        scales_out[i] = __float2bfloat16_rn(__half2float(scales[i]));
    }
}

// ------------------------------------------------------------------ AWQ -> bf16 for large batches
// Upstream runs dequantise-then-GEMM at >= 256 tokens (vllm awq.py:276-281).  Same here: from 256 rows on, every extra 128-row
// chunk of the dequant-fused kernel repeats the dequantisation of its weight tile, so t3b_decode_step / t3b_forward_tokens
// dequantise a layer's four projections ONCE into a bf16 scratch (nn.Linear layout, L2-resident for the GEMMs that follow) and
// run the bf16 kernels.  The arithmetic is the fused kernel's: bf16_rn((q - z) * s) through the 0x4300 magic.
struct AwqDeqJob {
    const uint32_t* qw;          // repacked nibbles [N/128][K/64][2][128][4 words]
    const unsigned char* zeros;  // u8 [K/128, N]
    const bf16* scales;          // bf16 [K/128, N]
    bf16* out;                   // bf16 [N, K]
    int N, K, gate_up;
    unsigned first_word;         // prefix sum of N * K / 8 over the jobs
};
struct AwqDeqArgs {
    AwqDeqJob job[4];
    unsigned total_words;
};

__global__ void __launch_bounds__(256) awq_dequant_kernel(const AwqDeqArgs a) {
    pdl_launch_dependents();
    pdl_wait();                      // the scratch may still be read by the previous layer's GEMMs
    for (unsigned gi = blockIdx.x * blockDim.x + threadIdx.x; gi < a.total_words; gi += gridDim.x * blockDim.x) {
        int j = 0;
#pragma unroll
        for (int t = 1; t < 4; ++t) if (gi >= a.job[t].first_word) j = t;
        const AwqDeqJob& jb = a.job[j];
        const unsigned i = gi - jb.first_word;
        const int w = (int)(i & 3), row = (int)((i >> 2) & 127), half = (int)((i >> 9) & 1);
        const unsigned blk = i >> 10;
        const int kbs = jb.K / 64;
        const int kb = (int)(blk % (unsigned)kbs), tile = (int)(blk / (unsigned)kbs);
        const int n = jb.gate_up ? (row < 64 ? tile * 64 + row : jb.N / 2 + tile * 64 + (row - 64)) : tile * 128 + row;
        const int k0 = kb * 64 + half * 32 + w * 8;
        const int g = k0 >> 7;
        const __nv_bfloat162 s2 = __bfloat162bfloat162(jb.scales[(size_t)g * jb.N + n]);
        const __nv_bfloat162 zb2 = __float2bfloat162_rn(128.f + (float)jb.zeros[(size_t)g * jb.N + n]);
        const uint32_t word = __ldg(jb.qw + i);
        uint32_t o[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint32_t t = ((word >> (4 * q)) & 0x000F000Fu) | 0x43004300u;        // bf16x2 (128 + q[2q], 128 + q[2q+1])
            __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&t);
            v = __hmul2(__hsub2(v, zb2), s2);
            o[q] = *reinterpret_cast<uint32_t*>(&v);
        }
        *reinterpret_cast<uint4*>(jb.out + (size_t)n * jb.K + k0) = make_uint4(o[0], o[1], o[2], o[3]);
    }
}

// dequantise up to 4 AWQ projections (t3b_linear_t, kind T3B_W_AWQ) into consecutive bf16 [N, K] matrices starting at `scratch`
int32_t t3b_awq_dequant_launch(const t3b_linear_t* const* lins, int n_lin, void* scratch, cudaStream_t stream) {
    T3B_CHECK_ARG(n_lin >= 1 && n_lin <= 4 && scratch, "awq dequant: bad arguments");
    AwqDeqArgs a = {};
    unsigned words = 0;
    char* out = static_cast<char*>(scratch);
    for (int i = 0; i < 4; ++i) {
        const t3b_linear_t* l = lins[i < n_lin ? i : n_lin - 1];
        AwqDeqJob& j = a.job[i];
        j.first_word = i < n_lin ? words : 0xffffffffu;
        if (i >= n_lin) continue;
        T3B_CHECK_ARG(l->kind == T3B_W_AWQ && l->w && l->scales && l->zeros, "awq dequant: projection %d is not AWQ", i);
        j.qw = (const uint32_t*)l->w; j.zeros = (const unsigned char*)l->zeros; j.scales = (const bf16*)l->scales;
        j.out = (bf16*)out; j.N = l->n; j.K = l->k; j.gate_up = (l->flags & T3B_LIN_GATE_UP_INTERLEAVED) ? 1 : 0;
        words += (unsigned)((size_t)l->n * l->k / 8);
        out += (size_t)l->n * l->k * 2;
    }
    a.total_words = words;
    int blocks = (int)((words + 255) / 256);
    const int cap = t3b_num_sms() * 16;
    if (blocks > cap) blocks = cap;
    T3B_CUDA(t3b_launch(awq_dequant_kernel, dim3((unsigned)blocks), dim3(256), 0, stream, a));
    return T3B_OK;
}

extern "C" int32_t t3b_repack_weights_awq(const int32_t* qweight, const int32_t* qzeros, const void* scales_fp16, int32_t n, int32_t k,
                                          int32_t gate_up_interleave, void* w_out, void* zeros_out, void* scales_out, void* stream) {
    T3B_CHECK_ARG(qweight && qzeros && scales_fp16 && w_out && zeros_out && scales_out, "t3b_repack_weights_awq: null pointer");
    T3B_CHECK_SHAPE(n > 0 && k > 0 && n % 128 == 0 && k % 128 == 0, "t3b_repack_weights_awq: n %d / k %d must be multiples of 128", n, k);
    T3B_CHECK_ALIGN(w_out);
    const size_t words = (size_t)n * k / 8;
    int blocks = (int)((words + 255) / 256);
    if (blocks > t3b_num_sms() * 16) blocks = t3b_num_sms() * 16;
    awq_repack_kernel<<<blocks, 256, 0, t3b_stream(stream)>>>(qweight, qzeros, (const __half*)scales_fp16, n, k, gate_up_interleave ? 1 : 0,
                                                             (uint32_t*)w_out, (unsigned char*)zeros_out, (bf16*)scales_out);
    T3B_LAUNCH_CHECK();
    return T3B_OK;
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

// fp32 [outer, inner] row-major, box = 64 x box_outer, no swizzle: the residual sums the normalising operand path stages
static int make_tmap_f32(CUtensorMap* m, const void* ptr, uint64_t inner, uint64_t outer, uint32_t box_outer) {
    auto enc = get_encode_fn();
    if (!enc) return -1;
    cuuint64_t gdim[2] = {inner, outer};
    cuuint64_t gstride[1] = {inner * 4};
    cuuint32_t box[2] = {UM_K, box_outer};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(ptr), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : (int)r;
}

int t3b_make_tmap_bf16(CUtensorMap* m, const void* ptr, uint64_t inner, uint64_t outer, uint32_t box_outer, bool weights) {
    return make_tmap(m, ptr, inner, outer, box_outer, weights);
}

int t3b_umma_available() { return get_encode_fn() != nullptr; }

// Split-K factor (cluster mode, DSMEM reduce): the largest power of two <= 8 that divides the k-blocks and whose grid still
// co-schedules in one wave.  The limits are measured (scripts/probe/split_sweep.py, profiles/r2t_split_sweep.log): clusters of 8
// need 8 free SMs inside one GPC, so ~13 of them run at once (o / down at 256 rows: 128 blocks as 16 clusters of 8 took 11.2 /
// 15.4 us, 64 blocks as 16 clusters of 4 take 5.6 / 9.9); the exchange of a 128-row partial tile also costs more per peer than
// the k-blocks it saves.  Depends on (N, K, 128-row chunks) only, never on the row count inside a chunk, so a request's
// arithmetic does not change with the batch it runs in.
int t3b_umma_pick_split(int rows, int N, int K) {
    const int tiles = (N + UM_M - 1) / UM_M * ((rows + 127) / 128);
    const int kb = K / UM_K;
    static const int cand[3] = {8, 4, 2};
    static const int limit[3] = {104, 128, 160};        // blocks of the whole grid at that cluster size
    for (int i = 0; i < 3; ++i) {
        const int s = cand[i];
        if (kb % s == 0 && kb / s >= 2 && tiles * s <= limit[i]) return s;
    }
    return 1;
}

template <int WSRC>
static int32_t launch_umma(const CUtensorMap& tw, const CUtensorMap& tx, const UmmaParams& p, int tiles, int chunks, bool cluster,
                           cudaStream_t stream) {
    const uint32_t stage_bytes = UM_A_BYTES + (uint32_t)p.Rp * 128u + (WSRC ? UM_RAW_BYTES : 0) + (p.xsrc ? (uint32_t)p.Rp * 256u : 0u);
    const size_t ring = (size_t)p.stages * stage_bytes;
    const size_t red = !cluster ? 0 : (p.n_split > 1 ? (size_t)p.Rp * UM_M * 4 : (size_t)16 * UM_M * 4);
    const size_t smem = ring + red + 1024 + 512 + (p.xsrc ? 512 + 2048 : 0);      // alignment slack, barriers [, rstd + norm weights]
    auto kern = linear_umma_kernel<WSRC>;
    T3B_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    T3B_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));   // see t3b_launch
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(tiles, p.n_split, chunks);
    cfg.blockDim = dim3(WSRC ? 512 : (p.xsrc ? 384 : 256));
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    int na = 1;
    if (cluster && p.n_split > 1) {
        attr[1].id = cudaLaunchAttributeClusterDimension;
        attr[1].val.clusterDim.x = 1; attr[1].val.clusterDim.y = p.n_split; attr[1].val.clusterDim.z = 1;
        na = 2;
    }
    cfg.attrs = attr; cfg.numAttrs = na;
    T3B_CUDA(cudaLaunchKernelEx(&cfg, kern, tw, tx, p));
    return T3B_OK;
}

// y != NULL: bf16 output (cluster split-K when n_split > 1); partials != NULL: legacy fp32 partials.
// epi: EPI_STORE / EPI_SILU (bf16 output only).  awq == NULL: bf16 weights `w` [N, K].
// stages_cap: ring depth cap (0 = 5); row_chunk_req: rows per thread block (0 = 128 with cluster split-K, 256 without).
int32_t t3b_linear_umma_launch(const void* x, const void* w, const t3b_linear_t* awq, int rows, int N, int K, void* y, float* partials,
                               int n_split, int epi, int stages_cap, int row_chunk_req, PfArgs pf, int w_wait, cudaStream_t stream,
                               const UmmaFuse* fuse) {
    T3B_CHECK_SHAPE(K % UM_K == 0 && K >= UM_K, "umma: K %d must be a multiple of 64", K);
    T3B_CHECK_SHAPE(rows >= 1, "umma: rows %d", rows);
    const int kb_total = K / UM_K;
    const bool legacy = partials != nullptr && n_split > 1;
    T3B_CHECK_ARG(n_split >= 1 && n_split <= kb_total, "umma: n_split %d not in [1, %d]", n_split, kb_total);
    T3B_CHECK_ARG(legacy || y != nullptr, "umma: y is null");
    T3B_CHECK_ARG(x != nullptr || (fuse && fuse->xs32), "umma: x is null");
    T3B_CHECK_ARG(legacy || n_split == 1 || n_split == 2 || n_split == 4 || n_split == 8, "umma: cluster split-K %d must be 1, 2, 4 or 8", n_split);
    T3B_CHECK_ARG(!(legacy && epi != EPI_STORE), "umma: fused epilogues need the bf16 output mode");
    T3B_CHECK_ARG(epi == EPI_STORE || epi == EPI_SILU, "umma: unknown epilogue %d", epi);
    const bool f_out = fuse && fuse->resid, f_in = fuse && fuse->xs32;
    if (f_out) {
        T3B_CHECK_ARG(!legacy && n_split > 1 && epi == EPI_STORE && N % UM_M == 0 && rows <= 128 && fuse->s32 && fuse->ssq,
                      "umma: the residual epilogue needs cluster split-K, <= 128 rows and N %d a multiple of 128", N);
        epi = EPI_RESID;
    }
    if (f_in)
        T3B_CHECK_ARG(!legacy && rows <= 128 && fuse->xssq && fuse->ln_w && fuse->ssq_tiles >= 1,
                      "umma: the normalising operand path needs <= 128 rows and the producer's sums");
    if (epi == EPI_SILU) T3B_CHECK_SHAPE(N % 128 == 0, "umma: gate_up N %d must be a multiple of 128", N);
    if (awq) {
        T3B_CHECK_SHAPE(N % 128 == 0 && K % 128 == 0, "umma: AWQ needs N %d and K %d to be multiples of 128", N, K);
        T3B_CHECK_ARG(awq->w && awq->scales && awq->zeros, "umma: AWQ tensors are null");
        T3B_CHECK_ARG(((awq->flags & T3B_LIN_GATE_UP_INTERLEAVED) != 0) == (epi == EPI_SILU),
                      "umma: AWQ gate_up weights repacked with gate_up_interleave must run the SiLU epilogue (and only they may)");
    }
    // Rows per block.  Cluster split-K keeps a [Rp][128] fp32 buffer per CTA: 128-row chunks.  Without split-K (measured table,
    // scripts/probe/split_sweep.py, profiles/r2u_split_sweep.log): decode batches of 129..1024 rows run 128-row chunks - twice the
    // blocks and a deeper ring than 256-row chunks (gate_up at 256 rows: 25 -> 14 us; qkv at 1024 rows: 15.9 -> 10.6) - except the
    // widest projection above 512 rows (gate_up at 1024 rows: 26 us with 256-row chunks, 31 with 128); prefill passes (thousands of
    // rows) keep 256-row chunks: they have blocks to spare and would only re-read the weights more often.
    const int n_tiles = (N + UM_M - 1) / UM_M;
    // (bf16 only: every extra row chunk of an AWQ tile repeats its dequantisation - 512 rows: 5.02 ms/step with 128-row chunks, 4.91 with 256)
    const bool mid = !legacy && !awq && n_split == 1 && rows > 128 && rows <= 1024 && !(n_tiles >= 64 && rows > 512);
    int max_chunk = (!legacy && n_split > 1) ? 128 : (mid ? 128 : 256);
    if (row_chunk_req > 0) {
        T3B_CHECK_ARG(row_chunk_req % 16 == 0 && row_chunk_req <= max_chunk, "umma: row_chunk %d must be a multiple of 16 <= %d", row_chunk_req, max_chunk);
        max_chunk = row_chunk_req;
    }
    const int row_chunk = rows > max_chunk ? max_chunk : rows;
    const int chunks = (rows + row_chunk - 1) / row_chunk;
    const int Rp = (row_chunk + 15) & ~15;
    if (!legacy && n_split > 1) T3B_CHECK_SHAPE(Rp % n_split == 0, "umma: %d rows per block not divisible by split-K %d", Rp, n_split);
    UmmaParams p = {};
    p.y = legacy ? nullptr : (bf16*)y;
    p.partials = legacy ? partials : nullptr;
    p.rows = rows; p.N = N; p.kb_total = kb_total; p.n_split = n_split; p.Rp = Rp; p.row_chunk = row_chunk; p.epi = epi;
    p.pf = pf;
    p.w_wait = w_wait;
    if (f_out) { p.resid = (bf16*)fuse->resid; p.s32 = fuse->s32; p.ssq = fuse->ssq; }
    if (f_in) { p.xsrc = 1; p.xs32 = fuse->xs32; p.xssq = fuse->xssq; p.ln_w = (const bf16*)fuse->ln_w; p.eps = fuse->eps; p.ssq_tiles = fuse->ssq_tiles; }
    p.tmem_cols = 32;
    while ((int)p.tmem_cols < Rp) p.tmem_cols <<= 1;
    const uint32_t stage_bytes = UM_A_BYTES + (uint32_t)Rp * 128u + (awq ? UM_RAW_BYTES : 0) + (f_in ? (uint32_t)Rp * 256u : 0u);
    const uint32_t red_b = legacy ? 0u : (n_split > 1 ? (uint32_t)Rp * UM_M * 4u : 16u * UM_M * 4u);
    const uint32_t budget = 220u * 1024u - red_b;
    int stages = (int)(budget / stage_bytes);
    const int nkb_max = (kb_total + n_split - 1) / n_split;
    if (stages > nkb_max) stages = nkb_max;
    if (stages > UM_MAX_STAGES) stages = UM_MAX_STAGES;
    // Ring depth cap (default 5): an 8-deep ring fills the SM's shared memory, so no block of the NEXT kernel (attention, norm) can
    // become resident beside a GEMM block and run its pre-wait part early.  Measured at 64 rows: 8 stages 1.460 ms/step, 4-6 stages
    // 1.397-1.402 (profiles/r01t_notes.md).
    const int cap = stages_cap > 0 ? stages_cap : 5;
    if (stages > cap) stages = cap;
    // prefill shapes (256-row chunks, no split-K): a 48 KB stage x 4 leaves one block per SM, so the long epilogue (256 TMEM
    // columns through the exchange buffer) never overlaps another block's main loop; 2 stages -> two blocks per SM
    if (Rp > 128 && stages > 2 && stages_cap <= 0) stages = 2;
    // 128-row chunks without split-K and more blocks than SMs: a 3-deep ring (96 KB) lets two blocks share an SM, so one block's
    // epilogue runs under the other's main loop and the grid is one wave (gate_up at 512 rows: 27.8 -> 15.6 us)
    if (mid && stages_cap <= 0 && n_tiles * chunks > t3b_num_sms() && stages > 3) stages = 3;
    if (stages < 2) stages = 2;
    p.stages = stages;
    CUtensorMap tw, tx;
    memset(&tw, 0, sizeof(tw));
    int e1 = awq ? 0 : make_tmap(&tw, w, (uint64_t)K, (uint64_t)N, 64, true);
    memset(&tx, 0, sizeof(tx));
    int e2 = f_in ? make_tmap_f32(&tx, fuse->xs32, (uint64_t)K, (uint64_t)rows, (uint32_t)Rp)
                  : make_tmap(&tx, x, (uint64_t)K, (uint64_t)rows, (uint32_t)Rp, false);
    if (e1 || e2) T3B_FAIL(T3B_ERR_CUDA, "umma: cuTensorMapEncodeTiled failed (%d, %d)", e1, e2);
    const int tiles = epi == EPI_SILU ? N / 128 : (N + UM_M - 1) / UM_M;
    if (awq) {
        p.qw = (const unsigned char*)awq->w; p.scales = (const bf16*)awq->scales; p.zeros = (const unsigned char*)awq->zeros;
        return launch_umma<1>(tw, tx, p, tiles, chunks, !legacy, stream);
    }
    return launch_umma<0>(tw, tx, p, tiles, chunks, !legacy, stream);
}
