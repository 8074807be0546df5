// F2.4 paged decode attention (llama.py:211-221,231), MHA (group = 1), head_dim 64, block 16.
//
// HBM-bound byte work: per (query, head) the kernel streams (ctx) x 128 B of K and of V exactly once.
// Pool layout [block][head][16 tok][64] makes one (block, head) slice 2 KB contiguous, so a warp's
// four 128-bit loads per slice are four fully coalesced 512 B segments.
//
// Decode steps run the FUSED variant: the CTA reads its head's q/k/v straight from the QKV projection
// output, applies llama3/neox RoPE (F2.3, bf16 roundings of rotary_embedding/common.py:144-174), writes the
// rotated k and v into the paged pool (reshape_and_cache) and then attends over the row's context including
// that token - one launch per layer instead of three.
//
// CTA = one (query, head); its NW warps take the row's KV pages round-robin two at a time (16 128-bit
// loads in flight per lane), keep a private online-softmax state and are merged through shared
// memory at the end (fixed order -> deterministic).  Dot products are reduced across the 8 lanes of
// a token with a 7-shuffle butterfly-transpose, softmax statistics with warp shuffles.
#include <cstdlib>

#include "t3b_common.cuh"

#define ATT_PAGES_PER_ITER 2

// 128-bit K/V load that never allocates in L1: a line of the pool is written once (by the step that appends the token) and no
// kernel keeps pool lines in L1, so there is nothing stale to hit (measured 1.6 us per launch faster than ld.global.cg)
__device__ __forceinline__ uint4 ldg_kv(const void* p) {
    uint4 r;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
}

// FUSED: q points at the QKV projection output bf16 [n_tok, 3*hidden]; k_pool / v_pool are written.
// NW = warps per CTA; the min-blocks bound trims registers so that 64 rows x 16 heads (4 warps) is ONE wave of 7 CTAs per SM.
template <bool FUSED, int NW>
__global__ void __launch_bounds__(NW * 32, NW == 2 ? 14 : (NW == 4 ? 7 : 3)) paged_attn_decode_kernel(const bf16* __restrict__ q, const int32_t* __restrict__ tok_row,
                                                                const int32_t* __restrict__ tok_idx,
                                                                const int32_t* __restrict__ block_tables, int max_blocks,
                                                                bf16* k_pool, bf16* v_pool, int n_heads, bf16* __restrict__ out,
                                                                const int32_t* __restrict__ rope_pos, const bf16* __restrict__ cos_sin,
                                                                int max_pos) {
    __shared__ float sm_m[8], sm_l[8];
    __shared__ float sm_acc[8][64];
    pdl_launch_dependents();       // early trigger: the next kernel's launch overlaps this one (it waits before reading)
    const int t = blockIdx.x / n_heads, h = blockIdx.x - t * n_heads;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = NW;
    const int t4 = lane >> 3, c = lane & 7;
    // Block-table entries of this warp's pages, one iteration per lane (lane i <-> pages (warp + i * NW) * 2 + {0, 1}), fetched once
    // before the dependency wait instead of one L2 round trip per iteration in front of the K/V loads.  The table is written by
    // the host (stream-ordered, long before this kernel); entries past the row's context are never used.
    int my_b0 = 0, my_b1 = 0;
    {
        const int trow = tok_row ? ld_act(tok_row + t) : t;
        const int32_t* btq = block_tables + (size_t)trow * max_blocks;
        const int pg = (warp + lane * NW) * ATT_PAGES_PER_ITER;
        if (pg < max_blocks) my_b0 = ld_act(btq + pg);
        my_b1 = pg + 1 < max_blocks ? ld_act(btq + pg + 1) : my_b0;
    }
    if (FUSED) {
        // Before the dependency wait: pull this (row, head)'s cached K/V lines from HBM into L2 while the QKV projection is
        // still running.  Hint only - the context length may be one step stale and nothing read here is used for arithmetic
        // (block-table entries are always valid block indices, so every prefetched address lies inside the pool).
        // (prefetching a second unit for blocks that are not resident yet was measured slower: 17.6 vs 16.3 us per launch)
        const int np = min((ld_act(tok_idx + t) + 16) >> 4, max_blocks);
        const int32_t* btp = block_tables + (size_t)t * max_blocks;
        for (int i = threadIdx.x; i < np * 32; i += NW * 32) {
            const int page = i >> 5, w = i & 31;
            const bf16* line = (w < 16 ? k_pool : v_pool) + (((size_t)ld_act(btp + page) * n_heads + h) * T3B_KV_BLOCK) * 64 + (w & 15) * 64;
            asm volatile("prefetch.global.L2 [%0];" :: "l"(line));
        }
    }
    pdl_wait();
    const int row = tok_row ? ld_act(tok_row + t) : t;
    const int ctx = ld_act(tok_idx + t) + 1;                      // keys 0..tok_idx[t] (own token is written first)
    const int n_pages = (ctx + 15) >> 4;
    const int32_t* bt = block_tables + (size_t)row * max_blocks;
    const float kScale = 0.125f * 1.4426950408889634f;   // head_dim^-0.5 * log2(e)

    float qf[8];
    if (FUSED) {
        const int hidden = n_heads * 64;
        const bf16* src = q + (size_t)t * 3 * hidden + h * 64 + c * 8;
        float kf[8], cs[8], sn[8];
        unpack8(ld_act(reinterpret_cast<const uint4*>(src)), qf);
        unpack8(ld_act(reinterpret_cast<const uint4*>(src + hidden)), kf);
        const uint4 vraw = ld_act(reinterpret_cast<const uint4*>(src + 2 * hidden));
        int pos = ld_act(rope_pos + t);
        pos = min(max(pos, 0), max_pos - 1);
        unpack8(*reinterpret_cast<const uint4*>(cos_sin + (size_t)pos * 64 + (c & 3) * 8), cs);
        unpack8(*reinterpret_cast<const uint4*>(cos_sin + (size_t)pos * 64 + 32 + (c & 3) * 8), sn);
        const bool hi = c >= 4;                           // lanes c < 4 hold x1 (dims 0..31), c >= 4 hold x2
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const float qp = __shfl_xor_sync(0xffffffffu, qf[e], 4), kp = __shfl_xor_sync(0xffffffffu, kf[e], 4);
            // o1 = x1*cos - x2*sin ; o2 = x2*cos + x1*sin, every product and the sum rounded to bf16
            qf[e] = hi ? bf16_round(bf16_round(qf[e] * cs[e]) + bf16_round(qp * sn[e])) : bf16_round(bf16_round(qf[e] * cs[e]) - bf16_round(qp * sn[e]));
            kf[e] = hi ? bf16_round(bf16_round(kf[e] * cs[e]) + bf16_round(kp * sn[e])) : bf16_round(bf16_round(kf[e] * cs[e]) - bf16_round(kp * sn[e]));
        }
        if (warp == 0 && t4 == 0) {
            const int idx = ctx - 1;
            const size_t slot = (((size_t)ld_act(bt + (idx >> 4)) * n_heads + h) * T3B_KV_BLOCK + (idx & 15)) * 64 + c * 8;
            *reinterpret_cast<uint4*>(k_pool + slot) = pack8(kf);
            *reinterpret_cast<uint4*>(v_pool + slot) = vraw;
        }
        __syncthreads();                                  // the new K/V line is visible to every warp of this CTA
    } else {
        unpack8(ld_act(reinterpret_cast<const uint4*>(q + ((size_t)t * n_heads + h) * 64 + c * 8)), qf);
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) qf[e] *= kScale;

    float m = -INFINITY, l = 0.f, acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;

    int iter = 0;
    for (int p0 = warp * ATT_PAGES_PER_ITER; p0 < n_pages; p0 += nw * ATT_PAGES_PER_ITER, ++iter) {
        uint4 kv[ATT_PAGES_PER_ITER][4], vv[ATT_PAGES_PER_ITER][4];
        int blk2[2];
        if (iter < 32) {                                                 // warp-uniform
            blk2[0] = __shfl_sync(0xffffffffu, my_b0, iter);
            blk2[1] = __shfl_sync(0xffffffffu, my_b1, iter);
        } else {
            blk2[0] = ld_act(bt + p0);
            blk2[1] = ld_act(bt + (p0 + 1 < n_pages ? p0 + 1 : p0));
        }
#pragma unroll
        for (int pp = 0; pp < ATT_PAGES_PER_ITER; ++pp) {
            const int page = p0 + pp;
            const int blk = page < n_pages ? blk2[pp] : blk2[0];         // tail page re-reads a valid page, masked below
            const size_t base = (((size_t)blk * n_heads + h) * T3B_KV_BLOCK) * 64 + (size_t)lane * 8;
#pragma unroll
            for (int j = 0; j < 4; ++j) kv[pp][j] = ldg_kv(k_pool + base + j * 256);
#pragma unroll
            for (int j = 0; j < 4; ++j) vv[pp][j] = ldg_kv(v_pool + base + j * 256);
        }
        // partial dot products: value i = pp*4 + j belongs to token p0*16 + pp*16 + 4j + t4
        float s[8];
#pragma unroll
        for (int pp = 0; pp < ATT_PAGES_PER_ITER; ++pp)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float kf[8];
                unpack8(kv[pp][j], kf);
                float d = 0.f;
#pragma unroll
                for (int e = 0; e < 8; ++e) d = fmaf(qf[e], kf[e], d);
                s[pp * 4 + j] = d;
            }
        // butterfly-transpose reduce over the 8 lanes (c) of a token: lane c ends with value i == c
        float w4[4], w2[2], sc;
        {
            const bool up = (c & 4) != 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float send = up ? s[i] : s[i + 4], keep = up ? s[i + 4] : s[i];
                w4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
            }
            const bool up2 = (c & 2) != 0;
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const float send = up2 ? w4[i] : w4[i + 2], keep = up2 ? w4[i + 2] : w4[i];
                w2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
            }
            const bool up1 = (c & 1) != 0;
            const float send = up1 ? w2[0] : w2[1], keep = up1 ? w2[1] : w2[0];
            sc = keep + __shfl_xor_sync(0xffffffffu, send, 1);
        }
        const int my_tok = (p0 + (c >> 2)) * 16 + 4 * (c & 3) + t4;
        if (my_tok >= ctx) sc = -INFINITY;
        const float m_new = fmaxf(m, warp_max(sc));     // finite: page p0 always holds >= 1 valid token
        const float corr = exp2f(m - m_new);
        const float pr = exp2f(sc - m_new);
        l = l * corr + pr;                              // per-lane partial of the row sum
        m = m_new;
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] *= corr;
#pragma unroll
        for (int pp = 0; pp < ATT_PAGES_PER_ITER; ++pp)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float pj = __shfl_sync(0xffffffffu, pr, t4 * 8 + pp * 4 + j);
                float vf[8];
                unpack8(vv[pp][j], vf);
#pragma unroll
                for (int e = 0; e < 8; ++e) acc[e] = fmaf(pj, vf[e], acc[e]);
            }
    }
    // fold the four token groups (t4) of the warp, then the row sum
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 8);
        acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 16);
    }
    l = warp_sum(l);
    if (lane < 8) {
#pragma unroll
        for (int e = 0; e < 8; ++e) sm_acc[warp][lane * 8 + e] = acc[e];
    }
    if (lane == 0) { sm_m[warp] = m; sm_l[warp] = l; }
    __syncthreads();
    if (threadIdx.x < 64) {
        float mm = -INFINITY;
        for (int w = 0; w < nw; ++w) mm = fmaxf(mm, sm_m[w]);
        float ll = 0.f, o = 0.f;
        for (int w = 0; w < nw; ++w) {
            const float f = sm_m[w] == -INFINITY ? 0.f : exp2f(sm_m[w] - mm);
            ll += sm_l[w] * f;
            o += sm_acc[w][threadIdx.x] * f;
        }
        out[((size_t)t * n_heads + h) * 64 + threadIdx.x] = __float2bfloat16_rn(o / ll);
    }
}

// warps = 0: pick by occupancy (enough CTAs x warps to cover the machine at small batch, fewer at large)
int32_t t3b_paged_attn_launch(const void* q, int n_tok, const int32_t* tok_row, const int32_t* tok_idx, const int32_t* block_tables,
                              int max_blocks, const void* k_pool, const void* v_pool, int n_heads, void* out, int warps, cudaStream_t s) {
    const long units = (long)n_tok * n_heads;
    const int sms = t3b_num_sms();
    int nw = warps;
    if (nw <= 0) {
        nw = 8;
        if (units >= (long)sms * 8) nw = 2;
        else if (units >= (long)sms * 4) nw = 4;
    }
    (void)sms;
    T3B_CHECK_ARG(nw == 2 || nw == 4 || nw == 8, "paged attention: warps %d must be 2, 4 or 8", nw);
    auto kern = nw == 2 ? paged_attn_decode_kernel<false, 2> : (nw == 4 ? paged_attn_decode_kernel<false, 4> : paged_attn_decode_kernel<false, 8>);
    T3B_CUDA(t3b_launch(kern, dim3((unsigned)units), dim3(nw * 32), 0, s, (const bf16*)q, tok_row, tok_idx,
                        block_tables, max_blocks, (bf16*)const_cast<void*>(k_pool), (bf16*)const_cast<void*>(v_pool), n_heads, (bf16*)out,
                        (const int32_t*)nullptr, (const bf16*)nullptr, 0));
    return T3B_OK;
}

static int pick_warps(long units) {
    static const int forced = []() { const char* e = getenv("T3B_ATTN_WARPS"); return e ? atoi(e) : 0; }();
    if (forced == 2 || forced == 4 || forced == 8) return forced;
    const int sms = t3b_num_sms();
    if (units >= (long)sms * 8) return 2;
    if (units >= (long)sms * 4) return 4;
    return 8;
}

// decode step: one token per KV row (token j <-> row j)
int32_t t3b_attn_fused_launch(const void* qkv, int n_tok, const int32_t* tok_idx, const int32_t* rope_pos, const void* rope_cos_sin, int max_pos,
                              const int32_t* block_tables, int max_blocks, void* k_pool, void* v_pool, int n_heads, void* out, cudaStream_t s) {
    const long units = (long)n_tok * n_heads;
    const int nw = pick_warps(units);
    auto kern = nw == 2 ? paged_attn_decode_kernel<true, 2> : (nw == 4 ? paged_attn_decode_kernel<true, 4> : paged_attn_decode_kernel<true, 8>);
    T3B_CUDA(t3b_launch(kern, dim3((unsigned)units), dim3(nw * 32), 0, s, (const bf16*)qkv,
                        (const int32_t*)nullptr, tok_idx, block_tables, max_blocks, (bf16*)k_pool, (bf16*)v_pool, n_heads, (bf16*)out, rope_pos,
                        (const bf16*)rope_cos_sin, max_pos));
    return T3B_OK;
}

extern "C" int32_t t3b_attn_decode_fused(const void* qkv, int32_t n_tok, const int32_t* tok_idx, const int32_t* rope_pos,
                                         const void* rope_cos_sin, int32_t max_pos, const int32_t* block_tables, int32_t max_blocks,
                                         void* k_pool, void* v_pool, int32_t n_heads, void* out, void* stream) {
    T3B_CHECK_ARG(qkv && tok_idx && rope_pos && rope_cos_sin && block_tables && k_pool && v_pool && out, "t3b_attn_decode_fused: null pointer");
    T3B_CHECK_ARG(n_tok >= 0 && max_blocks > 0 && max_pos > 0 && n_heads >= 1, "t3b_attn_decode_fused: bad sizes");
    T3B_CHECK_ALIGN(qkv); T3B_CHECK_ALIGN(rope_cos_sin); T3B_CHECK_ALIGN(k_pool); T3B_CHECK_ALIGN(v_pool); T3B_CHECK_ALIGN(out);
    if (n_tok == 0) return T3B_OK;
    return t3b_attn_fused_launch(qkv, n_tok, tok_idx, rope_pos, rope_cos_sin, max_pos, block_tables, max_blocks, k_pool, v_pool, n_heads, out,
                                 t3b_stream(stream));
}

extern "C" int32_t t3b_paged_attn_decode(const void* q, int32_t n_tok, const int32_t* tok_row, const int32_t* tok_idx,
                                         const int32_t* block_tables, int32_t max_blocks, const void* k_pool, const void* v_pool,
                                         int32_t n_heads, void* out, void* stream) {
    T3B_CHECK_ARG(q && tok_idx && block_tables && k_pool && v_pool && out, "t3b_paged_attn_decode: null pointer");
    T3B_CHECK_ARG(n_tok >= 0 && max_blocks > 0 && n_heads >= 1, "t3b_paged_attn_decode: bad sizes");
    T3B_CHECK_ALIGN(q); T3B_CHECK_ALIGN(k_pool); T3B_CHECK_ALIGN(v_pool); T3B_CHECK_ALIGN(out);
    if (n_tok == 0) return T3B_OK;
    return t3b_paged_attn_launch(q, n_tok, tok_row, tok_idx, block_tables, max_blocks, k_pool, v_pool, n_heads, out, 0, t3b_stream(stream));
}
