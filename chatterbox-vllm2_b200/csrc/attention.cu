// F2.4 paged decode attention (llama.py:211-221,231), MHA (group = 1), head_dim 64, block 16, with F2.3 (llama3 / neox RoPE,
// rotary_embedding/common.py:144-174) and the KV-cache write (reshape_and_cache) fused in for decode steps.
//
// HBM-bound byte work: per (query, head) item the kernel streams the row's cached K and V exactly once.  The pool layout
// [block][head][16 tok][64] makes one (block, head) slice ONE contiguous bulk copy (2 KB bf16 / 1088 B fp8).
//
// Persistent kernel, one CTA of AT_NW = 8 warps per SM.  (Two 4-warp CTAs with a 2-deep ring - small enough to sit beside a running
// QKV GEMM block - measured the same at 64 rows and 2-6 % slower at 256 / 512 rows: profiles/r2_notes.md r2h.)
// A TEAM of W warps (W = 1, 2, 4 or 8; AT_NW / W teams per CTA) owns a whole
// (token, head) item: the item's head - q/k/v load, RoPE, KV write - and its fold run once per item, and with W = 1 (every batch
// of >= 37 requests) a warp never meets another warp: no barrier, no exchange, the result goes from registers to global memory.
// Items are dealt round-robin to the teams of the grid; inside a team the item's pages are dealt round-robin two at a time
// ("pair" = pages 2p, 2p+1).  Every warp owns a private ring of S pair slots in shared memory that lane 0 fills with
// cp.async.bulk (K0 | K1 | V0 | V1 -> one mbarrier per slot) S pairs AHEAD of the pair the warp is computing - across item
// boundaries, so the head and the tail of an item never leave HBM idle.  The first ring fill is issued before
// griddepcontrol.wait: cached K/V, context lengths, RoPE positions and block tables were written by earlier steps (the sampler
// that advances them does not trigger its dependents early), so only q/k/v of the new token wait for the QKV GEMM.
// W > 1 (small batches, where one warp per item would leave most of the machine idle): the team's warps merge their
// online-softmax states through shared memory at a named barrier of their own, in warp order (deterministic).
// (The first version of this kernel split every item over all 8 warps of a CTA: 962 warp instructions per (item, warp) for a
// pair that needs ~250, and a CTA barrier per item - issue-latency bound at 0.36 of HBM, profiles/r2_notes.md.)
//
// The new token of a decode step never round-trips through the pool: warp 0 of the team seeds its state with (q.k_new, v_new)
// from registers and stores k (rotated) and v for the following steps.  A request's arithmetic depends on its own context and on
// W only (pair p -> warp p % W, merge in warp order); W is chosen from the item count, or fixed by the caller
// (t3b_options_t.attn_team_warps - an engine fixes it so that a request's tokens do not depend on the live batch).
//
// KVT = 1: FP8-e4m3 K/V with one fp32 scale per (token, head), slice = 16 x 64 B values | 16 x 4 B scales; the scale of a key is
// applied to its score and the scale of a value to its probability, so the inner products run on the raw fp8 values.
#include <cuda_fp8.h>

#include "t3b_common.cuh"

#define AT_NW 8                     // warps per CTA of the decode kernel
#define AT_CTAS 1                   // resident CTAs per SM the decode kernel is sized for
#define PF_NW 8                     // warps per CTA of the prefill kernel
#define AT_CAP 128                  // items per CTA (all its teams) whose metadata is staged in shared memory

struct AttnArgs {
    const bf16* q;                  // FUSED: qkv bf16 [n_tok, 3 * hidden]; else q bf16 [n_tok, hidden] (already rotated)
    const int32_t* row_order;       // deal order of the tokens (NULL = identity): item i = (token row_order[i / H], head i % H)
    const int32_t* tok_row;         // KV row of token j (NULL = j)
    const int32_t* tok_idx;         // index of token j in its row: FUSED = cached tokens before it; else it attends 0..idx
    const int32_t* rope_pos;
    const bf16* cos_sin;
    int max_pos;
    const int32_t* block_tables;
    int max_blocks;
    unsigned char* k_pool;          // layer base
    unsigned char* v_pool;
    int n_heads;
    bf16* out;
    int n_items;                    // n_tok * n_heads
    PfArgs pf;                      // weights of a later launch to pull into L2 (hint)
};

// Profiling build only (-DT3B_ATTN_TRACE, scripts/probe/attn_trace.py): lane 0 of warp 0 of every CTA stamps %globaltimer at the
// phases of the kernel into g_attn_trace[block][8].  Not compiled into the product library.
#ifdef T3B_ATTN_TRACE
__device__ unsigned long long* g_attn_trace = nullptr;
extern "C" int32_t t3b_debug_set_attn_trace(unsigned long long* buf) {
    return cudaMemcpyToSymbol(g_attn_trace, &buf, sizeof(buf)) == cudaSuccess ? 0 : -4;
}
// the buffer pointer is read ONCE per kernel (AT_TRACE_INIT): a stamp that loads it first would show - and add - an L2 round trip
#define AT_TRACE_INIT unsigned long long* const at_tr_ = threadIdx.x == 0 ? g_attn_trace : nullptr
#define AT_STAMP(i) do { if (at_tr_) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); \
                         at_tr_[(size_t)blockIdx.x * 16 + (i)] = t_; } } while (0)
// stamp that cannot be scheduled before `val` is available
#define AT_STAMP_DEP(i, val) do { if (at_tr_) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_) : "f"(val)); \
                                  at_tr_[(size_t)blockIdx.x * 16 + (i)] = t_; } } while (0)
#else
#define AT_TRACE_INIT do { } while (0)
#define AT_STAMP(i) do { } while (0)
#define AT_STAMP_DEP(i, val) do { } while (0)
#endif

template <int KVT> struct KvFmt;
template <> struct KvFmt<0> { static constexpr int SLICE = 2048; };                 // 16 x 64 bf16
template <> struct KvFmt<1> { static constexpr int SLICE = 1024 + 64; };            // 16 x 64 e4m3 | 16 fp32 scales

__device__ __forceinline__ void fp8x8_to_float(const uint2 v, float* f) {
    const uint32_t w[2] = {v.x, v.y};
#pragma unroll
    for (int i = 0; i < 2; ++i) {
#pragma unroll
        for (int hlf = 0; hlf < 2; ++hlf) {
            const __half2_raw hr = __nv_cvt_fp8x2_to_halfraw2((__nv_fp8x2_storage_t)((w[i] >> (16 * hlf)) & 0xffffu), __NV_E4M3);
            const float2 ff = __half22float2(*reinterpret_cast<const __half2*>(&hr));
            f[i * 4 + hlf * 2] = ff.x; f[i * 4 + hlf * 2 + 1] = ff.y;
        }
    }
}

// quantise 8 values of one (token, head) vector whose absolute maximum over all 64 dims is `amax`
__device__ __forceinline__ uint2 float8_to_fp8(const float* f, float scale) {
    uint32_t w[2] = {0u, 0u};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const __nv_fp8x2_storage_t p = __nv_cvt_float2_to_fp8x2(make_float2(f[2 * i] / scale, f[2 * i + 1] / scale), __NV_SATFINITE, __NV_E4M3);
        w[i >> 1] |= (uint32_t)p << (16 * (i & 1));
    }
    return make_uint2(w[0], w[1]);
}

// One warp, 32 tokens (pages k0/v0 and, if `two`, k1/v1; first token index tok0), online-softmax update of (m, l, acc).
// Lane (t4, c) = (lane >> 3, lane & 7) holds dims [8c, 8c+8) of tokens 4j + t4 of each page; qf = q * head_dim^-0.5 * log2(e);
// tokens >= nc are masked.  acc / l are per-lane partials (folded over t4 / the warp by the caller), m is warp-uniform.
template <int KVT>
__device__ __forceinline__ void attn_pair(const unsigned char* k0, const unsigned char* k1, const unsigned char* v0, const unsigned char* v1,
                                          bool two, int tok0, int nc, const float* qf, int lane, float& m, float& l, float* acc) {
    const int t4 = lane >> 3, c = lane & 7;
    float s[8];
#pragma unroll
    for (int pg = 0; pg < 2; ++pg) {
        if (pg == 1 && !two) {
#pragma unroll
            for (int j = 0; j < 4; ++j) s[4 + j] = 0.f;
            continue;
        }
        const unsigned char* kp = pg ? k1 : k0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float kf[8];
            if (KVT == 1) fp8x8_to_float(*reinterpret_cast<const uint2*>(kp + j * 256 + lane * 8), kf);
            else unpack8(*reinterpret_cast<const uint4*>(kp + j * 512 + lane * 16), kf);
            float d = 0.f;
#pragma unroll
            for (int e = 0; e < 8; ++e) d = fmaf(qf[e], kf[e], d);
            s[pg * 4 + j] = d;
        }
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
    // lane (t4, c) now owns token tok0 + (c >> 2) * 16 + 4 (c & 3) + t4
    const int tin = 4 * (c & 3) + t4;                  // token inside its page
    const int my_tok = tok0 + (c >> 2) * 16 + tin;
    float vsc = 1.f;
    if (KVT == 1) {
        if ((c >> 2) == 0 || two) {
            sc *= reinterpret_cast<const float*>(((c >> 2) ? k1 : k0) + 1024)[tin];
            vsc = reinterpret_cast<const float*>(((c >> 2) ? v1 : v0) + 1024)[tin];
        }
    }
    if (my_tok >= nc) sc = -INFINITY;
    const float m_new = fmaxf(m, warp_max(sc));         // finite: the pair holds >= 1 valid token
    const float corr = exp2f(m - m_new);
    float pr = exp2f(sc - m_new);
    l = l * corr + pr;                                  // per-lane partial of the row sum
    m = m_new;
    if (KVT == 1) pr = my_tok < nc ? pr * vsc : 0.f;    // the value's scale rides on its probability
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] *= corr;
#pragma unroll
    for (int pg = 0; pg < 2; ++pg) {
        if (pg == 1 && !two) continue;
        const unsigned char* vp = pg ? v1 : v0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float pj = __shfl_sync(0xffffffffu, pr, t4 * 8 + pg * 4 + j);
            float vf[8];
            if (KVT == 1) fp8x8_to_float(*reinterpret_cast<const uint2*>(vp + j * 256 + lane * 8), vf);
            else unpack8(*reinterpret_cast<const uint4*>(vp + j * 512 + lane * 16), vf);
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] = fmaf(pj, vf[e], acc[e]);
        }
    }
}

template <bool FUSED, int W, int S, int KVT>
__global__ void __launch_bounds__(AT_NW * 32, AT_CTAS) attn_ring_kernel(const AttnArgs a) {
    constexpr int SLICE = KvFmt<KVT>::SLICE;
    constexpr int PAIR = 4 * SLICE;                       // K0 | K1 | V0 | V1
    constexpr int T = AT_NW / W;                          // teams per CTA
    constexpr int CAP = AT_CAP / T;                       // items per team whose metadata is staged
    extern __shared__ __align__(128) unsigned char at_smem[];
    unsigned char* ring = at_smem;                        // [AT_NW][S][PAIR]
    uint64_t* full = reinterpret_cast<uint64_t*>(at_smem + (size_t)AT_NW * S * PAIR);      // [AT_NW * S]
    float* part_acc = reinterpret_cast<float*>(full + AT_NW * S);                          // [2][AT_NW][64]
    float* part_ml = part_acc + 2 * AT_NW * 64;                                            // [2][AT_NW][2]
    int* s_nc = reinterpret_cast<int*>(part_ml + 2 * AT_NW * 2);                           // [AT_CAP] cached tokens to stream
    int* s_row = s_nc + AT_CAP;                                                            // [AT_CAP] KV row
    int* s_pos = s_row + AT_CAP;                                                           // [AT_CAP] RoPE position (FUSED)
    int* s_tok = s_pos + AT_CAP;                                                           // [AT_CAP] token (= row of the step) of the item

    pdl_launch_dependents();       // the next kernel's launch overlaps this one (it waits before reading)
    AT_TRACE_INIT;
    AT_STAMP(0);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int team = warp / W, wt = warp - team * W;      // team = the warps that share one item
    const int t4 = lane >> 3, c = lane & 7;
    const int G = gridDim.x, H = a.n_heads;
    const int NT = G * T;                                 // teams in the grid; team-major so that every SM gets teams before any gets two
    const int gt = team * G + (int)blockIdx.x;
    const int n_my = gt < a.n_items ? (a.n_items - gt + NT - 1) / NT : 0;
    const int mb = team * CAP;                            // this team's slice of the metadata arrays
    for (int k = wt * 32 + lane; k < n_my; k += W * 32) {
        const int rank = (gt + k * NT) / H;
        const int t = a.row_order ? ld_act(a.row_order + rank) : rank;
        s_tok[mb + k] = t;
        const int idx = ld_act(a.tok_idx + t);
        s_nc[mb + k] = FUSED ? idx : idx + 1;
        s_row[mb + k] = a.tok_row ? ld_act(a.tok_row + t) : t;
        if (FUSED) s_pos[mb + k] = min(max(ld_act(a.rope_pos + t), 0), a.max_pos - 1);
    }
    if (threadIdx.x < AT_NW * S) mbar_init(&full[threadIdx.x], 1);
    mbar_fence_init();
    __syncthreads();

    unsigned char* my_ring = ring + (size_t)warp * S * PAIR;
    uint64_t* my_full = full + warp * S;

    // ---- producer cursor: (pk, pp) = next pair of this warp to fetch, (nb0, nb1) its two block ids - loaded when the cursor
    // moves, so that a block-table read is never right in front of the copy it feeds
    int pk = 0, pp = wt, pn = 0, nb0 = 0, nb1 = 0;
    const uint64_t kv_policy = l2_policy_evict_first();     // K/V is read once per step: it must not displace prefetched weights
    auto settle = [&]() {
        while (pk < n_my && pp >= ((s_nc[mb + pk] + 31) >> 5)) { ++pk; pp = wt; }
        nb0 = 0; nb1 = 0;
        if (pk < n_my) {
            const int32_t* bt = a.block_tables + (size_t)s_row[mb + pk] * a.max_blocks + 2 * pp;
            nb0 = ld_act(bt);
            if (2 * pp + 1 < ((s_nc[mb + pk] + 15) >> 4)) nb1 = ld_act(bt + 1);
        }
    };
    auto produce = [&]() {
        if (pk >= n_my) return;                                            // warp-uniform
        {
            // One cp.async.bulk costs the issuing THREAD ~600 cycles; four lanes issue the four copies of a pair with one warp
            // instruction instead (scripts/probe/bulk_bw.cu: 2,470 -> 725 cycles per pair, 5.4 -> 6.6 TB/s at 8 warps per SM).
            const int h = (gt + pk * NT) % H;
            const bool two = pp * 2 + 1 < ((s_nc[mb + pk] + 15) >> 4);
            const int slot = pn % S;
            unsigned char* dst = my_ring + (size_t)slot * PAIR;
            if (lane == 0) mbar_expect_tx(&my_full[slot], (two ? 4u : 2u) * SLICE);
            if (lane < 4 && (two || (lane & 1) == 0)) {
                const size_t off = ((size_t)((lane & 1) ? nb1 : nb0) * H + h) * SLICE;
                const unsigned char* src = ((lane & 2) ? a.v_pool : a.k_pool) + off;
                bulk_load_1d_hint(dst + lane * SLICE, src, SLICE, &my_full[slot], kv_policy);      // slot = K0 | K1 | V0 | V1
            }
        }
        ++pn; pp += W;
        settle();
    };
    // FUSED (decode): everything the ring reads was written by earlier steps -> the first fill goes out before the dependency
    // wait.  Otherwise (prefill) the K/V of this very pass were written by the kernel right in front of this one.
    if (!FUSED) pdl_wait();
    AT_STAMP(1);
    // First ring fill in ONE round trip: lane 4 j + c takes copy c (K0 | K1 | V0 | V1) of the warp's j-th pair of its first item,
    // loads that page's block id itself and issues its copy - one block-table latency and one bulk-copy issue for the whole ring
    // instead of S times (load ids -> issue -> load ids ...): 2.0 -> 0.7 us before the first byte is requested (r2l trace).
    while (pk < n_my && pp >= ((s_nc[mb + pk] + 31) >> 5)) { ++pk; pp = wt; }
    if (pk < n_my) {
        const int nc0 = s_nc[mb + pk], npg0 = (nc0 + 15) >> 4, np0 = (nc0 + 31) >> 5;
        const int cnt = min(S, (np0 - pp + W - 1) / W);
        const int j = lane >> 2, which = lane & 3;
        const int pr = pp + j * W, page = 2 * pr + (which & 1);
        const bool act = j < cnt && page < npg0;
        int blk = 0;
        if (act) blk = ld_act(a.block_tables + (size_t)s_row[mb + pk] * a.max_blocks + page);
        if (j < cnt && which == 0) mbar_expect_tx(&my_full[j], (2 * pr + 1 < npg0 ? 4u : 2u) * SLICE);
        if (act) {
            const int h = (gt + pk * NT) % H;
            bulk_load_1d_hint(my_ring + (size_t)j * PAIR + which * SLICE, ((which & 2) ? a.v_pool : a.k_pool) + ((size_t)blk * H + h) * SLICE, SLICE,
                              &my_full[j], kv_policy);
        }
        pn = cnt; pp += cnt * W;
    }
    settle();
#pragma unroll 1
    for (int i = pn; i < S; ++i) produce();
    AT_STAMP(2);
    if (warp == AT_NW - 1) pf_slice_warp(a.pf, (int)blockIdx.x, G, lane);      // behind the first ring fill, one piece per lane
    if (FUSED) pdl_wait();
    AT_STAMP(3);
    const float kScale = 0.125f * 1.4426950408889634f;   // head_dim^-0.5 * log2(e)
    const int hidden = H * 64;
    // head of an item: q (and k, v of the new token) with the RoPE table rows, fetched one item ahead
    uint4 hq, hk, hv, hcs, hsn;
    int hblk = 0;                                           // block that receives the new token
    auto load_head = [&](int k) {
        hq = hk = hv = hcs = hsn = make_uint4(0, 0, 0, 0);
        if (k >= n_my) return;
        const int item = gt + k * NT, h = item % H, t = s_tok[mb + k];
        if (FUSED) {
            const bf16* src = a.q + (size_t)t * 3 * hidden + h * 64 + c * 8;
            hq = ld_act(reinterpret_cast<const uint4*>(src));
            hk = ld_act(reinterpret_cast<const uint4*>(src + hidden));
            hv = ld_act(reinterpret_cast<const uint4*>(src + 2 * hidden));
            const int pos = s_pos[mb + k];
            hcs = __ldg(reinterpret_cast<const uint4*>(a.cos_sin + (size_t)pos * 64 + (c & 3) * 8));
            hsn = __ldg(reinterpret_cast<const uint4*>(a.cos_sin + (size_t)pos * 64 + 32 + (c & 3) * 8));
            if (wt == 0) hblk = ld_act(a.block_tables + (size_t)s_row[mb + k] * a.max_blocks + (s_nc[mb + k] >> 4));
        } else {
            hq = ld_act(reinterpret_cast<const uint4*>(a.q + ((size_t)t * H + h) * 64 + c * 8));
        }
    };
    load_head(0);

    int cn = 0;
#pragma unroll 1
    for (int k = 0; k < n_my; ++k) {
        const int item = gt + k * NT, h = item % H, t = s_tok[mb + k];
        const int nc = s_nc[mb + k], np = (nc + 31) >> 5, n_pages = (nc + 15) >> 4;
        float qf[8], m = -INFINITY, l = 0.f, acc[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = 0.f;
        unpack8(hq, qf);
        if (k == 0) AT_STAMP_DEP(5, qf[0]);                 // q of the first item has arrived
        if (FUSED) {
            float kf[8], cs[8], sn[8], vf[8];
            unpack8(hk, kf); unpack8(hcs, cs); unpack8(hsn, sn); unpack8(hv, vf);
            // o1 = x1*cos - x2*sin ; o2 = x2*cos + x1*sin, every product and the sum rounded to bf16.  Lanes c < 4 hold x1 (dims 0..31),
            // c >= 4 hold x2: the partner's value comes by shuffle and the sign rides on the (exactly negated) second product.  All
            // shuffles of the head are issued before any of the arithmetic: as one loop per element the compiler funnelled them through one
            // register - 16 shuffle -> round -> round -> add chains back to back, 1.8 us of a 4.2 us launch at small batch (r5p trace).
            const float sg = c >= 4 ? 1.f : -1.f;
            float qp[8], kp[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) qp[e] = __shfl_xor_sync(0xffffffffu, qf[e], 4);
            if (wt == 0) {
#pragma unroll
                for (int e = 0; e < 8; ++e) kp[e] = __shfl_xor_sync(0xffffffffu, kf[e], 4);
            }
#pragma unroll
            for (int e = 0; e < 8; ++e) qf[e] = bf16_round(bf16_round(qf[e] * cs[e]) + sg * bf16_round(qp[e] * sn[e]));
            if (k == 0) AT_STAMP_DEP(8, qf[0] + qf[7]);
            if (wt == 0) {
                // the new token: rotated, stored for the following steps, and attended to straight from registers
#pragma unroll
                for (int e = 0; e < 8; ++e) kf[e] = bf16_round(bf16_round(kf[e] * cs[e]) + sg * bf16_round(kp[e] * sn[e]));
                if (k == 0) AT_STAMP_DEP(9, kf[0] + kf[7]);
                const size_t sl = ((size_t)hblk * H + h) * SLICE;
                if (k == 0) AT_STAMP_DEP(10, (float)sl);
                if (KVT == 1) {
                    float ka = 0.f, va = 0.f;
#pragma unroll
                    for (int e = 0; e < 8; ++e) { ka = fmaxf(ka, fabsf(kf[e])); va = fmaxf(va, fabsf(vf[e])); }
#pragma unroll
                    for (int o = 1; o < 8; o <<= 1) { ka = fmaxf(ka, __shfl_xor_sync(0xffffffffu, ka, o)); va = fmaxf(va, __shfl_xor_sync(0xffffffffu, va, o)); }
                    const float ks = ka > 0.f ? ka / 448.f : 1.f, vs = va > 0.f ? va / 448.f : 1.f;
                    const uint2 kq = float8_to_fp8(kf, ks), vq = float8_to_fp8(vf, vs);
                    if (t4 == 0) {
                        *reinterpret_cast<uint2*>(a.k_pool + sl + (nc & 15) * 64 + c * 8) = kq;
                        *reinterpret_cast<uint2*>(a.v_pool + sl + (nc & 15) * 64 + c * 8) = vq;
                        if (c == 0) {
                            *reinterpret_cast<float*>(a.k_pool + sl + 1024 + (nc & 15) * 4) = ks;
                            *reinterpret_cast<float*>(a.v_pool + sl + 1024 + (nc & 15) * 4) = vs;
                        }
                    }
                    // attend to what the following steps will read back: the dequantised values
                    fp8x8_to_float(kq, kf); fp8x8_to_float(vq, vf);
#pragma unroll
                    for (int e = 0; e < 8; ++e) { kf[e] *= ks; vf[e] *= vs; }
                } else if (t4 == 0) {
                    *reinterpret_cast<uint4*>(a.k_pool + sl + ((nc & 15) * 64 + c * 8) * 2) = pack8(kf);
                    *reinterpret_cast<uint4*>(a.v_pool + sl + ((nc & 15) * 64 + c * 8) * 2) = hv;
                }
                float d = 0.f;
#pragma unroll
                for (int e = 0; e < 8; ++e) d = fmaf(qf[e], kf[e], d);
                d += __shfl_xor_sync(0xffffffffu, d, 1);
                d += __shfl_xor_sync(0xffffffffu, d, 2);
                d += __shfl_xor_sync(0xffffffffu, d, 4);
                m = d * kScale;
                if (k == 0) AT_STAMP_DEP(11, m);
                l = lane == 0 ? 1.f : 0.f;
#pragma unroll
                for (int e = 0; e < 8; ++e) acc[e] = t4 == 0 ? vf[e] : 0.f;
            }
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) qf[e] *= kScale;
        if (k == 0) AT_STAMP_DEP(2, qf[0] + m);            // head of the first item done (trace builds: overwrites ring_issued)
        load_head(k + 1);                                  // in flight while this item's pairs are processed

#pragma unroll 1
        for (int pr = wt; pr < np; pr += W) {
            const int slot = cn % S;
            mbar_wait(&my_full[slot], (uint32_t)((cn / S) & 1));
            if (cn == 0) AT_STAMP(4);
            const unsigned char* sl = my_ring + (size_t)slot * PAIR;
            const bool two = pr * 2 + 1 < n_pages;         // warp-uniform
            attn_pair<KVT>(sl, sl + SLICE, sl + 2 * SLICE, sl + 3 * SLICE, two, pr * 32, nc, qf, lane, m, l, acc);
            ++cn;
            __syncwarp();                                   // every lane is done reading the slot
            produce();                                      // refill it with the pair S ahead
        }
        if (k == 0) AT_STAMP(6);
        // fold the four token groups (t4) of the warp, then the row sum
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 8);
            acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 16);
        }
        l = warp_sum(l);
        if (W == 1) {
            // the warp owns the whole item: no exchange, lanes 0..7 store 8 dims each
            if (lane < 8) {
                const float inv = 1.f / l;
                float o[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) o[e] = acc[e] * inv;
                *reinterpret_cast<uint4*>(a.out + ((size_t)t * H + h) * 64 + lane * 8) = pack8(o);
            }
        } else {
            // park the warp's state; the team meets at its own named barrier and merges in warp order (deterministic)
            float* pa = part_acc + ((k & 1) * AT_NW + warp) * 64;
            if (lane < 8) {
                *reinterpret_cast<float4*>(pa + lane * 8) = make_float4(acc[0], acc[1], acc[2], acc[3]);
                *reinterpret_cast<float4*>(pa + lane * 8 + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
            }
            if (lane == 0) { part_ml[((k & 1) * AT_NW + warp) * 2] = m; part_ml[((k & 1) * AT_NW + warp) * 2 + 1] = l; }
            asm volatile("bar.sync %0, %1;" :: "r"(1 + team), "r"(W * 32) : "memory");
            const int tt = wt * 32 + lane;
            if (tt < 64) {
                const float* pm = part_ml + ((k & 1) * AT_NW + team * W) * 2;
                const float* pb = part_acc + ((k & 1) * AT_NW + team * W) * 64;
                float mm = -INFINITY;
#pragma unroll
                for (int w = 0; w < W; ++w) mm = fmaxf(mm, pm[w * 2]);
                float ll = 0.f, o = 0.f;
#pragma unroll
                for (int w = 0; w < W; ++w) {
                    const float f = pm[w * 2] == -INFINITY ? 0.f : exp2f(pm[w * 2] - mm);
                    ll += pm[w * 2 + 1] * f;
                    o += pb[w * 64 + tt] * f;
                }
                a.out[((size_t)t * H + h) * 64 + tt] = __float2bfloat16_rn(o / ll);
            }
        }
    }
    AT_STAMP(7);
}

// ------------------------------------------------------------------ prefill (causal) attention
// CTA = one (sequence, head): the prompt's K/V slices come from the pool into shared memory ONCE (bulk copies, one mbarrier);
// the CTA's warps then take the prompt's queries round-robin, each query a sequential sweep over its causal prefix with the
// same 32-token update as the decode kernel (a prompt's result depends on nothing but the prompt).
struct PrefillArgs {
    const bf16* q;                  // bf16 [n_tok, hidden], rotated
    const int32_t* seq_start;       // [n_seq + 1]
    const int32_t* seq_row;         // [n_seq]
    const int32_t* block_tables;
    int max_blocks;
    const unsigned char* k_pool;
    const unsigned char* v_pool;
    int n_heads;
    bf16* out;
    int max_pages;
};

template <int KVT>
__global__ void __launch_bounds__(PF_NW * 32) prefill_attn_kernel(const PrefillArgs a) {
    constexpr int SLICE = KvFmt<KVT>::SLICE;
    extern __shared__ __align__(128) unsigned char at_smem[];
    __shared__ uint64_t bar;
    pdl_launch_dependents();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, c = lane & 7;
    const int H = a.n_heads, seq = blockIdx.x / H, h = blockIdx.x - seq * H;
    const int t0 = a.seq_start[seq], P = a.seq_start[seq + 1] - t0;
    const int n_pages = min((P + 15) >> 4, a.max_pages);
    unsigned char* ks = at_smem;                                   // [max_pages][SLICE]
    unsigned char* vs = at_smem + (size_t)a.max_pages * SLICE;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    __syncthreads();
    pdl_wait();                                                    // the K/V of this pass were written by the kernel in front
    if (warp == 0) {
        if (lane == 0) mbar_expect_tx(&bar, (uint32_t)n_pages * 2u * SLICE);
        __syncwarp();
        const int32_t* bt = a.block_tables + (size_t)a.seq_row[seq] * a.max_blocks;
        for (int p = lane; p < n_pages; p += 32) {
            const size_t off = ((size_t)bt[p] * H + h) * SLICE;
            bulk_load_1d(ks + (size_t)p * SLICE, a.k_pool + off, SLICE, &bar);
            bulk_load_1d(vs + (size_t)p * SLICE, a.v_pool + off, SLICE, &bar);
        }
    }
    mbar_wait(&bar, 0);
    const float kScale = 0.125f * 1.4426950408889634f;
    // longest queries first: warp w takes queries P-1-w, P-1-w-NW, ...
    for (int j = P - 1 - warp; j >= 0; j -= PF_NW) {
        float qf[8], m = -INFINITY, l = 0.f, acc[8];
        unpack8(ld_act(reinterpret_cast<const uint4*>(a.q + ((size_t)(t0 + j) * H + h) * 64 + c * 8)), qf);
#pragma unroll
        for (int e = 0; e < 8; ++e) { qf[e] *= kScale; acc[e] = 0.f; }
        const int nc = j + 1, np = (nc + 31) >> 5, npg = (nc + 15) >> 4;
        for (int pr = 0; pr < np; ++pr) {
            const bool two = pr * 2 + 1 < npg;
            attn_pair<KVT>(ks + (size_t)(2 * pr) * SLICE, ks + (size_t)(2 * pr + 1) * SLICE, vs + (size_t)(2 * pr) * SLICE,
                           vs + (size_t)(2 * pr + 1) * SLICE, two, pr * 32, nc, qf, lane, m, l, acc);
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 8);
            acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 16);
        }
        l = warp_sum(l);
        if (lane < 8) {
            float o[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) o[e] = acc[e] / l;
            *reinterpret_cast<uint4*>(a.out + ((size_t)(t0 + j) * H + h) * 64 + lane * 8) = pack8(o);
        }
    }
}

// ------------------------------------------------------------------ prefill attention on the tensor cores (bf16 K/V)
// The warp-per-query kernel above spends ~250 instructions per (query, 32 keys): 110 us per layer for 64 prompts of 99 tokens,
// the largest piece of a prefill pass after the projections.  Here a warp owns a TILE of 16 queries of one (sequence, head) and
// runs S = Q K^T and O = P V as mma.sync.m16n8k16 (bf16 in, fp32 accumulate) over key chunks of 64 with the usual online softmax:
//   - K / V of the prompt are copied once into shared memory, rows of 128 B with their 16-byte chunks XOR-swizzled by (token & 7),
//     so the ldmatrix reads (K as the col-major B operand; V through ldmatrix.trans) are bank-conflict free;
//   - Q fragments come straight from global memory; the softmax scale is applied to the fp32 scores;
//   - P is split into bf16 hi + lo parts (two MMAs), so the probabilities enter P V with ~2^-17 relative error, not bf16's 2^-9:
//     the result stays within the 1-ulp band of the fp32-softmax tests.
// A prompt's result depends on nothing but the prompt (tile i of a sequence always sees keys 0..16i+15 in the same order).
__device__ __forceinline__ void ldsm_x4(uint32_t* r, uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t* r, uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

#define PM_NW 4
__global__ void __launch_bounds__(PM_NW * 32) prefill_attn_mma_kernel(const PrefillArgs a) {
    extern __shared__ __align__(128) unsigned char at_smem[];
    pdl_launch_dependents();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    const int H = a.n_heads, seq = blockIdx.x / H, h = blockIdx.x - seq * H;
    const int t0 = a.seq_start[seq], P = a.seq_start[seq + 1] - t0;
    const int P16 = (P + 15) & ~15;
    unsigned char* ks = at_smem;                                   // [P16][128 B], chunk c of token r at ((c ^ (r & 7)) * 16)
    unsigned char* vs = at_smem + (size_t)a.max_pages * 2048;
    pdl_wait();                                                    // the K/V of this pass were written by the kernel in front
    {
        const int32_t* bt = a.block_tables + (size_t)a.seq_row[seq] * a.max_blocks;
        for (int i = threadIdx.x; i < P16 * 8; i += PM_NW * 32) {
            const int r = i >> 3, c = i & 7;
            uint4 kq = make_uint4(0, 0, 0, 0), vq = kq;            // rows past the prompt are zero: they meet probability 0, never NaN
            if (r < P) {
                const size_t off = ((size_t)bt[r >> 4] * H + h) * 2048 + (size_t)(r & 15) * 128 + c * 16;
                kq = ld_act(reinterpret_cast<const uint4*>(a.k_pool + off));
                vq = ld_act(reinterpret_cast<const uint4*>(a.v_pool + off));
            }
            const int sw = r * 128 + ((c ^ (r & 7)) << 4);
            *reinterpret_cast<uint4*>(ks + sw) = kq;
            *reinterpret_cast<uint4*>(vs + sw) = vq;
        }
    }
    __syncthreads();
    const uint32_t ks_u = smem_u32(ks), vs_u = smem_u32(vs);
    const float kScale = 0.125f * 1.4426950408889634f;
    const int hidden = H * 64, n_tiles = P16 >> 4;
    // longest tiles first, round-robin over the warps
    for (int ti = n_tiles - 1 - warp; ti >= 0; ti -= PM_NW) {
        const int q0 = ti * 16;
        uint32_t qa[4][4];                                         // A fragments of Q for the 4 k-steps of 16 dims
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            const int r0 = min(q0 + g, P - 1), r1 = min(q0 + g + 8, P - 1);        // rows past the prompt repeat the last one (never stored)
            const bf16* p0 = a.q + (size_t)(t0 + r0) * hidden + h * 64 + kk * 16 + 2 * t;
            const bf16* p1 = a.q + (size_t)(t0 + r1) * hidden + h * 64 + kk * 16 + 2 * t;
            qa[kk][0] = ld_act(reinterpret_cast<const unsigned*>(p0));
            qa[kk][1] = ld_act(reinterpret_cast<const unsigned*>(p1));
            qa[kk][2] = ld_act(reinterpret_cast<const unsigned*>(p0 + 8));
            qa[kk][3] = ld_act(reinterpret_cast<const unsigned*>(p1 + 8));
        }
        float o[8][4], m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) { o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f; }
        const int k_end = q0 + 16;                                 // keys this tile may see (causal), a multiple of 16
        for (int kc = 0; kc < k_end; kc += 64) {
            const int nt = min(8, (k_end - kc) >> 3);              // 8-key tiles in this chunk (even)
            float sc[8][4];
#pragma unroll
            for (int j = 0; j < 8; ++j) { sc[j][0] = sc[j][1] = sc[j][2] = sc[j][3] = 0.f; }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (j < nt) {
                    const int key = kc + 8 * j + (lane & 7);
#pragma unroll
                    for (int k2 = 0; k2 < 2; ++k2) {               // two k-steps per ldmatrix.x4: chunks 4 k2 .. 4 k2 + 3 of the key rows
                        uint32_t b[4];
                        ldsm_x4(b, ks_u + key * 128 + ((((4 * k2 + (lane >> 3)) ^ (key & 7))) << 4));
                        mma_bf16(sc[j], qa[2 * k2], b[0], b[1]);
                        mma_bf16(sc[j], qa[2 * k2 + 1], b[2], b[3]);
                    }
                }
            }
            // scale, causal mask, online softmax (rows q0 + g and q0 + g + 8; a row's 4 lanes differ in t)
            float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (j < nt) {
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int key = kc + 8 * j + 2 * t + e;
                        sc[j][e] = key <= q0 + g ? sc[j][e] * kScale : -INFINITY;
                        sc[j][2 + e] = key <= q0 + g + 8 ? sc[j][2 + e] * kScale : -INFINITY;
                        mx0 = fmaxf(mx0, sc[j][e]); mx1 = fmaxf(mx1, sc[j][2 + e]);
                    }
                }
            }
            mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
            mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
            const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);             // finite: key kc <= every query of the tile
            const float c0 = exp2f(m0 - mn0), c1 = exp2f(m1 - mn1);
            m0 = mn0; m1 = mn1; l0 *= c0; l1 *= c1;
#pragma unroll
            for (int j = 0; j < 8; ++j) { o[j][0] *= c0; o[j][1] *= c0; o[j][2] *= c1; o[j][3] *= c1; }
            uint32_t ph[8][2], pl[8][2];                            // P as bf16 hi / lo pairs: [j][0] = row g, [j][1] = row g + 8
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (j < nt) {
                    const float p0 = exp2f(sc[j][0] - m0), p1 = exp2f(sc[j][1] - m0), p2 = exp2f(sc[j][2] - m1), p3 = exp2f(sc[j][3] - m1);
                    l0 += p0 + p1; l1 += p2 + p3;
                    const float h0 = bf16_round(p0), h1 = bf16_round(p1), h2 = bf16_round(p2), h3 = bf16_round(p3);
                    ph[j][0] = pack_bf16(h0, h1); ph[j][1] = pack_bf16(h2, h3);
                    pl[j][0] = pack_bf16(p0 - h0, p1 - h1); pl[j][1] = pack_bf16(p2 - h2, p3 - h3);
                } else {
                    ph[j][0] = ph[j][1] = pl[j][0] = pl[j][1] = 0u;
                }
            }
            // O += P V: k-step kk covers keys kc + 16 kk .. + 15 = score tiles 2 kk, 2 kk + 1
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                if (2 * kk < nt) {
                    const uint32_t ah[4] = {ph[2 * kk][0], ph[2 * kk][1], ph[2 * kk + 1][0], ph[2 * kk + 1][1]};
                    const uint32_t al[4] = {pl[2 * kk][0], pl[2 * kk][1], pl[2 * kk + 1][0], pl[2 * kk + 1][1]};
                    const int key = kc + 16 * kk + ((lane >> 3) & 1) * 8 + (lane & 7);
#pragma unroll
                    for (int d2 = 0; d2 < 4; ++d2) {               // two 8-dim tiles per ldmatrix.x4.trans: chunks 2 d2, 2 d2 + 1
                        uint32_t b[4];
                        ldsm_x4_t(b, vs_u + key * 128 + ((((2 * d2 + (lane >> 4)) ^ (key & 7))) << 4));
                        mma_bf16(o[2 * d2], ah, b[0], b[1]);
                        mma_bf16(o[2 * d2], al, b[0], b[1]);
                        mma_bf16(o[2 * d2 + 1], ah, b[2], b[3]);
                        mma_bf16(o[2 * d2 + 1], al, b[2], b[3]);
                    }
                }
            }
        }
        l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
        l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
        const float i0 = 1.f / l0, i1 = 1.f / l1;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (q0 + g < P)
                *reinterpret_cast<uint32_t*>(a.out + (size_t)(t0 + q0 + g) * hidden + h * 64 + 8 * j + 2 * t) = pack_bf16(o[j][0] * i0, o[j][1] * i0);
            if (q0 + g + 8 < P)
                *reinterpret_cast<uint32_t*>(a.out + (size_t)(t0 + q0 + g + 8) * hidden + h * 64 + 8 * j + 2 * t) = pack_bf16(o[j][2] * i1, o[j][3] * i1);
        }
    }
}

extern "C" int32_t t3b_prefill_attn_max_len(int32_t kv_dtype) {
    const int slice = kv_dtype == T3B_KV_FP8 ? KvFmt<1>::SLICE : KvFmt<0>::SLICE;
    return (220 * 1024 / (2 * slice)) * T3B_KV_BLOCK;
}

extern "C" int32_t t3b_prefill_attn(const void* q, int32_t n_seq, const int32_t* seq_start, const int32_t* seq_row,
                                    const int32_t* block_tables, int32_t max_blocks, const void* k_pool, const void* v_pool,
                                    int32_t n_heads, int32_t kv_dtype, int32_t max_seq_len, void* out, void* stream) {
    T3B_CHECK_ARG(q && seq_start && seq_row && block_tables && k_pool && v_pool && out, "t3b_prefill_attn: null pointer");
    T3B_CHECK_ARG(n_seq >= 0 && max_blocks > 0 && n_heads >= 1 && max_seq_len >= 1, "t3b_prefill_attn: bad sizes");
    T3B_CHECK_ARG(kv_dtype == T3B_KV_BF16 || kv_dtype == T3B_KV_FP8, "t3b_prefill_attn: unknown KV dtype %d", kv_dtype);
    T3B_CHECK_SHAPE(max_seq_len <= t3b_prefill_attn_max_len(kv_dtype), "t3b_prefill_attn: max_seq_len %d > %d", max_seq_len,
                    t3b_prefill_attn_max_len(kv_dtype));
    T3B_CHECK_ALIGN(q); T3B_CHECK_ALIGN(k_pool); T3B_CHECK_ALIGN(v_pool); T3B_CHECK_ALIGN(out);
    if (n_seq == 0) return T3B_OK;
    PrefillArgs a = {};
    a.q = (const bf16*)q; a.seq_start = seq_start; a.seq_row = seq_row; a.block_tables = block_tables; a.max_blocks = max_blocks;
    a.k_pool = (const unsigned char*)k_pool; a.v_pool = (const unsigned char*)v_pool; a.n_heads = n_heads; a.out = (bf16*)out;
    a.max_pages = (max_seq_len + 15) / 16;
    if (a.max_pages > max_blocks) a.max_pages = max_blocks;
    if (kv_dtype == T3B_KV_BF16) {                     // tensor-core kernel: a warp per 16-query tile
        const size_t smem_mma = (size_t)a.max_pages * 2 * 2048;
        T3B_CUDA(cudaFuncSetAttribute(prefill_attn_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_mma));
        T3B_CUDA(t3b_launch(prefill_attn_mma_kernel, dim3((unsigned)(n_seq * n_heads)), dim3(PM_NW * 32), smem_mma, t3b_stream(stream), a));
        return T3B_OK;
    }
    const int slice = kv_dtype == T3B_KV_FP8 ? KvFmt<1>::SLICE : KvFmt<0>::SLICE;
    const size_t smem = (size_t)a.max_pages * 2 * slice;
    auto kern = kv_dtype == T3B_KV_FP8 ? prefill_attn_kernel<1> : prefill_attn_kernel<0>;
    T3B_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    T3B_CUDA(t3b_launch(kern, dim3((unsigned)(n_seq * n_heads)), dim3(PF_NW * 32), smem, t3b_stream(stream), a));
    return T3B_OK;
}

// ------------------------------------------------------------------ host
// Warps per item.  0 = by item count: the largest power of two <= 8 that still gives every item its own team in one wave
// (items x W <= SMs x 8).  An engine passes a fixed value (t3b_options_t.attn_team_warps) so that a request's summation order
// does not change with the number of live requests.
static int at_team_warps(int n_items, int requested) {
    if (requested == 1 || requested == 2 || requested == 4 || requested == 8) return requested;
    const int slots = t3b_num_sms() * AT_NW * AT_CTAS;
    int w = AT_NW;
    while (w > 1 && (long)n_items * w > slots) w >>= 1;
    return w;
}

template <bool FUSED, int W, int S, int KVT>
static cudaError_t launch_ring(const AttnArgs& a, cudaStream_t s) {
    constexpr int PAIR = 4 * KvFmt<KVT>::SLICE;
    constexpr int T = AT_NW / W;
    const size_t smem = (size_t)AT_NW * S * PAIR + AT_NW * S * 8 + 2 * AT_NW * 66 * 4 + 4 * AT_CAP * 4;
    auto kern = attn_ring_kernel<FUSED, W, S, KVT>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    const int slots = t3b_num_sms() * AT_CTAS;            // resident CTAs
    // teams are dealt CTA-first (team j of every CTA before team j + 1 of any), so every SM pulls its share of the bytes:
    // one CTA per SM as soon as there are that many items (1024 items -> 148 CTAs with 6-7 busy warps, not 128 full ones)
    int grid = a.n_items < slots ? a.n_items : slots;
    const int need = (a.n_items + AT_CAP - 1) / AT_CAP;
    if (grid < need) grid = need;
    return t3b_launch(kern, dim3((unsigned)grid), dim3(AT_NW * 32), smem, s, a);
}

template <bool FUSED, int S, int KVT>
static cudaError_t launch_ring_w(const AttnArgs& a, int w, cudaStream_t s) {
    switch (w) {
        case 1: return launch_ring<FUSED, 1, S, KVT>(a, s);
        case 2: return launch_ring<FUSED, 2, S, KVT>(a, s);
        case 4: return launch_ring<FUSED, 4, S, KVT>(a, s);
        default: return launch_ring<FUSED, 8, S, KVT>(a, s);
    }
}

static int32_t attn_launch(bool fused, const AttnArgs& a, int kv_dtype, int team_warps, cudaStream_t s) {
    T3B_CHECK_ARG(kv_dtype == T3B_KV_BF16 || kv_dtype == T3B_KV_FP8, "attention: unknown KV dtype %d", kv_dtype);
    const int w = at_team_warps(a.n_items, team_warps);
    cudaError_t e;
    if (kv_dtype == T3B_KV_FP8) e = fused ? launch_ring_w<true, 5, 1>(a, w, s) : launch_ring_w<false, 5, 1>(a, w, s);
    // bf16: 2 pair slots per warp = 131 KB per block.  Bandwidth does not need a third slot (8 warps x 2 x 8 KB in flight per SM =
    // 19 MB over the chip, more than twice the ~8 MB latency-bandwidth product of HBM), and at 131 KB a
    // block of the o projection that follows (80 KB) fits on the SM beside it: 64 rows 1.337 -> 1.315 ms/step, 16 rows
    // 0.982 -> 0.931, no config slower (profiles/r2_notes.md r4j).  fp8: 5 slots of 4.25 KB (measured in round 2).
    else e = fused ? launch_ring_w<true, 2, 0>(a, w, s) : launch_ring_w<false, 2, 0>(a, w, s);
    T3B_CUDA(e);
    return T3B_OK;
}

int32_t t3b_paged_attn_launch(const void* q, int n_tok, const int32_t* tok_row, const int32_t* tok_idx, const int32_t* block_tables,
                              int max_blocks, const void* k_pool, const void* v_pool, int n_heads, int kv_dtype, void* out, int team_warps,
                              cudaStream_t s) {
    AttnArgs a = {};
    a.q = (const bf16*)q; a.tok_row = tok_row; a.tok_idx = tok_idx; a.block_tables = block_tables; a.max_blocks = max_blocks;
    a.k_pool = (unsigned char*)const_cast<void*>(k_pool); a.v_pool = (unsigned char*)const_cast<void*>(v_pool);
    a.n_heads = n_heads; a.out = (bf16*)out; a.n_items = n_tok * n_heads;
    return attn_launch(false, a, kv_dtype, team_warps, s);
}

// decode step: one token per KV row (token j <-> row j)
int32_t t3b_attn_fused_launch(const void* qkv, int n_tok, const int32_t* tok_idx, const int32_t* rope_pos, const void* rope_cos_sin, int max_pos,
                              const int32_t* block_tables, int max_blocks, void* k_pool, void* v_pool, int n_heads, int kv_dtype, void* out,
                              int team_warps, PfArgs pf, const int32_t* row_order, cudaStream_t s) {
    AttnArgs a = {};
    a.pf = pf;
    a.row_order = row_order;
    a.q = (const bf16*)qkv; a.tok_idx = tok_idx; a.rope_pos = rope_pos; a.cos_sin = (const bf16*)rope_cos_sin; a.max_pos = max_pos;
    a.block_tables = block_tables; a.max_blocks = max_blocks; a.k_pool = (unsigned char*)k_pool; a.v_pool = (unsigned char*)v_pool;
    a.n_heads = n_heads; a.out = (bf16*)out; a.n_items = n_tok * n_heads;
    return attn_launch(true, a, kv_dtype, team_warps, s);
}

extern "C" int32_t t3b_attn_decode_fused(const void* qkv, int32_t n_tok, const int32_t* tok_idx, const int32_t* rope_pos,
                                         const void* rope_cos_sin, int32_t max_pos, const int32_t* block_tables, int32_t max_blocks,
                                         void* k_pool, void* v_pool, int32_t n_heads, int32_t kv_dtype, void* out,
                                         const t3b_options_t* opt, void* stream) {
    T3B_CHECK_ARG(qkv && tok_idx && rope_pos && rope_cos_sin && block_tables && k_pool && v_pool && out, "t3b_attn_decode_fused: null pointer");
    T3B_CHECK_ARG(n_tok >= 0 && max_blocks > 0 && max_pos > 0 && n_heads >= 1, "t3b_attn_decode_fused: bad sizes");
    T3B_CHECK_ALIGN(qkv); T3B_CHECK_ALIGN(rope_cos_sin); T3B_CHECK_ALIGN(k_pool); T3B_CHECK_ALIGN(v_pool); T3B_CHECK_ALIGN(out);
    if (n_tok == 0) return T3B_OK;
    return t3b_attn_fused_launch(qkv, n_tok, tok_idx, rope_pos, rope_cos_sin, max_pos, block_tables, max_blocks, k_pool, v_pool, n_heads,
                                 kv_dtype, out, opt ? opt->attn_team_warps : 0, pf_none(), nullptr, t3b_stream(stream));
}

extern "C" int32_t t3b_paged_attn_decode(const void* q, int32_t n_tok, const int32_t* tok_row, const int32_t* tok_idx,
                                         const int32_t* block_tables, int32_t max_blocks, const void* k_pool, const void* v_pool,
                                         int32_t n_heads, int32_t kv_dtype, void* out, const t3b_options_t* opt, void* stream) {
    T3B_CHECK_ARG(q && tok_idx && block_tables && k_pool && v_pool && out, "t3b_paged_attn_decode: null pointer");
    T3B_CHECK_ARG(n_tok >= 0 && max_blocks > 0 && n_heads >= 1, "t3b_paged_attn_decode: bad sizes");
    T3B_CHECK_ALIGN(q); T3B_CHECK_ALIGN(k_pool); T3B_CHECK_ALIGN(v_pool); T3B_CHECK_ALIGN(out);
    if (n_tok == 0) return T3B_OK;
    return t3b_paged_attn_launch(q, n_tok, tok_row, tok_idx, block_tables, max_blocks, k_pool, v_pool, n_heads, kv_dtype, out,
                                 opt ? opt->attn_team_warps : 0, t3b_stream(stream));
}
