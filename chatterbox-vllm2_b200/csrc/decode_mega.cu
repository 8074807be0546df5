// The whole T3 CFG decode step (E1, F1, F2.*, C1 of SURVEY.md 8(a)) as ONE persistent kernel: one CTA per SM,
// every byte the step needs that does not depend on this step's activations - the repacked weights of all
// 30 layers + the speech head and the past KV pages of this CTA's (row, head) units - flows through a single
// shared-memory ring filled by one producer warp with 1-D bulk (TMA) copies.  The producer runs ahead of the
// compute across phase boundaries, so HBM keeps streaming while the CTAs synchronise; the dependent chain of
// ~150 small phases is paid in grid-barrier latency, not in kernel launches.
//
//   warp 0      producer: walks the static schedule [qkv | KV pages | o | gate_up | down] x layers | head and
//               copies 16 KB units into the ring (cp.async.bulk + mbarrier complete_tx).
//   warp 1      MMA issuer: one lane; per weight unit 4 x tcgen05.mma (M = 128 features, N = CFG rows padded
//               to 16, K = 16), accumulator in TMEM (two 128-column buffers); tcgen05.commit frees the ring
//               slot / the X stage and publishes the accumulator.
//   warps 2-9   workers: build the activation operand (RMSNorm fused in: bf16(bf16(s * rstd) * w), written
//               128B-swizzled K-major), drain TMEM into fp32 partial tiles, reduce the partials of a tile
//               across the CTAs that share it (fixed order -> deterministic) with the epilogue fused in
//               (residual add + sum of squares | SiLU * mul | plain store), and run the paged attention of the
//               CTA's (row, head) units out of the ring (RoPE + KV write fused, online softmax, warp shuffles).
//
// Work split: a GEMM phase has U = tiles x k-blocks units of 128 x 64 weights; CTA c owns the contiguous run
// [U c / G, U (c + 1) / G) (stream-K), so every SM streams the same number of bytes (+-1 unit) whatever the
// layer shape.  A tile shared by several CTAs is reduced through L2: each writes its fp32 partial, bumps the
// tile's counter, waits until all contributors have, then reduces a disjoint set of rows.  Phases are
// separated by a grid barrier (one atomic counter); only the workers ever wait on it.
//
// The split depends on (shape, SM count) only - never on the number of rows - so a request's arithmetic
// does not change with the batch it runs in.
#include <cuda_bf16.h>

#include "t3b_common.cuh"

#define MG_THREADS 320
#define MG_WORKERS 256
#define MG_SLOT 16384
#define MG_MAX_SLOTS 12
#define MG_MAX_XST 8
#define MG_MAXSEG 3
#define MG_CTR_TILES 64            // first tile counter (counter 0 = grid barrier)

struct MegaLayer { const unsigned char *qkv, *o, *gate_up, *down; const bf16 *ln1, *ln2; };
static_assert(sizeof(MegaLayer) == sizeof(t3b_mega_layer_t), "MegaLayer mirrors t3b_mega_layer_t");

struct MegaParams {
    const MegaLayer* layers;
    const unsigned char* head;
    const bf16 *final_norm, *emb, *pos_emb, *cos_sin, *x_in;
    const int32_t *tokens, *speech_pos, *rope_pos, *ctx_lens, *block_tables;
    bf16 *k_pool, *v_pool;
    long long layer_stride;
    int n_layers, D, H, F, V, emb_rows, pos_rows, max_pos, max_blocks;
    int rows, Rp, nslot, nx, max_tiles;
    float eps;
    float* s_f32;      // [rows][D]  fp32 residual stream s = x + res (layernorm.py:40-59 keeps it in fp32 for the norm)
    float* ssq;        // [D/128][rows]  per-tile sum of s^2
    bf16 *qkv, *attn, *act, *logits, *hidden_out;
    float* partials;   // [G][MG_MAXSEG][rows][128]
    unsigned int* counters;
    unsigned long long* trace;   // debug: [G][MG_TRACE] globaltimer stamps (NULL = off)
};
#define MG_TRACE 400

enum { XK_PLAIN = 0, XK_NORM = 1 };
enum { EP_STORE = 0, EP_RESID = 1, EP_SILU = 2 };

struct Gemm {
    const unsigned char* w;   // repacked [tile][kb][128 x 64 SW128]
    int ntiles, nkb, U, geff, seq;
    int xkind; const void* xsrc; const bf16* lnw; int K;
    int epi; void* y; int N;
};

__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void worker_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }
__device__ __forceinline__ unsigned int ld_acquire(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release(unsigned int* p) { asm volatile("red.release.gpu.global.add.u32 [%0], 1;" :: "l"(p) : "memory"); }
__device__ __forceinline__ unsigned long long gtimer() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ uint4 ldcg16(const void* p) { return __ldcg(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ float4 ldcgf4(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }

// spin until *ctr >= target (bounded: a protocol bug traps instead of hanging the box)
__device__ __forceinline__ void spin_ge(const unsigned int* ctr, unsigned int target) {
    if (ld_acquire(ctr) >= target) return;
    const long long t0 = clock64();
    while (ld_acquire(ctr) < target) {
        if (clock64() - t0 > 2000000000LL) { printf("t3b mega: counter timeout block %d target %u have %u\n", blockIdx.x, target, ld_acquire(ctr)); __trap(); }
    }
}

__device__ __forceinline__ Gemm make_gemm(const MegaParams& p, int layer, int which, int G) {
    Gemm g;
    const int D = p.D, F = p.F;
    if (layer < p.n_layers) {
        const MegaLayer& L = p.layers[layer];
        g.seq = layer * 4 + which;
        if (which == 0)      { g.w = L.qkv;     g.ntiles = 3 * D / 128; g.K = D; g.xkind = XK_NORM;  g.xsrc = p.s_f32; g.lnw = L.ln1; g.epi = EP_STORE; g.y = p.qkv; g.N = 3 * D; }
        else if (which == 1) { g.w = L.o;       g.ntiles = D / 128;     g.K = D; g.xkind = XK_PLAIN; g.xsrc = p.attn;  g.lnw = nullptr; g.epi = EP_RESID; g.y = nullptr; g.N = D; }
        else if (which == 2) { g.w = L.gate_up; g.ntiles = F / 64;      g.K = D; g.xkind = XK_NORM;  g.xsrc = p.s_f32; g.lnw = L.ln2; g.epi = EP_SILU;  g.y = p.act; g.N = 2 * F; }
        else                 { g.w = L.down;    g.ntiles = D / 128;     g.K = F; g.xkind = XK_PLAIN; g.xsrc = p.act;   g.lnw = nullptr; g.epi = EP_RESID; g.y = nullptr; g.N = D; }
    } else {
        g.seq = p.n_layers * 4;
        g.w = p.head; g.ntiles = (p.V + 127) / 128; g.K = D; g.xkind = XK_NORM; g.xsrc = p.s_f32; g.lnw = p.final_norm; g.epi = EP_STORE; g.y = p.logits; g.N = p.V;
    }
    g.nkb = g.K / 64;
    g.U = g.ntiles * g.nkb;
    g.geff = g.U < G ? g.U : G;
    return g;
}
__device__ __forceinline__ int unit_begin(const Gemm& g, int c) { return c >= g.geff ? g.U : (int)(((long long)g.U * c) / g.geff); }
__device__ __forceinline__ int unit_end(const Gemm& g, int c) { return c >= g.geff ? g.U : (int)(((long long)g.U * (c + 1)) / g.geff); }
__device__ __forceinline__ int unit_owner(const Gemm& g, int u) { return (int)((((long long)(u + 1)) * g.geff - 1) / g.U); }

// ring slots this CTA's attention units take in one layer (64 tokens of one (row, head) per slot)
__device__ __forceinline__ int att_slots(const MegaParams& p, int c, int G) {
    int n = 0;
    for (int u = c; u < p.rows * p.H; u += G) n += (__ldg(p.ctx_lens + u / p.H) + 63) >> 6;
    return n;
}

struct MegaSmem {
    uint64_t full[MG_MAX_SLOTS], w_empty[MG_MAX_SLOTS], kv_empty[MG_MAX_SLOTS];
    uint64_t x_full[MG_MAX_XST], x_empty[MG_MAX_XST];
    uint64_t tmem_full[2], tmem_empty[2];
    uint32_t tmem_ptr;
    uint32_t w_cnt[MG_MAX_SLOTS], kv_cnt[MG_MAX_SLOTS];
    uint8_t ptype[MG_MAX_SLOTS];
    float rstd[128];
    float att_m[2][8], att_l[2][8];
    float att_acc[2][8][64];
};

// ===================================================================== workers (warps 2..9)
__device__ __forceinline__ void mega_workers(const MegaParams& p, MegaSmem& sm, unsigned char* ring, unsigned char* xring, uint32_t tmem_base) {
    const int tid = threadIdx.x - 64, wid = tid >> 5, lane = tid & 31;
    const int c = blockIdx.x, G = gridDim.x;
    const int rows = p.rows, D = p.D, H = p.H;
    const int nslot = p.nslot, nx = p.nx;
    const uint32_t xst_bytes = (uint32_t)p.Rp * 128u;
    uint32_t q = 0, xq = 0, tseq = 0, bar_target = 0;
    unsigned int* gbar = p.counters;
    int tr = 0;
    auto stamp = [&]() { if (p.trace != nullptr && tid == 0 && tr < MG_TRACE) p.trace[(size_t)c * MG_TRACE + tr++] = gtimer(); };
    auto grid_arrive = [&]() { stamp(); __threadfence(); worker_sync(); if (tid == 0) red_release(gbar); };
    auto grid_wait = [&]() { bar_target += (uint32_t)G; if (tid == 0) spin_ge(gbar, bar_target); worker_sync(); stamp(); };

    // ---- E1 (t3.py:440,475-480): s = bf16(speech_emb[tok] + speech_pos_emb[p]); both rows of a request get the same vector
    for (int r = c; r < rows; r += G) {
        const bf16 *e0 = nullptr, *e1 = nullptr;
        if (p.x_in == nullptr) {
            int tok = __ldg(p.tokens + (r >> 1));
            tok = min(max(tok, 0), p.emb_rows - 1);
            int pp = __ldg(p.speech_pos + (r >> 1)) % p.pos_rows;
            if (pp < 0) pp += p.pos_rows;
            e0 = p.emb + (size_t)tok * D;
            e1 = p.pos_emb + (size_t)pp * D;
        }
        for (int t = wid; t < D / 128; t += 8) {
            const int n = t * 128 + lane * 4;
            float v[4];
            if (p.x_in != nullptr) {
                const uint2 raw = *reinterpret_cast<const uint2*>(p.x_in + (size_t)r * D + n);
                v[0] = bf16lo(raw.x); v[1] = bf16hi(raw.x); v[2] = bf16lo(raw.y); v[3] = bf16hi(raw.y);
            } else {
                const uint2 a = __ldg(reinterpret_cast<const uint2*>(e0 + n)), b = __ldg(reinterpret_cast<const uint2*>(e1 + n));
                v[0] = bf16_round(bf16lo(a.x) + bf16lo(b.x)); v[1] = bf16_round(bf16hi(a.x) + bf16hi(b.x));
                v[2] = bf16_round(bf16lo(a.y) + bf16lo(b.y)); v[3] = bf16_round(bf16hi(a.y) + bf16hi(b.y));
            }
            *reinterpret_cast<float4*>(p.s_f32 + (size_t)r * D + n) = make_float4(v[0], v[1], v[2], v[3]);
            const float ss = warp_sum(v[0] * v[0] + v[1] * v[1] + v[2] * v[2] + v[3] * v[3]);
            if (lane == 0) p.ssq[t * rows + r] = ss;
        }
    }
    grid_arrive();

    // ---- one projection: X operand -> (MMA warp) -> TMEM -> fp32 partials -> cross-CTA reduce + epilogue
    auto gemm_phase = [&](const Gemm& g, bool is_head) {
        grid_wait();
        if (g.xkind == XK_NORM) {
            if (tid < rows) {
                float tot = 0.f;
                for (int t = 0; t < D / 128; ++t) tot += __ldcg(p.ssq + t * rows + tid);      // fixed order
                sm.rstd[tid] = rsqrtf(tot / (float)D + p.eps);
            }
            worker_sync();
            if (is_head && p.hidden_out != nullptr) {
                for (int r = c; r < rows; r += G)
                    for (int i = tid; i < D / 8; i += MG_WORKERS) {
                        const float4 a = ldcgf4(p.s_f32 + (size_t)r * D + i * 8), b = ldcgf4(p.s_f32 + (size_t)r * D + i * 8 + 4);
                        const float s8[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
                        float w8[8], o[8];
                        unpack8(__ldg(reinterpret_cast<const uint4*>(g.lnw) + i), w8);
                        const float rs = sm.rstd[r];
#pragma unroll
                        for (int e = 0; e < 8; ++e) o[e] = bf16_round(s8[e] * rs) * w8[e];
                        *reinterpret_cast<uint4*>(p.hidden_out + (size_t)r * D + i * 8) = pack8(o);
                    }
            }
        }
        const int b = unit_begin(g, c), e = unit_end(g, c);
        for (int u = b; u < e; ++u) {
            const int kb = u % g.nkb;
            const uint32_t xs = xq % nx;
            if (xq >= (uint32_t)nx) mbar_wait(&sm.x_empty[xs], ((xq / nx) - 1) & 1);
            unsigned char* dst = xring + (size_t)xs * xst_bytes;
            for (int i = tid; i < rows * 8; i += MG_WORKERS) {
                const int r = i >> 3, ch = i & 7;
                uint4 v;
                if (g.xkind == XK_PLAIN) {
                    v = ldcg16(reinterpret_cast<const bf16*>(g.xsrc) + (size_t)r * g.K + kb * 64 + ch * 8);
                } else {
                    const float* sp = reinterpret_cast<const float*>(g.xsrc) + (size_t)r * g.K + kb * 64 + ch * 8;
                    const float4 a = ldcgf4(sp), bb = ldcgf4(sp + 4);
                    const float s8[8] = {a.x, a.y, a.z, a.w, bb.x, bb.y, bb.z, bb.w};
                    float w8[8], o[8];
                    unpack8(__ldg(reinterpret_cast<const uint4*>(g.lnw + kb * 64 + ch * 8)), w8);
                    const float rs = sm.rstd[r];
#pragma unroll
                    for (int e2 = 0; e2 < 8; ++e2) o[e2] = bf16_round(s8[e2] * rs) * w8[e2];   // x.to(weight.dtype) * weight
                    v = pack8(o);
                }
                *reinterpret_cast<uint4*>(dst + r * 128 + ((ch ^ (r & 7)) << 4)) = v;
            }
            fence_async_smem();
            worker_sync();
            if (tid == 0) mbar_arrive(&sm.x_full[xs]);
            ++xq;
        }
        const int t0 = b < e ? b / g.nkb : 0, t1 = b < e ? (e - 1) / g.nkb : -1;
        const int quad = (wid + 2) & 3, hf = wid >> 2, f = quad * 32 + lane;
        for (int t = t0; t <= t1; ++t) {
            const uint32_t buf = tseq & 1;
            mbar_wait(&sm.tmem_full[buf], (tseq >> 1) & 1);
            tc_fence_after();
            float* part = p.partials + ((size_t)(c * MG_MAXSEG + (t - t0)) * rows) * 128;
            for (int c0 = hf * 16; c0 < rows; c0 += 32) {
                uint32_t v[16];
                tmem_ld16(tmem_base + buf * 128 + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 16; ++j)
                    if (c0 + j < rows) part[(size_t)(c0 + j) * 128 + f] = __uint_as_float(v[j]);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.tmem_empty[buf]);
            ++tseq;
        }
        __threadfence();
        worker_sync();
        unsigned int* tctr = p.counters + MG_CTR_TILES + (size_t)g.seq * p.max_tiles;
        if (tid == 0)
            for (int t = t0; t <= t1; ++t) red_release(tctr + t);
        for (int t = t0; t <= t1; ++t) {
            const int cf = unit_owner(g, t * g.nkb), cl = unit_owner(g, (t + 1) * g.nkb - 1);
            const int nseg = cl - cf + 1, myseg = c - cf;
            if (tid == 0) spin_ge(tctr + t, (unsigned int)nseg);
            worker_sync();
            for (int r = myseg + nseg * wid; r < rows; r += nseg * 8) {
                float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int sg = 0; sg < nseg; ++sg) {                                   // fixed order -> deterministic
                    const int cc = cf + sg, slot = t - unit_begin(g, cc) / g.nkb;
                    const float4 v = ldcgf4(p.partials + ((size_t)(cc * MG_MAXSEG + slot) * rows + r) * 128 + lane * 4);
                    acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
                }
                if (g.epi == EP_STORE) {
                    const int n = t * 128 + lane * 4;
                    bf16* y = reinterpret_cast<bf16*>(g.y) + (size_t)r * g.N + n;
                    if (n < g.N) *reinterpret_cast<uint32_t*>(y) = pack_bf16(acc.x, acc.y);
                    if (n + 2 < g.N) *reinterpret_cast<uint32_t*>(y + 2) = pack_bf16(acc.z, acc.w);
                } else if (g.epi == EP_RESID) {
                    const int n = t * 128 + lane * 4;
                    float* sp = p.s_f32 + (size_t)r * D + n;
                    const float4 so = ldcgf4(sp);
                    float4 sn;                                                        // s = bf16(gemm) + res, res = bf16(s_old)
                    sn.x = bf16_round(acc.x) + bf16_round(so.x); sn.y = bf16_round(acc.y) + bf16_round(so.y);
                    sn.z = bf16_round(acc.z) + bf16_round(so.z); sn.w = bf16_round(acc.w) + bf16_round(so.w);
                    *reinterpret_cast<float4*>(sp) = sn;
                    const float ss = warp_sum(sn.x * sn.x + sn.y * sn.y + sn.z * sn.z + sn.w * sn.w);
                    if (lane == 0) p.ssq[t * rows + r] = ss;
                } else {
                    // lanes 0..15 hold gate features 4 lane.., lanes 16..31 the matching up features (llama.py:117-121)
                    const float ux = __shfl_down_sync(0xffffffffu, acc.x, 16), uy = __shfl_down_sync(0xffffffffu, acc.y, 16);
                    const float uz = __shfl_down_sync(0xffffffffu, acc.z, 16), uw = __shfl_down_sync(0xffffffffu, acc.w, 16);
                    if (lane < 16) {
                        const float gg[4] = {bf16_round(acc.x), bf16_round(acc.y), bf16_round(acc.z), bf16_round(acc.w)};
                        const float uu[4] = {bf16_round(ux), bf16_round(uy), bf16_round(uz), bf16_round(uw)};
                        float a4[4];
#pragma unroll
                        for (int e2 = 0; e2 < 4; ++e2) a4[e2] = bf16_round(gg[e2] / (1.0f + expf(-gg[e2]))) * uu[e2];
                        *reinterpret_cast<uint2*>(reinterpret_cast<bf16*>(g.y) + (size_t)r * p.F + t * 64 + lane * 4) =
                            make_uint2(pack_bf16(a4[0], a4[1]), pack_bf16(a4[2], a4[3]));
                    }
                }
            }
        }
        q += (uint32_t)(e - b);
        grid_arrive();
    };
    // ---- F2.3 + F2.4: RoPE, KV write and paged attention of this CTA's (row, head) units, past KV from the ring
    auto att_phase = [&](int layer) {
        grid_wait();
        bf16* kl = p.k_pool + (size_t)layer * p.layer_stride;
        bf16* vl = p.v_pool + (size_t)layer * p.layer_stride;
        const int t4 = lane >> 3, c8 = lane & 7;
        const float kScale = 0.125f * 1.4426950408889634f;   // head_dim^-0.5 * log2(e)
        uint32_t gi = 0;
        int ui = 0;
        for (int u = c; u < rows * H; u += G, ++ui) {
            const int r = u / H, h = u - r * H;
            const int L = __ldg(p.ctx_lens + r);               // tokens already in the cache; this token is number L
            float qf[8], kf[8], vnew[8];
            {
                const bf16* src = p.qkv + (size_t)r * 3 * D + h * 64 + c8 * 8;
                float cs[8], sn[8];
                unpack8(ldcg16(src), qf);
                unpack8(ldcg16(src + D), kf);
                const uint4 vraw = ldcg16(src + 2 * D);
                unpack8(vraw, vnew);
                int pos = __ldg(p.rope_pos + r);
                pos = min(max(pos, 0), p.max_pos - 1);
                unpack8(__ldg(reinterpret_cast<const uint4*>(p.cos_sin + (size_t)pos * 64 + (c8 & 3) * 8)), cs);
                unpack8(__ldg(reinterpret_cast<const uint4*>(p.cos_sin + (size_t)pos * 64 + 32 + (c8 & 3) * 8)), sn);
                const bool hi = c8 >= 4;                       // lanes c8 < 4 hold x1 (dims 0..31), c8 >= 4 hold x2
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    const float qp = __shfl_xor_sync(0xffffffffu, qf[e], 4), kp = __shfl_xor_sync(0xffffffffu, kf[e], 4);
                    qf[e] = hi ? bf16_round(bf16_round(qf[e] * cs[e]) + bf16_round(qp * sn[e])) : bf16_round(bf16_round(qf[e] * cs[e]) - bf16_round(qp * sn[e]));
                    kf[e] = hi ? bf16_round(bf16_round(kf[e] * cs[e]) + bf16_round(kp * sn[e])) : bf16_round(bf16_round(kf[e] * cs[e]) - bf16_round(kp * sn[e]));
                }
                if (wid == 0 && t4 == 0) {
                    const int blk = __ldg(p.block_tables + (size_t)r * p.max_blocks + (L >> 4));
                    const size_t off = (((size_t)blk * H + h) * T3B_KV_BLOCK + (L & 15)) * 64 + c8 * 8;
                    *reinterpret_cast<uint4*>(kl + off) = pack8(kf);
                    *reinterpret_cast<uint4*>(vl + off) = vraw;
                }
#pragma unroll
                for (int e = 0; e < 8; ++e) qf[e] *= kScale;
            }
            float m = -INFINITY, l = 0.f, acc[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] = 0.f;
            const int nchunk = (L + 63) >> 6, npages = (L + 15) >> 4;
            for (int ch = 0; ch < nchunk; ++ch) {
                const uint32_t g2 = gi + (uint32_t)ch;
                if ((int)(g2 & 3) != (wid >> 1)) continue;       // slot g2 belongs to warp pair g2 % 4
                const uint32_t qq = q + g2, s = qq % nslot;
                mbar_wait(&sm.full[s], (qq / nslot) & 1);
                const unsigned char* sl = ring + (size_t)s * MG_SLOT;
                const int pg0 = (wid & 1) * 2, page0 = ch * 4 + pg0;
                if (page0 < npages) {
                    uint4 kv[2][4], vv[2][4];
#pragma unroll
                    for (int pp = 0; pp < 2; ++pp) {
                        const bool valid = page0 + pp < npages;
                        const unsigned char* base = sl + (pg0 + pp) * 2048 + lane * 16;
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            kv[pp][j] = valid ? *reinterpret_cast<const uint4*>(base + j * 512) : make_uint4(0, 0, 0, 0);
                            vv[pp][j] = valid ? *reinterpret_cast<const uint4*>(base + 8192 + j * 512) : make_uint4(0, 0, 0, 0);
                        }
                    }
                    float sc8[8];
#pragma unroll
                    for (int pp = 0; pp < 2; ++pp)
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            float k8[8];
                            unpack8(kv[pp][j], k8);
                            float d = 0.f;
#pragma unroll
                            for (int e = 0; e < 8; ++e) d = fmaf(qf[e], k8[e], d);
                            sc8[pp * 4 + j] = d;
                        }
                    // butterfly-transpose reduce over the 8 lanes of a token: lane c8 ends with value c8
                    float w4[4], w2[2], sc;
                    {
                        const bool up = (c8 & 4) != 0;
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const float send = up ? sc8[i] : sc8[i + 4], keep = up ? sc8[i + 4] : sc8[i];
                            w4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
                        }
                        const bool up2 = (c8 & 2) != 0;
#pragma unroll
                        for (int i = 0; i < 2; ++i) {
                            const float send = up2 ? w4[i] : w4[i + 2], keep = up2 ? w4[i + 2] : w4[i];
                            w2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
                        }
                        const bool up1 = (c8 & 1) != 0;
                        const float send = up1 ? w2[0] : w2[1], keep = up1 ? w2[1] : w2[0];
                        sc = keep + __shfl_xor_sync(0xffffffffu, send, 1);
                    }
                    const int my_tok = (page0 + (c8 >> 2)) * 16 + 4 * (c8 & 3) + t4;
                    if (my_tok >= L) sc = -INFINITY;
                    const float m_new = fmaxf(m, warp_max(sc));       // finite: page0 holds >= 1 cached token
                    const float corr = exp2f(m - m_new), pr = exp2f(sc - m_new);
                    l = l * corr + pr;
                    m = m_new;
#pragma unroll
                    for (int e = 0; e < 8; ++e) acc[e] *= corr;
#pragma unroll
                    for (int pp = 0; pp < 2; ++pp)
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const float pj = __shfl_sync(0xffffffffu, pr, t4 * 8 + pp * 4 + j);
                            float v8[8];
                            unpack8(vv[pp][j], v8);
#pragma unroll
                            for (int e = 0; e < 8; ++e) acc[e] = fmaf(pj, v8[e], acc[e]);
                        }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&sm.kv_empty[s]);
            }
            gi += (uint32_t)nchunk;
            if (wid == 0) {                                        // this step's own token, straight from registers
                float d = 0.f;
#pragma unroll
                for (int e = 0; e < 8; ++e) d = fmaf(qf[e], kf[e], d);
                d += __shfl_xor_sync(0xffffffffu, d, 1);
                d += __shfl_xor_sync(0xffffffffu, d, 2);
                d += __shfl_xor_sync(0xffffffffu, d, 4);
                const float m_new = fmaxf(m, d), corr = exp2f(m - m_new), pr = exp2f(d - m_new);
                l = l * corr + (lane == 0 ? pr : 0.f);
                m = m_new;
#pragma unroll
                for (int e = 0; e < 8; ++e) acc[e] = acc[e] * corr + (t4 == 0 ? pr * vnew[e] : 0.f);
            }
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 8);
                acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 16);
            }
            l = warp_sum(l);
            const int mb = ui & 1;
            if (lane < 8) {
#pragma unroll
                for (int e = 0; e < 8; ++e) sm.att_acc[mb][wid][lane * 8 + e] = acc[e];
            }
            if (lane == 0) { sm.att_m[mb][wid] = m; sm.att_l[mb][wid] = l; }
            worker_sync();
            if (tid < 64) {
                float mm = -INFINITY;
                for (int w = 0; w < 8; ++w) mm = fmaxf(mm, sm.att_m[mb][w]);
                float ll = 0.f, o = 0.f;
                for (int w = 0; w < 8; ++w) {                       // fixed order -> deterministic
                    const float fct = sm.att_m[mb][w] == -INFINITY ? 0.f : exp2f(sm.att_m[mb][w] - mm);
                    ll += sm.att_l[mb][w] * fct;
                    o += sm.att_acc[mb][w][tid] * fct;
                }
                p.attn[(size_t)r * D + h * 64 + tid] = __float2bfloat16_rn(o / ll);
            }
        }
        q += gi;
        grid_arrive();
    };

    for (int layer = 0; layer < p.n_layers; ++layer) {
        gemm_phase(make_gemm(p, layer, 0, G), false);
        att_phase(layer);
        gemm_phase(make_gemm(p, layer, 1, G), false);
        gemm_phase(make_gemm(p, layer, 2, G), false);
        gemm_phase(make_gemm(p, layer, 3, G), false);
    }
    gemm_phase(make_gemm(p, p.n_layers, 0, G), true);
}

__global__ void __launch_bounds__(MG_THREADS, 1) decode_mega_kernel(const MegaParams p) {
    extern __shared__ unsigned char smem_dyn[];
    // 1024-byte alignment as an OFFSET into the __shared__ array: pointer arithmetic keeps the shared address space, so every
    // access below compiles to LDS / STS (a round trip through uintptr_t made them generic LD.E / ST.E)
    unsigned char* ring = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
    unsigned char* xring = ring + (size_t)p.nslot * MG_SLOT;
    const uint32_t xst_bytes = (uint32_t)p.Rp * 128u;
    MegaSmem& sm = *reinterpret_cast<MegaSmem*>(xring + (size_t)p.nx * xst_bytes);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int c = blockIdx.x, G = gridDim.x;
    const int nslot = p.nslot, nx = p.nx;

    if (threadIdx.x == 0) {
        for (int i = 0; i < nslot; ++i) {
            mbar_init(&sm.full[i], 1);
            mbar_init(&sm.w_empty[i], 1);
            mbar_init(&sm.kv_empty[i], 2);
            sm.w_cnt[i] = 0; sm.kv_cnt[i] = 0; sm.ptype[i] = 0;
        }
        for (int i = 0; i < nx; ++i) { mbar_init(&sm.x_full[i], 1); mbar_init(&sm.x_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&sm.tmem_full[i], 1); mbar_init(&sm.tmem_empty[i], 8); }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(&sm.tmem_ptr, 256);
    // padded rows of the activation operand stay zero for the whole kernel (workers only write rows < p.rows)
    for (uint32_t i = threadIdx.x; i < (uint32_t)nx * xst_bytes / 16; i += MG_THREADS) reinterpret_cast<uint4*>(xring)[i] = make_uint4(0, 0, 0, 0);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = sm.tmem_ptr;

    if (warp == 0) {
        // ================================================================= producer
        uint32_t q = 0;
        int s = 0;
        auto acquire = [&](uint8_t type) {          // lane 0 only: wait until slot s is free again, remember what goes in
            if (q >= (uint32_t)nslot) {
                if (sm.ptype[s] == 0) { mbar_wait(&sm.w_empty[s], sm.w_cnt[s] & 1); sm.w_cnt[s]++; }
                else                  { mbar_wait(&sm.kv_empty[s], sm.kv_cnt[s] & 1); sm.kv_cnt[s]++; }
            }
            sm.ptype[s] = type;
        };
        auto emit_gemm = [&](const Gemm& g) {
            if (lane == 0) {
                const int b = unit_begin(g, c), e = unit_end(g, c);
                for (int u = b; u < e; ++u) {
                    acquire(0);
                    mbar_expect_tx(&sm.full[s], MG_SLOT);
                    bulk_load_1d(ring + (size_t)s * MG_SLOT, g.w + (size_t)u * MG_SLOT, MG_SLOT, &sm.full[s]);
                    ++q; if (++s == nslot) s = 0;
                }
            }
            q = __shfl_sync(0xffffffffu, q, 0);
            s = __shfl_sync(0xffffffffu, s, 0);
        };
        for (int layer = 0; layer < p.n_layers; ++layer) {
            emit_gemm(make_gemm(p, layer, 0, G));
            // ---- past KV of this CTA's (row, head) units: K pages at slot + 2 KB * page, V pages 8 KB further
            const bf16* kl = p.k_pool + (size_t)layer * p.layer_stride;
            const bf16* vl = p.v_pool + (size_t)layer * p.layer_stride;
            for (int u = c; u < p.rows * p.H; u += G) {
                const int r = u / p.H, h = u - r * p.H;
                const int L = __ldg(p.ctx_lens + r);
                const int npages = (L + 15) >> 4;
                for (int ch = 0; ch * 64 < L; ++ch) {
                    const int np = min(4, npages - ch * 4);
                    if (lane == 0) { acquire(1); mbar_expect_tx(&sm.full[s], (uint32_t)np * 4096u); }
                    __syncwarp();
                    const int pg = lane & 3, isv = (lane >> 2) & 1;
                    if (lane < 8 && pg < np) {
                        const int blk = __ldg(p.block_tables + (size_t)r * p.max_blocks + ch * 4 + pg);
                        const bf16* src = (isv ? vl : kl) + ((size_t)blk * p.H + h) * (T3B_KV_BLOCK * 64);
                        bulk_load_1d(ring + (size_t)s * MG_SLOT + isv * 8192 + pg * 2048, src, 2048, &sm.full[s]);
                    }
                    ++q; if (++s == nslot) s = 0;
                }
            }
            emit_gemm(make_gemm(p, layer, 1, G));
            emit_gemm(make_gemm(p, layer, 2, G));
            emit_gemm(make_gemm(p, layer, 3, G));
        }
        emit_gemm(make_gemm(p, p.n_layers, 0, G));
    } else if (warp == 1) {
        // ================================================================= MMA issuer
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_bf16(128, (uint32_t)p.Rp);
            uint32_t q = 0, xq = 0, tseq = 0;
            auto run_gemm = [&](const Gemm& g) {
                const int b = unit_begin(g, c), e = unit_end(g, c);
                uint32_t buf = 0;
                for (int u = b; u < e; ++u) {
                    const int kb = u % g.nkb;
                    const bool first = (u == b) || kb == 0, last = (u == e - 1) || kb == g.nkb - 1;
                    if (first) {
                        buf = tseq & 1;
                        if (tseq >= 2) mbar_wait(&sm.tmem_empty[buf], ((tseq >> 1) - 1) & 1);
                    }
                    const uint32_t s = q % nslot, xs = xq % nx;
                    mbar_wait(&sm.full[s], (q / nslot) & 1);
                    mbar_wait(&sm.x_full[xs], (xq / nx) & 1);
                    tc_fence_after();
                    const uint64_t a_desc = umma_smem_desc_sw128(smem_u32(ring + (size_t)s * MG_SLOT));
                    const uint64_t b_desc = umma_smem_desc_sw128(smem_u32(xring + (size_t)xs * xst_bytes));
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_bf16(tmem_base + buf * 128, a_desc + (uint64_t)(2 * k), b_desc + (uint64_t)(2 * k), idesc, (uint32_t)(!first || k != 0));
                    umma_commit(&sm.w_empty[s]);
                    umma_commit(&sm.x_empty[xs]);
                    if (last) { umma_commit(&sm.tmem_full[buf]); ++tseq; }
                    ++q; ++xq;
                }
            };
            for (int layer = 0; layer < p.n_layers; ++layer) {
                run_gemm(make_gemm(p, layer, 0, G));
                q += (uint32_t)att_slots(p, c, G);
                run_gemm(make_gemm(p, layer, 1, G));
                run_gemm(make_gemm(p, layer, 2, G));
                run_gemm(make_gemm(p, layer, 3, G));
            }
            run_gemm(make_gemm(p, p.n_layers, 0, G));
        }
    } else {
        mega_workers(p, sm, ring, xring, tmem_base);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, 256);
}

// ------------------------------------------------------------------ one-off weight layout transform
// bf16 [n, k] row-major (nn.Linear) -> [tile][kb][128 rows x 64 k] blocks of 16 KB, each already in the
// 128B-swizzled K-major form tcgen05.mma reads (16-byte chunk c of row r sits at chunk c ^ (r & 7)), so a
// weight unit is ONE contiguous 16 KB bulk copy.  gate_up tiles pair 64 gate rows with their 64 up rows.
// Rows past n (the head's last tile) are zero.
__global__ void repack_bf16_kernel(const uint4* __restrict__ w, int n, int k, int gate_up, uint4* __restrict__ out, size_t n_chunks) {
    const int nkb = k / 64;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_chunks; i += (size_t)gridDim.x * blockDim.x) {
        const int pc = (int)(i & 7), row = (int)((i >> 3) & 127);
        const size_t blk = i >> 10;
        const int kb = (int)(blk % (size_t)nkb), tile = (int)(blk / (size_t)nkb);
        const int ch = pc ^ (row & 7);
        const int src_row = gate_up ? (row < 64 ? tile * 64 + row : n / 2 + tile * 64 + (row - 64)) : tile * 128 + row;
        out[i] = src_row < n ? w[((size_t)src_row * k + kb * 64 + ch * 8) / 8] : make_uint4(0, 0, 0, 0);
    }
}

static inline size_t mg_al(size_t x) { return (x + 255) & ~size_t(255); }

extern "C" size_t t3b_repacked_bytes(int32_t n, int32_t k) {
    if (n <= 0 || k <= 0) return 0;
    return (size_t)((n + 127) / 128) * 128 * (size_t)k * 2;
}

extern "C" int32_t t3b_repack_weights_bf16(const void* w, int32_t n, int32_t k, int32_t gate_up_interleave, void* out, void* stream) {
    T3B_CHECK_ARG(w && out, "t3b_repack_weights_bf16: null pointer");
    T3B_CHECK_SHAPE(n > 0 && k > 0 && k % 64 == 0, "t3b_repack_weights_bf16: k %d must be a multiple of 64", k);
    T3B_CHECK_SHAPE(!gate_up_interleave || n % 128 == 0, "t3b_repack_weights_bf16: gate_up n %d must be a multiple of 128", n);
    T3B_CHECK_ALIGN(w); T3B_CHECK_ALIGN(out);
    const size_t chunks = t3b_repacked_bytes(n, k) / 16;
    int blocks = (int)((chunks + 255) / 256);
    if (blocks > t3b_num_sms() * 16) blocks = t3b_num_sms() * 16;
    repack_bf16_kernel<<<blocks, 256, 0, t3b_stream(stream)>>>((const uint4*)w, n, k, gate_up_interleave ? 1 : 0, (uint4*)out, chunks);
    T3B_LAUNCH_CHECK();
    return T3B_OK;
}

// ------------------------------------------------------------------ workspace + launch
namespace {
struct MegaWs {
    char *s_f32, *ssq, *qkv, *attn, *act, *logits, *partials, *counters;
    size_t counter_bytes, total;
};
int mega_max_tiles(const t3b_model_t* m) {
    int t = 3 * m->hidden / 128;
    if (m->ffn / 64 > t) t = m->ffn / 64;
    if ((m->speech_vocab + 127) / 128 > t) t = (m->speech_vocab + 127) / 128;
    return t;
}
MegaWs mega_carve(const t3b_model_t* m, int rows, void* base) {
    MegaWs w;
    char* p = static_cast<char*>(base);
    const size_t D = m->hidden, F = m->ffn, V = m->speech_vocab, R = rows, G = t3b_num_sms();
    auto take = [&](size_t bytes) { char* r = p; p += mg_al(bytes); return r; };
    w.s_f32 = take(R * D * 4);
    w.ssq = take((D / 128) * R * 4);
    w.qkv = take(R * 3 * D * 2);
    w.attn = take(R * D * 2);
    w.act = take(R * F * 2);
    w.logits = take(R * V * 2 + 64);
    w.partials = take(G * MG_MAXSEG * R * 128 * 4);
    w.counter_bytes = (MG_CTR_TILES + (size_t)(4 * m->n_layers + 1) * mega_max_tiles(m)) * 4;
    w.counters = take(w.counter_bytes);
    w.total = static_cast<size_t>(p - static_cast<char*>(base));
    return w;
}
unsigned long long* g_mega_trace = nullptr;
}  // namespace

extern "C" void t3b_debug_set_mega_trace(void* p) { g_mega_trace = static_cast<unsigned long long*>(p); }
extern "C" int32_t t3b_mega_trace_stride(void) { return MG_TRACE; }

extern "C" int32_t t3b_mega_max_rows(void) { return 128; }

extern "C" size_t t3b_mega_workspace_bytes(const t3b_model_t* m, int32_t max_rows) {
    if (!m || max_rows <= 0) return 0;
    return mega_carve(m, max_rows, nullptr).total + 256;
}

static bool mega_shape_ok(int U, int nkb, int G) {
    const int geff = U < G ? U : G;
    for (int c = 0; c < geff; ++c) {
        const int b = (int)(((long long)U * c) / geff), e = (int)(((long long)U * (c + 1)) / geff);
        if (e > b && (e - 1) / nkb - b / nkb + 1 > MG_MAXSEG) return false;
    }
    return true;
}

// Whole decode step in one launch: embeddings (or x_in) -> 30 layers -> final norm (hidden_out, optional)
// -> speech head logits bf16 [rows, vocab] (logits_out).  rows <= t3b_mega_max_rows().
extern "C" int32_t t3b_decode_step_mega(const t3b_model_t* m, const t3b_mega_t* mw, const t3b_kv_t* kv, const t3b_state_t* st,
                                        const void* x_in, int32_t rows, void* hidden_out, void* logits_out, void* workspace,
                                        size_t workspace_bytes, void* stream) {
    T3B_CHECK_ARG(m && mw && mw->layers && mw->head && kv && kv->k && kv->v && st && workspace, "t3b_decode_step_mega: null pointer");
    T3B_CHECK_ARG(m->final_norm && m->speech_emb && m->speech_pos_emb && m->rope_cos_sin, "t3b_decode_step_mega: model tables are null");
    T3B_CHECK_ARG(st->rope_pos && st->ctx_lens && st->block_tables && st->max_blocks > 0, "t3b_decode_step_mega: state arrays are null");
    T3B_CHECK_ARG(x_in || (st->tokens && st->speech_pos), "t3b_decode_step_mega: no input (x_in and state tokens are null)");
    T3B_CHECK_ARG(rows >= 0 && rows % 2 == 0 && rows <= 128, "t3b_decode_step_mega: rows %d must be even and <= 128", rows);
    T3B_CHECK_SHAPE(m->head_dim == 64 && m->hidden == m->n_heads * 64 && m->hidden % 128 == 0 && m->ffn % 64 == 0 && m->speech_vocab % 2 == 0,
                    "t3b_decode_step_mega: unsupported model shape");
    T3B_CHECK_ALIGN(workspace);
    if (hidden_out) T3B_CHECK_ALIGN(hidden_out);
    if (rows == 0) return T3B_OK;
    T3B_CHECK_ARG(workspace_bytes >= t3b_mega_workspace_bytes(m, rows), "t3b_decode_step_mega: workspace %zu < %zu bytes", workspace_bytes,
                  t3b_mega_workspace_bytes(m, rows));
    const int G = t3b_num_sms();
    const int D = m->hidden, F = m->ffn, V = m->speech_vocab;
    T3B_CHECK_SHAPE(mega_shape_ok(3 * D / 128 * (D / 64), D / 64, G) && mega_shape_ok(D / 128 * (D / 64), D / 64, G) &&
                    mega_shape_ok(F / 64 * (D / 64), D / 64, G) && mega_shape_ok(D / 128 * (F / 64), F / 64, G) &&
                    mega_shape_ok((V + 127) / 128 * (D / 64), D / 64, G),
                    "t3b_decode_step_mega: a CTA would span more than %d tiles in one projection", MG_MAXSEG);
    const MegaWs ws = mega_carve(m, rows, workspace);
    cudaStream_t s = t3b_stream(stream);

    MegaParams p = {};
    p.layers = reinterpret_cast<const MegaLayer*>(mw->layers);
    p.head = static_cast<const unsigned char*>(mw->head);
    p.final_norm = (const bf16*)m->final_norm; p.emb = (const bf16*)m->speech_emb; p.pos_emb = (const bf16*)m->speech_pos_emb;
    p.cos_sin = (const bf16*)m->rope_cos_sin; p.x_in = (const bf16*)x_in;
    p.tokens = st->tokens; p.speech_pos = st->speech_pos; p.rope_pos = st->rope_pos; p.ctx_lens = st->ctx_lens; p.block_tables = st->block_tables;
    p.k_pool = (bf16*)kv->k; p.v_pool = (bf16*)kv->v; p.layer_stride = kv->layer_stride;
    p.n_layers = m->n_layers; p.D = D; p.H = m->n_heads; p.F = F; p.V = V; p.emb_rows = m->speech_vocab; p.pos_rows = m->speech_pos_rows;
    p.max_pos = m->max_pos; p.max_blocks = st->max_blocks; p.eps = m->rms_eps;
    p.rows = rows; p.Rp = (rows + 15) & ~15;
    p.nx = 65536 / (p.Rp * 128);
    if (p.nx > MG_MAX_XST) p.nx = MG_MAX_XST;
    const size_t x_bytes = (size_t)p.nx * p.Rp * 128;
    const size_t budget = 227 * 1024 - 1024 - sizeof(MegaSmem) - 64 - x_bytes;
    p.nslot = (int)(budget / MG_SLOT);
    if (p.nslot > MG_MAX_SLOTS) p.nslot = MG_MAX_SLOTS;
    p.max_tiles = mega_max_tiles(m);
    p.s_f32 = (float*)ws.s_f32; p.ssq = (float*)ws.ssq; p.qkv = (bf16*)ws.qkv; p.attn = (bf16*)ws.attn; p.act = (bf16*)ws.act;
    p.logits = logits_out ? (bf16*)logits_out : (bf16*)ws.logits;
    p.hidden_out = (bf16*)hidden_out;
    p.partials = (float*)ws.partials; p.counters = (unsigned int*)ws.counters;
    p.trace = g_mega_trace;

    const size_t smem = 1024 + (size_t)p.nslot * MG_SLOT + x_bytes + sizeof(MegaSmem) + 64;
    T3B_CUDA(cudaFuncSetAttribute(decode_mega_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    T3B_CUDA(cudaMemsetAsync(ws.counters, 0, ws.counter_bytes, s));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)G); cfg.blockDim = dim3(MG_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;       // all G CTAs must be co-resident: they meet at grid barriers
    attr[0].val.cooperative = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    T3B_CUDA(cudaLaunchKernelEx(&cfg, decode_mega_kernel, p));
    return T3B_OK;
}
