// The whole T3 CFG decode step (E1, F1, F2.*, C1 head of SURVEY.md 8(a)) as ONE persistent kernel with only
// TWO global exchange points per layer.
//
// Lesson of decode_mega.cu: a grid barrier costs ~1.1 us, but every *phase* behind it pays a
// chain of L2 round trips, so the step time is (number of global exchanges) x (phase latency).  This kernel
// cuts the exchanges from 5 to 2 per layer by choosing a work split in which every contraction whose input
// is produced by many CTAs is either cluster-local or a split-K whose partial sum merges into the next
// exchange:
//
//   grid = 32 clusters x 4 CTAs (one CTA per SM, 128 SMs; a B200 co-schedules at most 15 clusters of 8 but 33
//   of 4).  Inside a cluster CTAs talk through distributed shared memory.
//
//   phase A (cluster = (head h, row parity z): the cond rows of a head go to one cluster, the uncond rows to
//            the other - both rows of a request have the same context length, so the two are balanced):
//     qkv_h   CTA j computes 48 of the head's 192 q|k|v features over the FULL K = 1024 for its rows (tcgen05,
//             weights M-side, CFG rows N-side) and broadcasts them (bf16) to the 4 CTAs' smem.
//     attn_h  the (row, 64-token chunk) items are dealt round-robin to the 4 CTAs; KV pages come through the
//             TMA ring; partial (m, l, acc) go to the row's owner CTA, which adds this step's own token
//             (RoPE + KV write fused), normalises and broadcasts attn_h[row] (bf16).
//     o_h     CTA j computes output features [256 j, 256 j + 256) of  attn_h x Wo[:, 64 h .. 64 h + 64]^T
//             (a split-K over heads) -> fp32 partial P[h][row][1024] in global memory (L2).
//   exchange: s += bf16(sum_h P[h]);  xn = rmsnorm(s) * ln2       (row-parallel, fixed order -> deterministic)
//   phase B (cluster = FFN slice [128 c, 128 c + 128), all rows):
//     gate_up CTA j computes gate/up features [128 c + 32 j, +32) over the full K, SiLU * mul, broadcasts
//             act (bf16) to the cluster.
//     down    CTA j computes output features [256 j, +256) over K = the cluster's 128 act features
//             (a split-K over clusters) -> P[c][row][1024].
//   exchange: s += bf16(sum_c P[c]);  xn = rmsnorm(s) * ln1(next layer) | final norm
//   head:     72 vocabulary rows per CTA over the full K -> bf16 logits.
//
// Everything a step reads that does not depend on its activations (all weights, past KV pages) is fetched by
// one producer warp with 1-D bulk copies into a 16 KB-slot ring and runs ahead of the compute across phase and
// exchange boundaries.  Weights are repacked once (t3b_fused_repack) into per-CTA contiguous streams of
// UMMA-canonical tiles (K-major, SWIZZLE_128B), in the order the producer walks them.
//
//   warp 0      producer (TMA ring: weights + KV pages)
//   warp 1      MMA issuer (one lane): tcgen05.mma M = 128 (only the first 48 / 64 / 72 / 128 lanes carry
//               features; the tile rows past them read whatever follows in the ring and land in TMEM lanes
//               nobody reads), N = rows padded to 16, accumulators double-buffered in TMEM
//   warps 2-9   workers: exchange reduce + RMSNorm, TMA issue of the activation operand, TMEM drain,
//               DSMEM broadcasts, attention, epilogues
//
// Arithmetic is the bf16 pipeline of the oracle (rounding after every GEMM, norm, residual, RoPE, attention,
// SiLU*mul); reductions run in fixed orders that depend on shapes only, never on the batch: results are
// deterministic and a request's bits do not depend on the requests it shares a step with.
#include <cuda_bf16.h>

#include "t3b_common.cuh"


#define FZ_THREADS 320
#define FZ_WORKERS 256
#define FZ_SLOT 16384
#define FZ_MAX_SLOTS 12
#define FZ_CL 4                       // CTAs per cluster
#define FZ_NCL 32                     // clusters = (head, row parity) in phase A = FFN slices in phase B
#define FZ_NH 16                      // heads
#define FZ_NCTA (FZ_CL * FZ_NCL)
#define FZ_D 1024
#define FZ_F 4096
#define FZ_XBYTES 32768               // activation-operand ring
#define FZ_MAX_ROWS 128
#define FZ_QKV_ROWS 48
#define FZ_QKV_KB_BYTES (FZ_QKV_ROWS * 128)
#define FZ_QKV_BYTES 98304            // 16 k-blocks x 48 rows x 128 B
#define FZ_O_BYTES 32768              //  2 tiles x 128 rows (K = 64)
#define FZ_A_BYTES (FZ_QKV_BYTES + FZ_O_BYTES)        // per (head, rank): shared by the two clusters of a head
#define FZ_GU_BYTES 131072            // 16 k-blocks x 64 rows (32 gate + 32 up)
#define FZ_DN_BYTES 65536             //  2 tiles x 2 k-blocks x 128 rows
#define FZ_B_BYTES (FZ_GU_BYTES + FZ_DN_BYTES)        // per (cluster, rank)
#define FZ_LAYER_A_BYTES ((size_t)FZ_NH * FZ_CL * FZ_A_BYTES)      //  8 MB
#define FZ_LAYER_BYTES (FZ_LAYER_A_BYTES + (size_t)FZ_NCTA * FZ_B_BYTES)   // 32 MB = the layer's weights, no duplication
#define FZ_HEAD_ROWS 72
#define FZ_HEAD_KB_BYTES (FZ_HEAD_ROWS * 128)
#define FZ_HEAD_CTA_BYTES (16 * FZ_HEAD_KB_BYTES)
#define FZ_REC 68                     // floats per attention partial record: m, l, pad, pad, acc[64]
#define FZ_TRACE 2048                 // trace words per CTA: (id, globaltimer) pairs

struct FusedParams {
    const unsigned char* wl;          // [n_layers]{[16 heads][4][qkv | o], [32 clusters][4][gate_up | down]}
    const unsigned char* wh;          // [128 CTAs][144 KB]
    const bf16* norms;                // [2 n_layers + 1][1024]: ln1, ln2 per layer, final norm
    const bf16 *emb, *pos_emb, *cos_sin, *x_in;
    const int32_t *tokens, *speech_pos, *rope_pos, *ctx_lens, *block_tables;
    bf16 *k_pool, *v_pool;
    long long layer_stride;
    int n_layers, V, emb_rows, pos_rows, max_pos, max_blocks;
    int rows, Rp, Rh, Rph, nslot, nx, n_head_ctas;      // Rh = rows / 2 (one parity), Rp / Rph = padded to 16
    uint32_t off_qkvraw, off_attpart, off_attnx, off_actx, dsm_bytes;
    uint32_t meta_pages_max;          // capacity of the per-CTA KV page list (slots)
    float eps;
    float* s;                         // [2][rows][1024] fp32 residual stream (double buffered across exchanges)
    bf16* xn;                         // [rows][1024]    normalised activations handed to the next phase
    float* P;                         // [32][rows][1024] split-K partials of the o (16 used) / down (32) projections
    bf16 *logits, *hidden_out;
    unsigned int* counters;
    unsigned long long* trace;
    int dbg;                          // debug switches (t3b_debug_set_fused_flags)
};

struct FusedSmem {
    uint64_t full[FZ_MAX_SLOTS], w_empty[FZ_MAX_SLOTS], kv_empty[FZ_MAX_SLOTS];
    uint64_t x_full[16], x_empty[16];
    uint64_t tmem_full[2], tmem_empty[2];
    uint64_t attnx_full, actx_full;
    uint64_t cl_bar[2];
    uint32_t tmem_ptr;
    uint32_t w_cnt[FZ_MAX_SLOTS], kv_cnt[FZ_MAX_SLOTS];
    uint8_t ptype[FZ_MAX_SLOTS];
    float red[8];
    int n_groups, n_kv_slots;
    uint8_t grp_row[FZ_MAX_ROWS];
    uint16_t grp_slot0[FZ_MAX_ROWS];
};

// per-CTA copies of the step's request metadata (read once at kernel start: the producer and the attention warps
// would otherwise pay an L2 round trip per KV slot / per row and layer)
struct FusedMeta {
    int32_t* pages;       // [n_kv_slots][4] physical block of each 16-token page of a KV slot (ring order)
    int32_t* L;           // [Rh] cached tokens of local row lr (global row 2 lr + z)
    int32_t* blk_new;     // [Rh] physical block this step's token of the row is written to
    uint4* cs;            // [Rh][8] cos(32) | sin(32) bf16 of the row's RoPE position
};

namespace {
__device__ __forceinline__ void fz_fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fz_fence_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void fz_worker_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }
__device__ __forceinline__ unsigned int fz_ld_acquire(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void fz_red_release(unsigned int* p) { asm volatile("red.release.gpu.global.add.u32 [%0], 1;" :: "l"(p) : "memory"); }
__device__ __forceinline__ unsigned long long fz_gtimer() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ uint4 fz_ldcg16(const void* p) { return __ldcg(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ float4 fz_ldcgf4(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ uint32_t fz_cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t fz_clusterid() { uint32_t r; asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r)); return r; }
__device__ __forceinline__ void fz_cluster_barrier_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t fz_mapa(uint32_t saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
__device__ __forceinline__ void fz_st_cluster16(uint32_t raddr, const uint4& v) {
    asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" :: "r"(raddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void fz_st_cluster8(uint32_t raddr, float a, float b) {
    asm volatile("st.shared::cluster.v2.f32 [%0], {%1, %2};" :: "r"(raddr), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ void fz_remote_arrive(uint32_t raddr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" :: "r"(raddr) : "memory");
}
__device__ __forceinline__ bool fz_try_wait_cluster(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\t"
                 "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
                 "selp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
// one lane of a converged warp (the tcgen05 issue / commit instructions are single-thread; keeping the surrounding
// control flow warp-uniform lets ptxas hold descriptors in uniform registers instead of R2UR-ing them per MMA)
__device__ __forceinline__ bool fz_elect_one() {
    uint32_t p;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(p));
    return p != 0;
}
// spin until *ctr >= target (bounded: a protocol bug traps instead of hanging the box)
__device__ __forceinline__ void fz_spin_ge(const unsigned int* ctr, unsigned int target) {
    if (fz_ld_acquire(ctr) >= target) return;
    const long long t0 = clock64();
    while (fz_ld_acquire(ctr) < target) {
        if (clock64() - t0 > 2000000000LL) { printf("t3b fused: grid barrier timeout block %d target %u have %u\n", blockIdx.x, target, fz_ld_acquire(ctr)); __trap(); }
    }
}
}  // namespace


// ===================================================================== workers (warps 2..9)
__device__ __forceinline__ void fused_workers(const FusedParams& p, FusedSmem& sm, const FusedMeta& meta, unsigned char* ring,
                                              unsigned char* xring, unsigned char* dsm, uint32_t tmem_base) {
    const int tid = threadIdx.x - 64, wid = tid >> 5, lane = tid & 31;
    const int j = (int)fz_cluster_ctarank(), cl = (int)fz_clusterid(), b = cl * FZ_CL + j;
    const int h = cl >> 1, z = cl & 1;                               // phase A: head, row parity of this cluster
    const int rows = p.rows, Rp = p.Rp, Rh = p.Rh, Rph = p.Rph, nslot = p.nslot, nx = p.nx;
    const uint32_t xst = (uint32_t)Rp * 128u;
    const uint32_t dsm_a = smem_u32(dsm);
    const bf16* qraw = reinterpret_cast<const bf16*>(dsm + p.off_qkvraw);          // [192 features][Rph local rows]
    const float* attpart = reinterpret_cast<const float*>(dsm + p.off_attpart);    // [Rh / 4][4 src][FZ_REC]
    uint32_t q = 0, xq = 0, tseq = 0, bar_target = 0, ncl = 0;
    unsigned int* gbar = p.counters;
    int tr = 0;
    auto stamp = [&](int id) {
        if (p.trace != nullptr && tid == 0 && tr + 2 <= FZ_TRACE) { p.trace[(size_t)b * FZ_TRACE + tr] = (unsigned long long)id; p.trace[(size_t)b * FZ_TRACE + tr + 1] = fz_gtimer(); tr += 2; }
    };
    auto grid_arrive = [&]() { stamp(1); __threadfence(); fz_worker_sync(); if (tid == 0) fz_red_release(gbar); };
    auto grid_wait = [&]() { bar_target += (uint32_t)FZ_NCTA; if (tid == 0) fz_spin_ge(gbar, bar_target); fz_worker_sync(); stamp(2); };
    // barrier of the 4 x 256 worker threads of the cluster (producer / MMA warps run ahead, so the hardware
    // barrier.cluster cannot be used): remote mbarrier arrivals, two alternating barriers
    auto cluster_sync = [&]() {
        asm volatile("fence.acq_rel.cluster;" ::: "memory");
        fz_worker_sync();
        uint64_t* bar = &sm.cl_bar[ncl & 1];
        if (tid < FZ_CL) fz_remote_arrive(fz_mapa(smem_u32(bar), (uint32_t)tid));
        const uint32_t par = (ncl >> 1) & 1;
        if (!fz_try_wait_cluster(bar, par)) {
            const long long t0 = clock64();
            while (!fz_try_wait_cluster(bar, par))
                if (clock64() - t0 > 2000000000LL) { printf("t3b fused: cluster barrier timeout block %d sync %u\n", blockIdx.x, ncl); __trap(); }
        }
        ++ncl;
    };

    // ---- exchange e: s_new = (e == 0 ? embedding : bf16(sum_i P[i]) + bf16(s_old)); xn = bf16(bf16(s_new * rstd) * norm_e)
    //      row-parallel: CTA b owns row b.  E1 (t3.py:440,475-480), F2.1 (layernorm.py:40-59)
    auto reduce_row = [&](int e) {
        const int r = b, n = tid * 4;
        float4 sn;
        if (e == 0) {
            if (p.x_in != nullptr) {
                const uint2 raw = *reinterpret_cast<const uint2*>(p.x_in + (size_t)r * FZ_D + n);
                sn = make_float4(bf16lo(raw.x), bf16hi(raw.x), bf16lo(raw.y), bf16hi(raw.y));
            } else {
                int tok = __ldg(p.tokens + (r >> 1));
                tok = min(max(tok, 0), p.emb_rows - 1);
                int pp = __ldg(p.speech_pos + (r >> 1)) % p.pos_rows;
                if (pp < 0) pp += p.pos_rows;
                const uint2 a = __ldg(reinterpret_cast<const uint2*>(p.emb + (size_t)tok * FZ_D + n));
                const uint2 bb = __ldg(reinterpret_cast<const uint2*>(p.pos_emb + (size_t)pp * FZ_D + n));
                sn = make_float4(bf16_round(bf16lo(a.x) + bf16lo(bb.x)), bf16_round(bf16hi(a.x) + bf16hi(bb.x)),
                                 bf16_round(bf16lo(a.y) + bf16lo(bb.y)), bf16_round(bf16hi(a.y) + bf16hi(bb.y)));
            }
        } else {
            const int np = (e & 1) ? FZ_NH : FZ_NCL;              // after phase A: one partial per head; after phase B: per cluster
            const float4 so = fz_ldcgf4(p.s + ((size_t)((e + 1) & 1) * rows + r) * FZ_D + n);
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int i0 = 0; i0 < np; i0 += 16) {
                float4 v[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = fz_ldcgf4(p.P + ((size_t)(i0 + i) * rows + r) * FZ_D + n);
#pragma unroll
                for (int i = 0; i < 16; ++i) { acc.x += v[i].x; acc.y += v[i].y; acc.z += v[i].z; acc.w += v[i].w; }   // fixed order
            }
            sn.x = bf16_round(acc.x) + bf16_round(so.x); sn.y = bf16_round(acc.y) + bf16_round(so.y);
            sn.z = bf16_round(acc.z) + bf16_round(so.z); sn.w = bf16_round(acc.w) + bf16_round(so.w);
        }
        *reinterpret_cast<float4*>(p.s + ((size_t)(e & 1) * rows + r) * FZ_D + n) = sn;
        const float ss = warp_sum(sn.x * sn.x + sn.y * sn.y + sn.z * sn.z + sn.w * sn.w);
        if (lane == 0) sm.red[wid] = ss;
        fz_worker_sync();
        float tot = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) tot += sm.red[w];
        const float rs = rsqrtf(tot / (float)FZ_D + p.eps);
        const uint2 wr = __ldg(reinterpret_cast<const uint2*>(p.norms + (size_t)e * FZ_D + n));
        const uint2 o = make_uint2(pack_bf16(bf16_round(sn.x * rs) * bf16lo(wr.x), bf16_round(sn.y * rs) * bf16hi(wr.x)),
                                   pack_bf16(bf16_round(sn.z * rs) * bf16lo(wr.y), bf16_round(sn.w * rs) * bf16hi(wr.y)));
        *reinterpret_cast<uint2*>(p.xn + (size_t)r * FZ_D + n) = o;
        if (e == 2 * p.n_layers && p.hidden_out != nullptr) *reinterpret_cast<uint2*>(p.hidden_out + (size_t)r * FZ_D + n) = o;
    };
    auto exchange = [&](int e) {
        if (e > 0) grid_wait();
        if (b < rows) reduce_row(e);
        grid_arrive();
        grid_wait();
    };
    // activation operand: 16 k-blocks of xn rows {rmul i + radd, i < nrows} -> [N x 64] K-major SWIZZLE_128B stages, built
    // by the workers nx stages at a time (one L2 round trip per batch; rows past nrows are never read back)
    auto build_x = [&](int nrows, int rmul, int radd) {
        const int lnx = 31 - __clz(nx);                              // nx is a power of two dividing 16
        const int total = nrows * 8 * nx;
        for (int kb0 = 0; kb0 < 16; kb0 += nx) {
            const uint32_t use = (xq + (uint32_t)kb0) >> lnx;        // how often every stage has been filled before
            for (int i0 = tid; i0 < total; i0 += 4 * FZ_WORKERS) {
                uint4 v[4];
                uint32_t dst[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int i = i0 + u * FZ_WORKERS;
                    dst[u] = 0xffffffffu;
                    if (i < total) {
                        const int ch = i & 7, xs = (i >> 3) & (nx - 1), r = i >> (3 + lnx);
                        if (use > 0) mbar_wait(&sm.x_empty[xs], (use - 1) & 1);
                        v[u] = fz_ldcg16(p.xn + (size_t)(rmul * r + radd) * FZ_D + (kb0 + xs) * 64 + ch * 8);
                        dst[u] = (uint32_t)xs * xst + (uint32_t)(r * 128 + ((ch ^ (r & 7)) << 4));
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (dst[u] != 0xffffffffu) *reinterpret_cast<uint4*>(xring + dst[u]) = v[u];
            }
            fz_fence_async_smem();
            fz_worker_sync();
            if (tid < nx) mbar_arrive(&sm.x_full[tid]);
        }
        xq += 16;
    };
    const int quad = (wid + 2) & 3, hf = wid >> 2;                  // TMEM lane quadrant of this warp, column half
    uint32_t buf = 0;
    auto drain_begin = [&]() { buf = tseq & 1; mbar_wait(&sm.tmem_full[buf], (tseq >> 1) & 1); tc_fence_after(); };
    auto drain_end = [&]() { tc_fence_before(); __syncwarp(); if (lane == 0) mbar_arrive(&sm.tmem_empty[buf]); ++tseq; };
    // two fp32 accumulator tiles [128 features x ncols] -> P[part][row][256 j + 128 t + feature], row = rmul * col + radd
    auto drain_to_P = [&](int part, int ncols, int rmul, int radd) {
        drain_begin();
        stamp(3);
#pragma unroll 1
        for (int t = 0; t < 2; ++t) {
            float* dst = p.P + (size_t)part * rows * FZ_D + 256 * j + 128 * t + quad * 32 + lane;
            for (int c0 = hf * 16; c0 < ncols; c0 += 32) {
                uint32_t v[16];
                tmem_ld16(tmem_base + buf * 256 + t * 128 + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 16; ++i)
                    if (c0 + i < ncols) dst[(size_t)(rmul * (c0 + i) + radd) * FZ_D] = __uint_as_float(v[i]);
            }
        }
        drain_end();
    };
    // q|k|v features of one local row in the (t4, c8) lane layout: lane holds dims [8 c8, 8 c8 + 8); neox RoPE with the
    // bf16 roundings of rotary_embedding/common.py:144-174
    const int t4 = lane >> 3, c8 = lane & 7;
    auto load8 = [&](int f0, int lr, float* o) {
        const bf16* base = qraw + (size_t)(f0 + c8 * 8) * Rph + lr;
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = __bfloat162float(base[(size_t)e * Rph]);
    };
    auto rope8 = [&](float* x, const float* cs, const float* sn) {
        const bool hi = c8 >= 4;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const float xp = __shfl_xor_sync(0xffffffffu, x[e], 4);
            x[e] = hi ? bf16_round(bf16_round(x[e] * cs[e]) + bf16_round(xp * sn[e])) : bf16_round(bf16_round(x[e] * cs[e]) - bf16_round(xp * sn[e]));
        }
    };
    auto cos_sin8 = [&](int lr, float* cs, float* sn) {
        unpack8(meta.cs[lr * 8 + (c8 & 3)], cs);
        unpack8(meta.cs[lr * 8 + 4 + (c8 & 3)], sn);
    };
    const float kScale = 0.125f * 1.4426950408889634f;              // head_dim^-0.5 * log2(e)

    for (int layer = 0; layer < p.n_layers; ++layer) {
        bf16* kl = p.k_pool + (size_t)layer * p.layer_stride;
        bf16* vl = p.v_pool + (size_t)layer * p.layer_stride;
        // =========================================================== phase A
        exchange(2 * layer);
        build_x(Rh, 2, z);
        stamp(10);
        // ---- qkv_h: lanes 0..47 of the accumulator = features [48 j, 48 j + 48) of q|k|v of head h, columns = local rows
        //      -> bf16 -> feature-major [192][Rph] in all 4 CTAs
        drain_begin();
        stamp(11);
        if (quad < 2) {
            const int fl = quad * 32 + lane;
            for (int c0 = hf * 16; c0 < Rh; c0 += 32) {
                uint32_t v[16];
                tmem_ld16(tmem_base + buf * 256 + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
                tmem_ld_wait();
                if (fl < FZ_QKV_ROWS) {
                    uint4 lo, hi;
                    lo.x = pack_bf16(__uint_as_float(v[0]), __uint_as_float(v[1]));   lo.y = pack_bf16(__uint_as_float(v[2]), __uint_as_float(v[3]));
                    lo.z = pack_bf16(__uint_as_float(v[4]), __uint_as_float(v[5]));   lo.w = pack_bf16(__uint_as_float(v[6]), __uint_as_float(v[7]));
                    hi.x = pack_bf16(__uint_as_float(v[8]), __uint_as_float(v[9]));   hi.y = pack_bf16(__uint_as_float(v[10]), __uint_as_float(v[11]));
                    hi.z = pack_bf16(__uint_as_float(v[12]), __uint_as_float(v[13])); hi.w = pack_bf16(__uint_as_float(v[14]), __uint_as_float(v[15]));
                    const uint32_t off = dsm_a + p.off_qkvraw + (uint32_t)(((FZ_QKV_ROWS * j + fl) * Rph + c0) * 2);
#pragma unroll
                    for (int peer = 0; peer < FZ_CL; ++peer) {
                        const uint32_t ra = fz_mapa(off, (uint32_t)peer);
                        fz_st_cluster16(ra, lo);
                        fz_st_cluster16(ra + 16, hi);
                    }
                }
            }
        }
        drain_end();
        q += 8;
        stamp(12);
        cluster_sync();                                              // #1: q|k|v of the cluster's rows are in every CTA
        stamp(13);
        // ---- attention over the cached tokens: group g = this CTA's 64-token chunks of one row, one warp per group
        for (int g = wid; g < sm.n_groups; g += 8) {
            const int lr = sm.grp_row[g];
            const int L = meta.L[lr];
            const int nch = (L + 63) >> 6, npages = (L + 15) >> 4;
            float qf[8], cs[8], sn[8];
            load8(0, lr, qf);
            cos_sin8(lr, cs, sn);
            rope8(qf, cs, sn);
#pragma unroll
            for (int e = 0; e < 8; ++e) qf[e] *= kScale;
            float m = -INFINITY, l = 0.f, acc[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] = 0.f;
            const uint32_t qq = q + sm.grp_slot0[g];
            uint32_t s = qq % nslot, sph = (qq / nslot) & 1;
            for (int ch = (j - lr) & 3; ch < nch; ch += 4) {
                mbar_wait(&sm.full[s], sph);
                const unsigned char* sl = ring + (size_t)s * FZ_SLOT;
#pragma unroll 1
                for (int half = 0; half < 2; ++half) {
                    const int pg0 = half * 2, page0 = ch * 4 + pg0;
                    if (page0 >= npages) break;
                    uint4 kv[2][4], vv[2][4];
#pragma unroll
                    for (int pp = 0; pp < 2; ++pp) {
                        const bool valid = page0 + pp < npages;
                        const unsigned char* base = sl + (pg0 + pp) * 2048 + lane * 16;
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            kv[pp][i] = valid ? *reinterpret_cast<const uint4*>(base + i * 512) : make_uint4(0, 0, 0, 0);
                            vv[pp][i] = valid ? *reinterpret_cast<const uint4*>(base + 8192 + i * 512) : make_uint4(0, 0, 0, 0);
                        }
                    }
                    float sc8[8];
#pragma unroll
                    for (int pp = 0; pp < 2; ++pp)
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            float k8[8];
                            unpack8(kv[pp][i], k8);
                            float d = 0.f;
#pragma unroll
                            for (int e = 0; e < 8; ++e) d = fmaf(qf[e], k8[e], d);
                            sc8[pp * 4 + i] = d;
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
                        for (int i = 0; i < 4; ++i) {
                            const float pj = __shfl_sync(0xffffffffu, pr, t4 * 8 + pp * 4 + i);
                            float v8[8];
                            unpack8(vv[pp][i], v8);
#pragma unroll
                            for (int e = 0; e < 8; ++e) acc[e] = fmaf(pj, v8[e], acc[e]);
                        }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&sm.kv_empty[s]);
                if (++s == (uint32_t)nslot) { s = 0; sph ^= 1; }
            }
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 8);
                acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 16);
            }
            l = warp_sum(l);
            // partial (m, l, acc[64]) -> record [lr / 4][src = j] in the smem of the row's owner CTA (lr % 4)
            const uint32_t rec = fz_mapa(dsm_a + p.off_attpart + (uint32_t)((((lr >> 2) * FZ_CL + j) * FZ_REC) * 4), (uint32_t)(lr & 3));
            if (lane == 0) fz_st_cluster8(rec, m, l);
            if (lane < 8) {
                fz_st_cluster16(rec + 16 + lane * 32, make_uint4(__float_as_uint(acc[0]), __float_as_uint(acc[1]), __float_as_uint(acc[2]), __float_as_uint(acc[3])));
                fz_st_cluster16(rec + 32 + lane * 32, make_uint4(__float_as_uint(acc[4]), __float_as_uint(acc[5]), __float_as_uint(acc[6]), __float_as_uint(acc[7])));
            }
        }
        q += (uint32_t)sm.n_kv_slots;
        stamp(14);
        cluster_sync();                                              // #2: every partial sits with its row's owner
        stamp(15);
        // ---- owner (local rows lr = j mod 4): merge the <= 4 partials in chunk order, add this step's own token, write
        //      K/V, broadcast attn_h[lr] (bf16) into every CTA's o-projection operand
        for (int lr = j + 4 * wid; lr < Rh; lr += 32) {
            const int L = meta.L[lr];
            const int nch = (L + 63) >> 6;
            float qf[8], kf[8], vf[8], cs[8], sn[8];
            load8(0, lr, qf); load8(64, lr, kf); load8(128, lr, vf);
            cos_sin8(lr, cs, sn);
            rope8(qf, cs, sn);
            rope8(kf, cs, sn);
#pragma unroll
            for (int e = 0; e < 8; ++e) qf[e] *= kScale;
            float m = -INFINITY, l = 0.f, acc[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] = 0.f;
            for (int ch0 = 0; ch0 < FZ_CL && ch0 < nch; ++ch0) {   // fixed (chunk) order -> deterministic, slot invariant
                const float* rec = attpart + (size_t)(((lr >> 2) * FZ_CL + ((ch0 + lr) & 3)) * FZ_REC);
                const float pm = rec[0], pl = rec[1];
                const float4 a0 = *reinterpret_cast<const float4*>(rec + 4 + c8 * 8), a1 = *reinterpret_cast<const float4*>(rec + 8 + c8 * 8);
                const float m_new = fmaxf(m, pm), c1 = exp2f(m - m_new), c2 = exp2f(pm - m_new);
                l = l * c1 + pl * c2;
                acc[0] = acc[0] * c1 + a0.x * c2; acc[1] = acc[1] * c1 + a0.y * c2; acc[2] = acc[2] * c1 + a0.z * c2; acc[3] = acc[3] * c1 + a0.w * c2;
                acc[4] = acc[4] * c1 + a1.x * c2; acc[5] = acc[5] * c1 + a1.y * c2; acc[6] = acc[6] * c1 + a1.z * c2; acc[7] = acc[7] * c1 + a1.w * c2;
                m = m_new;
            }
            {
                float d = 0.f;
#pragma unroll
                for (int e = 0; e < 8; ++e) d = fmaf(qf[e], kf[e], d);
                d += __shfl_xor_sync(0xffffffffu, d, 1);
                d += __shfl_xor_sync(0xffffffffu, d, 2);
                d += __shfl_xor_sync(0xffffffffu, d, 4);
                const float m_new = fmaxf(m, d), corr = exp2f(m - m_new), pr = exp2f(d - m_new);
                l = l * corr + pr;
#pragma unroll
                for (int e = 0; e < 8; ++e) acc[e] = acc[e] * corr + pr * vf[e];
            }
            float o8[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) o8[e] = acc[e] / l;
            const uint4 ov = pack8(o8);
            const uint32_t off = dsm_a + p.off_attnx + (uint32_t)((lr >> 3) * 1024 + (lr & 7) * 128 + ((c8 ^ (lr & 7)) << 4));
            fz_st_cluster16(fz_mapa(off, (uint32_t)t4), ov);       // lane group t4 -> peer t4
            if (t4 == 0) {                                          // reshape_and_cache: this token is number L of row r
                const int blk = meta.blk_new[lr];
                const size_t koff = (((size_t)blk * FZ_NH + h) * T3B_KV_BLOCK + (L & 15)) * 64 + c8 * 8;
                *reinterpret_cast<uint4*>(kl + koff) = pack8(kf);
                *reinterpret_cast<uint4*>(vl + koff) = pack8(vf);
            }
        }
        stamp(16);
        fz_fence_async_all();
        cluster_sync();                                              // #3: attn_h of the cluster's rows is in every CTA
        if (tid == 0) { fz_fence_async_all(); mbar_arrive(&sm.attnx_full); }
        stamp(17);
        // ---- o_h partial -> P[h][r = 2 lr + z]
        drain_to_P(h, Rh, 2, z);
        q += 2;
        grid_arrive();
        // =========================================================== phase B
        exchange(2 * layer + 1);
        build_x(rows, 1, 0);
        stamp(20);
        // ---- gate_up: lanes 0..31 gate, 32..63 up of features [128 cl + 32 j, +32) -> fp32 scratch (the X ring is idle now)
        drain_begin();
        stamp(21);
        float* gu = reinterpret_cast<float*>(xring);                 // [64][Rp]
        if (quad < 2) {
            for (int c0 = hf * 16; c0 < rows; c0 += 32) {
                uint32_t v[16];
                tmem_ld16(tmem_base + buf * 256 + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
                tmem_ld_wait();
                float* dst = gu + (size_t)(quad * 32 + lane) * Rp + c0;
#pragma unroll
                for (int i = 0; i < 16; i += 4) *reinterpret_cast<float4*>(dst + i) = make_float4(__uint_as_float(v[i]), __uint_as_float(v[i + 1]), __uint_as_float(v[i + 2]), __uint_as_float(v[i + 3]));
            }
        }
        drain_end();
        q += 8;
        fz_worker_sync();
        // act = bf16(bf16(silu(bf16 g)) * bf16 u)  (llama.py:117-121, activation.py:138-141) -> every CTA's down operand
        for (int i = tid; i < rows * 4; i += FZ_WORKERS) {
            const int r = i >> 2, fb = i & 3;
            float a8[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const float gg = bf16_round(gu[(size_t)(fb * 8 + e) * Rp + r]), uu = bf16_round(gu[(size_t)(32 + fb * 8 + e) * Rp + r]);
                a8[e] = bf16_round(gg / (1.0f + expf(-gg))) * uu;
            }
            const uint4 av = pack8(a8);
            const uint32_t off = dsm_a + p.off_actx + (uint32_t)((j >> 1) * (Rp * 128) + (r >> 3) * 1024 + (r & 7) * 128 + ((((j & 1) * 4 + fb) ^ (r & 7)) << 4));
#pragma unroll
            for (int peer = 0; peer < FZ_CL; ++peer) fz_st_cluster16(fz_mapa(off, (uint32_t)peer), av);
        }
        stamp(22);
        fz_fence_async_all();
        cluster_sync();                                              // #4: the cluster's 128 act features are in every CTA
        if (tid == 0) { fz_fence_async_all(); mbar_arrive(&sm.actx_full); }
        stamp(23);
        // ---- down partial -> P[cl]
        drain_to_P(cl, rows, 1, 0);
        q += 4;
        grid_arrive();
    }
    // =============================================================== final norm + speech head
    exchange(2 * p.n_layers);
    if (b < p.n_head_ctas) {
        build_x(rows, 1, 0);
        drain_begin();
        const int fl = quad * 32 + lane, n = b * FZ_HEAD_ROWS + fl;
        if (quad < 3) {
            for (int c0 = hf * 16; c0 < rows; c0 += 32) {
                uint32_t v[16];
                tmem_ld16(tmem_base + buf * 256 + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
                tmem_ld_wait();
                if (fl < FZ_HEAD_ROWS && n < p.V) {
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        if (c0 + i < rows) p.logits[(size_t)(c0 + i) * p.V + n] = __float2bfloat16_rn(__uint_as_float(v[i]));
                }
            }
        }
        drain_end();
    }
    stamp(30);
}

__global__ void __cluster_dims__(FZ_CL, 1, 1) __launch_bounds__(FZ_THREADS, 1)
decode_fused_kernel(const FusedParams p) {
    extern __shared__ unsigned char smem_dyn[];
    // 1024-byte alignment as an OFFSET into the __shared__ array: pointer arithmetic keeps the shared address space, so every
    // access below compiles to LDS / STS (a round trip through uintptr_t made them generic LD.E / ST.E)
    unsigned char* ring = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
    unsigned char* xring = ring + (size_t)p.nslot * FZ_SLOT;
    unsigned char* dsm = xring + FZ_XBYTES;
    FusedSmem& sm = *reinterpret_cast<FusedSmem*>(dsm + p.dsm_bytes);
    FusedMeta meta;
    meta.cs = reinterpret_cast<uint4*>(dsm + p.dsm_bytes + ((sizeof(FusedSmem) + 15) & ~size_t(15)));
    meta.pages = reinterpret_cast<int32_t*>(meta.cs + (size_t)p.Rh * 8);
    meta.L = meta.pages + (size_t)p.meta_pages_max * 4;
    meta.blk_new = meta.L + p.Rh;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int j = (int)fz_cluster_ctarank(), cl = (int)fz_clusterid(), b = cl * FZ_CL + j;
    const int h = cl >> 1, z = cl & 1;
    const int nslot = p.nslot, nx = p.nx, Rp = p.Rp, Rh = p.Rh;
    const uint32_t xst = (uint32_t)Rp * 128u;

    if (threadIdx.x == 0) {
        for (int i = 0; i < nslot; ++i) {
            mbar_init(&sm.full[i], 1);
            mbar_init(&sm.w_empty[i], 1);
            mbar_init(&sm.kv_empty[i], 1);
            sm.w_cnt[i] = 0; sm.kv_cnt[i] = 0; sm.ptype[i] = 0;
        }
        for (int i = 0; i < 16; ++i) { mbar_init(&sm.x_full[i], 1); mbar_init(&sm.x_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&sm.tmem_full[i], 1); mbar_init(&sm.tmem_empty[i], 8); mbar_init(&sm.cl_bar[i], FZ_CL); }
        mbar_init(&sm.attnx_full, 1);
        mbar_init(&sm.actx_full, 1);
        mbar_fence_init();
    }
    if (threadIdx.x == 64) {
        // this CTA's attention groups: local row lr (global row 2 lr + z) contributes its 64-token chunks
        // ch = (j - lr) mod 4, +4, ...
        int n = 0, slots = 0;
        for (int lr = 0; lr < Rh; ++lr) {
            const int L = __ldg(p.ctx_lens + 2 * lr + z);
            const int nch = (L + 63) >> 6, ch0 = (j - lr) & 3;
            if (ch0 < nch) { sm.grp_row[n] = (uint8_t)lr; sm.grp_slot0[n] = (uint16_t)slots; slots += (nch - ch0 + 3) >> 2; ++n; }
        }
        sm.n_groups = n; sm.n_kv_slots = slots;
    }
    if (warp == 1) tmem_alloc(&sm.tmem_ptr, 512);
    // row metadata of this cluster's rows: context length, the block this step's token goes to, cos|sin of its position
    for (int i = threadIdx.x; i < Rh * 8; i += FZ_THREADS) {
        const int lr = i >> 3, r = 2 * lr + z;
        int pos = __ldg(p.rope_pos + r);
        pos = min(max(pos, 0), p.max_pos - 1);
        meta.cs[i] = __ldg(reinterpret_cast<const uint4*>(p.cos_sin + (size_t)pos * 64) + (i & 7));
        if ((i & 7) == 0) {
            const int L = __ldg(p.ctx_lens + r);
            meta.L[lr] = L;
            meta.blk_new[lr] = __ldg(p.block_tables + (size_t)r * p.max_blocks + (L >> 4));
        }
    }
    tc_fence_before();
    __syncthreads();
    // physical blocks of the 4 pages of every KV slot this CTA streams per layer (ring order)
    for (int i = threadIdx.x; i < sm.n_groups * 32; i += FZ_THREADS) {       // <= 8 slots (2048 tokens) per group
        const int g = i >> 5, sl = (i >> 2) & 7, pq = i & 3;
        const int lr = sm.grp_row[g], r = 2 * lr + z;
        const int npages = (meta.L[lr] + 15) >> 4, ch0 = (j - lr) & 3;
        const int nsl = (((npages + 3) >> 2) - ch0 + 3) >> 2;
        if (sl < nsl) {
            const int pg = (ch0 + 4 * sl) * 4 + pq;
            meta.pages[(sm.grp_slot0[g] + sl) * 4 + pq] = pg < npages ? __ldg(p.block_tables + (size_t)r * p.max_blocks + pg) : 0;
        }
    }
    __syncthreads();
    tc_fence_after();
    fz_cluster_barrier_all();          // every CTA of the cluster is running and has initialised its barriers
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
        auto emit_w = [&](const unsigned char* src, uint32_t bytes) {
            if (lane == 0) {
                acquire(0);
                mbar_expect_tx(&sm.full[s], bytes);
                bulk_load_1d(ring + (size_t)s * FZ_SLOT, src, bytes, &sm.full[s]);
            }
            ++q; if (++s == nslot) s = 0;
        };
        for (int layer = 0; layer < p.n_layers; ++layer) {
            const unsigned char* wa = p.wl + (size_t)layer * FZ_LAYER_BYTES + (size_t)(h * FZ_CL + j) * FZ_A_BYTES;
            const unsigned char* wb = p.wl + (size_t)layer * FZ_LAYER_BYTES + FZ_LAYER_A_BYTES + (size_t)b * FZ_B_BYTES;
            for (int i = 0; i < 8; ++i) emit_w(wa + i * 2 * FZ_QKV_KB_BYTES, 2 * FZ_QKV_KB_BYTES);
            // ---- past KV of this CTA's chunks: K pages at slot + 2 KB * page, V pages 8 KB further
            const bf16* kl = p.k_pool + (size_t)layer * p.layer_stride;
            const bf16* vl = p.v_pool + (size_t)layer * p.layer_stride;
            int kvs = 0;
            for (int g = 0; g < sm.n_groups; ++g) {
                const int lr = sm.grp_row[g];
                const int L = meta.L[lr];
                const int nch = (L + 63) >> 6, npages = (L + 15) >> 4;
                for (int ch = (j - lr) & 3; ch < nch; ch += 4, ++kvs) {
                    const int np = min(4, npages - ch * 4);
                    if (lane == 0) { acquire(1); mbar_expect_tx(&sm.full[s], (uint32_t)np * 4096u); }
                    __syncwarp();
                    const int pg = lane & 3, isv = (lane >> 2) & 1;
                    if (lane < 8 && pg < np) {
                        const int blk = meta.pages[kvs * 4 + pg];
                        const bf16* src = (isv ? vl : kl) + ((size_t)blk * FZ_NH + h) * (T3B_KV_BLOCK * 64);
                        bulk_load_1d(ring + (size_t)s * FZ_SLOT + isv * 8192 + pg * 2048, src, 2048, &sm.full[s]);
                    }
                    ++q; if (++s == nslot) s = 0;
                }
            }
            for (int i = 0; i < 2; ++i) emit_w(wa + FZ_QKV_BYTES + i * FZ_SLOT, FZ_SLOT);
            for (int i = 0; i < 8; ++i) emit_w(wb + i * FZ_SLOT, FZ_SLOT);
            for (int i = 0; i < 4; ++i) emit_w(wb + FZ_GU_BYTES + i * FZ_SLOT, FZ_SLOT);
        }
        if (b < p.n_head_ctas)
            for (int i = 0; i < 16; ++i) emit_w(p.wh + (size_t)b * FZ_HEAD_CTA_BYTES + i * FZ_HEAD_KB_BYTES, FZ_HEAD_KB_BYTES);
    } else if (warp == 1) {
        // ================================================================= MMA issuer (warp-uniform; one elected lane issues)
        {
            const uint32_t idesc_a = umma_idesc_bf16(128, (uint32_t)p.Rph), idesc_b = umma_idesc_bf16(128, (uint32_t)Rp);
            // ring / X-stage / accumulator cursors kept incrementally (index + phase bit): no divisions on the issue path
            uint32_t s = 0, sph = 0, xs = 0, xph = 0, tseq = 0, buf = 0;
            const uint32_t ring_a = smem_u32(ring), x_a = smem_u32(xring);
            const uint64_t desc0 = umma_smem_desc_sw128(0);
            auto desc = [&](uint32_t addr) { return desc0 | (uint64_t)((addr & 0x3FFFFu) >> 4); };
            auto slot_next = [&]() { if (++s == (uint32_t)nslot) { s = 0; sph ^= 1; } };
            auto x_next = [&]() { if (++xs == (uint32_t)nx) { xs = 0; xph ^= 1; } };
            auto gemm_begin = [&]() {
                buf = tseq & 1;
                if (tseq >= 2) mbar_wait(&sm.tmem_empty[buf], ((tseq >> 1) - 1) & 1);
            };
            auto commit = [&](uint64_t* bar) { if (fz_elect_one()) umma_commit(bar); __syncwarp(); };
            auto gemm_end = [&]() { commit(&sm.tmem_full[buf]); ++tseq; };
            auto mma4 = [&](uint32_t d_col, uint32_t a_addr, uint32_t b_addr, uint32_t idesc, bool first) {
                const uint64_t a_desc = desc(a_addr), b_desc = desc(b_addr);
                const uint32_t d = tmem_base + d_col;
                if (fz_elect_one()) {
                    umma_bf16(d, a_desc, b_desc, idesc, first ? 0u : 1u);
                    umma_bf16(d, a_desc + 2, b_desc + 2, idesc, 1u);
                    umma_bf16(d, a_desc + 4, b_desc + 4, idesc, 1u);
                    umma_bf16(d, a_desc + 6, b_desc + 6, idesc, 1u);
                }
                __syncwarp();
            };
            int mtr = 0;
            auto mstamp = [&](int id) {
                if (p.trace != nullptr && lane == 0 && mtr + 2 <= FZ_TRACE) {
                    unsigned long long* t = p.trace + (size_t)(FZ_NCTA + b) * FZ_TRACE;
                    t[mtr] = (unsigned long long)id; t[mtr + 1] = fz_gtimer(); mtr += 2;
                }
            };
            // full-K projection: n_slots ring slots holding kb_per_slot k-blocks of kb_bytes each, B operand from the X ring
            auto gemm_fullk = [&](int n_slots, int kb_per_slot, uint32_t kb_bytes, uint32_t idesc) {
                gemm_begin();
                mstamp(40);
                for (int i = 0; i < n_slots; ++i) {
                    mbar_wait(&sm.full[s], sph);
                    if (i == 0) mstamp(41);
                    if (i == n_slots - 1) mstamp(44);
                    for (int kk = 0; kk < kb_per_slot; ++kk) {
                        mbar_wait(&sm.x_full[xs], xph);
                        if (i == 0 && kk == 0) mstamp(42);
                        tc_fence_after();
                        mma4(buf * 256, ring_a + s * FZ_SLOT + kk * kb_bytes, x_a + xs * xst, idesc, i == 0 && kk == 0);
                        commit(&sm.x_empty[xs]);
                        x_next();
                    }
                    commit(&sm.w_empty[s]);
                    slot_next();
                }
                mstamp(43);
                gemm_end();
            };
            // 2 output tiles x kbs k-blocks (one 16 KB ring slot each, tile-major); B tiles [N x 64] sit in the DSMEM region
            auto gemm_tiles = [&](int kbs, uint64_t* ready, uint32_t parity, const unsigned char* bsrc, uint32_t idesc) {
                gemm_begin();
                mbar_wait(ready, parity);
                const uint32_t b_a = smem_u32(bsrc);
                for (int t = 0; t < 2; ++t)
                    for (int kb = 0; kb < kbs; ++kb) {
                        mbar_wait(&sm.full[s], sph);
                        tc_fence_after();
                        mma4(buf * 256 + t * 128, ring_a + s * FZ_SLOT, b_a + kb * xst, idesc, kb == 0);
                        commit(&sm.w_empty[s]);
                        slot_next();
                    }
                gemm_end();
            };
            for (int layer = 0; layer < p.n_layers; ++layer) {
                gemm_fullk(8, 2, FZ_QKV_KB_BYTES, idesc_a);                                     // qkv_h
                for (int n = __shfl_sync(0xffffffffu, sm.n_kv_slots, 0); n > 0; --n) slot_next();   // the KV slots belong to the workers
                gemm_tiles(1, &sm.attnx_full, layer & 1, dsm + p.off_attnx, idesc_a);           // o_h
                gemm_fullk(8, 2, 8192, idesc_b);                                                // gate_up
                gemm_tiles(2, &sm.actx_full, layer & 1, dsm + p.off_actx, idesc_b);             // down
            }
            if (b < p.n_head_ctas) gemm_fullk(16, 1, FZ_HEAD_KB_BYTES, idesc_b);                // speech head
        }
    } else {
        fused_workers(p, sm, meta, ring, xring, dsm, tmem_base);
    }
    tc_fence_before();
    __syncthreads();
    fz_cluster_barrier_all();          // nobody exits while a peer may still address its shared memory
    if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// ------------------------------------------------------------------ one-off weight layout transform
// Tiles are K-major and 128B-swizzled (16-byte chunk ch of tile row rr is stored at chunk ch ^ (rr & 7)).
// Section A, per (head h, rank j), 128 KB (read by both clusters of the head):
//   qkv      16 k-blocks x 48 rows: features 48 j + rr of head h's q(64) | k(64) | v(64)
//   o        2 tiles x 128 rows: output features 256 j + 128 t + rr, K = head h's 64 dims
// Section B, per (cluster c, rank j), 192 KB:
//   gate_up  16 k-blocks x 64 rows: gate rows 128 c + 32 j + rr (rr < 32), then the matching up rows
//   down     2 tiles x 2 k-blocks x 128 rows: output features 256 j + 128 t + rr, K = act features [128 c + 64 kb, +64)
__global__ void fused_repack_layer_kernel(const uint4* __restrict__ wqkv, const uint4* __restrict__ wo, const uint4* __restrict__ wgu,
                                          const uint4* __restrict__ wd, uint4* __restrict__ out) {
    const size_t n_a = FZ_LAYER_A_BYTES / 16, n_chunks = FZ_LAYER_BYTES / 16;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_chunks; i += (size_t)gridDim.x * blockDim.x) {
        uint4 v;
        if (i < n_a) {
            const int hj = (int)(i / (FZ_A_BYTES / 16)), h = hj >> 2, j = hj & 3;
            int o = (int)(i % (FZ_A_BYTES / 16));
            const int pc = o & 7;
            if (o < FZ_QKV_BYTES / 16) {
                const int kb = o / (FZ_QKV_KB_BYTES / 16), rr = (o % (FZ_QKV_KB_BYTES / 16)) >> 3, ch = pc ^ (rr & 7);
                const int f = FZ_QKV_ROWS * j + rr, row = (f >> 6) * FZ_D + h * 64 + (f & 63);
                v = wqkv[((size_t)row * FZ_D + kb * 64 + ch * 8) / 8];
            } else {
                o -= FZ_QKV_BYTES / 16;
                const int t = o / 1024, rr = (o % 1024) >> 3, ch = pc ^ (rr & 7);
                v = wo[((size_t)(256 * j + 128 * t + rr) * FZ_D + h * 64 + ch * 8) / 8];
            }
        } else {
            const size_t i2 = i - n_a;
            const int cj = (int)(i2 / (FZ_B_BYTES / 16)), c = cj >> 2, j = cj & 3;
            int o = (int)(i2 % (FZ_B_BYTES / 16));
            const int pc = o & 7;
            if (o < FZ_GU_BYTES / 16) {
                const int kb = o / 512, rr = (o % 512) >> 3, ch = pc ^ (rr & 7);
                const int f = 128 * c + 32 * j + (rr & 31), row = rr < 32 ? f : FZ_F + f;
                v = wgu[((size_t)row * FZ_D + kb * 64 + ch * 8) / 8];
            } else {
                o -= FZ_GU_BYTES / 16;
                const int idx = o / 1024, t = idx >> 1, kb = idx & 1, rr = (o % 1024) >> 3, ch = pc ^ (rr & 7);
                v = wd[((size_t)(256 * j + 128 * t + rr) * FZ_F + 128 * c + kb * 64 + ch * 8) / 8];
            }
        }
        out[i] = v;
    }
}
// speech head: CTA b holds vocabulary rows [72 b, 72 b + 72) as 16 k-blocks x 72 rows (rows >= V are zero)
__global__ void fused_repack_head_kernel(const uint4* __restrict__ wh, int V, uint4* __restrict__ out) {
    const size_t n_chunks = (size_t)FZ_NCTA * FZ_HEAD_CTA_BYTES / 16;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_chunks; i += (size_t)gridDim.x * blockDim.x) {
        const int cta = (int)(i / (FZ_HEAD_CTA_BYTES / 16)), o = (int)(i % (FZ_HEAD_CTA_BYTES / 16));
        const int kb = o / (FZ_HEAD_KB_BYTES / 16), rr = (o % (FZ_HEAD_KB_BYTES / 16)) >> 3, pc = o & 7, ch = pc ^ (rr & 7);
        const int n = cta * FZ_HEAD_ROWS + rr;
        out[i] = n < V ? wh[((size_t)n * FZ_D + kb * 64 + ch * 8) / 8] : make_uint4(0, 0, 0, 0);
    }
}

// ------------------------------------------------------------------ host
namespace {
size_t fz_al(size_t x) { return (x + 255) & ~size_t(255); }
bool fz_shape_ok(const t3b_model_t* m) {
    return m->hidden == FZ_D && m->n_heads == FZ_NH && m->head_dim == 64 && m->ffn == FZ_F && m->speech_vocab > 0 &&
           (m->speech_vocab + FZ_HEAD_ROWS - 1) / FZ_HEAD_ROWS <= FZ_NCTA;
}
struct FusedBlob { size_t layers, head, norms, total; };
FusedBlob fz_blob(const t3b_model_t* m) {
    FusedBlob b;
    b.layers = 0;
    b.head = (size_t)m->n_layers * FZ_LAYER_BYTES;
    b.norms = b.head + (size_t)FZ_NCTA * FZ_HEAD_CTA_BYTES;
    b.total = b.norms + (size_t)(2 * m->n_layers + 1) * FZ_D * 2;
    return b;
}
struct FusedWs { char *s, *xn, *P, *logits, *counters; size_t total; };
FusedWs fz_carve(const t3b_model_t* m, int rows, void* base) {
    FusedWs w;
    char* p = static_cast<char*>(base);
    auto take = [&](size_t bytes) { char* r = p; p += fz_al(bytes); return r; };
    w.s = take((size_t)2 * rows * FZ_D * 4);
    w.xn = take((size_t)rows * FZ_D * 2);
    w.P = take((size_t)FZ_NCL * rows * FZ_D * 4);
    w.logits = take((size_t)rows * m->speech_vocab * 2 + 64);
    w.counters = take(256);
    w.total = static_cast<size_t>(p - static_cast<char*>(base));
    return w;
}
unsigned long long* g_fused_trace = nullptr;
int g_fused_dbg = 0;
}  // namespace
extern "C" void t3b_debug_set_fused_flags(int32_t f) { g_fused_dbg = f; }

extern "C" void t3b_debug_set_fused_trace(void* p) { g_fused_trace = static_cast<unsigned long long*>(p); }
extern "C" int32_t t3b_fused_trace_stride(void) { return FZ_TRACE; }
extern "C" int32_t t3b_fused_max_rows(void) { return FZ_MAX_ROWS; }

extern "C" size_t t3b_fused_weight_bytes(const t3b_model_t* m) {
    if (!m || !fz_shape_ok(m)) return 0;
    return fz_blob(m).total;
}

extern "C" size_t t3b_fused_workspace_bytes(const t3b_model_t* m, int32_t max_rows) {
    if (!m || max_rows <= 0) return 0;
    return fz_carve(m, max_rows, nullptr).total + 256;
}

extern "C" int32_t t3b_fused_repack(const t3b_model_t* m, void* out, void* stream) {
    T3B_CHECK_ARG(m && out && m->layers && m->final_norm && m->head.w, "t3b_fused_repack: null pointer");
    T3B_CHECK_SHAPE(fz_shape_ok(m), "t3b_fused_repack: the fused step kernel is built for hidden 1024 / 16 heads x 64 / ffn 4096 (got %d / %d x %d / %d)",
                    m->hidden, m->n_heads, m->head_dim, m->ffn);
    T3B_CHECK_ALIGN(out);
    cudaStream_t s = t3b_stream(stream);
    const FusedBlob bl = fz_blob(m);
    char* o = static_cast<char*>(out);
    const int blocks = t3b_num_sms() * 8;
    for (int l = 0; l < m->n_layers; ++l) {
        const t3b_layer_t& L = m->layers[l];
        T3B_CHECK_ARG(L.qkv.kind == T3B_W_BF16 && L.o.kind == T3B_W_BF16 && L.gate_up.kind == T3B_W_BF16 && L.down.kind == T3B_W_BF16,
                      "t3b_fused_repack: layer %d is not bf16 (the fused step kernel streams bf16 weights)", l);
        T3B_CHECK_ARG(L.qkv.w && L.o.w && L.gate_up.w && L.down.w && L.ln1 && L.ln2, "t3b_fused_repack: layer %d has null weights", l);
        fused_repack_layer_kernel<<<blocks, 256, 0, s>>>((const uint4*)L.qkv.w, (const uint4*)L.o.w, (const uint4*)L.gate_up.w, (const uint4*)L.down.w,
                                                         (uint4*)(o + bl.layers + (size_t)l * FZ_LAYER_BYTES));
        T3B_LAUNCH_CHECK();
        T3B_CUDA(cudaMemcpyAsync(o + bl.norms + (size_t)(2 * l) * FZ_D * 2, L.ln1, FZ_D * 2, cudaMemcpyDeviceToDevice, s));
        T3B_CUDA(cudaMemcpyAsync(o + bl.norms + (size_t)(2 * l + 1) * FZ_D * 2, L.ln2, FZ_D * 2, cudaMemcpyDeviceToDevice, s));
    }
    T3B_CUDA(cudaMemcpyAsync(o + bl.norms + (size_t)(2 * m->n_layers) * FZ_D * 2, m->final_norm, FZ_D * 2, cudaMemcpyDeviceToDevice, s));
    fused_repack_head_kernel<<<blocks, 256, 0, s>>>((const uint4*)m->head.w, m->speech_vocab, (uint4*)(o + bl.head));
    T3B_LAUNCH_CHECK();
    return T3B_OK;
}

// Whole decode step in one launch: embeddings (or x_in) -> 30 layers -> final norm (hidden_out, optional)
// -> speech head logits bf16 [rows, vocab] (logits_out).  rows even, <= t3b_fused_max_rows().
extern "C" int32_t t3b_decode_step_fused(const t3b_model_t* m, const void* fused_weights, const t3b_kv_t* kv, const t3b_state_t* st,
                                         const void* x_in, int32_t rows, void* hidden_out, void* logits_out, void* workspace,
                                         size_t workspace_bytes, void* stream) {
    T3B_CHECK_ARG(m && fused_weights && kv && kv->k && kv->v && st && workspace, "t3b_decode_step_fused: null pointer");
    T3B_CHECK_ARG(m->speech_emb && m->speech_pos_emb && m->rope_cos_sin, "t3b_decode_step_fused: model tables are null");
    T3B_CHECK_ARG(st->rope_pos && st->ctx_lens && st->block_tables && st->max_blocks > 0, "t3b_decode_step_fused: state arrays are null");
    T3B_CHECK_ARG(x_in || (st->tokens && st->speech_pos), "t3b_decode_step_fused: no input (x_in and state tokens are null)");
    T3B_CHECK_ARG(rows >= 0 && rows % 2 == 0 && rows <= FZ_MAX_ROWS, "t3b_decode_step_fused: rows %d must be even and <= %d", rows, FZ_MAX_ROWS);
    T3B_CHECK_SHAPE(fz_shape_ok(m), "t3b_decode_step_fused: unsupported model shape");
    T3B_CHECK_ALIGN(workspace); T3B_CHECK_ALIGN(fused_weights);
    if (hidden_out) T3B_CHECK_ALIGN(hidden_out);
    if (rows == 0) return T3B_OK;
    T3B_CHECK_ARG(workspace_bytes >= t3b_fused_workspace_bytes(m, rows), "t3b_decode_step_fused: workspace %zu < %zu bytes", workspace_bytes,
                  t3b_fused_workspace_bytes(m, rows));
    T3B_CHECK_ARG(t3b_num_sms() >= FZ_NCTA, "t3b_decode_step_fused: needs %d SMs", FZ_NCTA);
    const FusedWs ws = fz_carve(m, rows, workspace);
    const FusedBlob bl = fz_blob(m);
    cudaStream_t s = t3b_stream(stream);
    const char* fw = static_cast<const char*>(fused_weights);

    FusedParams p = {};
    p.wl = (const unsigned char*)(fw + bl.layers); p.wh = (const unsigned char*)(fw + bl.head); p.norms = (const bf16*)(fw + bl.norms);
    p.emb = (const bf16*)m->speech_emb; p.pos_emb = (const bf16*)m->speech_pos_emb; p.cos_sin = (const bf16*)m->rope_cos_sin; p.x_in = (const bf16*)x_in;
    p.tokens = st->tokens; p.speech_pos = st->speech_pos; p.rope_pos = st->rope_pos; p.ctx_lens = st->ctx_lens; p.block_tables = st->block_tables;
    p.k_pool = (bf16*)kv->k; p.v_pool = (bf16*)kv->v; p.layer_stride = kv->layer_stride;
    p.n_layers = m->n_layers; p.V = m->speech_vocab; p.emb_rows = m->speech_vocab; p.pos_rows = m->speech_pos_rows;
    p.max_pos = m->max_pos; p.max_blocks = st->max_blocks; p.eps = m->rms_eps;
    p.rows = rows; p.Rp = (rows + 15) & ~15;
    p.Rh = rows / 2; p.Rph = (p.Rh + 15) & ~15;
    p.nx = 16;                                           // X stages: a power of two (dividing the 16 k-blocks) that fits 32 KB
    while (p.nx * p.Rp * 128 > FZ_XBYTES) p.nx >>= 1;
    p.n_head_ctas = (m->speech_vocab + FZ_HEAD_ROWS - 1) / FZ_HEAD_ROWS;
    // DSMEM region: the act operand of the down projection (phase B) aliases the q|k|v / partial / attn buffers (phase A)
    const uint32_t r4 = (uint32_t)(p.Rh + 3) / 4;
    p.off_actx = 0;
    p.off_qkvraw = 0;
    p.off_attpart = (uint32_t)fz_al((size_t)192 * p.Rph * 2);
    p.off_attnx = (uint32_t)((p.off_attpart + r4 * FZ_CL * FZ_REC * 4 + 1023) & ~1023u);
    const uint32_t a_bytes = p.off_attnx + (uint32_t)p.Rph * 128u, b_bytes = (uint32_t)p.Rp * 256u;
    p.dsm_bytes = ((a_bytes > b_bytes ? a_bytes : b_bytes) + 1023) & ~1023u;
    // per-CTA metadata: cos|sin 128 B + L + block per local row, 4 page ids per KV slot (a row's chunks are dealt over 4 CTAs)
    const uint32_t max_ch = (uint32_t)(st->max_blocks + 3) / 4;
    p.meta_pages_max = (uint32_t)p.Rh * ((max_ch + 3) / 4);
    T3B_CHECK_SHAPE(max_ch <= 32, "t3b_decode_step_fused: max_blocks %d (> 2048 tokens per row) not supported", st->max_blocks);
    const size_t meta_bytes = (size_t)p.Rh * 128 + (size_t)p.meta_pages_max * 16 + (size_t)p.Rh * 8 + 64;
    const size_t fixed = 1024 + FZ_XBYTES + p.dsm_bytes + sizeof(FusedSmem) + 64 + meta_bytes;
    p.nslot = (int)((227 * 1024 - fixed) / FZ_SLOT);
    if (p.nslot > FZ_MAX_SLOTS) p.nslot = FZ_MAX_SLOTS;
    T3B_CHECK_SHAPE(p.nslot >= 4, "t3b_decode_step_fused: %d rows leave room for only %d ring slots", rows, p.nslot);
    p.s = (float*)ws.s; p.xn = (bf16*)ws.xn; p.P = (float*)ws.P;
    p.logits = logits_out ? (bf16*)logits_out : (bf16*)ws.logits;
    p.hidden_out = (bf16*)hidden_out;
    p.counters = (unsigned int*)ws.counters;
    p.trace = g_fused_trace;
    p.dbg = g_fused_dbg;

    const size_t smem = (size_t)p.nslot * FZ_SLOT + fixed;
    T3B_CUDA(cudaFuncSetAttribute(decode_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    T3B_CUDA(cudaMemsetAsync(ws.counters, 0, 256, s));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(FZ_NCTA); cfg.blockDim = dim3(FZ_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;        // all 32 clusters must be co-resident: they meet at grid barriers
    attr[0].val.cooperative = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    static int coop = -1;                                // cooperative + cluster launch: fall back to a plain launch if refused
    if (coop != 0) {
        cudaError_t e = cudaLaunchKernelEx(&cfg, decode_fused_kernel, p);
        if (e == cudaSuccess) { coop = 1; return T3B_OK; }
        if (coop == 1) T3B_FAIL(T3B_ERR_CUDA, "t3b_decode_step_fused: launch failed: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        coop = 0;
    }
    cfg.numAttrs = 0;
    int ncl = 0;
    T3B_CUDA(cudaOccupancyMaxActiveClusters(&ncl, decode_fused_kernel, &cfg));
    T3B_CHECK_ARG(ncl >= FZ_NCL, "t3b_decode_step_fused: only %d of %d clusters can be co-resident", ncl, FZ_NCL);
    T3B_CUDA(cudaLaunchKernelEx(&cfg, decode_fused_kernel, p));
    return T3B_OK;
}
