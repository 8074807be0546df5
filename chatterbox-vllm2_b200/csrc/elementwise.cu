// E1 embedding gather, F2.1 fused add+RMSNorm, F2.3 RoPE + paged-KV write, F2.6 SiLU*mul.
// All are HBM/L2-bound byte work: 128-bit accesses, one pass, bf16 rounding at the points where
// the reference's bf16 pipeline materialises a tensor (see oracle/t3_ref.py header).
#include <cuda_fp8.h>

#include "t3b_common.cuh"

// ------------------------------------------------------------------ E1
// t3.py:440,475-480: out[dup*b+j] = bf16(emb[tok] + pos_emb[p % pos_rows])
__global__ void __launch_bounds__(128) embed_gather_kernel(const uint4* __restrict__ emb, const uint4* __restrict__ pos_emb,
                                                           const int32_t* __restrict__ tokens, const int32_t* __restrict__ speech_pos,
                                                           int hidden8, int emb_rows, int pos_rows, int dup, uint4* __restrict__ out) {
    pdl_launch_dependents();       // early trigger: the next kernel's launch overlaps this one (it waits before reading)
    pdl_wait();
    const int b = blockIdx.x;
    int tok = ld_act(tokens + b);
    tok = min(max(tok, 0), emb_rows - 1);
    int p = ld_act(speech_pos + b) % pos_rows;
    if (p < 0) p += pos_rows;
    for (int i = threadIdx.x; i < hidden8; i += blockDim.x) {
        float a[8], c[8];
        unpack8(emb[(size_t)tok * hidden8 + i], a);
        unpack8(pos_emb[(size_t)p * hidden8 + i], c);
#pragma unroll
        for (int e = 0; e < 8; ++e) a[e] += c[e];
        const uint4 v = pack8(a);
        for (int j = 0; j < dup; ++j) out[(size_t)(b * dup + j) * hidden8 + i] = v;
    }
}

// ------------------------------------------------------------------ F2.1
// vllm ir/ops/layernorm.py:40-59.  One CTA per row; each thread owns up to 2 chunks of 8 elements
// (hidden <= 2048*... generic loop kept in registers for hidden == 1024 with 128 threads).
template <int MAX_CHUNKS>
__global__ void __launch_bounds__(128) add_rmsnorm_kernel(const uint4* __restrict__ x, const float4* __restrict__ partials, int n_split,
                                                          size_t split_stride4,  // rows*hidden/4
                                                          uint4* __restrict__ residual, const uint4* __restrict__ weight,
                                                          uint4* __restrict__ y, int hidden8, float eps, int has_residual) {
    __shared__ float red[4];
    pdl_launch_dependents();       // early trigger: the next kernel's launch overlaps this one (it waits before reading)
    uint4 wraw[MAX_CHUNKS];        // norm weights never depend on the previous kernel: fetched before the wait
#pragma unroll
    for (int c = 0; c < MAX_CHUNKS; ++c) wraw[c] = threadIdx.x + c * 128 < hidden8 ? __ldg(weight + threadIdx.x + c * 128) : make_uint4(0, 0, 0, 0);
    pdl_wait();
    const int row = blockIdx.x;
    float s[MAX_CHUNKS][8];
    float ss = 0.f;
#pragma unroll
    for (int c = 0; c < MAX_CHUNKS; ++c) {
        const int i = threadIdx.x + c * 128;
        if (i < hidden8) {
            float xv[8];
            if (x != nullptr) {
                unpack8(ld_act(x + (size_t)row * hidden8 + i), xv);
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) xv[e] = 0.f;
                for (int sp = 0; sp < n_split; ++sp) {            // fixed order -> deterministic
                    const float4* p = partials + sp * split_stride4 + ((size_t)row * hidden8 + i) * 2;
                    const float4 a = ld_act(p), b = ld_act(p + 1);
                    xv[0] += a.x; xv[1] += a.y; xv[2] += a.z; xv[3] += a.w;
                    xv[4] += b.x; xv[5] += b.y; xv[6] += b.z; xv[7] += b.w;
                }
#pragma unroll
                for (int e = 0; e < 8; ++e) xv[e] = bf16_round(xv[e]);  // the GEMM's bf16 output
            }
            if (has_residual) {
                float r[8];
                unpack8(ld_act(residual + (size_t)row * hidden8 + i), r);
#pragma unroll
                for (int e = 0; e < 8; ++e) xv[e] += r[e];
            }
            residual[(size_t)row * hidden8 + i] = pack8(xv);
#pragma unroll
            for (int e = 0; e < 8; ++e) { s[c][e] = xv[e]; ss += xv[e] * xv[e]; }
        }
    }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    const float tot = (red[0] + red[1]) + (red[2] + red[3]);
    const float rstd = rsqrtf(tot / (float)(hidden8 * 8) + eps);
#pragma unroll
    for (int c = 0; c < MAX_CHUNKS; ++c) {
        const int i = threadIdx.x + c * 128;
        if (i < hidden8) {
            float w[8], o[8];
            unpack8(wraw[c], w);
#pragma unroll
            for (int e = 0; e < 8; ++e) o[e] = bf16_round(s[c][e] * rstd) * w[e];   // x.to(weight.dtype) * weight
            y[(size_t)row * hidden8 + i] = pack8(o);
        }
    }
}

// F2.1 over a row whose residual sum already exists in fp32 (left by the o / down projection's epilogue, linear_umma.cu EPI_RESID):
// the last normalisation of the stack when the per-layer ones ride inside the projections.  Same arithmetic and the same
// reduction order as add_rmsnorm_kernel.
__global__ void __launch_bounds__(128) rmsnorm_s32_kernel(const float4* __restrict__ s32, const uint4* __restrict__ weight, uint4* __restrict__ y,
                                                          int hidden8, float eps) {
    __shared__ float red[4];
    pdl_launch_dependents();
    const int i = threadIdx.x;
    const uint4 wraw = i < hidden8 ? __ldg(weight + i) : make_uint4(0, 0, 0, 0);
    pdl_wait();
    const int row = blockIdx.x;
    float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float ss = 0.f;
    if (i < hidden8) {
        const float4* p = s32 + ((size_t)row * hidden8 + i) * 2;
        const float4 a = ld_act(p), b = ld_act(p + 1);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
#pragma unroll
        for (int e = 0; e < 8; ++e) ss += v[e] * v[e];
    }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    const float tot = (red[0] + red[1]) + (red[2] + red[3]);
    const float rstd = rsqrtf(tot / (float)(hidden8 * 8) + eps);
    if (i < hidden8) {
        float w[8], o[8];
        unpack8(wraw, w);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = bf16_round(v[e] * rstd) * w[e];
        y[(size_t)row * hidden8 + i] = pack8(o);
    }
}
int32_t t3b_rmsnorm_s32_launch(const float* s32, const void* weight, void* y, int rows, int hidden, float eps, cudaStream_t stream) {
    T3B_CHECK_SHAPE(hidden % 8 == 0 && hidden <= 1024, "rmsnorm_s32: hidden %d unsupported", hidden);
    T3B_CUDA(t3b_launch(rmsnorm_s32_kernel, dim3(rows), dim3(128), 0, stream, (const float4*)s32, (const uint4*)weight, (uint4*)y, hidden / 8, eps));
    return T3B_OK;
}

// ------------------------------------------------------------------ F2.3 + KV write
// One CTA per token, one warp per head: lane j rotates the pair (j, j+32) of q and k
// (rotary_embedding/common.py:144-174 in bf16 tensor arithmetic: each product and the sum round).
// KVT = 1: K (rotated) and V are stored as e4m3 with one fp32 scale per (token, head) = max|x| / 448 (T3B_KV_FP8).
template <int KVT>
__global__ void __launch_bounds__(512) rope_kv_write_kernel(const bf16* __restrict__ qkv, const float* __restrict__ partials, int n_split,
                                                            size_t split_stride, const int32_t* __restrict__ tok_row,
                                                            const int32_t* __restrict__ tok_idx, const int32_t* __restrict__ rope_pos,
                                                            const bf16* __restrict__ cos_sin, int max_pos,
                                                            const int32_t* __restrict__ block_tables, int max_blocks,
                                                            unsigned char* __restrict__ k_pool, unsigned char* __restrict__ v_pool, int n_heads,
                                                            bf16* __restrict__ q_out) {
    pdl_launch_dependents();       // early trigger: the next kernel's launch overlaps this one (it waits before reading)
    pdl_wait();
    const int t = blockIdx.x;
    const int lane = threadIdx.x & 31;
    const int hidden = n_heads * 64;
    const int row = tok_row ? ld_act(tok_row + t) : t;
    const int idx = ld_act(tok_idx + t);
    int pos = ld_act(rope_pos + t);
    pos = min(max(pos, 0), max_pos - 1);
    const float c = __bfloat162float(cos_sin[(size_t)pos * 64 + lane]);
    const float s = __bfloat162float(cos_sin[(size_t)pos * 64 + 32 + lane]);
    const int blk = ld_act(block_tables + (size_t)row * max_blocks + (idx >> 4));
    for (int h = threadIdx.x >> 5; h < n_heads; h += blockDim.x >> 5) {
        float v[6];   // q1 q2 k1 k2 v1 v2
#pragma unroll
        for (int part = 0; part < 3; ++part) {
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const size_t col = (size_t)part * hidden + h * 64 + half * 32 + lane;
                float a;
                if (qkv != nullptr) {
                    a = ld_act(qkv + (size_t)t * 3 * hidden + col);
                } else {
                    a = 0.f;
                    for (int sp = 0; sp < n_split; ++sp) a += ld_act(partials + sp * split_stride + (size_t)t * 3 * hidden + col);
                    a = bf16_round(a);
                }
                v[part * 2 + half] = a;
            }
        }
        const float q1 = bf16_round(bf16_round(v[0] * c) - bf16_round(v[1] * s));
        const float q2 = bf16_round(bf16_round(v[1] * c) + bf16_round(v[0] * s));
        const float k1 = bf16_round(bf16_round(v[2] * c) - bf16_round(v[3] * s));
        const float k2 = bf16_round(bf16_round(v[3] * c) + bf16_round(v[2] * s));
        bf16* qo = q_out + (size_t)t * hidden + h * 64;
        qo[lane] = __float2bfloat16_rn(q1);
        qo[32 + lane] = __float2bfloat16_rn(q2);
        if (KVT == 1) {
            const float ka = warp_max(fmaxf(fabsf(k1), fabsf(k2))), va = warp_max(fmaxf(fabsf(v[4]), fabsf(v[5])));
            const float ks = ka > 0.f ? ka / 448.f : 1.f, vs = va > 0.f ? va / 448.f : 1.f;
            const size_t sl = ((size_t)blk * n_heads + h) * 1088;
            unsigned char* kp = k_pool + sl + (idx & 15) * 64;
            unsigned char* vp = v_pool + sl + (idx & 15) * 64;
            kp[lane] = (unsigned char)__nv_cvt_float_to_fp8(k1 / ks, __NV_SATFINITE, __NV_E4M3);
            kp[32 + lane] = (unsigned char)__nv_cvt_float_to_fp8(k2 / ks, __NV_SATFINITE, __NV_E4M3);
            vp[lane] = (unsigned char)__nv_cvt_float_to_fp8(v[4] / vs, __NV_SATFINITE, __NV_E4M3);
            vp[32 + lane] = (unsigned char)__nv_cvt_float_to_fp8(v[5] / vs, __NV_SATFINITE, __NV_E4M3);
            if (lane == 0) {
                *reinterpret_cast<float*>(k_pool + sl + 1024 + (idx & 15) * 4) = ks;
                *reinterpret_cast<float*>(v_pool + sl + 1024 + (idx & 15) * 4) = vs;
            }
        } else {
            const size_t slot = (((size_t)blk * n_heads + h) * T3B_KV_BLOCK + (idx & 15)) * 64;
            bf16* kp = reinterpret_cast<bf16*>(k_pool) + slot;
            bf16* vp = reinterpret_cast<bf16*>(v_pool) + slot;
            kp[lane] = __float2bfloat16_rn(k1);
            kp[32 + lane] = __float2bfloat16_rn(k2);
            vp[lane] = __float2bfloat16_rn(v[4]);
            vp[32 + lane] = __float2bfloat16_rn(v[5]);
        }
    }
}

// The same operation on bf16 qkv with 128-bit accesses (the prefill pass: thousands of tokens): 8 lanes per head, lane c holds dims
// [8c, 8c+8) (c < 4: x1, c >= 4: x2 - the rotation partner sits 4 lanes away); 128 threads per token, 4 tokens per block.  Same
// roundings as the scalar kernel (each product and the sum round to bf16) - bit-identical; 56 -> ~15 us per layer at 6,336 tokens.
template <int KVT>
__global__ void __launch_bounds__(512) rope_kv_write_vec_kernel(const bf16* __restrict__ qkv, int n_tok, const int32_t* __restrict__ tok_row,
                                                                const int32_t* __restrict__ tok_idx, const int32_t* __restrict__ rope_pos,
                                                                const bf16* __restrict__ cos_sin, int max_pos,
                                                                const int32_t* __restrict__ block_tables, int max_blocks,
                                                                unsigned char* __restrict__ k_pool, unsigned char* __restrict__ v_pool,
                                                                bf16* __restrict__ q_out) {
    pdl_launch_dependents();
    pdl_wait();
    const int t = blockIdx.x * 4 + (threadIdx.x >> 7);
    if (t >= n_tok) return;
    const int u = threadIdx.x & 127, h = u >> 3, c = u & 7;          // 16 heads x 8 lanes
    const int hidden = 1024;
    const int row = tok_row ? ld_act(tok_row + t) : t;
    const int idx = ld_act(tok_idx + t);
    int pos = ld_act(rope_pos + t);
    pos = min(max(pos, 0), max_pos - 1);
    const int blk = ld_act(block_tables + (size_t)row * max_blocks + (idx >> 4));
    const bf16* src = qkv + (size_t)t * 3 * hidden + h * 64 + c * 8;
    float qf[8], kf[8], vf[8], cs[8], sn[8];
    unpack8(ld_act(reinterpret_cast<const uint4*>(src)), qf);
    unpack8(ld_act(reinterpret_cast<const uint4*>(src + hidden)), kf);
    const uint4 vraw = ld_act(reinterpret_cast<const uint4*>(src + 2 * hidden));
    unpack8(__ldg(reinterpret_cast<const uint4*>(cos_sin + (size_t)pos * 64 + (c & 3) * 8)), cs);
    unpack8(__ldg(reinterpret_cast<const uint4*>(cos_sin + (size_t)pos * 64 + 32 + (c & 3) * 8)), sn);
    // all shuffles first, the sign of the second product by an exact +-1 (same arithmetic as the decode attention's item head, where one
    // shuffle -> round chain per element measured 1.8 us per launch: profiles/r2_notes.md r5p)
    const float sg = c >= 4 ? 1.f : -1.f;
    float qp[8], kp[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) { qp[e] = __shfl_xor_sync(0xffffffffu, qf[e], 4); kp[e] = __shfl_xor_sync(0xffffffffu, kf[e], 4); }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        qf[e] = bf16_round(bf16_round(qf[e] * cs[e]) + sg * bf16_round(qp[e] * sn[e]));
        kf[e] = bf16_round(bf16_round(kf[e] * cs[e]) + sg * bf16_round(kp[e] * sn[e]));
    }
    *reinterpret_cast<uint4*>(q_out + (size_t)t * hidden + h * 64 + c * 8) = pack8(qf);
    if (KVT == 1) {
        unpack8(vraw, vf);
        float ka = 0.f, va = 0.f;
#pragma unroll
        for (int e = 0; e < 8; ++e) { ka = fmaxf(ka, fabsf(kf[e])); va = fmaxf(va, fabsf(vf[e])); }
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) { ka = fmaxf(ka, __shfl_xor_sync(0xffffffffu, ka, o)); va = fmaxf(va, __shfl_xor_sync(0xffffffffu, va, o)); }
        const float ks = ka > 0.f ? ka / 448.f : 1.f, vs = va > 0.f ? va / 448.f : 1.f;
        uint32_t kw[2] = {0u, 0u}, vw[2] = {0u, 0u};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            kw[i >> 1] |= (uint32_t)__nv_cvt_float2_to_fp8x2(make_float2(kf[2 * i] / ks, kf[2 * i + 1] / ks), __NV_SATFINITE, __NV_E4M3) << (16 * (i & 1));
            vw[i >> 1] |= (uint32_t)__nv_cvt_float2_to_fp8x2(make_float2(vf[2 * i] / vs, vf[2 * i + 1] / vs), __NV_SATFINITE, __NV_E4M3) << (16 * (i & 1));
        }
        const size_t sl = ((size_t)blk * 16 + h) * 1088;
        *reinterpret_cast<uint2*>(k_pool + sl + (idx & 15) * 64 + c * 8) = make_uint2(kw[0], kw[1]);
        *reinterpret_cast<uint2*>(v_pool + sl + (idx & 15) * 64 + c * 8) = make_uint2(vw[0], vw[1]);
        if (c == 0) {
            *reinterpret_cast<float*>(k_pool + sl + 1024 + (idx & 15) * 4) = ks;
            *reinterpret_cast<float*>(v_pool + sl + 1024 + (idx & 15) * 4) = vs;
        }
    } else {
        const size_t sl = ((size_t)blk * 16 + h) * 2048 + ((idx & 15) * 64 + c * 8) * 2;
        *reinterpret_cast<uint4*>(k_pool + sl) = pack8(kf);
        *reinterpret_cast<uint4*>(v_pool + sl) = vraw;
    }
}

// ------------------------------------------------------------------ prefill embedding assembly (t3.py:542-561)
// One CTA per sequence (2 per request: cond, uncond); 128 threads = one 1024-wide row per pass.
__global__ void __launch_bounds__(256) prefill_embed_kernel(const uint4* __restrict__ conds, const int32_t* __restrict__ text_ids,
                                                            const int32_t* __restrict__ text_start, const int32_t* __restrict__ out_start,
                                                            const int32_t* __restrict__ slots, const uint4* __restrict__ text_emb,
                                                            const uint4* __restrict__ text_pos_emb, int text_rows, int text_pos_rows,
                                                            const uint4* __restrict__ speech_emb, const uint4* __restrict__ speech_pos_emb,
                                                            int bos_token, int hidden8, int cond_len, uint4* __restrict__ x_out,
                                                            int32_t* __restrict__ tok_row, int32_t* __restrict__ tok_idx) {
    pdl_launch_dependents();
    pdl_wait();
    const int seq = blockIdx.x, r = seq >> 1, half = seq & 1;
    const int t0 = text_start[r], n_text = text_start[r + 1] - t0;
    const int o0 = out_start[seq], P = cond_len + n_text + 1;
    const int row = 2 * slots[r] + half;
    for (int i = threadIdx.x; i < P; i += blockDim.x) { tok_row[o0 + i] = row; tok_idx[o0 + i] = i; }
    const int per = blockDim.x / hidden8 > 0 ? blockDim.x / hidden8 : 1;       // rows per pass
    const int sub = threadIdx.x / hidden8, col = threadIdx.x - sub * hidden8;
    if (sub >= per) return;
    for (int i = sub; i < P; i += per) {
        uint4 v;
        if (i < cond_len) {
            v = conds[((size_t)r * cond_len + i) * hidden8 + col];
        } else if (i < P - 1) {
            if (half == 0) {
                int id = text_ids[t0 + i - cond_len];
                id = min(max(id, 0), text_rows - 1);
                const int p = min(i - cond_len, text_pos_rows - 1);
                float a[8], b[8];
                unpack8(text_emb[(size_t)id * hidden8 + col], a);
                unpack8(text_pos_emb[(size_t)p * hidden8 + col], b);
#pragma unroll
                for (int e = 0; e < 8; ++e) a[e] += b[e];
                v = pack8(a);
            } else {
                v = make_uint4(0, 0, 0, 0);               // t3.py:556: the uncond row zeroes the text embeddings
            }
        } else {
            float a[8], b[8];                             // BOS = speech_emb[start] + speech_pos_emb[0] (t3.py:550-551)
            unpack8(speech_emb[(size_t)bos_token * hidden8 + col], a);
            unpack8(speech_pos_emb[col], b);
#pragma unroll
            for (int e = 0; e < 8; ++e) a[e] += b[e];
            v = pack8(a);
        }
        x_out[(size_t)(o0 + i) * hidden8 + col] = v;
    }
}

// ------------------------------------------------------------------ S3Gen hand-off (tts.py:492-514)
// order-preserving compaction of the ids whose un-offset value is a valid S3 code; one CTA, ballot scan
__global__ void __launch_bounds__(1024) filter_speech_tokens_kernel(const int32_t* __restrict__ ids, int n, int offset, int n_valid,
                                                                     int32_t* __restrict__ out, int32_t* __restrict__ n_out) {
    __shared__ int warp_cnt[32];
    __shared__ int base;
    pdl_launch_dependents();
    pdl_wait();
    if (threadIdx.x == 0) base = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i0 = 0; i0 < n; i0 += 1024) {
        const int i = i0 + threadIdx.x;
        const int v = i < n ? ids[i] - offset : -1;
        const bool keep = i < n && v >= 0 && v < n_valid;
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) warp_cnt[warp] = __popc(m);
        __syncthreads();
        int off = base;
        for (int w = 0; w < warp; ++w) off += warp_cnt[w];
        const int pos = off + __popc(m & ((1u << lane) - 1u));
        __syncthreads();                                   // every thread has read `base` / warp_cnt and ids[i] (out may alias ids)
        if (keep) out[pos] = v;
        if (threadIdx.x == 1023) base = pos + (keep ? 1 : 0);
        __syncthreads();
    }
    if (threadIdx.x == 0 && n_out) *n_out = base;
}

// ------------------------------------------------------------------ F2.6
// activation.py:138-141: F.silu(gate) * up on bf16 tensors -> two roundings.
__global__ void __launch_bounds__(256) silu_mul_kernel(const uint4* __restrict__ gu, const float4* __restrict__ partials, int n_split,
                                                       size_t split_stride4, int ffn8, uint4* __restrict__ act, int total) {
    pdl_launch_dependents();       // early trigger: the next kernel's launch overlaps this one (it waits before reading)
    pdl_wait();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int row = i / ffn8, c = i - row * ffn8;
        float g[8], u[8];
        if (gu != nullptr) {
            unpack8(ld_act(gu + (size_t)row * 2 * ffn8 + c), g);
            unpack8(ld_act(gu + (size_t)row * 2 * ffn8 + ffn8 + c), u);
        } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) { g[e] = 0.f; u[e] = 0.f; }
            for (int sp = 0; sp < n_split; ++sp) {
                const float4* pg = partials + sp * split_stride4 + ((size_t)row * 2 * ffn8 + c) * 2;
                const float4* pu = pg + (size_t)ffn8 * 2;
                const float4 a = ld_act(pg), b = ld_act(pg + 1), d = ld_act(pu), f = ld_act(pu + 1);
                g[0] += a.x; g[1] += a.y; g[2] += a.z; g[3] += a.w; g[4] += b.x; g[5] += b.y; g[6] += b.z; g[7] += b.w;
                u[0] += d.x; u[1] += d.y; u[2] += d.z; u[3] += d.w; u[4] += f.x; u[5] += f.y; u[6] += f.z; u[7] += f.w;
            }
#pragma unroll
            for (int e = 0; e < 8; ++e) { g[e] = bf16_round(g[e]); u[e] = bf16_round(u[e]); }
        }
        float o[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = bf16_round(g[e] / (1.0f + expf(-g[e]))) * u[e];
        act[(size_t)row * ffn8 + c] = pack8(o);
    }
}

// ------------------------------------------------------------------ C ABI
extern "C" int32_t t3b_embed_gather(const void* emb, const void* pos_emb, const int32_t* tokens, const int32_t* speech_pos,
                                    int32_t n_req, int32_t hidden, int32_t emb_rows, int32_t pos_rows, int32_t dup, void* out,
                                    void* stream) {
    T3B_CHECK_ARG(emb && pos_emb && tokens && speech_pos && out, "t3b_embed_gather: null pointer");
    T3B_CHECK_ARG(n_req >= 0 && dup >= 1 && emb_rows > 0 && pos_rows > 0, "t3b_embed_gather: bad sizes");
    T3B_CHECK_SHAPE(hidden > 0 && hidden % 8 == 0, "t3b_embed_gather: hidden %d not a multiple of 8", hidden);
    T3B_CHECK_ALIGN(emb); T3B_CHECK_ALIGN(pos_emb); T3B_CHECK_ALIGN(out);
    if (n_req == 0) return T3B_OK;
    T3B_CUDA(t3b_launch(embed_gather_kernel, dim3(n_req), dim3(128), 0, t3b_stream(stream), (const uint4*)emb, (const uint4*)pos_emb,
                        tokens, speech_pos, hidden / 8, emb_rows, pos_rows, dup, (uint4*)out));
    return T3B_OK;
}

extern "C" int32_t t3b_add_rmsnorm(const void* x, const float* partials, int32_t n_split, void* residual, const void* weight,
                                   void* y, int32_t rows, int32_t hidden, float eps, int32_t has_residual, void* stream) {
    T3B_CHECK_ARG((x != nullptr) != (partials != nullptr), "t3b_add_rmsnorm: exactly one of x / partials");
    T3B_CHECK_ARG(residual && weight && y && rows >= 0, "t3b_add_rmsnorm: null pointer");
    T3B_CHECK_ARG(x || n_split >= 1, "t3b_add_rmsnorm: n_split %d", n_split);
    T3B_CHECK_SHAPE(hidden > 0 && hidden % 8 == 0 && hidden <= 4096, "t3b_add_rmsnorm: hidden %d unsupported", hidden);
    if (x) T3B_CHECK_ALIGN(x);
    if (partials) T3B_CHECK_ALIGN(partials);
    T3B_CHECK_ALIGN(residual); T3B_CHECK_ALIGN(weight); T3B_CHECK_ALIGN(y);
    if (rows == 0) return T3B_OK;
    const size_t stride4 = (size_t)rows * hidden / 4;
    if (hidden <= 1024) {
        T3B_CUDA(t3b_launch(add_rmsnorm_kernel<1>, dim3(rows), dim3(128), 0, t3b_stream(stream), (const uint4*)x, (const float4*)partials,
                            n_split, stride4, (uint4*)residual, (const uint4*)weight, (uint4*)y, hidden / 8, eps, has_residual));
    } else {
        T3B_CUDA(t3b_launch(add_rmsnorm_kernel<4>, dim3(rows), dim3(128), 0, t3b_stream(stream), (const uint4*)x, (const float4*)partials,
                            n_split, stride4, (uint4*)residual, (const uint4*)weight, (uint4*)y, hidden / 8, eps, has_residual));
    }
    return T3B_OK;
}

extern "C" int32_t t3b_rope_kv_write(const void* qkv, const float* partials, int32_t n_split, int32_t n_tok, const int32_t* tok_row,
                                     const int32_t* tok_idx, const int32_t* rope_pos, const void* rope_cos_sin, int32_t max_pos,
                                     const int32_t* block_tables, int32_t max_blocks, void* k_pool, void* v_pool, int32_t n_heads,
                                     int32_t kv_dtype, void* q_out, void* stream) {
    T3B_CHECK_ARG((qkv != nullptr) != (partials != nullptr), "t3b_rope_kv_write: exactly one of qkv / partials");
    T3B_CHECK_ARG(tok_idx && rope_pos && rope_cos_sin && block_tables && k_pool && v_pool && q_out, "t3b_rope_kv_write: null pointer");
    T3B_CHECK_ARG(n_tok >= 0 && max_blocks > 0 && max_pos > 0 && (qkv || n_split >= 1), "t3b_rope_kv_write: bad sizes");
    T3B_CHECK_SHAPE(n_heads >= 1 && n_heads <= 64, "t3b_rope_kv_write: n_heads %d", n_heads);
    T3B_CHECK_ARG(kv_dtype == T3B_KV_BF16 || kv_dtype == T3B_KV_FP8, "t3b_rope_kv_write: unknown KV dtype %d", kv_dtype);
    if (n_tok == 0) return T3B_OK;
    if (qkv != nullptr && n_heads == 16 && n_tok >= 64) {              // the prefill pass: 128-bit accesses, 4 tokens per block
        auto vk = kv_dtype == T3B_KV_FP8 ? rope_kv_write_vec_kernel<1> : rope_kv_write_vec_kernel<0>;
        T3B_CUDA(t3b_launch(vk, dim3((n_tok + 3) / 4), dim3(512), 0, t3b_stream(stream), (const bf16*)qkv, n_tok, tok_row, tok_idx, rope_pos,
                            (const bf16*)rope_cos_sin, max_pos, block_tables, max_blocks, (unsigned char*)k_pool, (unsigned char*)v_pool,
                            (bf16*)q_out));
        return T3B_OK;
    }
    const int threads = 32 * (n_heads < 16 ? n_heads : 16);
    auto kern = kv_dtype == T3B_KV_FP8 ? rope_kv_write_kernel<1> : rope_kv_write_kernel<0>;
    T3B_CUDA(t3b_launch(kern, dim3(n_tok), dim3(threads), 0, t3b_stream(stream), (const bf16*)qkv, partials, n_split,
                        (size_t)n_tok * 3 * n_heads * 64, tok_row, tok_idx, rope_pos, (const bf16*)rope_cos_sin, max_pos, block_tables,
                        max_blocks, (unsigned char*)k_pool, (unsigned char*)v_pool, n_heads, (bf16*)q_out));
    return T3B_OK;
}

extern "C" int32_t t3b_prefill_embed(const void* conds, const int32_t* text_ids, const int32_t* text_start, const int32_t* out_start,
                                     const int32_t* slots, int32_t n_req, const void* text_emb, const void* text_pos_emb, int32_t text_rows,
                                     int32_t text_pos_rows, const void* speech_emb, const void* speech_pos_emb, int32_t bos_token,
                                     int32_t hidden, int32_t cond_len, void* x_out, int32_t* tok_row, int32_t* tok_idx, void* stream) {
    T3B_CHECK_ARG(conds && text_ids && text_start && out_start && slots && text_emb && text_pos_emb && speech_emb && speech_pos_emb && x_out &&
                  tok_row && tok_idx, "t3b_prefill_embed: null pointer");
    T3B_CHECK_ARG(n_req >= 0 && text_rows > 0 && text_pos_rows > 0 && bos_token >= 0 && cond_len >= 0, "t3b_prefill_embed: bad sizes");
    T3B_CHECK_SHAPE(hidden > 0 && hidden % 8 == 0 && hidden / 8 <= 256, "t3b_prefill_embed: hidden %d unsupported", hidden);
    T3B_CHECK_ALIGN(conds); T3B_CHECK_ALIGN(text_emb); T3B_CHECK_ALIGN(text_pos_emb); T3B_CHECK_ALIGN(speech_emb); T3B_CHECK_ALIGN(speech_pos_emb);
    T3B_CHECK_ALIGN(x_out);
    if (n_req == 0) return T3B_OK;
    T3B_CUDA(t3b_launch(prefill_embed_kernel, dim3(2 * n_req), dim3(256), 0, t3b_stream(stream), (const uint4*)conds, text_ids, text_start,
                        out_start, slots, (const uint4*)text_emb, (const uint4*)text_pos_emb, text_rows, text_pos_rows, (const uint4*)speech_emb,
                        (const uint4*)speech_pos_emb, bos_token, hidden / 8, cond_len, (uint4*)x_out, tok_row, tok_idx));
    return T3B_OK;
}

extern "C" int32_t t3b_filter_speech_tokens(const int32_t* ids, int32_t n, int32_t offset, int32_t n_valid, int32_t* out, int32_t* n_out,
                                            void* stream) {
    T3B_CHECK_ARG(ids && out && n_out && n >= 0 && n_valid > 0, "t3b_filter_speech_tokens: bad arguments");
    T3B_CUDA(t3b_launch(filter_speech_tokens_kernel, dim3(1), dim3(1024), 0, t3b_stream(stream), ids, n, offset, n_valid, out, n_out));
    return T3B_OK;
}

extern "C" int32_t t3b_silu_mul(const void* gate_up, const float* partials, int32_t n_split, int32_t rows, int32_t ffn, void* act,
                                void* stream) {
    T3B_CHECK_ARG((gate_up != nullptr) != (partials != nullptr), "t3b_silu_mul: exactly one of gate_up / partials");
    T3B_CHECK_ARG(act && rows >= 0 && (gate_up || n_split >= 1), "t3b_silu_mul: bad arguments");
    T3B_CHECK_SHAPE(ffn > 0 && ffn % 8 == 0, "t3b_silu_mul: ffn %d not a multiple of 8", ffn);
    if (gate_up) T3B_CHECK_ALIGN(gate_up);
    if (partials) T3B_CHECK_ALIGN(partials);
    T3B_CHECK_ALIGN(act);
    if (rows == 0) return T3B_OK;
    const int total = rows * (ffn / 8);
    int blocks = (total + 255) / 256;
    const int cap = t3b_num_sms() * 8;
    if (blocks > cap) blocks = cap;
    T3B_CUDA(t3b_launch(silu_mul_kernel, dim3(blocks), dim3(256), 0, t3b_stream(stream), (const uint4*)gate_up, (const float4*)partials,
                        n_split, (size_t)rows * 2 * ffn / 4, ffn / 8, (uint4*)act, total));
    return T3B_OK;
}

