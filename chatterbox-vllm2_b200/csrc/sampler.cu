// C1 (CFG combine) + S1..S4 (penalty, greedy | temperature -> min-p -> top-p -> exponential race)
// in one pass over the 8194-entry speech vocabulary, one CTA per request, logits resident in shared
// memory.  Follows vllm v1/sample/sampler.py:67-307, logits_processor/builtin.py:100-114,
// ops/topk_topp_sampler.py:328-416, layers/utils.py:51-89 and t3.py:650-673.
//
// No sort: top-p needs the ascending-order cumulative mass only to find ONE threshold, which a
// 4-pass 8-bit radix select over the fp32 bit patterns of exp(l - max) finds exactly.  Masses are
// accumulated as 2^-40 fixed point integers, so the result is independent of summation order
// (deterministic across runs and launch shapes); ties at the threshold are resolved in index order
// like the reference's stable ascending sort.
//
// RNG contract (identical in oracle/sampler_ref.py): x = Philox4x32-10(ctr=(i, step, 0, 0),
// key=(seed_lo, seed_hi))[0]; u = ((x >> 9) + 0.5) * 2^-23; q = -log(u); token = argmax p_i / q_i.
#include "t3b_common.cuh"

#define SMP_THREADS 256
#define SMP_MAX_V 8448
#define SMP_EPS 1e-5f

struct SamplerArgs {
    const bf16* head;            // bf16 [2B, V] head outputs (cond, uncond interleaved) or NULL
    const float* head_partials;  // fp32 [n_split, 2B, V] or NULL
    int n_split;
    const float* logits_in;      // fp32 [B, V] already combined (t3b_sample_logits) or NULL
    int V;
    int n_req;
    t3b_sampling_t sp;
    t3b_state_t st;
    int has_state;
    int advance;
    int32_t* out_tokens;
    float* logits_out;
};

__device__ __forceinline__ uint32_t philox_x0(uint32_t c0, uint32_t c1, uint32_t k0, uint32_t k1) {
    uint32_t c2 = 0, c3 = 0;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        c0 = hi1 ^ c1 ^ k0; c1 = lo1; c2 = hi0 ^ c3 ^ k1; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return c0;
}

// ---- block reductions over SMP_THREADS threads (fixed tree -> deterministic) ----
__device__ __forceinline__ float block_max(float v, float* red) {
    v = warp_max(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = red[0];
#pragma unroll
    for (int i = 1; i < SMP_THREADS / 32; ++i) r = fmaxf(r, red[i]);
    return r;
}
__device__ __forceinline__ unsigned long long block_sum_u64(unsigned long long v, unsigned long long* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    unsigned long long r = 0;
#pragma unroll
    for (int i = 0; i < SMP_THREADS / 32; ++i) r += red[i];
    return r;
}
// argmax with the lowest index on ties (torch.argmax convention)
__device__ __forceinline__ int block_argmax(float v, int idx, float* redv, int* redi) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, v, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (ov > v || (ov == v && oi < idx)) { v = ov; idx = oi; }
    }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) { redv[threadIdx.x >> 5] = v; redi[threadIdx.x >> 5] = idx; }
    __syncthreads();
    float bv = redv[0];
    int bi = redi[0];
#pragma unroll
    for (int i = 1; i < SMP_THREADS / 32; ++i)
        if (redv[i] > bv || (redv[i] == bv && redi[i] < bi)) { bv = redv[i]; bi = redi[i]; }
    return bi;
}

__device__ __forceinline__ unsigned long long to_fixed(float e) { return (unsigned long long)((double)e * 1099511627776.0); }  // 2^40

__global__ void __launch_bounds__(SMP_THREADS) cfg_sample_kernel(const SamplerArgs a) {
    __shared__ float lg[SMP_MAX_V];
    __shared__ uint32_t seen[(SMP_MAX_V + 31) / 32];
    __shared__ unsigned long long hist[256];
    __shared__ unsigned long long red64[SMP_THREADS / 32];
    __shared__ float redf[SMP_THREADS / 32];
    __shared__ int redi[SMP_THREADS / 32];
    __shared__ unsigned int hcnt[256];
    __shared__ unsigned long long s_below;
    __shared__ unsigned int s_digit, s_tie_rank_base;

    // No early griddepcontrol.launch_dependents here: this kernel advances the decode state (ctx_lens, positions, history), and the
    // kernels of the next step read that state BEFORE their own dependency wait (attention prefetches K/V by it), so they must not
    // start until this grid has completed (the implicit trigger at exit).
    pdl_wait();
    const int b = blockIdx.x, tid = threadIdx.x, V = a.V;
    const bool active = a.has_state ? (a.st.active == nullptr || a.st.active[b] != 0) : true;

    // ---- C1: load + CFG combine (t3.py:659-662) -------------------------------------------------
    if (a.logits_in != nullptr) {
        for (int i = tid; i < V; i += SMP_THREADS) lg[i] = ld_act(a.logits_in + (size_t)b * V + i);
    } else {
        const float w = a.sp.cfg_scale ? a.sp.cfg_scale[b] : 0.5f;
        const int R = 2 * a.n_req;
        // bf16 head rows: two logits per 32-bit load when the rows are 4-byte aligned (V even: 8194), half the load instructions
        const bool pairs = a.head != nullptr && (V & 1) == 0 && (reinterpret_cast<uintptr_t>(a.head) & 3u) == 0;
        if (pairs) {
            const unsigned* hc = reinterpret_cast<const unsigned*>(a.head + (size_t)(2 * b) * V);
            const unsigned* hu = reinterpret_cast<const unsigned*>(a.head + (size_t)(2 * b + 1) * V);
            // one CTA per request and nothing else to hide latency behind: 8 x 2 loads in flight per thread (as a plain loop the 17
            // iterations were 17 L2 round trips back to back)
            constexpr int UN = 8;
            for (int i0 = tid; i0 < V / 2; i0 += SMP_THREADS * UN) {
                unsigned wc[UN], wu[UN];
#pragma unroll
                for (int q = 0; q < UN; ++q) {
                    const int i2 = i0 + q * SMP_THREADS;
                    wc[q] = i2 < V / 2 ? ld_act(hc + i2) : 0u;
                    wu[q] = i2 < V / 2 ? ld_act(hu + i2) : 0u;
                }
#pragma unroll
                for (int q = 0; q < UN; ++q) {
                    const int i2 = i0 + q * SMP_THREADS;
                    if (i2 >= V / 2) break;
#pragma unroll
                    for (int k = 0; k < 2; ++k) {
                        const float c = k ? bf16hi(wc[q]) : bf16lo(wc[q]), u = k ? bf16hi(wu[q]) : bf16lo(wu[q]);
                        float l;
                        if (a.sp.combine == T3B_COMBINE_BF16) l = bf16_round(c + bf16_round(w * bf16_round(c - u)));
                        else l = c + w * (c - u);
                        lg[2 * i2 + k] = l;
                        if (a.logits_out) a.logits_out[(size_t)b * V + 2 * i2 + k] = l;
                    }
                }
            }
        }
        for (int i = tid; i < (pairs ? 0 : V); i += SMP_THREADS) {
            float c, u;
            if (a.head != nullptr) {
                c = ld_act(a.head + (size_t)(2 * b) * V + i);
                u = ld_act(a.head + (size_t)(2 * b + 1) * V + i);
            } else {
                c = 0.f; u = 0.f;
                for (int s = 0; s < a.n_split; ++s) {
                    c += ld_act(a.head_partials + ((size_t)s * R + 2 * b) * V + i);
                    u += ld_act(a.head_partials + ((size_t)s * R + 2 * b + 1) * V + i);
                }
                c = bf16_round(c); u = bf16_round(u);              // ParallelLMHead output dtype
            }
            float l;
            if (a.sp.combine == T3B_COMBINE_BF16) l = bf16_round(c + bf16_round(w * bf16_round(c - u)));
            else l = c + w * (c - u);
            lg[i] = l;
            if (a.logits_out) a.logits_out[(size_t)b * V + i] = l;
        }
    }
    for (int i = tid; i < (V + 31) / 32; i += SMP_THREADS) seen[i] = 0u;
    __syncthreads();
    if (!active) { if (tid == 0 && a.out_tokens) a.out_tokens[b] = -1; return; }

    // ---- token heuristics of alignment_stream_analyzer.py:143-195, run in the loop: the call that shapes the logits of token i
    // (i = tokens generated so far) is the analyzer's (i+1)-th step, after it has seen tokens 0..i-1:
    //   curr_frame_pos = i + 1; text_position = min(frame // 2, n_text - 1); complete from the first frame with
    //   text_position >= n_text - 3 (a pure function of n_text: frame max(1, 2 (n_text - 3)));
    //   3 equal tokens in a row or >= 10 frames after completion => every logit -2^15 except EOS = +2^15 (the EOS is then the
    //   only possible outcome of greedy and of sampling); else EOS = -2^15 while text_position < n_text - 3 and n_text > 5.
    const int stop_tok = (a.has_state && a.st.stop_tokens) ? a.st.stop_tokens[b] : -1;
    bool force_eos = false;
    if (a.sp.analyzer_text_tokens != nullptr && a.has_state && a.st.hist_len != nullptr && stop_tok >= 0 && stop_tok < V) {
        const int n_text = a.sp.analyzer_text_tokens[b];
        if (n_text > 0) {
            const int i = a.st.hist_len[b], frame = i + 1;
            bool rep = false;
            if (i >= 3 && i <= a.st.hist_stride) {
                const int32_t* hh = a.st.history + (size_t)b * a.st.hist_stride;
                rep = hh[i - 1] == hh[i - 2] && hh[i - 2] == hh[i - 3];
            }
            const int f_star = max(1, 2 * (n_text - 3));
            const bool long_tail = frame - f_star >= 10;
            const int text_pos = min(frame / 2, n_text - 1);
            force_eos = rep || long_tail;
            if (!force_eos && text_pos < n_text - 3 && n_text > 5) {
                if (tid == 0) lg[stop_tok] = -32768.f;
                __syncthreads();
            }
        }
    }

    // ---- S2: repetition penalty over the generated tokens (utils.py:51-89) -----------------------
    const float rho = a.sp.repetition_penalty ? a.sp.repetition_penalty[b] : 1.0f;
    if (a.has_state && a.st.history != nullptr && rho != 1.0f) {
        int hl = a.st.hist_len[b];
        hl = min(hl, a.st.hist_stride);
        for (int i = tid; i < hl; i += SMP_THREADS) {
            const int t = a.st.history[(size_t)b * a.st.hist_stride + i];
            if (t >= 0 && t < V) atomicOr(&seen[t >> 5], 1u << (t & 31));
        }
        __syncthreads();
        for (int i = tid; i < V; i += SMP_THREADS)
            if (seen[i >> 5] & (1u << (i & 31))) { const float l = lg[i]; lg[i] = l > 0.f ? l / rho : l * rho; }
        __syncthreads();
    }

    const float temp = a.sp.temperature ? a.sp.temperature[b] : 0.f;
    int token;
    if (force_eos) {
        token = stop_tok;
    } else if (temp < SMP_EPS) {
        // ---- S3 greedy: argmax, first index on ties --------------------------------------------
        float bv = -INFINITY; int bi = 0x7fffffff;
        for (int i = tid; i < V; i += SMP_THREADS) { const float l = lg[i]; if (l > bv) { bv = l; bi = i; } }
        token = block_argmax(bv, bi, redf, redi);
    } else {
        // ---- S3 temperature / min-p / top-p ----------------------------------------------------
        float mx = -INFINITY;
        for (int i = tid; i < V; i += SMP_THREADS) { const float l = lg[i] / temp; lg[i] = l; mx = fmaxf(mx, l); }
        mx = block_max(mx, redf);
        // lg <- e_i = exp(l_i - max) (softmax numerators); masked entries become 0
        unsigned long long zs = 0;
        for (int i = tid; i < V; i += SMP_THREADS) { const float e = expf(lg[i] - mx); lg[i] = e; zs += to_fixed(e); }
        unsigned long long Z = block_sum_u64(zs, red64);
        const float min_p = a.sp.min_p ? a.sp.min_p[b] : 0.f;
        if (min_p > 0.f) {
            // p_i < min_p * p_max  with p = e / Z, p_max = 1 / Z  (builtin.py:100-114)
            const float Zf = (float)((double)Z * (1.0 / 1099511627776.0));
            const float thr = (1.0f / Zf) * min_p;
            zs = 0;
            for (int i = tid; i < V; i += SMP_THREADS) {
                float e = lg[i];
                if (e / Zf < thr) { e = 0.f; lg[i] = 0.f; }
                zs += to_fixed(e);
            }
            Z = block_sum_u64(zs, red64);
        }
        const float top_p = a.sp.top_p ? a.sp.top_p[b] : 1.f;
        if (top_p < 1.f) {
            // ascending cumulative mass <= (1 - top_p) is dropped (topk_topp_sampler.py:356-363).
            // budget in fixed point; find the key T with mass{key < T} <= budget < mass{key <= T}.
            // top_p is validated on the host to (0, 1]; the clamp keeps the crossing bin defined for any input
            unsigned long long budget = (unsigned long long)((double)fmaxf(1.0f - top_p, 0.f) * (double)Z);
            if (budget >= Z) budget = Z > 0 ? Z - 1 : 0;
            uint32_t prefix = 0;
            if (tid == 0) { s_below = 0; s_digit = 0; }
            for (int pass = 0; pass < 4; ++pass) {
                const int shift = 24 - 8 * pass;
                hist[tid] = 0; hcnt[tid] = 0;          // SMP_THREADS == 256 bins
                __syncthreads();
                for (int i = tid; i < V; i += SMP_THREADS) {
                    const float e = lg[i];
                    if (e > 0.f) {
                        const uint32_t key = __float_as_uint(e);
                        if (pass == 0 || (key >> (shift + 8)) == (prefix >> (shift + 8))) {
                            atomicAdd(&hist[(key >> shift) & 255u], to_fixed(e));
                            atomicAdd(&hcnt[(key >> shift) & 255u], 1u);
                        }
                    }
                }
                __syncthreads();
                // inclusive scan of the 256 bin masses (one bin per thread)
                unsigned long long v = hist[tid];
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const unsigned long long n = __shfl_up_sync(0xffffffffu, v, o);
                    if ((tid & 31) >= o) v += n;
                }
                if ((tid & 31) == 31) red64[tid >> 5] = v;
                __syncthreads();
                unsigned long long off = 0;
                for (int w = 0; w < (tid >> 5); ++w) off += red64[w];
                const unsigned long long incl = v + off, excl = incl - hist[tid];
                const unsigned long long below = s_below;
                __syncthreads();
                // the crossing bin: first non-empty bin whose inclusive mass exceeds the budget
                const bool cross = hcnt[tid] > 0 && below + excl <= budget && below + incl > budget;
                if (cross) { s_digit = (unsigned)tid; s_below = below + excl; }
                __syncthreads();
                prefix |= s_digit << shift;
            }
            const uint32_t T = prefix;
            const float eT = __uint_as_float(T);
            const unsigned long long fT = to_fixed(eT);
            // ties at T: the first m (index order) still fit in the budget and are dropped
            unsigned long long m_drop = fT > 0 ? (budget - s_below) / fT : 0ull;
            // rank of each tied element in index order: chunked exclusive count
            unsigned int cnt = 0;
            const int per = (V + SMP_THREADS - 1) / SMP_THREADS;
            const int i0 = tid * per, i1 = min(V, i0 + per);
            for (int i = i0; i < i1; ++i) cnt += (lg[i] > 0.f && __float_as_uint(lg[i]) == T) ? 1u : 0u;
            unsigned int incl = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned int n = __shfl_up_sync(0xffffffffu, incl, o);
                if ((tid & 31) >= o) incl += n;
            }
            __syncthreads();
            if ((tid & 31) == 31) redi[tid >> 5] = (int)incl;
            __syncthreads();
            unsigned int off = 0;
            for (int w = 0; w < (tid >> 5); ++w) off += (unsigned)redi[w];
            unsigned int rank = off + incl - cnt;
            unsigned int n_tied = 0;
            for (int w = 0; w < SMP_THREADS / 32; ++w) n_tied += (unsigned)redi[w];
            if (m_drop >= n_tied) m_drop = n_tied > 0 ? n_tied - 1 : 0;   // "at least one" (:362): the last sorted entry stays
            __syncthreads();
            for (int i = i0; i < i1; ++i) {
                const float e = lg[i];
                if (e > 0.f) {
                    const uint32_t key = __float_as_uint(e);
                    if (key < T) lg[i] = 0.f;
                    else if (key == T) { if (rank < m_drop) lg[i] = 0.f; ++rank; }
                }
            }
            __syncthreads();
            zs = 0;
            for (int i = tid; i < V; i += SMP_THREADS) zs += to_fixed(lg[i]);
            Z = block_sum_u64(zs, red64);
        }
        // ---- S4: exponential race (topk_topp_sampler.py:395-416) --------------------------------
        const float Zf = (float)((double)Z * (1.0 / 1099511627776.0));
        const unsigned long long seed = a.sp.seed ? a.sp.seed[b] : 0ull;
        const uint32_t step = (a.has_state && a.st.step) ? (uint32_t)a.st.step[b] : 0u;
        float bv = -INFINITY; int bi = 0x7fffffff;
        for (int i = tid; i < V; i += SMP_THREADS) {
            const float e = lg[i];
            if (e > 0.f) {
                const uint32_t x = philox_x0((uint32_t)i, step, (uint32_t)seed, (uint32_t)(seed >> 32));
                const float u = ((float)(x >> 9) + 0.5f) * 1.1920928955078125e-07f;   // 2^-23
                const float q = -logf(u);
                const float sc = (e / Zf) / q;
                if (sc > bv) { bv = sc; bi = i; }
            }
        }
        token = block_argmax(bv, bi, redf, redi);
    }

    if (tid == 0) {
        if (a.out_tokens) a.out_tokens[b] = token;
        if (a.has_state && a.advance) {
            const t3b_state_t& st = a.st;
            st.tokens[b] = token;
            if (st.history && st.hist_len[b] < st.hist_stride) st.history[(size_t)b * st.hist_stride + st.hist_len[b]] = token;
            if (st.hist_len) st.hist_len[b] += 1;
            if (st.step) st.step[b] += 1;
            st.speech_pos[b] += 1;
            for (int r = 2 * b; r < 2 * b + 2; ++r)
                if (st.ctx_lens[r] + 1 < st.max_ctx) { st.ctx_lens[r] += 1; st.rope_pos[r] += 1; }
            if (stop_tok >= 0 && token == stop_tok && st.active) st.active[b] = 0;
        }
    }
}

int32_t t3b_sampler_launch(const void* head, const float* head_partials, int n_split, const float* logits_in, int V, int n_req,
                           const t3b_sampling_t* sp, const t3b_state_t* st, int advance, int32_t* out_tokens, float* logits_out,
                           cudaStream_t stream) {
    T3B_CHECK_SHAPE(V >= 2 && V <= SMP_MAX_V, "sampler: vocab %d not in [2, %d]", V, SMP_MAX_V);
    T3B_CHECK_ARG(sp != nullptr, "sampler: sampling params are null");
    T3B_CHECK_ARG(!advance || st != nullptr, "sampler: advance needs a state");
    if (st && advance)
        T3B_CHECK_ARG(st->tokens && st->speech_pos && st->rope_pos && st->ctx_lens, "sampler: state arrays are null");
    if (n_req == 0) return T3B_OK;
    SamplerArgs a;
    a.head = (const bf16*)head; a.head_partials = head_partials; a.n_split = n_split; a.logits_in = logits_in;
    a.V = V; a.n_req = n_req; a.sp = *sp; a.has_state = st != nullptr; a.advance = advance;
    if (st) a.st = *st; else memset(&a.st, 0, sizeof(a.st));
    a.out_tokens = out_tokens; a.logits_out = logits_out;
    T3B_CUDA(t3b_launch(cfg_sample_kernel, dim3(n_req), dim3(SMP_THREADS), 0, stream, a));
    return T3B_OK;
}
