/*
 * t3b200.h - C ABI of libt3b200.so: the B200 (sm_100a) implementation of the Chatterbox T3
 * CFG decode step.
 *
 * Nothing comparable exists in the reference (groxaxo/chatterbox-vllm2 is pure Python and all of
 * its arithmetic runs inside the vllm==0.10.0 wheel); every entry point below cites the reference
 * interface (file:line, relative to the reference repo; "vllm:" = the engine it drives) whose
 * arithmetic it replaces.  INTEGRATION.md shows the ctypes binding a maintainer adds on the
 * reference side.
 *
 * Conventions
 *   - every function returns int32_t: 0 = T3B_OK, <0 = T3B_ERR_*; t3b_last_error() gives the
 *     thread-local message of the last failure on this host thread.
 *   - all data pointers are DEVICE pointers (tensor.data_ptr()) unless the name says host_*;
 *     the last argument is the cudaStream_t to launch on (passed as void*).
 *   - no allocation, no ownership transfer, no host synchronisation inside any entry point;
 *     scratch comes from the caller (t3b_workspace_bytes()).  All launches are CUDA-graph
 *     capturable.
 *   - bf16 = 16-bit bfloat16, row-major.  R = CFG rows (row 2i = cond, 2i+1 = uncond of request i;
 *     t3.py:697-708 batches the two halves the same way), B = R/2 requests.
 *   - thread model: one host thread per GPU context; re-entrant across streams given disjoint
 *     workspaces.
 */
#ifndef T3B200_H
#define T3B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define T3B_OK               0
#define T3B_ERR_ARG         -1   /* null / out-of-range argument */
#define T3B_ERR_SHAPE       -2   /* shape the kernels do not support */
#define T3B_ERR_ALIGN       -3   /* pointer not 16-byte aligned */
#define T3B_ERR_CUDA        -4   /* a CUDA runtime / driver call failed */
#define T3B_ERR_UNSUPPORTED -5   /* device is not sm_100 */

#define T3B_KV_BLOCK 16          /* tokens per paged-KV block (BASELINE.json configs[3]) */
#define T3B_HEAD_DIM 64

/* linear-layer weight formats */
#define T3B_W_BF16 0             /* w = bf16 [N, K] row-major (nn.Linear layout) */
#define T3B_W_AWQ  1             /* AutoAWQ GEMM, group 128, repacked by t3b_repack_weights_awq */

/* kernel selection for t3b_linear_* (0 lets the library choose by row count) */
#define T3B_IMPL_AUTO 0
#define T3B_IMPL_GEMV 1          /* SIMT 128-bit-load GEMV, rows <= 4, bf16 weights  */
#define T3B_IMPL_UMMA 2          /* TMA-fed tcgen05.mma, accumulators in TMEM        */

/* fused epilogues of t3b_linear_* */
#define T3B_EPI_NONE     0       /* y[rows, n] = bf16(x W^T)                                             */
#define T3B_EPI_SILU_MUL 1       /* w = gate_up [2*ffn, k]: y[rows, ffn] = silu(gate) * up (llama.py:117-121) */

/* paged-KV element formats (t3b_kv_t.dtype) */
#define T3B_KV_BF16 0            /* (block, head) slice = 16 tok x 64 bf16 = 2048 B                        */
#define T3B_KV_FP8  1            /* slice = 16 tok x 64 e4m3 (1024 B) | 16 fp32 scales, one per (token, head): value = e4m3 * scale,
                                    scale = max|x| / 448 over the head's 64 dims (1 if the vector is zero) = 1088 B         */

/* t3b_linear_t.flags */
#define T3B_LIN_GATE_UP_INTERLEAVED 1   /* AWQ tiles repacked as 64 gate rows + 64 up rows (SILU_MUL epilogue only) */

/* CFG combine arithmetic (t3.py:662 runs in bf16 tensors; SURVEY.md 8(a) also defines fp32) */
#define T3B_COMBINE_BF16 0
#define T3B_COMBINE_FP32 1

typedef struct t3b_linear {
    int32_t kind;                /* T3B_W_* */
    int32_t n, k;                /* out features, in features */
    const void* w;               /* bf16 [n,k]  |  AWQ repacked nibbles */
    const void* scales;          /* AWQ: bf16 [k/128, n].  bf16: NULL, or the copy of w in the fragment order of the tensor-core GEMV
                                    (t3b_repack_weights_gemv) - with it and t3b_options_t.gemv_mma, <= 16 rows run that kernel   */
    const void* zeros;           /* AWQ: uint8 [k/128, n]         (else NULL) */
    int32_t flags;               /* T3B_LIN_* */
} t3b_linear_t;

typedef struct t3b_layer {       /* one LlamaDecoderLayer (vllm:model_executor/models/llama.py:316-333) */
    t3b_linear_t qkv;            /* [3*hidden, hidden]  q|k|v rows stacked (llama.py:437-443) */
    t3b_linear_t o;              /* [hidden, hidden] */
    t3b_linear_t gate_up;        /* [2*ffn, hidden]     gate rows first */
    t3b_linear_t down;           /* [hidden, ffn] */
    const void* ln1;             /* bf16 [hidden] input_layernorm.weight */
    const void* ln2;             /* bf16 [hidden] post_attention_layernorm.weight */
} t3b_layer_t;

typedef struct t3b_model {
    int32_t n_layers, hidden, n_heads, head_dim, ffn, speech_vocab;
    int32_t speech_pos_rows;     /* rows of speech_pos_emb (4100, t3.py:283) */
    int32_t max_pos;             /* rows of rope_cos_sin */
    float   rms_eps;
    const t3b_layer_t* layers;   /* HOST array [n_layers] of structs holding device pointers */
    const void* final_norm;      /* bf16 [hidden]  tfmr.norm.weight */
    t3b_linear_t head;           /* speech_head [speech_vocab, hidden], bf16 only */
    const void* speech_emb;      /* bf16 [speech_vocab, hidden] */
    const void* speech_pos_emb;  /* bf16 [speech_pos_rows, hidden] */
    const void* rope_cos_sin;    /* bf16 [max_pos, head_dim]  cos(32) | sin(32) */
} t3b_model_t;

/* Paged KV pool.  Layout per layer: [num_blocks][n_heads] slices of T3B_KV_BLOCK tokens x head_dim (T3B_KV_BF16:
 * [16][64] bf16; T3B_KV_FP8: see above), K and V separate.  (vLLM's flash layout is [block][token][head][dim]; heads are
 * hoisted here so one (block, head) slice is one contiguous bulk copy.)  Block 0 is reserved as a scratch block for idle
 * rows.  Several rows may name the same block in their tables (shared conditioning prefix, t3.py:554-556). */
typedef struct t3b_kv {
    void* k; void* v;            /* [n_layers][num_blocks][n_heads][slice] */
    int64_t layer_stride;        /* BYTES between layers = num_blocks * n_heads * slice bytes */
    int32_t num_blocks;
    int32_t dtype;               /* T3B_KV_* */
} t3b_kv_t;

/* Per-step device-resident request state; the sampler advances it so a captured graph can be
 * replayed with no host writes between steps. */
typedef struct t3b_state {
    int32_t* tokens;             /* [B]  in: last sampled speech token (un-offset); out: new token */
    int32_t* speech_pos;         /* [B]  index into speech_pos_emb for `tokens` (t3.py:468-478)   */
    int32_t* rope_pos;           /* [R]  RoPE position of the token being fed (= ctx_lens)        */
    int32_t* ctx_lens;           /* [R]  tokens already in the row's KV cache                     */
    const int32_t* block_tables; /* [R, max_blocks] physical block ids                            */
    int32_t  max_blocks;
    int32_t* history;            /* [B, hist_stride] generated tokens (penalty history, S2)       */
    int32_t* hist_len;           /* [B] */
    int32_t  hist_stride;
    int32_t* step;               /* [B]  sampler step counter (Philox counter word 1)             */
    int32_t* active;             /* [B]  1 = decoding, 0 = idle slot / finished                   */
    const int32_t* stop_tokens;  /* [B]  per request (tts.py:457 stop_token_ids[0] - 2500 = 6562, t3_config.py:9); sampling it
                                         clears active[b]; < 0 or NULL = never stops (ignore_eos)  */
    int32_t  max_ctx;            /* capacity guard: a row never advances past this many tokens     */
    const int32_t* row_order;    /* [R] or NULL: a permutation of the step's rows, the order in which the decode attention deals
                                    (row, head) items to its warps.  Longest context first balances ragged batches (1024 rows of
                                    100..1100 tokens: 5.5 -> 6.4 TB/s); a scheduling hint only - results do not depend on it   */
} t3b_state_t;

typedef struct t3b_sampling {    /* per-request device arrays [B]; tts.py:455-464 SamplingParams */
    const float* cfg_scale;      /* CHATTERBOX_CFG_SCALE (t3.py:296), per request here             */
    const float* repetition_penalty;
    const float* temperature;    /* < 1e-5 = greedy (vllm:v1/sample/sampler.py _SAMPLING_EPS)      */
    const float* min_p;
    const float* top_p;
    const uint64_t* seed;
    const int32_t* analyzer_text_tokens;  /* [B] or NULL: > 0 runs the token heuristics of alignment_stream_analyzer.py:143-195
                                             in the loop with text_tokens_count = this value (early-EOS suppression, 3x repeat
                                             => EOS, long tail => EOS); <= 0 = off                    */
    int32_t combine;             /* T3B_COMBINE_* */
} t3b_sampling_t;

/* Launch-shape knobs of the decode path (NULL / 0 fields = the library's measured defaults).  Everything that used to be an
 * environment variable inside the library is an explicit argument here. */
typedef struct t3b_options {
    int32_t gemv_max_rows;       /* bf16 projections use the GEMV kernels up to this many rows (0 = default 2, max 4 - 16 with gemv_mma; -1 = never) */
    int32_t gemv_no_norm;        /* 1 = keep add+RMSNorm as separate launches at GEMV row counts (default: folded in)     */
    int32_t umma_stages;         /* tcgen05 GEMM: shared-memory ring depth cap (0 = default 5)                             */
    int32_t umma_row_chunk;      /* tcgen05 GEMM: CFG rows per thread block, multiple of 16 (0 = default)                  */
    int32_t umma_split[4];       /* cluster split-K of the qkv / o / gate_up / down projections (0 = automatic)            */
    int32_t attn_team_warps;     /* decode attention: warps that share one (row, head) item: 1, 2, 4 or 8 (0 = by item count) */
    int32_t weight_prefetch;     /* 1 = every launch of t3b_decode_step pulls the weights of the projection two launches ahead into
                                    L2; 2 (GEMV chains only) = the same projection of the NEXT layer, a whole layer ahead.  No gain where
                                    the chain is latency-bound (tcgen05 / SIMT chains, profiles/r2_notes.md); the tensor-core GEMV chain
                                    gains 3-5 % (what an engine of <= 4 requests sets).  Default off                                */
    int32_t umma_big_min_rows;   /* bf16 projections without split-K run the persistent double-buffered tcgen05 kernel (csrc/linear_big.cu)
                                    from this many rows on (0 = default 1024: prefill passes and decode batches of >= 512 requests; < 0 = never)                       */
    int32_t norm_fusion;         /* t3b_decode_step can fold add+RMSNorm into the tcgen05 projections either side of it (5 launches per layer
                                    instead of 7; up to 32 rows).  0 = where it is measured faster (9..32 rows), 1 = wherever the shapes allow
                                    (3..32 rows: what an engine picks so that every batch size it runs uses the same arithmetic), -1 = never.
                                    The folded chain adds a row's sum of squares in a different order than the norm kernel: results agree
                                    to a bf16 step, not bit for bit                                                                */
    int32_t gemv_mma;            /* 1 = the GEMV row counts run on the tensor cores (mma.sync m16n8k16 on weight fragments loaded straight
                                    from global memory, csrc/linear_gemv.cu) for every bf16 projection that carries its fragment-ordered
                                    copy (t3b_linear_t.scales): up to 16 rows (gemv_max_rows may then be up to 16), K = 1024 or 4096.  fp32 accumulation in another order than the SIMT GEMV: results agree to a bf16 step.
                                    0 = library default, -1 = the SIMT kernel                                                        */
} t3b_options_t;

const char* t3b_last_error(void);
int32_t     t3b_version(void);
/* 0 if the current device is sm_100 and the driver entry points resolve, else T3B_ERR_UNSUPPORTED */
int32_t     t3b_check_device(void);
/* scratch needed by t3b_decode_step / t3b_forward_tokens / t3b_head_cfg_sample for up to max_rows rows (tokens) */
size_t      t3b_workspace_bytes(const t3b_model_t* m, int32_t max_rows);
/* bytes of one (block, head) K or V slice for a T3B_KV_* format (0 = unknown format) */
int32_t     t3b_kv_slice_bytes(int32_t kv_dtype);

/* E1  t3.py:440,475-480 (+ learned_pos_emb.py:21-32): out[dup*b + j] = emb[tokens[b]] +
 *     pos_emb[speech_pos[b] % pos_rows], j < dup.  out bf16 [n_req*dup, hidden]. */
int32_t t3b_embed_gather(const void* emb, const void* pos_emb, const int32_t* tokens,
                         const int32_t* speech_pos, int32_t n_req, int32_t hidden, int32_t emb_rows,
                         int32_t pos_rows, int32_t dup, void* out, void* stream);

/* F2.1 vllm:ir/ops/layernorm.py:9-21,40-59 (RMSNorm / fused_add_rms_norm of llama.py:311-331).
 *     x comes either as bf16 [rows,hidden] (x != NULL) or as n_split fp32 split-K partials
 *     [n_split, rows, hidden] that are summed in order and rounded to bf16 first.
 *     has_residual = 0: residual := x (first layer).  residual is updated in place. */
int32_t t3b_add_rmsnorm(const void* x, const float* partials, int32_t n_split, void* residual,
                        const void* weight, void* y, int32_t rows, int32_t hidden, float eps,
                        int32_t has_residual, void* stream);

/* F2.2 / F2.5 / F2.6 / C1  the projections (llama.py:164-180,228-232,117-121; t3.py:659-660):
 *     y[r, n] = sum_k x[r,k] * w[n,k], fp32 accumulate, bf16 output [rows, n] (T3B_EPI_NONE) or the fused
 *     SiLU*mul of the gate_up projection [rows, n/2] (T3B_EPI_SILU_MUL, activation.py:138-141).
 *     impl: T3B_IMPL_AUTO picks the SIMT GEMV for rows <= opt->gemv_max_rows (default 2) and the tcgen05 kernel above.
 *     Split-K: n_split <= 0 lets the library pick (returned through *n_split_out, may be NULL); the
 *     partial tiles are reduced on-chip through the thread-block cluster's distributed shared memory in a
 *     fixed order (deterministic) and y is written.  Only when partials != NULL AND n_split > 1 is given
 *     explicitly are the fp32 partials written to global memory instead ([n_split, rows, n], y untouched)
 *     for a consumer that sums them (t3b_add_rmsnorm / t3b_rope_kv_write / t3b_silu_mul accept them).
 *     opt may be NULL. */
int32_t t3b_linear_bf16(const void* x, const t3b_linear_t* w, int32_t rows, void* y, float* partials,
                        int32_t n_split, int32_t* n_split_out, int32_t impl, int32_t epilogue,
                        const t3b_options_t* opt, void* stream);

/* Q   vllm:model_executor/layers/quantization/awq.py:214-283: same contract as t3b_linear_bf16 with
 *     W[k,n] = bf16((nib - zero) * scale), group 128, dequantised on the fly into the tcgen05 operand tile
 *     (every row count runs T3B_IMPL_UMMA; activations are bf16, scales are converted fp16 -> bf16 by
 *     t3b_repack_weights_awq). */
int32_t t3b_linear_awq(const void* x, const t3b_linear_t* w, int32_t rows, void* y, float* partials,
                       int32_t n_split, int32_t* n_split_out, int32_t impl, int32_t epilogue,
                       const t3b_options_t* opt, void* stream);

/* one-off layout transform of AutoAWQ tensors (qweight int32 [k, n/8], qzeros int32 [k/128, n/8],
 * scales fp16 [k/128, n]; nibble order 0,2,4,6,1,3,5,7 - awq_triton.py:56-62) into the kernel
 * layout: w_out uint8 [n/128][k/64][2][128][16] (tile, k-block, k half, feature row, 32 nibbles; inside
 * a 32-bit word the nibble positions hold k offsets 0,2,4,6,1,3,5,7), zeros_out uint8 [k/128, n],
 * scales_out bf16 [k/128, n].  gate_up_interleave != 0: tile t = gate rows [64t, 64t+64) then up rows
 * [n/2 + 64t, ...) (set T3B_LIN_GATE_UP_INTERLEAVED in the t3b_linear_t; needs T3B_EPI_SILU_MUL). */
int32_t t3b_repack_weights_awq(const int32_t* qweight, const int32_t* qzeros, const void* scales_fp16,
                               int32_t n, int32_t k, int32_t gate_up_interleave, void* w_out, void* zeros_out,
                               void* scales_out, void* stream);

/* Small-batch layout of a bf16 projection (no counterpart in the reference: vLLM keeps nn.Linear layout): w bf16 [n, k] -> out,
 * t3b_gemv_frag_bytes(n, k, gate_up) bytes (0 = shape not supported: k must be 1024 or 4096, gate_up needs k = 1024 and n % 32 == 0),
 * in the order the tensor-core GEMV reads its mma.sync A fragments ([16-feature tile][k step of 32][lo | hi][lane] x 16 bytes;
 * gate_up != 0: a tile = 8 gate features + the 8 matching up features, for the T3B_EPI_SILU_MUL epilogue).  One-off, at load; the
 * pointer goes into t3b_linear_t.scales of the bf16 projection. */
size_t  t3b_gemv_frag_bytes(int32_t n, int32_t k, int32_t gate_up);
int32_t t3b_repack_weights_gemv(const void* w, int32_t n, int32_t k, int32_t gate_up, void* out, void* stream);

/* F2.3 + KV write  (rotary_embedding/common.py:144-174, llama3 table; reshape_and_cache of
 *     llama.py:231): qkv as bf16 [n_tok, 3*hidden] or split-K partials.  Token j belongs to KV row
 *     tok_row[j] (NULL = j), sits at index tok_idx[j] of that row's sequence and is rotated by
 *     rope_pos[j].  Writes q (rotated) bf16 [n_tok, hidden] and k (rotated), v into the pool (kv_dtype = T3B_KV_*). */
int32_t t3b_rope_kv_write(const void* qkv, const float* partials, int32_t n_split, int32_t n_tok,
                          const int32_t* tok_row, const int32_t* tok_idx, const int32_t* rope_pos,
                          const void* rope_cos_sin, int32_t max_pos, const int32_t* block_tables,
                          int32_t max_blocks, void* k_pool, void* v_pool, int32_t n_heads, int32_t kv_dtype,
                          void* q_out, void* stream);

/* F2.4 paged decode attention (llama.py:211-221,231): o[j] = softmax(q[j] K^T / 8) V over the
 *     first tok_idx[j]+1 tokens of row tok_row[j], per head.  out bf16 [n_tok, hidden].  opt (may be NULL): attn_team_warps. */
int32_t t3b_paged_attn_decode(const void* q, int32_t n_tok, const int32_t* tok_row, const int32_t* tok_idx,
                              const int32_t* block_tables, int32_t max_blocks, const void* k_pool,
                              const void* v_pool, int32_t n_heads, int32_t kv_dtype, void* out, const t3b_options_t* opt,
                              void* stream);

/* F2.3 + KV write + F2.4 in one launch for a decode step (token j <-> KV row j): reads q/k/v of every head
 *     from qkv bf16 [n_tok, 3*hidden], rotates q and k by rope_pos[j], stores k (rotated) and v at index
 *     tok_idx[j] of row j and attends over tokens 0..tok_idx[j].  out bf16 [n_tok, hidden].
 *     Persistent kernel: K/V slices staged through shared memory by bulk copies (csrc/attention.cu). */
int32_t t3b_attn_decode_fused(const void* qkv, int32_t n_tok, const int32_t* tok_idx, const int32_t* rope_pos,
                              const void* rope_cos_sin, int32_t max_pos, const int32_t* block_tables,
                              int32_t max_blocks, void* k_pool, void* v_pool, int32_t n_heads, int32_t kv_dtype,
                              void* out, const t3b_options_t* opt, void* stream);

/* Prefill attention (the causal pass of llama.py:231 over whole prompts, t3.py:542-561): sequence s is the tokens
 *     [seq_start[s], seq_start[s+1]) of q (bf16 [n_tok, hidden], already rotated), belongs to KV row seq_row[s] and starts
 *     at index 0 of that row; its K/V must already be in the pool (t3b_rope_kv_write).  Every (sequence, head) reads the
 *     prompt's K/V from the pool ONCE into shared memory and serves all of its queries from there:
 *     out[j] = softmax(q[j] K[0..j]^T / 8) V[0..j].  out bf16 [n_tok, hidden].  Prompts up to 432 tokens (bf16 KV) /
 *     800 tokens (fp8 KV); longer prompts run t3b_paged_attn_decode per token. */
int32_t t3b_prefill_attn(const void* q, int32_t n_seq, const int32_t* seq_start, const int32_t* seq_row,
                         const int32_t* block_tables, int32_t max_blocks, const void* k_pool, const void* v_pool,
                         int32_t n_heads, int32_t kv_dtype, int32_t max_seq_len, void* out, void* stream);
int32_t t3b_prefill_attn_max_len(int32_t kv_dtype);

/* Prefill embedding assembly (t3.py:542-561) for n_req requests in one launch:
 *     cond row   = [cond_emb(34) | text_emb[id] + text_pos_emb[i] | speech_emb[6561] + speech_pos_emb[0]]
 *     uncond row = [cond_emb(34) | zeros                          | the same BOS embedding]
 *     conds bf16 [n_req, 34, hidden]; text ids of request r are text_ids[text_start[r] .. text_start[r+1]).
 *     Request r's cond sequence starts at token out_start[2r], its uncond sequence at out_start[2r+1]
 *     (out_start int32 [2 n_req + 1], the caller's prefix sums of P_r = 34 + n_text_r + 1 taken twice).
 *     Writes x bf16 [n_tok, hidden] and the per-token index arrays tok_row[j] = 2 * slot[r] + half, tok_idx[j] = position. */
int32_t t3b_prefill_embed(const void* conds, const int32_t* text_ids, const int32_t* text_start, const int32_t* out_start,
                          const int32_t* slots, int32_t n_req, const void* text_emb, const void* text_pos_emb,
                          int32_t text_rows, int32_t text_pos_rows, const void* speech_emb, const void* speech_pos_emb,
                          int32_t bos_token, int32_t hidden, int32_t cond_len, void* x_out, int32_t* tok_row,
                          int32_t* tok_idx, void* stream);

/* F2.6 activation (vllm:model_executor/layers/activation.py:138-141): act = silu(g) * u with the
 *     bf16 roundings of the reference; gate_up as bf16 [rows, 2*ffn] or split-K partials. */
int32_t t3b_silu_mul(const void* gate_up, const float* partials, int32_t n_split, int32_t rows, int32_t ffn,
                     void* act, void* stream);

/* F2   LlamaModel.forward (llama.py:395-432) over n_tok tokens given as embeddings x_in (bf16
 *     [n_tok, hidden]): token j belongs to KV row tok_row[j], is the tok_idx[j]-th token of
 *     that row (its K/V are written there and it attends causally to tokens 0..tok_idx[j]) and is
 *     rotated by rope_pos[j].  One call with every prompt token of the new rows is the prefill of
 *     t3.py:542-561.  seq_start / seq_row (n_seq sequences, see t3b_prefill_attn; NULL / 0 = none) let the
 *     attention read each prompt's K/V once; without them every token runs the paged decode kernel.
 *     hidden_out bf16 [n_tok, hidden].  Launch shapes are fixed (no split-K), so a prompt's result does not
 *     depend on the prompts it is batched with. */
int32_t t3b_forward_tokens(const t3b_model_t* m, const t3b_kv_t* kv, const void* x_in, int32_t n_tok,
                           const int32_t* tok_row, const int32_t* tok_idx, const int32_t* rope_pos,
                           const int32_t* block_tables, int32_t max_blocks, int32_t n_seq, const int32_t* seq_start,
                           const int32_t* seq_row, int32_t max_seq_len, void* hidden_out, void* workspace,
                           size_t workspace_bytes, const t3b_options_t* opt, void* stream);

/* F1 + F2  T3VllmModel.forward (t3.py:676-713) over LlamaModel.forward (llama.py:395-432) for one
 *     decode token per CFG row.  x_in (bf16 [rows, hidden]) != NULL feeds given embeddings (prefill-
 *     by-decode, t3.py:542-561 rows); x_in == NULL gathers E1 from st->tokens/speech_pos.  Uses
 *     st->rope_pos / ctx_lens / block_tables.  hidden_out bf16 [rows, hidden] = final-norm output.
 *     opt (may be NULL) = launch-shape knobs. */
int32_t t3b_decode_step(const t3b_model_t* m, const t3b_kv_t* kv, const t3b_state_t* st, const void* x_in,
                        int32_t rows, void* hidden_out, void* workspace, size_t workspace_bytes,
                        const t3b_options_t* opt, void* stream);

/* C1 + S1..S4  compute_logits (t3.py:650-673) fused with the sampler the engine runs next
 *     (vllm:v1/sample/sampler.py:67-307, ops/topk_topp_sampler.py:328-416, layers/utils.py:51-89):
 *     head on both rows, l = c + w (c - u), the analyzer's token heuristics when sp->analyzer_text_tokens
 *     is set (alignment_stream_analyzer.py:143-195), repetition penalty over st->history, greedy or
 *     temperature/min-p/top-p + exponential race (Philox4x32-10 keyed by seed, counter (index, step)).
 *     Writes out_tokens[b] (un-offset id; caller adds SPEECH_TOKEN_OFFSET = 2500, t3.py:666-672),
 *     optionally the combined fp32 logits [B, speech_vocab], and - if advance != 0 - appends to
 *     history and advances tokens/speech_pos/rope_pos/ctx_lens/step, clearing active[b] on stop. */
int32_t t3b_head_cfg_sample(const t3b_model_t* m, const void* hidden, int32_t rows, const t3b_sampling_t* sp,
                            const t3b_state_t* st, int32_t advance, int32_t* out_tokens, float* logits_out,
                            void* workspace, size_t workspace_bytes, const t3b_options_t* opt, void* stream);

/* t3b_head_cfg_sample without any allocation by anyone.  NO COUNTERPART IN THE REFERENCE: added
 *     for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 *     forbid it).  Same arguments, outputs and state advance as t3b_head_cfg_sample, minus the workspace: the head logits
 *     (bf16 [rows, speech_vocab], the scratch between the head projection and the sampler) live in a file-scope
 *     `static __device__` array inside the library, T3B_STATIC_MAX_ROWS x T3B_STATIC_MAX_VOCAB bf16 + 256 B (~1.03 MiB), which
 *     every GPU that loads the library holds until the process exits.  The call makes no cudaMalloc / cudaFree and calls no
 *     allocating library routine; the array's device address is looked up once per device (cudaGetSymbolAddress, with that
 *     device current - make the first call per device outside graph capture).
 *     Limits: rows <= T3B_STATIC_MAX_ROWS (32 requests) and m->speech_vocab <= T3B_STATIC_MAX_VOCAB, else T3B_ERR_ARG with a
 *     message naming the limit.  ALL calls on a device share the one array: it keeps the previous call's logits (nothing needs
 *     to start at zero - the projection overwrites every element the sampler reads), and two calls that overlap on the GPU
 *     (different streams, or graph branches without an edge between them) RACE; serialise them or use t3b_head_cfg_sample
 *     with separate workspaces. */
#define T3B_STATIC_MAX_ROWS  64
#define T3B_STATIC_MAX_VOCAB 8448
int32_t t3b_head_cfg_sample_static(const t3b_model_t* m, const void* hidden, int32_t rows, const t3b_sampling_t* sp,
                                   const t3b_state_t* st, int32_t advance, int32_t* out_tokens, float* logits_out,
                                   const t3b_options_t* opt, void* stream);

/* S1..S4 alone on caller-provided combined fp32 logits [n_req, vocab] (parity tests of the sampler). */
int32_t t3b_sample_logits(const float* logits, int32_t n_req, int32_t vocab, const t3b_sampling_t* sp,
                          const t3b_state_t* st, int32_t* out_tokens, void* stream);

/* C1(combine) + S1..S4 on head logits bf16 [rows, speech_vocab] (row 2b = cond, 2b+1 = uncond), the
 * second half of t3b_head_cfg_sample.  Same outputs and state advance. */
int32_t t3b_cfg_sample_head_logits(const void* head_logits, int32_t rows, int32_t vocab, const t3b_sampling_t* sp,
                                   const t3b_state_t* st, int32_t advance, int32_t* out_tokens, float* logits_out, void* stream);

/* S3Gen hand-off (tts.py:492-514): out[i] = ids[i] - 2500 for the ids with 0 <= ids[i] - 2500 < 6561, in order;
 * *n_out (device int32) receives the count.  ids / out int32 [n] device arrays (out may alias ids). */
int32_t t3b_filter_speech_tokens(const int32_t* ids, int32_t n, int32_t offset, int32_t n_valid, int32_t* out,
                                 int32_t* n_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* T3B200_H */
