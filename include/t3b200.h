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
#define T3B_EPI_RESIDUAL 2       /* t3b_linear_fused only: out_z = bf16(x W^T) + bf16(res_z) (llama.py:316-333)   */

/* t3b_linear_t.flags */
#define T3B_LIN_GATE_UP_INTERLEAVED 1   /* AWQ tiles repacked as 64 gate rows + 64 up rows (SILU_MUL epilogue only) */

/* CFG combine arithmetic (t3.py:662 runs in bf16 tensors; SURVEY.md 8(a) also defines fp32) */
#define T3B_COMBINE_BF16 0
#define T3B_COMBINE_FP32 1

typedef struct t3b_linear {
    int32_t kind;                /* T3B_W_* */
    int32_t n, k;                /* out features, in features */
    const void* w;               /* bf16 [n,k]  |  AWQ repacked nibbles */
    const void* scales;          /* AWQ: bf16 [k/128, n]          (else NULL) */
    const void* zeros;           /* AWQ: uint8 [k/128, n]         (else NULL) */
    int32_t flags;               /* T3B_LIN_* */
} t3b_linear_t;

/* F2.1 folded into the projections around it (t3b_linear_fused).  The residual stream of a decode step is
 * kept as z = fp32 [rows, hidden] (the unrounded sum fused_add_rms_norm normalises, layernorm.py:40-59)
 * together with ss = fp32 [hidden/128][rows], the sum of squares of every 128-feature slice of a row. */
typedef struct t3b_linear_fuse {
    const float* x_z;            /* != NULL: the projection's input is rmsnorm(x_z) * norm_w, built inside the kernel */
    const float* x_ss;           /*          [x_ss_tiles][rows] partial sums of squares of x_z's rows                */
    int32_t      x_ss_tiles;
    const void*  norm_w;         /*          bf16 [k]                                                                */
    float        eps;
    const float* res_z;          /* T3B_EPI_RESIDUAL: residual in, fp32 [rows, n]                                     */
    float*       out_z;          /*                   residual out, fp32 [rows, n] (may alias res_z)                  */
    float*       out_ss;         /*                   [n/128][rows] sums of squares of out_z                          */
    /* tail norm (T3B_EPI_NONE, n == 1024, cluster split-K): the fused_add_rms_norm that FOLLOWS the projection runs in
     * its tail - the last thread block to finish a group of rows does  resid += y;  out = rmsnorm(resid) * norm  with the
     * arithmetic of t3b_add_rmsnorm (bit-identical), saving the separate launch.  y is still written.              */
    void*        tail_resid;     /* bf16 [rows, n], updated in place                                                  */
    const void*  tail_norm_w;    /* bf16 [n]   (eps = the eps field above)                                            */
    void*        tail_out;       /* bf16 [rows, n]; NULL = no tail norm                                               */
    uint32_t*    tail_counters;  /* 64 x uint32, zero before the first use; the kernel leaves them zero              */
} t3b_linear_fuse_t;

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

/* Paged KV pool.  Layout per layer: [num_blocks][n_heads][T3B_KV_BLOCK][head_dim] bf16, K and V
 * separate.  (vLLM's flash layout is [block][token][head][dim]; heads are hoisted here so one
 * (block, head) slice is 2 KB contiguous.)  Block 0 is reserved as a scratch block for idle rows. */
typedef struct t3b_kv {
    void* k; void* v;            /* bf16 [n_layers][num_blocks][n_heads][16][64] */
    int64_t layer_stride;        /* elements between layers = num_blocks*n_heads*16*64 */
    int32_t num_blocks;
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
    int32_t  stop_token;         /* 6562 (t3_config.py:9); sampling it clears active[b]           */
    int32_t  max_ctx;            /* capacity guard: a row never advances past this many tokens     */
} t3b_state_t;

typedef struct t3b_sampling {    /* per-request device arrays [B]; tts.py:455-464 SamplingParams */
    const float* cfg_scale;      /* CHATTERBOX_CFG_SCALE (t3.py:296), per request here             */
    const float* repetition_penalty;
    const float* temperature;    /* < 1e-5 = greedy (vllm:v1/sample/sampler.py _SAMPLING_EPS)      */
    const float* min_p;
    const float* top_p;
    const uint64_t* seed;
    int32_t combine;             /* T3B_COMBINE_* */
} t3b_sampling_t;

const char* t3b_last_error(void);
int32_t     t3b_version(void);
/* 0 if the current device is sm_100 and the driver entry points resolve, else T3B_ERR_UNSUPPORTED */
int32_t     t3b_check_device(void);
/* scratch needed by t3b_decode_step / t3b_head_cfg_sample for up to max_rows CFG rows */
size_t      t3b_workspace_bytes(const t3b_model_t* m, int32_t max_rows);

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
 *     Split-K: n_split <= 0 lets the library pick (returned through *n_split_out, may be NULL); the
 *     partial tiles are reduced on-chip through the thread-block cluster's distributed shared memory in a
 *     fixed order (deterministic) and y is written.  Only when partials != NULL AND n_split > 1 is given
 *     explicitly are the fp32 partials written to global memory instead ([n_split, rows, n], y untouched)
 *     for a consumer that sums them (t3b_add_rmsnorm / t3b_rope_kv_write / t3b_silu_mul accept them). */
int32_t t3b_linear_bf16(const void* x, const t3b_linear_t* w, int32_t rows, void* y, float* partials,
                        int32_t n_split, int32_t* n_split_out, int32_t impl, int32_t epilogue, void* stream);

/* F2.1 + F2.2/F2.5/F2.6 in one launch (tcgen05 path, any row count; bf16 or AWQ weights):
 *     input  = x (bf16 [rows,k]) or, when f->x_z is set, bf16(bf16(x_z * rstd) * norm_w) with
 *              rstd_r = rsqrt(sum_t x_ss[t][r] / k + eps)   (vllm:ir/ops/layernorm.py:40-59);
 *     output = y per `epilogue`, or for T3B_EPI_RESIDUAL the updated residual stream out_z / out_ss
 *              (the residual add of llama.py:316-333; y may be NULL).
 *     n_split <= 0 lets the library pick the cluster split-K.  Sums run in fixed orders that depend on the
 *     shapes only. */
int32_t t3b_linear_fused(const void* x, const t3b_linear_t* w, int32_t rows, void* y, const t3b_linear_fuse_t* f,
                         int32_t n_split, int32_t epilogue, void* stream);
/* start / end of a step's fp32 residual stream: z = x (bf16 [rows,hidden]) with its ss; and
 * y = bf16(bf16(z * rstd) * weight) (the final norm, llama.py:430-431). */
int32_t t3b_z_init(const void* x, int32_t rows, int32_t hidden, float* z, float* ss, void* stream);
int32_t t3b_z_rmsnorm(const float* z, const float* ss, const void* weight, int32_t rows, int32_t hidden, float eps,
                      void* y, void* stream);

/* Q   vllm:model_executor/layers/quantization/awq.py:214-283: same contract as t3b_linear_bf16 with
 *     W[k,n] = bf16((nib - zero) * scale), group 128, dequantised on the fly into the tcgen05 operand tile
 *     (every row count runs T3B_IMPL_UMMA; activations are bf16, scales are converted fp16 -> bf16 by
 *     t3b_repack_weights_awq). */
int32_t t3b_linear_awq(const void* x, const t3b_linear_t* w, int32_t rows, void* y, float* partials,
                       int32_t n_split, int32_t* n_split_out, int32_t impl, int32_t epilogue, void* stream);

/* one-off layout transform of AutoAWQ tensors (qweight int32 [k, n/8], qzeros int32 [k/128, n/8],
 * scales fp16 [k/128, n]; nibble order 0,2,4,6,1,3,5,7 - awq_triton.py:56-62) into the kernel
 * layout: w_out uint8 [n/128][k/64][2][128][16] (tile, k-block, k half, feature row, 32 nibbles; inside
 * a 32-bit word the nibble positions hold k offsets 0,2,4,6,1,3,5,7), zeros_out uint8 [k/128, n],
 * scales_out bf16 [k/128, n].  gate_up_interleave != 0: tile t = gate rows [64t, 64t+64) then up rows
 * [n/2 + 64t, ...) (set T3B_LIN_GATE_UP_INTERLEAVED in the t3b_linear_t; needs T3B_EPI_SILU_MUL). */
int32_t t3b_repack_weights_awq(const int32_t* qweight, const int32_t* qzeros, const void* scales_fp16,
                               int32_t n, int32_t k, int32_t gate_up_interleave, void* w_out, void* zeros_out,
                               void* scales_out, void* stream);

/* F2.3 + KV write  (rotary_embedding/common.py:144-174, llama3 table; reshape_and_cache of
 *     llama.py:231): qkv as bf16 [n_tok, 3*hidden] or split-K partials.  Token j belongs to KV row
 *     tok_row[j] (NULL = j), sits at index tok_idx[j] of that row's sequence and is rotated by
 *     rope_pos[j].  Writes q (rotated) bf16 [n_tok, hidden] and k (rotated), v into the pool. */
int32_t t3b_rope_kv_write(const void* qkv, const float* partials, int32_t n_split, int32_t n_tok,
                          const int32_t* tok_row, const int32_t* tok_idx, const int32_t* rope_pos,
                          const void* rope_cos_sin, int32_t max_pos, const int32_t* block_tables,
                          int32_t max_blocks, void* k_pool, void* v_pool, int32_t n_heads, void* q_out,
                          void* stream);

/* F2.4 paged decode attention (llama.py:211-221,231): o[j] = softmax(q[j] K^T / 8) V over the
 *     first tok_idx[j]+1 tokens of row tok_row[j], per head.  out bf16 [n_tok, hidden]. */
int32_t t3b_paged_attn_decode(const void* q, int32_t n_tok, const int32_t* tok_row, const int32_t* tok_idx,
                              const int32_t* block_tables, int32_t max_blocks, const void* k_pool,
                              const void* v_pool, int32_t n_heads, void* out, void* stream);

/* F2.3 + KV write + F2.4 in one launch for a decode step (token j <-> KV row j): reads q/k/v of every head
 *     from qkv bf16 [n_tok, 3*hidden], rotates q and k by rope_pos[j], stores k (rotated) and v at index
 *     tok_idx[j] of row j and attends over tokens 0..tok_idx[j].  out bf16 [n_tok, hidden]. */
int32_t t3b_attn_decode_fused(const void* qkv, int32_t n_tok, const int32_t* tok_idx, const int32_t* rope_pos,
                              const void* rope_cos_sin, int32_t max_pos, const int32_t* block_tables,
                              int32_t max_blocks, void* k_pool, void* v_pool, int32_t n_heads, void* out,
                              void* stream);

/* F2.6 activation (vllm:model_executor/layers/activation.py:138-141): act = silu(g) * u with the
 *     bf16 roundings of the reference; gate_up as bf16 [rows, 2*ffn] or split-K partials. */
int32_t t3b_silu_mul(const void* gate_up, const float* partials, int32_t n_split, int32_t rows, int32_t ffn,
                     void* act, void* stream);

/* F2   LlamaModel.forward (llama.py:395-432) over n_tok tokens given as embeddings x_in (bf16
 *     [n_tok, hidden]): token j belongs to KV row tok_row[j] (NULL = j), is the tok_idx[j]-th token of
 *     that row (its K/V are written there and it attends causally to tokens 0..tok_idx[j]) and is
 *     rotated by rope_pos[j].  One call with every prompt token of the new rows is the prefill of
 *     t3.py:542-561; one token per row is a decode step.  hidden_out bf16 [n_tok, hidden]. */
int32_t t3b_forward_tokens(const t3b_model_t* m, const t3b_kv_t* kv, const void* x_in, int32_t n_tok,
                           const int32_t* tok_row, const int32_t* tok_idx, const int32_t* rope_pos,
                           const int32_t* block_tables, int32_t max_blocks, void* hidden_out,
                           void* workspace, size_t workspace_bytes, void* stream);

/* F1 + F2  T3VllmModel.forward (t3.py:676-713) over LlamaModel.forward (llama.py:395-432) for one
 *     decode token per CFG row.  x_in (bf16 [rows, hidden]) != NULL feeds given embeddings (prefill-
 *     by-decode, t3.py:542-561 rows); x_in == NULL gathers E1 from st->tokens/speech_pos.  Uses
 *     st->rope_pos / ctx_lens / block_tables.  hidden_out bf16 [rows, hidden] = final-norm output. */
int32_t t3b_decode_step(const t3b_model_t* m, const t3b_kv_t* kv, const t3b_state_t* st, const void* x_in,
                        int32_t rows, void* hidden_out, void* workspace, size_t workspace_bytes, void* stream);

/* C1 + S1..S4  compute_logits (t3.py:650-673) fused with the sampler the engine runs next
 *     (vllm:v1/sample/sampler.py:67-307, ops/topk_topp_sampler.py:328-416, layers/utils.py:51-89):
 *     head on both rows, l = c + w (c - u), repetition penalty over st->history, greedy or
 *     temperature/min-p/top-p + exponential race (Philox4x32-10 keyed by seed, counter (index, step)).
 *     Writes out_tokens[b] (un-offset id; caller adds SPEECH_TOKEN_OFFSET = 2500, t3.py:666-672),
 *     optionally the combined fp32 logits [B, speech_vocab], and - if advance != 0 - appends to
 *     history and advances tokens/speech_pos/rope_pos/ctx_lens/step, clearing active[b] on stop. */
int32_t t3b_head_cfg_sample(const t3b_model_t* m, const void* hidden, int32_t rows, const t3b_sampling_t* sp,
                            const t3b_state_t* st, int32_t advance, int32_t* out_tokens, float* logits_out,
                            void* workspace, size_t workspace_bytes, void* stream);

/* S1..S4 alone on caller-provided combined fp32 logits [n_req, vocab] (parity tests of the sampler). */
int32_t t3b_sample_logits(const float* logits, int32_t n_req, int32_t vocab, const t3b_sampling_t* sp,
                          const t3b_state_t* st, int32_t* out_tokens, void* stream);

/* ---- the persistent whole-step kernel -------------------------------------------------------------
 * Same arithmetic as t3b_decode_step + the speech head of t3b_head_cfg_sample (t3.py:676-713,
 * llama.py:395-432, t3.py:650-660), executed by ONE cooperative launch (one CTA per SM) that streams
 * weights and past KV pages through a shared-memory ring; see csrc/decode_mega.cu. */

/* device pointers to the weights of one LlamaDecoderLayer in the unit layout of
 * t3b_repack_weights_bf16 (gate_up with gate_up_interleave = 1) + the two norm weights (bf16 [hidden]) */
typedef struct t3b_mega_layer {
    const void* qkv; const void* o; const void* gate_up; const void* down;
    const void* ln1; const void* ln2;
} t3b_mega_layer_t;

typedef struct t3b_mega {
    const t3b_mega_layer_t* layers;   /* DEVICE array [n_layers] */
    const void* head;                 /* speech_head repacked (rows past speech_vocab zero-filled) */
} t3b_mega_t;

/* bytes of the repacked form of a bf16 [n, k] weight: ceil(n/128) * 128 * k * 2 */
size_t  t3b_repacked_bytes(int32_t n, int32_t k);
/* one-off layout transform: bf16 [n, k] -> [tile][k/64] blocks
 * of 128 x 64, 128B-swizzled K-major, 16 KB each; gate_up_interleave pairs 64 gate rows with 64 up rows */
int32_t t3b_repack_weights_bf16(const void* w, int32_t n, int32_t k, int32_t gate_up_interleave, void* out, void* stream);
int32_t t3b_mega_max_rows(void);      /* largest `rows` t3b_decode_step_mega accepts (128) */
size_t  t3b_mega_workspace_bytes(const t3b_model_t* m, int32_t max_rows);
/* F1 + F2 + C1(head): one decode token per CFG row, rows even and <= t3b_mega_max_rows().  Inputs as
 * t3b_decode_step.  hidden_out (optional) bf16 [rows, hidden] = final-norm output; logits_out bf16
 * [rows, speech_vocab] = speech_head on every row (NULL = kept in the workspace). */
int32_t t3b_decode_step_mega(const t3b_model_t* m, const t3b_mega_t* mw, const t3b_kv_t* kv, const t3b_state_t* st,
                             const void* x_in, int32_t rows, void* hidden_out, void* logits_out, void* workspace,
                             size_t workspace_bytes, void* stream);
/* C1(combine) + S1..S4 on head logits bf16 [rows, speech_vocab] (row 2b = cond, 2b+1 = uncond), the
 * second half of t3b_head_cfg_sample.  Same outputs and state advance. */
int32_t t3b_cfg_sample_head_logits(const void* head_logits, int32_t rows, int32_t vocab, const t3b_sampling_t* sp,
                                   const t3b_state_t* st, int32_t advance, int32_t* out_tokens, float* logits_out, void* stream);

/* ---- the fused persistent step kernel (two global exchanges per layer) --------------------------------
 * Same arithmetic and interface as t3b_decode_step_mega; see csrc/decode_fused.cu.  16 clusters x 8 CTAs:
 * cluster = attention head = FFN slice, q|k|v -> attention -> o-partial and gate_up -> SiLU*mul ->
 * down-partial stay inside a cluster (distributed shared memory); the only global exchanges are the two
 * split-K reductions + RMSNorm per layer.  Built for hidden 1024, 16 heads x 64, ffn 4096, bf16 weights. */

/* bytes of the repacked weight blob (0 if the model shape is not supported) */
size_t  t3b_fused_weight_bytes(const t3b_model_t* m);
/* one-off layout transform of every bf16 projection, the speech head and the norm weights of `m` into the
 * per-CTA streams the kernel's producer warp walks (`out`: t3b_fused_weight_bytes(m) bytes, 16-byte aligned) */
int32_t t3b_fused_repack(const t3b_model_t* m, void* out, void* stream);
int32_t t3b_fused_max_rows(void);     /* largest `rows` t3b_decode_step_fused accepts (128) */
size_t  t3b_fused_workspace_bytes(const t3b_model_t* m, int32_t max_rows);
/* F1 + F2 + C1(head) (t3.py:676-713, llama.py:395-432, t3.py:650-660): one decode token per CFG row, rows
 * even and <= t3b_fused_max_rows().  Inputs as t3b_decode_step.  hidden_out (optional) bf16 [rows, hidden] =
 * final-norm output; logits_out bf16 [rows, speech_vocab] = speech_head on every row (NULL = kept in the
 * workspace).  Follow with t3b_cfg_sample_head_logits. */
int32_t t3b_decode_step_fused(const t3b_model_t* m, const void* fused_weights, const t3b_kv_t* kv, const t3b_state_t* st,
                              const void* x_in, int32_t rows, void* hidden_out, void* logits_out, void* workspace,
                              size_t workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* T3B200_H */
