// Library plumbing + the whole-step entry points (t3b_decode_step, t3b_head_cfg_sample).
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "t3b_common.cuh"

// implemented in the kernel translation units
int32_t t3b_linear_gemv_launch(const void* x, const void* w, int rows, int N, int K, void* y, int mode, cudaStream_t s,
                               const void* res_in, void* res_out, const void* norm_w, float eps, int has_res, PfArgs pf, const void* wf);
bool t3b_gemv_mma_ok(int N, int K, int mode);
size_t t3b_gemv_frag_bytes_impl(int N, int K, int gate_up);
int32_t t3b_gemv_frag_launch(const void* w, int N, int K, int gate_up, void* out, cudaStream_t s);
int32_t t3b_linear_umma_launch(const void* x, const void* w, const t3b_linear_t* awq, int rows, int N, int K, void* y, float* partials,
                               int n_split, int epi, int stages_cap, int row_chunk, PfArgs pf, int w_wait, cudaStream_t s,
                               const UmmaFuse* fuse);
int32_t t3b_rmsnorm_s32_launch(const float* s32, const void* weight, void* y, int rows, int hidden, float eps, cudaStream_t stream);
int32_t t3b_awq_dequant_launch(const t3b_linear_t* const* lins, int n_lin, void* scratch, cudaStream_t stream);
int32_t t3b_linear_big_launch(const void* x, const void* w, int rows, int N, int K, void* y, int epi, cudaStream_t stream);
int t3b_umma_pick_split(int rows, int N, int K);
int t3b_umma_available();
int32_t t3b_sampler_launch(const void* head, const float* head_partials, int n_split, const float* logits_in, int V, int n_req,
                           const t3b_sampling_t* sp, const t3b_state_t* st, int advance, int32_t* out_tokens, float* logits_out,
                           cudaStream_t stream);
int32_t t3b_attn_fused_launch(const void* qkv, int n_tok, const int32_t* tok_idx, const int32_t* rope_pos, const void* rope_cos_sin, int max_pos,
                              const int32_t* block_tables, int max_blocks, void* k_pool, void* v_pool, int n_heads, int kv_dtype, void* out,
                              int team_warps, PfArgs pf, const int32_t* row_order, cudaStream_t s);
int32_t t3b_paged_attn_launch(const void* q, int n_tok, const int32_t* tok_row, const int32_t* tok_idx, const int32_t* block_tables,
                              int max_blocks, const void* k_pool, const void* v_pool, int n_heads, int kv_dtype, void* out, int team_warps,
                              cudaStream_t s);

#define NORM_FUSE_MIN_ROWS 9          // below: the two norm launches are cheaper than the operand build (4 rows: 0.877 -> 0.903 ms/step)
#define NORM_FUSE_MAX_ROWS 32         // decode steps of 3..32 rows run F2.1 inside the projections (measured: 1.157 -> 1.076 ms/step at 32 rows;
                                      // at 64 rows the operand build costs what the two launches saved: 1.351 -> 1.369, profiles/r2_notes.md r4)
#define AWQ_DEQUANT_MIN_ROWS 256      // from here on AWQ layers are dequantised once per layer into a bf16 scratch (vllm awq.py:276-281 does the same)
#define GEMV_MMA_DEFAULT 0            // t3b_options_t.gemv_mma = 0
#define GEMV_MAX_ROWS_DEFAULT 2       // library default without options; engines fix their own (16 with the tensor-core GEMV chain)
static thread_local char g_err[512] = "";

void t3b_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int t3b_num_sms() {
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
            sms = 148;
    }
    return sms;
}

extern "C" const char* t3b_last_error(void) { return g_err; }
extern "C" int32_t t3b_version(void) { return 200; }
extern "C" int32_t t3b_kv_slice_bytes(int32_t kv_dtype) { return kv_dtype == T3B_KV_BF16 ? 2048 : (kv_dtype == T3B_KV_FP8 ? 1088 : 0); }

extern "C" int32_t t3b_check_device(void) {
    int dev = 0, major = 0;
    T3B_CUDA(cudaGetDevice(&dev));
    T3B_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    if (major != 10) T3B_FAIL(T3B_ERR_UNSUPPORTED, "libt3b200 is built for sm_100a only; device %d is sm_%dx", dev, major);
    if (!t3b_umma_available()) T3B_FAIL(T3B_ERR_UNSUPPORTED, "cuTensorMapEncodeTiled driver entry point not found");
    return T3B_OK;
}

// ------------------------------------------------------------------ workspace carving
namespace {
inline size_t al(size_t x) { return (x + 255) & ~size_t(255); }

struct Workspace {
    char *x_emb, *resid, *normed, *qkv, *q, *attn, *act, *lin, *head, *resid2, *deq, *s32, *ssq;
    size_t total;
};

Workspace carve(const t3b_model_t* m, int rows, void* base) {
    Workspace w;
    char* p = static_cast<char*>(base);
    const size_t D = m->hidden, F = m->ffn, V = m->speech_vocab, R = rows;
    auto take = [&](size_t bytes) { char* r = p; p += al(bytes); return r; };
    w.x_emb = take(R * D * 2);
    w.resid = take(R * D * 2);
    w.normed = take(R * D * 2);
    w.qkv = take(R * 3 * D * 2);
    w.q = take(R * D * 2);
    w.attn = take(R * D * 2);
    w.act = take(R * F * 2);
    w.lin = take(R * D * 2);
    w.head = take(R * V * 2 + 64);
    w.resid2 = take((R < 16 ? R : 16) * D * 2);  // ping-pong residual of the GEMV chains (rows <= 16)
    // bf16 copy of ONE layer's projections for AWQ models at large row counts
    const bool awq_big = m->layers && m->layers[0].qkv.kind == T3B_W_AWQ && rows >= AWQ_DEQUANT_MIN_ROWS;
    w.deq = take(awq_big ? (3 * D * D + D * D + 2 * F * D + D * F) * 2 : 0);
    // decode chain at <= 128 rows: fp32 residual sum + per-tile sums of squares handed from o / down to gate_up / qkv
    const size_t Rf = R < NORM_FUSE_MAX_ROWS ? R : NORM_FUSE_MAX_ROWS;
    w.s32 = take(Rf * D * 4);
    w.ssq = take(Rf * (D / 128) * 4);
    w.total = static_cast<size_t>(p - static_cast<char*>(base));
    return w;
}

}  // namespace

extern "C" size_t t3b_workspace_bytes(const t3b_model_t* m, int32_t max_rows) {
    if (!m || max_rows <= 0) return 0;
    return carve(m, max_rows, nullptr).total + 256;
}

struct Opt {
    int gemv_max_rows, gemv_norm, stages, row_chunk, split[4], attn_team, prefetch, w_wait, big_min_rows, norm_fuse, gemv_mma;
};
static Opt resolve(const t3b_options_t* o) {
    Opt r = {GEMV_MAX_ROWS_DEFAULT, 1, 0, 0, {0, 0, 0, 0}, 0, 0, 0, 1024, 1, GEMV_MMA_DEFAULT};      // norm_fuse: 0 never, 1 where faster, 2 wherever possible
    if (o) {
        if (o->gemv_mma != 0) r.gemv_mma = o->gemv_mma > 0 ? 1 : 0;
        const int cap = r.gemv_mma ? 16 : 4;
        if (o->gemv_max_rows != 0) r.gemv_max_rows = o->gemv_max_rows < 0 ? 0 : (o->gemv_max_rows > cap ? cap : o->gemv_max_rows);
        r.gemv_norm = o->gemv_no_norm ? 0 : 1;
        r.stages = o->umma_stages > 0 ? o->umma_stages : 0;
        r.row_chunk = o->umma_row_chunk > 0 ? o->umma_row_chunk : 0;
        for (int i = 0; i < 4; ++i) r.split[i] = o->umma_split[i] > 0 ? o->umma_split[i] : 0;
        r.attn_team = o->attn_team_warps;
        r.prefetch = o->weight_prefetch > 0 ? o->weight_prefetch : 0;
        r.norm_fuse = o->norm_fusion < 0 ? 0 : (o->norm_fusion > 0 ? 2 : 1);
        if (o->umma_big_min_rows != 0) r.big_min_rows = o->umma_big_min_rows < 0 ? 0x7fffffff : o->umma_big_min_rows;
    }
    return r;
}

// ------------------------------------------------------------------ linear dispatch
// epi: 0 = y[rows, n]; 1 = gate_up with the SiLU*mul epilogue -> y[rows, n/2].
static int32_t linear_any(const void* x, const t3b_linear_t* w, int rows, void* y, float* partials, int n_split, int* n_split_out, int impl,
                          int epi, const Opt& o, PfArgs pf, cudaStream_t s, const UmmaFuse* fuse = nullptr) {
    T3B_CHECK_ARG(x && w && w->w, "linear: null pointer");
    T3B_CHECK_ARG(rows >= 0 && w->n > 0 && w->k > 0, "linear: bad sizes rows=%d n=%d k=%d", rows, w->n, w->k);
    T3B_CHECK_ALIGN(x); T3B_CHECK_ALIGN(w->w);
    if (y) T3B_CHECK_ALIGN(y);
    if (partials) T3B_CHECK_ALIGN(partials);
    if (rows == 0) { if (n_split_out) *n_split_out = 1; return T3B_OK; }
    const bool awq = w->kind == T3B_W_AWQ;
    T3B_CHECK_ARG(awq || w->kind == T3B_W_BF16, "linear: unknown weight kind %d", w->kind);
    // the SIMT GEMV takes <= 4 rows of any K % 1024 == 0, the tensor-core GEMV <= 8 rows of the model's own K (1024 / 4096)
    const bool mma = o.gemv_mma && !awq && w->scales != nullptr && t3b_gemv_mma_ok(w->n, w->k, epi);     // scales = the fragment-ordered copy
    const bool gemv_fits = mma ? rows <= 16 : (rows <= 4 && w->k % 1024 == 0 && w->k <= 8192);
    if (impl == T3B_IMPL_AUTO) impl = (!awq && rows <= o.gemv_max_rows && gemv_fits) ? T3B_IMPL_GEMV : T3B_IMPL_UMMA;
    if (impl == T3B_IMPL_GEMV) {
        T3B_CHECK_ARG(fuse == nullptr, "linear: the fused normalisation runs on the tcgen05 kernel only");
        T3B_CHECK_ARG(!awq, "linear: AWQ weights run on the dequant-fused tcgen05 kernel at every row count (no GEMV)");
        T3B_CHECK_ARG(n_split <= 1, "linear: the GEMV path has no split-K (n_split %d)", n_split);
        T3B_CHECK_ARG(y != nullptr, "linear: y is null");
        if (n_split_out) *n_split_out = 1;
        return t3b_linear_gemv_launch(x, w->w, rows, w->n, w->k, y, epi, s, nullptr, nullptr, nullptr, 0.f, 0, pf, mma ? w->scales : nullptr);
    }
    T3B_CHECK_ARG(impl == T3B_IMPL_UMMA, "linear: unknown impl %d", impl);
    const bool legacy = partials != nullptr && n_split > 1;      // explicit request for global fp32 partials
    if (n_split <= 0) {
        const int which = (w->n == 3072 && w->k == 1024) ? 0 : (w->n == 1024 && w->k == 1024) ? 1 : (w->n == 8192 && w->k == 1024) ? 2
                        : (w->n == 1024 && w->k == 4096) ? 3 : -1;
        n_split = (which >= 0 && o.split[which] > 0 && rows <= 128) ? o.split[which] : t3b_umma_pick_split(rows, w->n, w->k);
    }
    if (!legacy && y == nullptr) T3B_FAIL(T3B_ERR_ARG, "linear: y is null (partials are only written for an explicit n_split > 1)");
    if (n_split_out) *n_split_out = n_split;
    // thousands of rows, no split-K, bf16 (AWQ layers arrive as their dequantised copy): the persistent double-buffered kernel.
    // (it requests weights after its dependency wait, so it is also safe on the AWQ scratch)
    if (!awq && !legacy && n_split == 1 && rows >= o.big_min_rows && !fuse)
        return t3b_linear_big_launch(x, w->w, rows, w->n, w->k, y, epi, s);
    return t3b_linear_umma_launch(x, w->w, awq ? w : nullptr, rows, w->n, w->k, y, legacy ? partials : nullptr, n_split, epi, o.stages,
                                  o.row_chunk, pf, o.w_wait, s, fuse);
}

extern "C" int32_t t3b_linear_bf16(const void* x, const t3b_linear_t* w, int32_t rows, void* y, float* partials, int32_t n_split,
                                   int32_t* n_split_out, int32_t impl, int32_t epilogue, const t3b_options_t* opt, void* stream) {
    T3B_CHECK_ARG(w && w->kind == T3B_W_BF16, "t3b_linear_bf16: weight kind must be T3B_W_BF16");
    T3B_CHECK_ARG(epilogue == T3B_EPI_NONE || epilogue == T3B_EPI_SILU_MUL, "t3b_linear_bf16: unknown epilogue %d", epilogue);
    return linear_any(x, w, rows, y, partials, n_split, n_split_out, impl, epilogue, resolve(opt), pf_none(), t3b_stream(stream));
}

extern "C" int32_t t3b_linear_awq(const void* x, const t3b_linear_t* w, int32_t rows, void* y, float* partials, int32_t n_split,
                                  int32_t* n_split_out, int32_t impl, int32_t epilogue, const t3b_options_t* opt, void* stream) {
    T3B_CHECK_ARG(w && w->kind == T3B_W_AWQ, "t3b_linear_awq: weight kind must be T3B_W_AWQ");
    T3B_CHECK_ARG(w->scales && w->zeros, "t3b_linear_awq: scales / zeros are null");
    T3B_CHECK_ARG(epilogue == T3B_EPI_NONE || epilogue == T3B_EPI_SILU_MUL, "t3b_linear_awq: unknown epilogue %d", epilogue);
    return linear_any(x, w, rows, y, partials, n_split, n_split_out, impl, epilogue, resolve(opt), pf_none(), t3b_stream(stream));
}

extern "C" size_t t3b_gemv_frag_bytes(int32_t n, int32_t k, int32_t gate_up) {
    return t3b_gemv_mma_ok(n, k, gate_up) ? t3b_gemv_frag_bytes_impl(n, k, gate_up) : 0;
}
extern "C" int32_t t3b_repack_weights_gemv(const void* w, int32_t n, int32_t k, int32_t gate_up, void* out, void* stream) {
    T3B_CHECK_ARG(w && out, "t3b_repack_weights_gemv: null pointer");
    T3B_CHECK_ALIGN(w); T3B_CHECK_ALIGN(out);
    return t3b_gemv_frag_launch(w, n, k, gate_up, out, t3b_stream(stream));
}

// ------------------------------------------------------------------ whole step
static int32_t check_model(const t3b_model_t* m) {
    T3B_CHECK_ARG(m && m->layers && m->final_norm && m->speech_emb && m->speech_pos_emb && m->rope_cos_sin && m->head.w,
                  "model: null pointer");
    T3B_CHECK_SHAPE(m->head_dim == T3B_HEAD_DIM, "model: head_dim %d (kernels are specialised for 64)", m->head_dim);
    T3B_CHECK_SHAPE(m->hidden == m->n_heads * m->head_dim, "model: hidden %d != heads %d x 64", m->hidden, m->n_heads);
    T3B_CHECK_SHAPE(m->hidden % 128 == 0 && m->ffn % 128 == 0 && m->hidden <= 4096, "model: hidden %d / ffn %d unsupported", m->hidden, m->ffn);
    T3B_CHECK_SHAPE(m->n_layers >= 1 && m->speech_vocab >= 2 && m->speech_pos_rows >= 1 && m->max_pos >= 1, "model: bad sizes");
    return T3B_OK;
}
static int32_t check_kv(const t3b_kv_t* kv, const t3b_model_t* m) {
    T3B_CHECK_ARG(kv && kv->k && kv->v, "kv: null pointer");
    const int slice = t3b_kv_slice_bytes(kv->dtype);
    T3B_CHECK_ARG(slice > 0, "kv: unknown dtype %d", kv->dtype);
    T3B_CHECK_ARG(kv->layer_stride >= (int64_t)kv->num_blocks * m->n_heads * slice && kv->layer_stride % 16 == 0,
                  "kv: layer_stride %lld bytes < %d blocks x %d heads x %d", (long long)kv->layer_stride, kv->num_blocks, m->n_heads, slice);
    return T3B_OK;
}

struct PrefillSeqs {          // sequences of a multi-token pass (t3b_prefill_attn); n_seq == 0: every token runs the paged decode kernel
    int n_seq;
    const int32_t* seq_start;
    const int32_t* seq_row;
    int max_seq_len;
};

// F2: the 30-layer stack over n_tok tokens (decode: one token per CFG row; prefill: the whole prompt of
// each new row, causal through tok_idx).  x0 = bf16 [n_tok, hidden] input embeddings.
// Decode layer = 7 launches: add+norm, qkv GEMM, attention (RoPE + KV write fused), o GEMM, add+norm,
// gate_up GEMM (SiLU*mul epilogue), down GEMM; at GEMV row counts (SIMT <= 4 rows, tensor cores <= 16 rows) the two add+norms run
// inside the qkv / gate_up GEMVs (5 launches), at <= 32 rows on the tcgen05 kernel inside the projections (5 launches).
static int32_t forward_impl(const t3b_model_t* m, const t3b_kv_t* kv, const void* x0, int n_tok, const int32_t* tok_row,
                            const int32_t* tok_idx, const int32_t* rope_pos, const int32_t* block_tables, int max_blocks,
                            const PrefillSeqs* ps, void* hidden_out, const Workspace& ws, const Opt& o, void* stream,
                            const int32_t* row_order = nullptr) {
    cudaStream_t s = t3b_stream(stream);
    const int D = m->hidden, H = m->n_heads;
    int32_t rc;
    // Multi-token passes (tok_row given = prefill) fix every launch shape that changes fp32 summation order (split-K = 1) so a
    // request's prefill is bit-identical no matter which other requests are admitted with it.  Decode steps pick shapes by row
    // count (identical within a slot bucket).
    const bool prefill = tok_row != nullptr;
    // AWQ models at >= 256 rows: the bf16 kernels read a scratch the dequant launch of the same layer wrote, so they must not
    // request weights before their dependency wait (every other weight is constant and IS requested before it)
    const bool deq = n_tok >= AWQ_DEQUANT_MIN_ROWS && m->layers[0].qkv.kind == T3B_W_AWQ;
    Opt ol = o;
    ol.w_wait = deq ? 1 : 0;
    // 17..32 rows: a 3-deep ring leaves room for a block of the NEXT projection on the same SM, whose weight loads (issued before
    // its dependency wait) then overlap this one (32 rows: 1.157 -> 1.115 ms/step without the folded norms, 1.134 -> 1.076 with)
    if (!prefill && o.stages == 0 && n_tok > 16 && n_tok <= 32) ol.stages = 3;
    auto lin = [&](const void* x, const t3b_linear_t* w, void* y, int epi, PfArgs pf) -> int32_t {
        return linear_any(x, w, n_tok, y, nullptr, prefill ? 1 : 0, nullptr, prefill ? T3B_IMPL_UMMA : T3B_IMPL_AUTO, epi, ol, pf, s);
    };
    // Decode steps are HBM-bound and every kernel of the chain is short: each launch pulls the weights of the projection two
    // launches ahead into L2 (qkv -> o, attention -> gate_up, o -> down, gate_up -> the next layer's qkv / the speech head), so
    // the weight stream never waits for the dependent chain.  Prefill passes are compute-bound: no prefetch.
    const bool pfo = !prefill && o.prefetch;
    bool frag_pf = false;          // set below: the chain reads the fragment-ordered copies, so those are what gets prefetched
    auto pfw = [&](const t3b_linear_t* w) {
        if (!pfo) return pf_none();
        if (frag_pf && w->kind == T3B_W_BF16 && w->scales != nullptr) {
            PfArgs a;
            a.p = static_cast<const unsigned char*>(w->scales);
            a.bytes = t3b_gemv_frag_bytes_impl(w->n, w->k, 0) & ~15ull;        // same size for a gate_up copy
            return a;
        }
        return pf_linear(w);
    };
    auto pf_after = [&](int li) { return li + 1 < m->n_layers ? pfw(&m->layers[li + 1].qkv) : pfw(&m->head); };
    auto kpool = [&](int li) { return static_cast<char*>(kv->k) + (size_t)li * kv->layer_stride; };
    auto vpool = [&](int li) { return static_cast<char*>(kv->v) + (size_t)li * kv->layer_stride; };
    // Small batch (rows <= gemv_max_rows, bf16): F2.1 runs in front of the qkv / gate_up GEMVs (each block normalises the <= 4 rows
    // it needs anyway; block 0 writes the residual to the other ping-pong buffer) - 5 launches per layer, bit-identical to the
    // 7-launch path.
    // tensor-core GEMV (<= 8 rows): every layer carries the fragment-ordered copies of its projections (t3b_repack_weights_gemv)
    bool small_mma = o.gemv_mma && t3b_gemv_mma_ok(2 * m->ffn, D, 1);
    for (int li = 0; small_mma && li < m->n_layers; ++li)
        small_mma = m->layers[li].qkv.kind == T3B_W_BF16 && m->layers[li].gate_up.kind == T3B_W_BF16 && m->layers[li].qkv.scales != nullptr &&
                    m->layers[li].gate_up.scales != nullptr;
    const bool small = !prefill && n_tok <= o.gemv_max_rows && o.gemv_norm && D == 1024 && m->layers[0].qkv.kind == T3B_W_BF16 &&
                       m->layers[0].gate_up.kind == T3B_W_BF16 && m->ffn % 1024 == 0 && (small_mma || n_tok <= 4);
    if (small) {
        frag_pf = small_mma;
        char* rcur = ws.resid;
        char* roth = ws.resid2;
        const void* xin = x0;
        for (int li = 0; li < m->n_layers; ++li) {
            const t3b_layer_t* L = &m->layers[li];
            T3B_CHECK_ARG(L->ln1 && L->ln2, "forward: layer %d norm weights are null", li);
            // weight_prefetch 1: every launch pulls the projection two launches ahead into L2; 2: the same projection of the NEXT layer
            // (a whole layer ahead; gate_up rides on the attention launch, the speech head on the last layer's qkv)
            const bool far = o.prefetch >= 2;
            const t3b_layer_t* Ln = li + 1 < m->n_layers ? &m->layers[li + 1] : nullptr;
            auto nxt = [&](const t3b_linear_t* w) { return w ? pfw(w) : pf_none(); };
            rc = t3b_linear_gemv_launch(xin, L->qkv.w, n_tok, L->qkv.n, L->qkv.k, ws.qkv, 0, s, rcur, roth, L->ln1, m->rms_eps, li > 0,
                                        far ? (Ln ? pfw(&Ln->qkv) : pfw(&m->head)) : pfw(&L->o), small_mma ? L->qkv.scales : nullptr);
            if (rc) return rc;
            { char* t = rcur; rcur = roth; roth = t; }
            rc = t3b_attn_fused_launch(ws.qkv, n_tok, tok_idx, rope_pos, m->rope_cos_sin, m->max_pos, block_tables, max_blocks, kpool(li),
                                       vpool(li), H, kv->dtype, ws.attn, o.attn_team, far ? nxt(Ln ? &Ln->gate_up : nullptr) : pfw(&L->gate_up),
                                       row_order, s);
            if (rc) return rc;
            rc = lin(ws.attn, &L->o, ws.lin, T3B_EPI_NONE, far ? nxt(Ln ? &Ln->o : nullptr) : pfw(&L->down));
            if (rc) return rc;
            rc = t3b_linear_gemv_launch(ws.lin, L->gate_up.w, n_tok, L->gate_up.n, L->gate_up.k, ws.act, 1, s, rcur, roth, L->ln2, m->rms_eps, 1,
                                        far ? pf_none() : pf_after(li), small_mma ? L->gate_up.scales : nullptr);
            if (rc) return rc;
            { char* t = rcur; rcur = roth; roth = t; }
            rc = lin(ws.act, &L->down, ws.lin, T3B_EPI_NONE, far ? nxt(Ln ? &Ln->down : nullptr) : pf_none());
            if (rc) return rc;
            xin = ws.lin;
        }
        return t3b_add_rmsnorm(xin, nullptr, 1, rcur, m->final_norm, hidden_out, n_tok, D, m->rms_eps, 1, stream);
    }
    // Decode steps at <= 128 rows (every projection runs the cluster split-K kernel): F2.1 is folded into the projections either
    // side of it.  o / down add the residual in their epilogue and leave the fp32 sum plus per-tile sums of squares; gate_up / the
    // next layer's qkv normalise their activation operand from those while the weight tiles are in flight.  5 launches per layer
    // instead of 7; only layer 0's first normalisation and the final one stay kernels.
    const bool fuse = !prefill && (o.norm_fuse == 2 || (o.norm_fuse == 1 && n_tok >= NORM_FUSE_MIN_ROWS)) && n_tok <= NORM_FUSE_MAX_ROWS &&
                      n_tok > o.gemv_max_rows && D == 1024 && !deq &&
                      t3b_umma_pick_split(n_tok, D, D) > 1 && t3b_umma_pick_split(n_tok, D, m->ffn) > 1 && o.split[1] != 1 && o.split[3] != 1;
    if (fuse) {
        float* s32 = reinterpret_cast<float*>(ws.s32);
        float* ssq = reinterpret_cast<float*>(ws.ssq);
        UmmaFuse f_out = {};
        f_out.resid = ws.resid; f_out.s32 = s32; f_out.ssq = ssq;
        auto lin_f = [&](const void* x, const t3b_linear_t* w, void* y, int epi, PfArgs pf, const UmmaFuse* f) -> int32_t {
            return linear_any(x, w, n_tok, y, nullptr, 0, nullptr, T3B_IMPL_UMMA, epi, ol, pf, s, f);
        };
        for (int li = 0; li < m->n_layers; ++li) {
            const t3b_layer_t* L = &m->layers[li];
            T3B_CHECK_ARG(L->ln1 && L->ln2, "forward: layer %d norm weights are null", li);
            UmmaFuse f_in = {};
            f_in.xs32 = s32; f_in.xssq = ssq; f_in.eps = m->rms_eps; f_in.ssq_tiles = D / 128;
            if (li == 0) {
                rc = t3b_add_rmsnorm(x0, nullptr, 1, ws.resid, L->ln1, ws.normed, n_tok, D, m->rms_eps, 0, stream);
                if (rc) return rc;
                rc = lin_f(ws.normed, &L->qkv, ws.qkv, T3B_EPI_NONE, pfw(&L->o), nullptr);
            } else {
                f_in.ln_w = L->ln1;
                rc = lin_f(ws.normed, &L->qkv, ws.qkv, T3B_EPI_NONE, pfw(&L->o), &f_in);
            }
            if (rc) return rc;
            rc = t3b_attn_fused_launch(ws.qkv, n_tok, tok_idx, rope_pos, m->rope_cos_sin, m->max_pos, block_tables, max_blocks, kpool(li),
                                       vpool(li), H, kv->dtype, ws.attn, o.attn_team, pfw(&L->gate_up), row_order, s);
            if (rc) return rc;
            rc = lin_f(ws.attn, &L->o, ws.lin, T3B_EPI_NONE, pfw(&L->down), &f_out);
            if (rc) return rc;
            f_in.ln_w = L->ln2;
            rc = lin_f(ws.normed, &L->gate_up, ws.act, T3B_EPI_SILU_MUL, pf_after(li), &f_in);
            if (rc) return rc;
            rc = lin_f(ws.act, &L->down, ws.lin, T3B_EPI_NONE, pf_none(), &f_out);
            if (rc) return rc;
        }
        return t3b_rmsnorm_s32_launch(s32, m->final_norm, hidden_out, n_tok, D, m->rms_eps, s);
    }
    const void* prev = x0;
    t3b_layer_t tmp = {};
    for (int li = 0; li < m->n_layers; ++li) {
        const t3b_layer_t* L = &m->layers[li];
        T3B_CHECK_ARG(L->ln1 && L->ln2, "forward: layer %d norm weights are null", li);
        if (deq) {
            // one launch dequantises the layer's four projections into the scratch; the bf16 kernels run on the copy
            const t3b_linear_t* lins[4] = {&L->qkv, &L->o, &L->gate_up, &L->down};
            rc = t3b_awq_dequant_launch(lins, 4, ws.deq, s);
            if (rc) return rc;
            tmp = *L;
            char* p = ws.deq;
            t3b_linear_t* dst[4] = {&tmp.qkv, &tmp.o, &tmp.gate_up, &tmp.down};
            for (int j = 0; j < 4; ++j) {
                dst[j]->kind = T3B_W_BF16; dst[j]->w = p; dst[j]->scales = nullptr; dst[j]->zeros = nullptr; dst[j]->flags = 0;
                p += (size_t)dst[j]->n * dst[j]->k * 2;
            }
            L = &tmp;
        }
        rc = t3b_add_rmsnorm(prev, nullptr, 1, ws.resid, L->ln1, ws.normed, n_tok, D, m->rms_eps, li > 0, stream);
        if (rc) return rc;
        rc = lin(ws.normed, &L->qkv, ws.qkv, T3B_EPI_NONE, pfw(&L->o));
        if (rc) return rc;
        if (prefill) {
            rc = t3b_rope_kv_write(ws.qkv, nullptr, 1, n_tok, tok_row, tok_idx, rope_pos, m->rope_cos_sin, m->max_pos, block_tables,
                                   max_blocks, kpool(li), vpool(li), H, kv->dtype, ws.q, stream);
            if (rc) return rc;
            if (ps && ps->n_seq > 0 && ps->max_seq_len <= t3b_prefill_attn_max_len(kv->dtype))
                rc = t3b_prefill_attn(ws.q, ps->n_seq, ps->seq_start, ps->seq_row, block_tables, max_blocks, kpool(li), vpool(li), H,
                                      kv->dtype, ps->max_seq_len, ws.attn, stream);
            else
                rc = t3b_paged_attn_launch(ws.q, n_tok, tok_row, tok_idx, block_tables, max_blocks, kpool(li), vpool(li), H, kv->dtype, ws.attn, 1, s);   // one warp per (token, head): fixed, so a prompt's result does not depend on its batch
        } else {
            rc = t3b_attn_fused_launch(ws.qkv, n_tok, tok_idx, rope_pos, m->rope_cos_sin, m->max_pos, block_tables, max_blocks, kpool(li),
                                       vpool(li), H, kv->dtype, ws.attn, o.attn_team, pfw(&L->gate_up), row_order, s);
        }
        if (rc) return rc;
        rc = lin(ws.attn, &L->o, ws.lin, T3B_EPI_NONE, pfw(&L->down));
        if (rc) return rc;
        rc = t3b_add_rmsnorm(ws.lin, nullptr, 1, ws.resid, L->ln2, ws.normed, n_tok, D, m->rms_eps, 1, stream);
        if (rc) return rc;
        rc = lin(ws.normed, &L->gate_up, ws.act, T3B_EPI_SILU_MUL, pf_after(li));
        if (rc) return rc;
        rc = lin(ws.act, &L->down, ws.lin, T3B_EPI_NONE, pf_none());
        if (rc) return rc;
        prev = ws.lin;
    }
    return t3b_add_rmsnorm(prev, nullptr, 1, ws.resid, m->final_norm, hidden_out, n_tok, D, m->rms_eps, 1, stream);
}

extern "C" int32_t t3b_forward_tokens(const t3b_model_t* m, const t3b_kv_t* kv, const void* x_in, int32_t n_tok,
                                      const int32_t* tok_row, const int32_t* tok_idx, const int32_t* rope_pos,
                                      const int32_t* block_tables, int32_t max_blocks, int32_t n_seq, const int32_t* seq_start,
                                      const int32_t* seq_row, int32_t max_seq_len, void* hidden_out, void* workspace,
                                      size_t workspace_bytes, const t3b_options_t* opt, void* stream) {
    int32_t rc = check_model(m);
    if (rc) return rc;
    rc = check_kv(kv, m);
    if (rc) return rc;
    T3B_CHECK_ARG(x_in && hidden_out && workspace, "t3b_forward_tokens: null pointer");
    T3B_CHECK_ARG(tok_row && tok_idx && rope_pos && block_tables && max_blocks > 0 && n_tok >= 0, "t3b_forward_tokens: bad token arrays");
    T3B_CHECK_ARG(n_seq == 0 || (seq_start && seq_row && max_seq_len > 0), "t3b_forward_tokens: n_seq %d needs seq_start, seq_row, max_seq_len", n_seq);
    T3B_CHECK_ALIGN(workspace); T3B_CHECK_ALIGN(hidden_out); T3B_CHECK_ALIGN(x_in);
    if (n_tok == 0) return T3B_OK;
    T3B_CHECK_ARG(workspace_bytes >= t3b_workspace_bytes(m, n_tok), "t3b_forward_tokens: workspace %zu < %zu bytes", workspace_bytes,
                  t3b_workspace_bytes(m, n_tok));
    const PrefillSeqs ps = {n_seq, seq_start, seq_row, max_seq_len};
    return forward_impl(m, kv, x_in, n_tok, tok_row, tok_idx, rope_pos, block_tables, max_blocks, &ps, hidden_out,
                        carve(m, n_tok, workspace), resolve(opt), stream);
}

extern "C" int32_t t3b_decode_step(const t3b_model_t* m, const t3b_kv_t* kv, const t3b_state_t* st, const void* x_in, int32_t rows,
                                   void* hidden_out, void* workspace, size_t workspace_bytes, const t3b_options_t* opt, void* stream) {
    int32_t rc = check_model(m);
    if (rc) return rc;
    rc = check_kv(kv, m);
    if (rc) return rc;
    T3B_CHECK_ARG(st && hidden_out && workspace, "t3b_decode_step: null pointer");
    T3B_CHECK_ARG(st->rope_pos && st->ctx_lens && st->block_tables && st->max_blocks > 0, "t3b_decode_step: state arrays are null");
    T3B_CHECK_ARG(rows >= 0 && rows % 2 == 0, "t3b_decode_step: rows %d must be even (cond/uncond pairs)", rows);
    T3B_CHECK_ARG(x_in || (st->tokens && st->speech_pos), "t3b_decode_step: no input (x_in and state tokens are null)");
    T3B_CHECK_ALIGN(workspace); T3B_CHECK_ALIGN(hidden_out);
    if (rows == 0) return T3B_OK;
    T3B_CHECK_ARG(workspace_bytes >= t3b_workspace_bytes(m, rows), "t3b_decode_step: workspace %zu < %zu bytes", workspace_bytes,
                  t3b_workspace_bytes(m, rows));
    const Workspace ws = carve(m, rows, workspace);
    const void* x0 = x_in;
    if (!x0) {
        rc = t3b_embed_gather(m->speech_emb, m->speech_pos_emb, st->tokens, st->speech_pos, rows / 2, m->hidden, m->speech_vocab,
                              m->speech_pos_rows, 2, ws.x_emb, stream);
        if (rc) return rc;
        x0 = ws.x_emb;
    }
    return forward_impl(m, kv, x0, rows, nullptr, st->ctx_lens, st->rope_pos, st->block_tables, st->max_blocks, nullptr, hidden_out, ws,
                        resolve(opt), stream, st->row_order);
}

extern "C" int32_t t3b_head_cfg_sample(const t3b_model_t* m, const void* hidden, int32_t rows, const t3b_sampling_t* sp,
                                       const t3b_state_t* st, int32_t advance, int32_t* out_tokens, float* logits_out, void* workspace,
                                       size_t workspace_bytes, const t3b_options_t* opt, void* stream) {
    int32_t rc = check_model(m);
    if (rc) return rc;
    T3B_CHECK_ARG(hidden && sp && workspace, "t3b_head_cfg_sample: null pointer");
    T3B_CHECK_ARG(rows >= 0 && rows % 2 == 0, "t3b_head_cfg_sample: rows %d must be even", rows);
    T3B_CHECK_ARG(m->head.kind == T3B_W_BF16, "t3b_head_cfg_sample: speech_head must be bf16");
    if (rows == 0) return T3B_OK;
    T3B_CHECK_ARG(workspace_bytes >= t3b_workspace_bytes(m, rows), "t3b_head_cfg_sample: workspace too small");
    const Workspace ws = carve(m, rows, workspace);
    rc = linear_any(hidden, &m->head, rows, ws.head, nullptr, 0, nullptr, T3B_IMPL_AUTO, T3B_EPI_NONE, resolve(opt), pf_none(), t3b_stream(stream));
    if (rc) return rc;
    return t3b_sampler_launch(ws.head, nullptr, 1, nullptr, m->speech_vocab, rows / 2, sp, st, advance, out_tokens, logits_out,
                              t3b_stream(stream));
}

// ------------------------------------------------------------------ no-allocation variant of t3b_head_cfg_sample
// The head logits of a step (bf16 [rows, speech_vocab], scratch between the head projection and the sampler) live in a file-scope
// device array instead of a caller-provided workspace, so the call needs no allocation at all.  Every GPU that loads the library
// holds the array (T3B_STATIC_SCRATCH_BYTES, ~1.03 MiB) until exit.  In host code `g_static_head` is only a stand-in for the
// symbol, NOT a device address: the address comes from cudaGetSymbolAddress, once per device, with that device current.
#define T3B_STATIC_SCRATCH_BYTES ((size_t)T3B_STATIC_MAX_ROWS * T3B_STATIC_MAX_VOCAB * 2 + 256)
static __device__ __align__(256) unsigned char g_static_head[T3B_STATIC_SCRATCH_BYTES];

#define T3B_MAX_DEVICES 64
static void* g_static_head_ptr[T3B_MAX_DEVICES];      // per device; nullptr = not looked up yet
static std::mutex g_static_mu;

static int32_t static_head_scratch(void** out) {
    int dev = -1;
    T3B_CUDA(cudaGetDevice(&dev));
    T3B_CHECK_ARG(dev >= 0 && dev < T3B_MAX_DEVICES, "static scratch: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (g_static_head_ptr[dev] == nullptr) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_head);
        if (e != cudaSuccess || p == nullptr)
            T3B_FAIL(T3B_ERR_CUDA, "static scratch: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
        g_static_head_ptr[dev] = p;
    }
    *out = g_static_head_ptr[dev];
    return T3B_OK;
}

extern "C" int32_t t3b_head_cfg_sample_static(const t3b_model_t* m, const void* hidden, int32_t rows, const t3b_sampling_t* sp,
                                              const t3b_state_t* st, int32_t advance, int32_t* out_tokens, float* logits_out,
                                              const t3b_options_t* opt, void* stream) {
    int32_t rc = check_model(m);
    if (rc) return rc;
    T3B_CHECK_ARG(hidden && sp, "t3b_head_cfg_sample_static: null pointer");
    T3B_CHECK_ARG(rows >= 0 && rows % 2 == 0, "t3b_head_cfg_sample_static: rows %d must be even", rows);
    T3B_CHECK_ARG(m->head.kind == T3B_W_BF16, "t3b_head_cfg_sample_static: speech_head must be bf16");
    T3B_CHECK_ARG(rows <= T3B_STATIC_MAX_ROWS, "t3b_head_cfg_sample_static: rows %d exceed the static scratch (T3B_STATIC_MAX_ROWS = %d); "
                  "use t3b_head_cfg_sample with a workspace", rows, T3B_STATIC_MAX_ROWS);
    T3B_CHECK_ARG(m->speech_vocab <= T3B_STATIC_MAX_VOCAB, "t3b_head_cfg_sample_static: speech_vocab %d exceeds the static scratch "
                  "(T3B_STATIC_MAX_VOCAB = %d); use t3b_head_cfg_sample with a workspace", m->speech_vocab, T3B_STATIC_MAX_VOCAB);
    T3B_CHECK_ALIGN(hidden);
    if (rows == 0) return T3B_OK;
    void* head = nullptr;
    rc = static_head_scratch(&head);
    if (rc) return rc;
    // nothing in the scratch has to start at zero: the projection writes every one of its rows x speech_vocab elements before
    // the sampler reads them (the array keeps the previous call's logits until then)
    rc = linear_any(hidden, &m->head, rows, head, nullptr, 0, nullptr, T3B_IMPL_AUTO, T3B_EPI_NONE, resolve(opt), pf_none(), t3b_stream(stream));
    if (rc) return rc;
    return t3b_sampler_launch(head, nullptr, 1, nullptr, m->speech_vocab, rows / 2, sp, st, advance, out_tokens, logits_out,
                              t3b_stream(stream));
}

extern "C" int32_t t3b_sample_logits(const float* logits, int32_t n_req, int32_t vocab, const t3b_sampling_t* sp, const t3b_state_t* st,
                                     int32_t* out_tokens, void* stream) {
    T3B_CHECK_ARG(logits && sp && out_tokens && n_req >= 0, "t3b_sample_logits: null pointer");
    return t3b_sampler_launch(nullptr, nullptr, 1, logits, vocab, n_req, sp, st, 0, out_tokens, nullptr, t3b_stream(stream));
}

extern "C" int32_t t3b_cfg_sample_head_logits(const void* head_logits, int32_t rows, int32_t vocab, const t3b_sampling_t* sp,
                                              const t3b_state_t* st, int32_t advance, int32_t* out_tokens, float* logits_out, void* stream) {
    T3B_CHECK_ARG(head_logits && sp && rows >= 0 && rows % 2 == 0 && vocab >= 2, "t3b_cfg_sample_head_logits: bad arguments");
    if (rows == 0) return T3B_OK;
    return t3b_sampler_launch(head_logits, nullptr, 1, nullptr, vocab, rows / 2, sp, st, advance, out_tokens, logits_out, t3b_stream(stream));
}
