// Library plumbing + the whole-step entry points (t3b_decode_step, t3b_head_cfg_sample).
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "t3b_common.cuh"

// implemented in the kernel translation units
int32_t t3b_linear_gemv_launch(const void* x, const void* w, int rows, int N, int K, void* y, int mode, cudaStream_t s,
                               const void* res_in, void* res_out, const void* norm_w, float eps, int has_res);
int32_t t3b_linear_umma_launch(const void* x, const void* w, const t3b_linear_t* awq, int rows, int N, int K, void* y, float* partials,
                               int n_split, int epi, cudaStream_t s, const t3b_linear_fuse_t* fz);
int t3b_umma_pick_split(int rows, int N, int K);
void t3b_umma_set_prefetch(const void* pool, const int32_t* ctx, const int32_t* bt, int rows, int heads, int max_blocks);
int t3b_umma_pick_split_partials(int rows, int N, int K);
int t3b_umma_available();
int32_t t3b_sampler_launch(const void* head, const float* head_partials, int n_split, const float* logits_in, int V, int n_req,
                           const t3b_sampling_t* sp, const t3b_state_t* st, int advance, int32_t* out_tokens, float* logits_out,
                           cudaStream_t stream);
int32_t t3b_attn_fused_launch(const void* qkv, int n_tok, const int32_t* tok_idx, const int32_t* rope_pos, const void* rope_cos_sin, int max_pos,
                              const int32_t* block_tables, int max_blocks, void* k_pool, void* v_pool, int n_heads, void* out, cudaStream_t s);
int32_t t3b_paged_attn_launch(const void* q, int n_tok, const int32_t* tok_row, const int32_t* tok_idx, const int32_t* block_tables,
                              int max_blocks, const void* k_pool, const void* v_pool, int n_heads, void* out, int warps, cudaStream_t s);

#define GEMV_MAX_ROWS_DEFAULT 2       // measured: 4 rows run 0.87 ms/step on the tcgen05 path, 0.96 on the GEMV (profiles/r02_notes.md)
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
extern "C" int32_t t3b_version(void) { return 100; }

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
    char *x_emb, *resid, *normed, *qkv, *q, *attn, *act, *lin, *head, *z, *zss, *counters, *resid2;
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
    w.z = take(R * D * 4);                       // fp32 residual stream of the fused-norm decode layers
    w.zss = take((D / 128) * R * 4);
    w.resid2 = take((R < 4 ? R : 4) * D * 2);    // ping-pong residual of the small-batch path (rows <= 4)
    w.counters = take(256);                      // tail-norm arrival counters (zeroed at the start of every decode step)
    w.total = static_cast<size_t>(p - static_cast<char*>(base));
    return w;
}

}  // namespace

extern "C" size_t t3b_workspace_bytes(const t3b_model_t* m, int32_t max_rows) {
    if (!m || max_rows <= 0) return 0;
    return carve(m, max_rows, nullptr).total + 256;
}

// rows up to which the SIMT GEMV is picked over the tcgen05 GEMM (T3B_GEMV_MAX_ROWS, 0..4)
static int gemv_max_rows() {
    static const int v = []() { const char* e = getenv("T3B_GEMV_MAX_ROWS"); const int r = e ? atoi(e) : GEMV_MAX_ROWS_DEFAULT; return r < 0 ? 0 : (r > 4 ? 4 : r); }();
    return v;
}

// ------------------------------------------------------------------ linear dispatch
// epi: 0 = y[rows, n]; 1 = gate_up with the SiLU*mul epilogue -> y[rows, n/2].
static int32_t linear_any(const void* x, const t3b_linear_t* w, int rows, void* y, float* partials, int n_split, int* n_split_out, int impl,
                          int epi, cudaStream_t s) {
    T3B_CHECK_ARG(x && w && w->w, "linear: null pointer");
    T3B_CHECK_ARG(rows >= 0 && w->n > 0 && w->k > 0, "linear: bad sizes rows=%d n=%d k=%d", rows, w->n, w->k);
    T3B_CHECK_ALIGN(x); T3B_CHECK_ALIGN(w->w);
    if (y) T3B_CHECK_ALIGN(y);
    if (partials) T3B_CHECK_ALIGN(partials);
    if (rows == 0) { if (n_split_out) *n_split_out = 1; return T3B_OK; }
    const bool awq = w->kind == T3B_W_AWQ;
    T3B_CHECK_ARG(awq || w->kind == T3B_W_BF16, "linear: unknown weight kind %d", w->kind);
    if (impl == T3B_IMPL_AUTO) impl = (!awq && rows <= gemv_max_rows() && w->k % 1024 == 0 && w->k <= 8192) ? T3B_IMPL_GEMV : T3B_IMPL_UMMA;
    if (impl == T3B_IMPL_GEMV) {
        T3B_CHECK_ARG(!awq, "linear: AWQ weights run on the dequant-fused tcgen05 kernel at every row count (no GEMV)");
        T3B_CHECK_ARG(n_split <= 1, "linear: the GEMV path has no split-K (n_split %d)", n_split);
        T3B_CHECK_ARG(y != nullptr, "linear: y is null");
        if (n_split_out) *n_split_out = 1;
        return t3b_linear_gemv_launch(x, w->w, rows, w->n, w->k, y, epi, s, nullptr, nullptr, nullptr, 0.f, 0);
    }
    T3B_CHECK_ARG(impl == T3B_IMPL_UMMA, "linear: unknown impl %d", impl);
    const bool legacy = partials != nullptr && n_split > 1;      // explicit request for global fp32 partials
    if (n_split <= 0) n_split = t3b_umma_pick_split(rows, w->n, w->k);
    if (!legacy && y == nullptr) T3B_FAIL(T3B_ERR_ARG, "linear: y is null (partials are only written for an explicit n_split > 1)");
    if (n_split_out) *n_split_out = n_split;
    return t3b_linear_umma_launch(x, w->w, awq ? w : nullptr, rows, w->n, w->k, y, legacy ? partials : nullptr, n_split, epi, s, nullptr);
}

extern "C" int32_t t3b_linear_fused(const void* x, const t3b_linear_t* w, int32_t rows, void* y, const t3b_linear_fuse_t* f, int32_t n_split,
                                    int32_t epilogue, void* stream) {
    T3B_CHECK_ARG(w && w->w && (w->kind == T3B_W_BF16 || w->kind == T3B_W_AWQ), "t3b_linear_fused: bad weights");
    T3B_CHECK_ARG(rows >= 0 && w->n > 0 && w->k > 0, "t3b_linear_fused: bad sizes rows=%d n=%d k=%d", rows, w->n, w->k);
    T3B_CHECK_ARG(epilogue >= T3B_EPI_NONE && epilogue <= T3B_EPI_RESIDUAL, "t3b_linear_fused: unknown epilogue %d", epilogue);
    T3B_CHECK_ARG(x || (f && f->x_z), "t3b_linear_fused: no input (x and f->x_z are null)");
    if (x) T3B_CHECK_ALIGN(x);
    if (y) T3B_CHECK_ALIGN(y);
    if (f && f->x_z) { T3B_CHECK_ALIGN(f->x_z); T3B_CHECK_ALIGN(f->norm_w); }
    if (rows == 0) return T3B_OK;
    const bool awq = w->kind == T3B_W_AWQ;
    if (awq) T3B_CHECK_ARG(w->scales && w->zeros, "t3b_linear_fused: scales / zeros are null");
    if (n_split <= 0) n_split = t3b_umma_pick_split(rows, w->n, w->k);
    return t3b_linear_umma_launch(x, w->w, awq ? w : nullptr, rows, w->n, w->k, y, nullptr, n_split, epilogue, t3b_stream(stream), f);
}

extern "C" int32_t t3b_linear_bf16(const void* x, const t3b_linear_t* w, int32_t rows, void* y, float* partials, int32_t n_split,
                                   int32_t* n_split_out, int32_t impl, int32_t epilogue, void* stream) {
    T3B_CHECK_ARG(w && w->kind == T3B_W_BF16, "t3b_linear_bf16: weight kind must be T3B_W_BF16");
    T3B_CHECK_ARG(epilogue == T3B_EPI_NONE || epilogue == T3B_EPI_SILU_MUL, "t3b_linear_bf16: unknown epilogue %d", epilogue);
    return linear_any(x, w, rows, y, partials, n_split, n_split_out, impl, epilogue, t3b_stream(stream));
}

extern "C" int32_t t3b_linear_awq(const void* x, const t3b_linear_t* w, int32_t rows, void* y, float* partials, int32_t n_split,
                                  int32_t* n_split_out, int32_t impl, int32_t epilogue, void* stream) {
    T3B_CHECK_ARG(w && w->kind == T3B_W_AWQ, "t3b_linear_awq: weight kind must be T3B_W_AWQ");
    T3B_CHECK_ARG(w->scales && w->zeros, "t3b_linear_awq: scales / zeros are null");
    T3B_CHECK_ARG(epilogue == T3B_EPI_NONE || epilogue == T3B_EPI_SILU_MUL, "t3b_linear_awq: unknown epilogue %d", epilogue);
    return linear_any(x, w, rows, y, partials, n_split, n_split_out, impl, epilogue, t3b_stream(stream));
}

// ------------------------------------------------------------------ whole step
static int g_fuse_norm = -1;        // -1: from the environment (T3B_FUSE_NORM, default off)
extern "C" void t3b_debug_set_fuse_norm(int v) { g_fuse_norm = v; }
static int g_gemv_norm = -1;        // -1: from the environment (T3B_GEMV_NORM, default on)
extern "C" void t3b_debug_set_gemv_norm(int v) { g_gemv_norm = v; }
static int g_kv_prefetch = -1;      // -1: from the environment (T3B_KV_PREFETCH, default off)
static int g_tail_norm = -1;        // -1: from the environment (T3B_TAIL_NORM, default off)
extern "C" void t3b_debug_set_tail_norm(int v) { g_tail_norm = v; }
static int32_t check_model(const t3b_model_t* m) {
    T3B_CHECK_ARG(m && m->layers && m->final_norm && m->speech_emb && m->speech_pos_emb && m->rope_cos_sin && m->head.w,
                  "model: null pointer");
    T3B_CHECK_SHAPE(m->head_dim == T3B_HEAD_DIM, "model: head_dim %d (kernels are specialised for 64)", m->head_dim);
    T3B_CHECK_SHAPE(m->hidden == m->n_heads * m->head_dim, "model: hidden %d != heads %d x 64", m->hidden, m->n_heads);
    T3B_CHECK_SHAPE(m->hidden % 128 == 0 && m->ffn % 128 == 0 && m->hidden <= 4096, "model: hidden %d / ffn %d unsupported", m->hidden, m->ffn);
    T3B_CHECK_SHAPE(m->n_layers >= 1 && m->speech_vocab >= 2 && m->speech_pos_rows >= 1 && m->max_pos >= 1, "model: bad sizes");
    return T3B_OK;
}

// F2: the 30-layer stack over n_tok tokens (decode: one token per CFG row; prefill: the whole prompt of
// each new row, causal through tok_idx).  x0 = bf16 [n_tok, hidden] input embeddings.
// Decode layer = 7 launches: add+norm, qkv GEMM, attention (RoPE + KV write fused), o GEMM, add+norm,
// gate_up GEMM (SiLU*mul epilogue), down GEMM.
static int32_t forward_impl(const t3b_model_t* m, const t3b_kv_t* kv, const void* x0, int n_tok, const int32_t* tok_row,
                            const int32_t* tok_idx, const int32_t* rope_pos, const int32_t* block_tables, int max_blocks,
                            void* hidden_out, const Workspace& ws, void* stream) {
    cudaStream_t s = t3b_stream(stream);
    const int D = m->hidden, H = m->n_heads;
    int32_t rc;
    // Multi-token passes (tok_row given = prefill) fix every launch shape that changes fp32 summation order
    // (split-K = 1, 2 warps per attention CTA) so a request's prefill is bit-identical no matter which other
    // requests are admitted with it.  Decode steps pick shapes by row count (identical within a slot bucket).
    const bool prefill = tok_row != nullptr;
    auto lin = [&](const void* x, const t3b_linear_t* w, void* y, int epi) -> int32_t {
        return linear_any(x, w, n_tok, y, nullptr, prefill ? 1 : 0, nullptr, prefill ? T3B_IMPL_UMMA : T3B_IMPL_AUTO, epi, s);
    };
    // Decode steps on the tcgen05 path (rows > 4): F2.1 lives inside the projections (t3b_linear_fused) - 5 launches per
    // layer: qkv (norm fused into the operand build), attention, o (+ residual), gate_up (norm, SiLU*mul), down (+ residual).
    // opt-in (T3B_FUSE_NORM=1): measured on B200 at 64 rows the 5-launch layer is slower than the 7-launch one (18.0k vs 19.9k
    // tokens/s, profiles/r01j_notes.md) - the SIMT operand build sits on the GEMM's critical path where the TMA load did not
    if (g_fuse_norm < 0) { const char* e = getenv("T3B_FUSE_NORM"); g_fuse_norm = (e && e[0] == '1') ? 1 : 0; }
    if (!prefill && n_tok > 4 && g_fuse_norm) {
        float* z = reinterpret_cast<float*>(ws.z);
        float* zss = reinterpret_cast<float*>(ws.zss);
        rc = t3b_z_init(x0, n_tok, D, z, zss, stream);
        if (rc) return rc;
        for (int li = 0; li < m->n_layers; ++li) {
            const t3b_layer_t* L = &m->layers[li];
            T3B_CHECK_ARG(L->ln1 && L->ln2, "forward: layer %d norm weights are null", li);
            t3b_linear_fuse_t fin = {z, zss, D / 128, L->ln1, m->rms_eps, nullptr, nullptr, nullptr};
            t3b_linear_fuse_t fout = {nullptr, nullptr, 0, nullptr, 0.f, z, z, zss};
            rc = t3b_linear_fused(nullptr, &L->qkv, n_tok, ws.qkv, &fin, 0, T3B_EPI_NONE, stream);
            if (rc) return rc;
            bf16* kp = static_cast<bf16*>(kv->k) + (size_t)li * kv->layer_stride;
            bf16* vp = static_cast<bf16*>(kv->v) + (size_t)li * kv->layer_stride;
            rc = t3b_attn_fused_launch(ws.qkv, n_tok, tok_idx, rope_pos, m->rope_cos_sin, m->max_pos, block_tables, max_blocks, kp, vp, H,
                                       ws.attn, s);
            if (rc) return rc;
            rc = t3b_linear_fused(ws.attn, &L->o, n_tok, nullptr, &fout, 0, T3B_EPI_RESIDUAL, stream);
            if (rc) return rc;
            fin.norm_w = L->ln2;
            rc = t3b_linear_fused(nullptr, &L->gate_up, n_tok, ws.act, &fin, 0, T3B_EPI_SILU_MUL, stream);
            if (rc) return rc;
            rc = t3b_linear_fused(ws.act, &L->down, n_tok, nullptr, &fout, 0, T3B_EPI_RESIDUAL, stream);
            if (rc) return rc;
        }
        return t3b_z_rmsnorm(z, zss, m->final_norm, n_tok, D, m->rms_eps, hidden_out, stream);
    }
    // Decode steps on the tcgen05 path, optional: the add + RMSNorm after the o and down projections runs in the tail of those
    // GEMMs (t3b_linear_fuse_t tail_*), bit-identical to t3b_add_rmsnorm: 5 launches per layer instead of 7.
    // opt-in (T3B_TAIL_NORM=1): bit-identical but measured slower on B200 (18.4k vs 21.1k tokens/s at 64 rows, profiles/r01k_notes.md):
    // fence + arrival atomic + a single block's norm cost more than a PDL-chained 64-block kernel whose launch is already hidden
    if (g_tail_norm < 0) { const char* e = getenv("T3B_TAIL_NORM"); g_tail_norm = (e && e[0] == '1') ? 1 : 0; }
    const bool tail = !prefill && n_tok > 4 && g_tail_norm && D == 1024 && m->layers[0].o.kind == T3B_W_BF16 &&
                      t3b_umma_pick_split(n_tok, D, D) > 1 && t3b_umma_pick_split(n_tok, D, m->ffn) > 1 && n_tok <= 1024;
    if (tail) {
        T3B_CUDA(cudaMemsetAsync(ws.counters, 0, 256, s));
        rc = t3b_add_rmsnorm(x0, nullptr, 1, ws.resid, m->layers[0].ln1, ws.normed, n_tok, D, m->rms_eps, 0, stream);
        if (rc) return rc;
        for (int li = 0; li < m->n_layers; ++li) {
            const t3b_layer_t* L = &m->layers[li];
            T3B_CHECK_ARG(L->ln1 && L->ln2, "forward: layer %d norm weights are null", li);
            const bool last = li + 1 == m->n_layers;
            t3b_linear_fuse_t fo = {};
            fo.eps = m->rms_eps; fo.tail_resid = ws.resid; fo.tail_norm_w = L->ln2; fo.tail_out = ws.normed;
            fo.tail_counters = reinterpret_cast<uint32_t*>(ws.counters);
            t3b_linear_fuse_t fd = fo;
            fd.tail_norm_w = last ? m->final_norm : m->layers[li + 1].ln1;
            fd.tail_out = last ? hidden_out : ws.normed;
            rc = lin(ws.normed, &L->qkv, ws.qkv, T3B_EPI_NONE);
            if (rc) return rc;
            bf16* kp = static_cast<bf16*>(kv->k) + (size_t)li * kv->layer_stride;
            bf16* vp = static_cast<bf16*>(kv->v) + (size_t)li * kv->layer_stride;
            rc = t3b_attn_fused_launch(ws.qkv, n_tok, tok_idx, rope_pos, m->rope_cos_sin, m->max_pos, block_tables, max_blocks, kp, vp, H,
                                       ws.attn, s);
            if (rc) return rc;
            rc = t3b_linear_fused(ws.attn, &L->o, n_tok, ws.lin, &fo, 0, T3B_EPI_NONE, stream);
            if (rc) return rc;
            rc = lin(ws.normed, &L->gate_up, ws.act, T3B_EPI_SILU_MUL);
            if (rc) return rc;
            rc = t3b_linear_fused(ws.act, &L->down, n_tok, ws.lin, &fd, 0, T3B_EPI_NONE, stream);
            if (rc) return rc;
        }
        return T3B_OK;
    }
    // opt-in (T3B_KV_PREFETCH=1): measured slower (17.3k vs 21.4k tokens/s, profiles/r01n_notes.md) - the prefetch warp keeps
    // every GEMM block alive for microseconds after its epilogue and competes with the weight stream
    // Small batch (rows <= 4, bf16): F2.1 runs in front of the qkv / gate_up GEMVs (each block normalises the <= 4 rows it needs
    // anyway; block 0 writes the residual to the other ping-pong buffer) - 5 launches per layer, bit-identical to the 7-launch path.
    if (g_gemv_norm < 0) { const char* e = getenv("T3B_GEMV_NORM"); g_gemv_norm = (e && e[0] == '0') ? 0 : 1; }
    const bool small = !prefill && n_tok <= gemv_max_rows() && g_gemv_norm && D == 1024 && m->layers[0].qkv.kind == T3B_W_BF16 &&
                       m->layers[0].gate_up.kind == T3B_W_BF16 && m->ffn % 1024 == 0;
    if (small) {
        char* rcur = ws.resid;
        char* roth = ws.resid2;
        const void* xin = x0;
        for (int li = 0; li < m->n_layers; ++li) {
            const t3b_layer_t* L = &m->layers[li];
            T3B_CHECK_ARG(L->ln1 && L->ln2, "forward: layer %d norm weights are null", li);
            rc = t3b_linear_gemv_launch(xin, L->qkv.w, n_tok, L->qkv.n, L->qkv.k, ws.qkv, 0, s, rcur, roth, L->ln1, m->rms_eps, li > 0);
            if (rc) return rc;
            { char* t = rcur; rcur = roth; roth = t; }
            bf16* kp = static_cast<bf16*>(kv->k) + (size_t)li * kv->layer_stride;
            bf16* vp = static_cast<bf16*>(kv->v) + (size_t)li * kv->layer_stride;
            rc = t3b_attn_fused_launch(ws.qkv, n_tok, tok_idx, rope_pos, m->rope_cos_sin, m->max_pos, block_tables, max_blocks, kp, vp, H,
                                       ws.attn, s);
            if (rc) return rc;
            rc = lin(ws.attn, &L->o, ws.lin, T3B_EPI_NONE);
            if (rc) return rc;
            rc = t3b_linear_gemv_launch(ws.lin, L->gate_up.w, n_tok, L->gate_up.n, L->gate_up.k, ws.act, 1, s, rcur, roth, L->ln2, m->rms_eps, 1);
            if (rc) return rc;
            { char* t = rcur; rcur = roth; roth = t; }
            rc = lin(ws.act, &L->down, ws.lin, T3B_EPI_NONE);
            if (rc) return rc;
            xin = ws.lin;
        }
        return t3b_add_rmsnorm(xin, nullptr, 1, rcur, m->final_norm, hidden_out, n_tok, D, m->rms_eps, 1, stream);
    }
    if (g_kv_prefetch < 0) { const char* e = getenv("T3B_KV_PREFETCH"); g_kv_prefetch = (e && e[0] == '1') ? 1 : 0; }
    const void* prev = x0;
    for (int li = 0; li < m->n_layers; ++li) {
        const t3b_layer_t* L = &m->layers[li];
        T3B_CHECK_ARG(L->ln1 && L->ln2, "forward: layer %d norm weights are null", li);
        rc = t3b_add_rmsnorm(prev, nullptr, 1, ws.resid, L->ln1, ws.normed, n_tok, D, m->rms_eps, li > 0, stream);
        if (rc) return rc;
        rc = lin(ws.normed, &L->qkv, ws.qkv, T3B_EPI_NONE);
        if (rc) return rc;
        bf16* kp = static_cast<bf16*>(kv->k) + (size_t)li * kv->layer_stride;
        bf16* vp = static_cast<bf16*>(kv->v) + (size_t)li * kv->layer_stride;
        if (prefill) {
            rc = t3b_rope_kv_write(ws.qkv, nullptr, 1, n_tok, tok_row, tok_idx, rope_pos, m->rope_cos_sin, m->max_pos, block_tables,
                                   max_blocks, kp, vp, H, ws.q, stream);
            if (rc) return rc;
            rc = t3b_paged_attn_launch(ws.q, n_tok, tok_row, tok_idx, block_tables, max_blocks, kp, vp, H, ws.attn, 2, s);
        } else {
            rc = t3b_attn_fused_launch(ws.qkv, n_tok, tok_idx, rope_pos, m->rope_cos_sin, m->max_pos, block_tables, max_blocks, kp, vp, H,
                                       ws.attn, s);
        }
        if (rc) return rc;
        rc = lin(ws.attn, &L->o, ws.lin, T3B_EPI_NONE);
        if (rc) return rc;
        rc = t3b_add_rmsnorm(ws.lin, nullptr, 1, ws.resid, L->ln2, ws.normed, n_tok, D, m->rms_eps, 1, stream);
        if (rc) return rc;
        // decode: the idle warp of the FFN projections pulls the NEXT layer's K (gate_up) and V (down) pages into L2
        const bool pf = !prefill && g_kv_prefetch && n_tok > 4 && li + 1 < m->n_layers;
        if (pf) t3b_umma_set_prefetch(static_cast<bf16*>(kv->k) + (size_t)(li + 1) * kv->layer_stride, tok_idx, block_tables, n_tok, H, max_blocks);
        rc = lin(ws.normed, &L->gate_up, ws.act, T3B_EPI_SILU_MUL);
        if (rc) return rc;
        if (pf) t3b_umma_set_prefetch(static_cast<bf16*>(kv->v) + (size_t)(li + 1) * kv->layer_stride, tok_idx, block_tables, n_tok, H, max_blocks);
        rc = lin(ws.act, &L->down, ws.lin, T3B_EPI_NONE);
        if (rc) return rc;
        prev = ws.lin;
    }
    return t3b_add_rmsnorm(prev, nullptr, 1, ws.resid, m->final_norm, hidden_out, n_tok, D, m->rms_eps, 1, stream);
}

extern "C" int32_t t3b_forward_tokens(const t3b_model_t* m, const t3b_kv_t* kv, const void* x_in, int32_t n_tok,
                                      const int32_t* tok_row, const int32_t* tok_idx, const int32_t* rope_pos,
                                      const int32_t* block_tables, int32_t max_blocks, void* hidden_out, void* workspace,
                                      size_t workspace_bytes, void* stream) {
    int32_t rc = check_model(m);
    if (rc) return rc;
    T3B_CHECK_ARG(kv && kv->k && kv->v && x_in && hidden_out && workspace, "t3b_forward_tokens: null pointer");
    T3B_CHECK_ARG(tok_idx && rope_pos && block_tables && max_blocks > 0 && n_tok >= 0, "t3b_forward_tokens: bad token arrays");
    T3B_CHECK_ALIGN(workspace); T3B_CHECK_ALIGN(hidden_out); T3B_CHECK_ALIGN(x_in);
    if (n_tok == 0) return T3B_OK;
    T3B_CHECK_ARG(workspace_bytes >= t3b_workspace_bytes(m, n_tok), "t3b_forward_tokens: workspace %zu < %zu bytes", workspace_bytes,
                  t3b_workspace_bytes(m, n_tok));
    return forward_impl(m, kv, x_in, n_tok, tok_row, tok_idx, rope_pos, block_tables, max_blocks, hidden_out,
                        carve(m, n_tok, workspace), stream);
}

extern "C" int32_t t3b_decode_step(const t3b_model_t* m, const t3b_kv_t* kv, const t3b_state_t* st, const void* x_in, int32_t rows,
                                   void* hidden_out, void* workspace, size_t workspace_bytes, void* stream) {
    int32_t rc = check_model(m);
    if (rc) return rc;
    T3B_CHECK_ARG(kv && kv->k && kv->v && st && hidden_out && workspace, "t3b_decode_step: null pointer");
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
    return forward_impl(m, kv, x0, rows, nullptr, st->ctx_lens, st->rope_pos, st->block_tables, st->max_blocks, hidden_out, ws, stream);
}

extern "C" int32_t t3b_head_cfg_sample(const t3b_model_t* m, const void* hidden, int32_t rows, const t3b_sampling_t* sp,
                                       const t3b_state_t* st, int32_t advance, int32_t* out_tokens, float* logits_out, void* workspace,
                                       size_t workspace_bytes, void* stream) {
    int32_t rc = check_model(m);
    if (rc) return rc;
    T3B_CHECK_ARG(hidden && sp && workspace, "t3b_head_cfg_sample: null pointer");
    T3B_CHECK_ARG(rows >= 0 && rows % 2 == 0, "t3b_head_cfg_sample: rows %d must be even", rows);
    T3B_CHECK_ARG(m->head.kind == T3B_W_BF16, "t3b_head_cfg_sample: speech_head must be bf16");
    if (rows == 0) return T3B_OK;
    T3B_CHECK_ARG(workspace_bytes >= t3b_workspace_bytes(m, rows), "t3b_head_cfg_sample: workspace too small");
    const Workspace ws = carve(m, rows, workspace);
    rc = linear_any(hidden, &m->head, rows, ws.head, nullptr, 0, nullptr, T3B_IMPL_AUTO, T3B_EPI_NONE, t3b_stream(stream));
    if (rc) return rc;
    return t3b_sampler_launch(ws.head, nullptr, 1, nullptr, m->speech_vocab, rows / 2, sp, st, advance, out_tokens, logits_out,
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
