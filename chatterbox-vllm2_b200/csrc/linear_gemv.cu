// F2.2 / F2.5 / F2.6 / C1 at small batch (rows <= 4): SIMT GEMV, HBM-bound.
//
// Each warp owns FPW output features and streams their weight rows with 128-bit non-allocating
// loads, FPW*KSTEP loads (16) in flight per lane; the activations (<= 4 rows) sit in shared memory
// as bf16 and are read conflict-free.  fp32 accumulate, one warp-shuffle tree per output, bf16
// store.  The first batch of weight loads is issued BEFORE griddepcontrol.wait: weights do not
// depend on the previous kernel, so under programmatic dependent launch the HBM stream keeps
// running across the kernel boundary.
//
// mode 1 (gate_up): the warp takes FPW gate rows j and the matching up rows j+ffn and writes
// act[r][j] = bf16(bf16(silu(bf16 g)) * bf16 u)  (llama.py:117-121 + activation.py:138-141).
#include "t3b_common.cuh"

// F2.1 in front of the GEMV (rows <= 4, K == 1024): x is the previous projection's output, the kernel adds the residual, normalises
// (the arithmetic and summation order of add_rmsnorm_kernel<1>: bit-identical) into its shared-memory operand, and block 0 writes
// the updated residual to res_out (a different buffer than res_in: the other blocks still read res_in).
struct GemvNorm {
    const bf16* res_in;
    bf16* res_out;
    const bf16* norm_w;      // NULL = plain GEMV
    float eps;
    int has_res;
};

// ONE: K == KSTEP * 256, the warp's whole weight slice is the batch requested before the wait - no second register buffer (the
// double-buffered loop costs 170 registers for gate_up: one block per SM, two waves; single-shot 2 blocks per SM, one wave)
template <int R, int FPW, int KSTEP, int MODE, bool ONE>
__global__ void __launch_bounds__(256, 2) linear_gemv_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w, int rows, int N, int K,
                                                          bf16* __restrict__ y, const GemvNorm nz, const PfArgs pf) {
    __shared__ float s_red[4][4];
    extern __shared__ __align__(16) unsigned char smem_raw[];
    pdl_launch_dependents();       // early trigger: the next kernel's launch overlaps this one (it waits before reading)
    uint4* xs = reinterpret_cast<uint4*>(smem_raw);                 // [R][K/8]
    constexpr int NW = (MODE == 1) ? 2 * FPW : FPW;                 // weight rows per warp
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n_out = (MODE == 1) ? N / 2 : N;                      // features this kernel writes per row
    const int f0 = (blockIdx.x * 8 + warp) * FPW;
    const bool live = f0 < n_out;
    const int K8 = K / 8;

    const uint4* wrow[NW];
#pragma unroll
    for (int f = 0; f < NW; ++f) {
        int n = (MODE == 1) ? ((f < FPW) ? f0 + f : n_out + f0 + (f - FPW)) : f0 + f;
        n = min(n, N - 1);
        wrow[f] = reinterpret_cast<const uint4*>(w + (size_t)n * K);
    }
    uint4 wv[NW][KSTEP];
    if (live) {
#pragma unroll
        for (int f = 0; f < NW; ++f)
#pragma unroll
            for (int c = 0; c < KSTEP; ++c) wv[f][c] = ldg_stream(wrow[f] + c * 32 + lane);
    }
    if (threadIdx.x == 0) pf_slice(pf, blockIdx.x, gridDim.x);      // behind this block's own weight loads
    uint4 nwraw = make_uint4(0, 0, 0, 0);
    if (nz.norm_w != nullptr) nwraw = __ldg(reinterpret_cast<const uint4*>(nz.norm_w) + (threadIdx.x & 127));     // before the wait: constant
    pdl_wait();
    if (nz.norm_w != nullptr) {
        // K == 1024: 128 chunks of 8 per row; threads [128 g, 128 g + 128) take rows g, g + 2 (like add_rmsnorm's 128 threads per row)
        const int g = threadIdx.x >> 7, i = threadIdx.x & 127, wq = (threadIdx.x >> 5) & 3;
        float xv[2][8];
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int r = g + 2 * q;
            if (r < R) {
                float ss = 0.f;
                if (r < rows) {
                    unpack8(ld_act(reinterpret_cast<const uint4*>(x) + (size_t)r * K8 + i), xv[q]);
                    if (nz.has_res) {
                        float rv[8];
                        unpack8(ld_act(reinterpret_cast<const uint4*>(nz.res_in) + (size_t)r * K8 + i), rv);
#pragma unroll
                        for (int e = 0; e < 8; ++e) xv[q][e] += rv[e];
                    }
                    if (blockIdx.x == 0) reinterpret_cast<uint4*>(nz.res_out)[(size_t)r * K8 + i] = pack8(xv[q]);
#pragma unroll
                    for (int e = 0; e < 8; ++e) ss += xv[q][e] * xv[q][e];
                }
                ss = warp_sum(ss);
                if (lane == 0) s_red[r][wq] = ss;
            }
        }
        __syncthreads();
        float wv8[8];
        unpack8(nwraw, wv8);
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int r = g + 2 * q;
            if (r < R) {
                uint4 o4 = make_uint4(0, 0, 0, 0);
                if (r < rows) {
                    const float tot = (s_red[r][0] + s_red[r][1]) + (s_red[r][2] + s_red[r][3]);
                    const float rstd = rsqrtf(tot / (float)K + nz.eps);
                    float o[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) o[e] = bf16_round(xv[q][e] * rstd) * wv8[e];
                    o4 = pack8(o);
                }
                xs[r * K8 + i] = o4;
            }
        }
    } else {
        for (int i = threadIdx.x; i < R * K8; i += 256) {
            const int r = i / K8;
            xs[i] = r < rows ? ld_act(reinterpret_cast<const uint4*>(x) + i) : make_uint4(0, 0, 0, 0);
        }
    }
    __syncthreads();
    if (!live) return;

    float acc[NW][R];
#pragma unroll
    for (int f = 0; f < NW; ++f)
#pragma unroll
        for (int r = 0; r < R; ++r) acc[f][r] = 0.f;

    if (ONE) {
#pragma unroll
        for (int c = 0; c < KSTEP; ++c) {
            float xf[R][8];
#pragma unroll
            for (int r = 0; r < R; ++r) unpack8(xs[r * K8 + c * 32 + lane], xf[r]);
#pragma unroll
            for (int f = 0; f < NW; ++f) {
                float wf[8];
                unpack8(wv[f][c], wf);
#pragma unroll
                for (int r = 0; r < R; ++r)
#pragma unroll
                    for (int e = 0; e < 8; ++e) acc[f][r] = fmaf(wf[e], xf[r][e], acc[f][r]);
            }
        }
    } else {
        for (int k0 = 0; k0 < K8; k0 += KSTEP * 32) {
            uint4 cur[NW][KSTEP];
    #pragma unroll
            for (int f = 0; f < NW; ++f)
    #pragma unroll
                for (int c = 0; c < KSTEP; ++c) cur[f][c] = wv[f][c];
            if (k0 + KSTEP * 32 < K8) {
    #pragma unroll
                for (int f = 0; f < NW; ++f)
    #pragma unroll
                    for (int c = 0; c < KSTEP; ++c) wv[f][c] = ldg_stream(wrow[f] + k0 + KSTEP * 32 + c * 32 + lane);
            }
    #pragma unroll
            for (int c = 0; c < KSTEP; ++c) {
                float xf[R][8];
    #pragma unroll
                for (int r = 0; r < R; ++r) unpack8(xs[r * K8 + k0 + c * 32 + lane], xf[r]);
    #pragma unroll
                for (int f = 0; f < NW; ++f) {
                    float wf[8];
                    unpack8(cur[f][c], wf);
    #pragma unroll
                    for (int r = 0; r < R; ++r)
    #pragma unroll
                        for (int e = 0; e < 8; ++e) acc[f][r] = fmaf(wf[e], xf[r][e], acc[f][r]);
                }
            }
        }
    }
#pragma unroll
    for (int f = 0; f < NW; ++f)
#pragma unroll
        for (int r = 0; r < R; ++r) acc[f][r] = warp_sum(acc[f][r]);
    if (lane == 0) {
#pragma unroll
        for (int f = 0; f < FPW; ++f) {
            if (f0 + f >= n_out) break;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if (r >= rows) break;
                float v;
                if (MODE == 1) {
                    const float g = bf16_round(acc[f][r]), u = bf16_round(acc[FPW + f][r]);
                    v = bf16_round(g / (1.0f + expf(-g))) * u;
                } else {
                    v = acc[f][r];
                }
                y[(size_t)r * n_out + f0 + f] = __float2bfloat16_rn(v);
            }
        }
    }
}

template <int R, int FPW, int KSTEP, int MODE, bool ONE>
static cudaError_t launch_gemv1(const void* x, const void* w, int rows, int N, int K, void* y, cudaStream_t s, const GemvNorm& nz,
                               const PfArgs& pf) {
    const int n_out = MODE == 1 ? N / 2 : N;
    const int warps = (n_out + FPW - 1) / FPW;
    const size_t smem = (size_t)R * K * 2;
    auto kern = linear_gemv_kernel<R, FPW, KSTEP, MODE, ONE>;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    return t3b_launch(kern, dim3((warps + 7) / 8), dim3(256), smem, s, (const bf16*)x, (const bf16*)w, rows, N, K, (bf16*)y, nz, pf);
}

template <int R, int FPW, int KSTEP, int MODE>
static cudaError_t launch_gemv(const void* x, const void* w, int rows, int N, int K, void* y, cudaStream_t s, const GemvNorm& nz,
                               const PfArgs& pf) {
    if (K == KSTEP * 256) return launch_gemv1<R, FPW, KSTEP, MODE, true>(x, w, rows, N, K, y, s, nz, pf);
    return launch_gemv1<R, FPW, KSTEP, MODE, false>(x, w, rows, N, K, y, s, nz, pf);
}

template <int R, int MODE>
static cudaError_t dispatch_gemv_r(const void* x, const void* w, int rows, int N, int K, void* y, cudaStream_t s, const GemvNorm& nz,
                                   const PfArgs& pf) {
    const int n_out = MODE == 1 ? N / 2 : N;
    const int sms = t3b_num_sms();
    if (MODE == 1) return launch_gemv<R, 2, 4, 1>(x, w, rows, N, K, y, s, nz, pf);
    // few, long rows (o / down with K = 4096): one feature per warp, 16 chunks in flight
    if (K % 4096 == 0 && n_out / 4 < sms * 8) return launch_gemv<R, 1, 16, 0>(x, w, rows, N, K, y, s, nz, pf);
    if (n_out % 4 == 0 && n_out / 4 >= sms * 8) return launch_gemv<R, 4, 4, 0>(x, w, rows, N, K, y, s, nz, pf);
    if (n_out % 2 == 0 && n_out / 2 >= sms * 8) return launch_gemv<R, 2, 4, 0>(x, w, rows, N, K, y, s, nz, pf);
    return launch_gemv<R, 1, 4, 0>(x, w, rows, N, K, y, s, nz, pf);
}

// mode: 0 = y[rows, N]; 1 = gate_up with fused SiLU*mul -> y[rows, N/2]
int32_t t3b_linear_gemv_launch(const void* x, const void* w, int rows, int N, int K, void* y, int mode, cudaStream_t s,
                               const void* res_in, void* res_out, const void* norm_w, float eps, int has_res, PfArgs pf) {
    GemvNorm nz = {(const bf16*)res_in, (bf16*)res_out, (const bf16*)norm_w, eps, has_res};
    if (norm_w != nullptr) {
        T3B_CHECK_SHAPE(K == 1024, "gemv: the fused norm needs K == 1024 (got %d)", K);
        T3B_CHECK_ARG(res_out != nullptr && (!has_res || res_in != nullptr) && res_in != res_out, "gemv: fused norm needs res_out != res_in");
    }
    T3B_CHECK_SHAPE(rows >= 1 && rows <= 4, "gemv: rows %d not in [1,4]", rows);
    T3B_CHECK_SHAPE(K % 1024 == 0 && K <= 8192, "gemv: K %d must be a multiple of 1024 (<= 8192)", K);
    T3B_CHECK_SHAPE(mode == 0 || N % 4 == 0, "gemv: gate_up N %d", N);
    cudaError_t e;
    if (rows == 1) e = mode ? dispatch_gemv_r<1, 1>(x, w, rows, N, K, y, s, nz, pf) : dispatch_gemv_r<1, 0>(x, w, rows, N, K, y, s, nz, pf);
    else if (rows == 2) e = mode ? dispatch_gemv_r<2, 1>(x, w, rows, N, K, y, s, nz, pf) : dispatch_gemv_r<2, 0>(x, w, rows, N, K, y, s, nz, pf);
    else e = mode ? dispatch_gemv_r<4, 1>(x, w, rows, N, K, y, s, nz, pf) : dispatch_gemv_r<4, 0>(x, w, rows, N, K, y, s, nz, pf);
    T3B_CUDA(e);
    return T3B_OK;
}
