// F2.2 / F2.5 / F2.6 / C1 at small batch.  Two kernels: the SIMT GEMV below (rows <= 4, reads the nn.Linear layout; what a model
// without fragment-ordered weight copies runs) and the tensor-core GEMV further down (rows <= 16, linear_gemv_mma_kernel: what engines
// of <= 8 requests run).
//
// SIMT kernel:
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

// ------------------------------------------------------------------ rows <= 8 on the tensor cores (t3b_options_t.gemv_mma)
// The SIMT kernel above spends 2 R + 1 ALU instructions per 8 weight bytes: gate_up is issue-bound from R = 2 on (4.2 us with
// L2-resident weights, scripts/probe/gemv_probe.py) and R = 8 would be four times that.  Here a warp holds a 16-feature x 32 * STEPS
// slice of W as mma.sync.m16n8k16 A fragments loaded STRAIGHT from global memory, before the dependency wait: lane (g, t) =
// (lane >> 2, lane & 3) holds 16 bytes of feature g and of feature g + 8 at k = 32 s + 8 t per step s.  Read from the nn.Linear
// layout that is 16 rows x 64 bytes per load instruction (measured: qkv 4.3 us vs 2.5 for the SIMT kernel on HBM-resident weights,
// profiles/r5d_gemv_probe.log), so the kernel reads a copy of W in fragment order (t3b_repack_weights_gemv, 512 contiguous bytes
// per load instruction; t3b_linear_t.scales of a bf16 projection).  The k order inside an MMA is
// permuted (k 2t, 2t+1, 2t+8, 2t+9 of the instruction <- 8t + {0,1,2,3} / 8t + {4,5,6,7} of the slice), which a dot product does not
// see as long as the B fragment (the activations, 8 rows = the n dimension, from padded shared memory) uses the same permutation.
// The WARPS warps of a block split K; their 16 x 8 fp32 partials meet in shared memory and are summed in warp order
// (deterministic, independent of the row count: a request's numbers do not depend on how many rows ride along).
// MODE 1 (gate_up): tile rows 0..7 = gate features, 8..15 = the matching up features, so a thread owns both halves of a SiLU * mul.
// TPB = 16-feature tiles per block (2 for gate_up: half as many blocks re-run the fused normalisation).
__device__ __forceinline__ void gemv_mma16816(float* c, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

template <int STEPS, int WARPS, int TPB, int NT, int MODE>
__global__ void __launch_bounds__(WARPS * 32) linear_gemv_mma_kernel(const bf16* __restrict__ x, const uint4* __restrict__ wf, int rows, int N,
                                                                     bf16* __restrict__ y, const GemvNorm nz, const PfArgs pf) {
    constexpr int K = STEPS * 32 * WARPS, K8 = K / 8;
    constexpr int XROW = K * 2 + 64;                               // bytes per activation row: +64 keeps the B-fragment loads conflict-free
    __shared__ float s_red[8][4];
    extern __shared__ __align__(16) unsigned char smem_mm[];
    unsigned char* xs = smem_mm;                                    // [rows][XROW]
    float* part = reinterpret_cast<float*>(smem_mm + (size_t)rows * XROW);      // [WARPS][TPB][NT][128]
    pdl_launch_dependents();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    const int n_out = (MODE == 1) ? N / 2 : N;
    const int kb = warp * STEPS * 32 + t * 8;                       // this lane's first k (of the activations; the weights come pre-ordered)
    uint4 wlo[TPB][STEPS], whi[TPB][STEPS];
#pragma unroll
    for (int tl = 0; tl < TPB; ++tl) {
        // fragment-ordered copy (gemv_frag_kernel): [tile][k-step of 32][lo | hi][lane] x 16 bytes - a warp's load is 512 contiguous
        // bytes, a warp's slice STEPS KB, a block's tile 32 K bytes
        const uint4* p = wf + ((size_t)(blockIdx.x * TPB + tl) * (K / 32) + warp * STEPS) * 64 + lane;
#pragma unroll
        for (int st = 0; st < STEPS; ++st) { wlo[tl][st] = ldg_stream(p + st * 64); whi[tl][st] = ldg_stream(p + st * 64 + 32); }
    }
    // behind this block's own weight loads: a later launch's weights -> L2, one piece per lane of the last warp.  (This branch also keeps
    // ptxas from hoisting the dependency wait above the weight loads - see the guard in tests/test_host_cpu.py.)
    if (warp == WARPS - 1) pf_slice_warp(pf, blockIdx.x, gridDim.x, lane);
    uint4 nwraw = make_uint4(0, 0, 0, 0);
    if (nz.norm_w != nullptr) nwraw = __ldg(reinterpret_cast<const uint4*>(nz.norm_w) + (threadIdx.x & 127));     // before the wait: constant
    pdl_wait();
    if (nz.norm_w != nullptr) {
        // F2.1 in front (K == 1024, 256 threads): threads [128 h, 128 h + 128) take rows h, h + 2, h + 4, h + 6 - per row the arithmetic
        // and summation order of add_rmsnorm_kernel<1> (128 threads x 8 elements, warp sums, (s0 + s1) + (s2 + s3))
        const int h = threadIdx.x >> 7, i = threadIdx.x & 127, wq = (threadIdx.x >> 5) & 3;
        float wv8[8];
        unpack8(nwraw, wv8);
        auto stage_group = [&](const int r0) {
        // every load of the up to four rows goes out first: written as one loop per row, the store of the updated residual (which
        // may alias the inputs as far as the compiler knows) pinned each row's loads behind the previous row's store - four L2 round
        // trips back to back at 8 rows (profiles/r2_notes.md r5r); the four warp reductions are interleaved for the same reason
        uint4 xr[4], rr[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int r = r0 + h + 2 * q;
            xr[q] = rr[q] = make_uint4(0, 0, 0, 0);
            if (r < rows) {
                xr[q] = ld_act(reinterpret_cast<const uint4*>(x) + (size_t)r * K8 + i);
                if (nz.has_res) rr[q] = ld_act(reinterpret_cast<const uint4*>(nz.res_in) + (size_t)r * K8 + i);
            }
        }
        float xv[4][8], ss[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float rv[8];
            unpack8(xr[q], xv[q]);
            unpack8(rr[q], rv);
            ss[q] = 0.f;
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                if (nz.has_res) xv[q][e] += rv[e];
                ss[q] += xv[q][e] * xv[q][e];
            }
        }
        if (blockIdx.x == 0) {
#pragma unroll
            for (int q = 0; q < 4; ++q)
                if (r0 + h + 2 * q < rows) reinterpret_cast<uint4*>(nz.res_out)[(size_t)(r0 + h + 2 * q) * K8 + i] = pack8(xv[q]);
        }
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1)
#pragma unroll
            for (int q = 0; q < 4; ++q) ss[q] += __shfl_xor_sync(0xffffffffu, ss[q], o);
        if (lane == 0) {
#pragma unroll
            for (int q = 0; q < 4; ++q) s_red[h + 2 * q][wq] = ss[q];
        }
        __syncthreads();
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int r = r0 + h + 2 * q;
            if (r < rows) {
                const float* sr = s_red[h + 2 * q];
                const float tot = (sr[0] + sr[1]) + (sr[2] + sr[3]);
                const float rstd = rsqrtf(tot / (float)K + nz.eps);
                float o[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) o[e] = bf16_round(xv[q][e] * rstd) * wv8[e];
                *reinterpret_cast<uint4*>(xs + (size_t)r * XROW + i * 16) = pack8(o);
            }
        }
        };
        stage_group(0);
        if (NT > 1 && rows > 8) {                                   // NT = 2: a second group of up to 8 rows
            __syncthreads();                                        // s_red of the first group has been read
            stage_group(8);
        }
    } else {
        for (int i = threadIdx.x; i < rows * K8; i += WARPS * 32) {
            const int r = i / K8, col = i - r * K8;
            *reinterpret_cast<uint4*>(xs + (size_t)r * XROW + col * 16) = ld_act(reinterpret_cast<const uint4*>(x) + i);
        }
    }
    __syncthreads();

    float c[TPB][NT][4];
#pragma unroll
    for (int tl = 0; tl < TPB; ++tl)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) c[tl][nt][0] = c[tl][nt][1] = c[tl][nt][2] = c[tl][nt][3] = 0.f;
    const unsigned char* xrow = xs + (size_t)g * XROW + (size_t)kb * 2;      // activation row 8 nt + g = column g of B operand nt
#pragma unroll
    for (int st = 0; st < STEPS; ++st) {
        uint4 xb[NT];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            xb[nt] = make_uint4(0, 0, 0, 0);
            if (nt * 8 + g < rows) xb[nt] = *reinterpret_cast<const uint4*>(xrow + (size_t)nt * 8 * XROW + st * 64);
        }
#pragma unroll
        for (int tl = 0; tl < TPB; ++tl)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                gemv_mma16816(c[tl][nt], wlo[tl][st].x, whi[tl][st].x, wlo[tl][st].y, whi[tl][st].y, xb[nt].x, xb[nt].y);
                gemv_mma16816(c[tl][nt], wlo[tl][st].z, whi[tl][st].z, wlo[tl][st].w, whi[tl][st].w, xb[nt].z, xb[nt].w);
            }
    }
    // c[0..3] = (feature g, rows 2t / 2t+1), (feature g + 8, rows 2t / 2t+1) of row group nt: slot (g * 4 + t) * 4 + j of a partial tile
#pragma unroll
    for (int tl = 0; tl < TPB; ++tl)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
            *reinterpret_cast<float4*>(part + ((size_t)((warp * TPB + tl) * NT + nt) * 32 + lane) * 4) =
                make_float4(c[tl][nt][0], c[tl][nt][1], c[tl][nt][2], c[tl][nt][3]);
    __syncthreads();
    if (MODE == 1) {
        // thread -> (tile, row, gate feature f): 8 adjacent features of a row are adjacent in y
        for (int i = threadIdx.x; i < TPB * NT * 64; i += WARPS * 32) {
            const int f = i & 7, r = (i >> 3) % (8 * NT), tl = i / (64 * NT);
            if (r >= rows) continue;
            const int nt = r >> 3, rr = r & 7;
            const int slot = ((f * 4 + (rr >> 1)) * 4) + (rr & 1);
            float gs = 0.f, us = 0.f;
#pragma unroll
            for (int wp = 0; wp < WARPS; ++wp) {
                gs += part[(size_t)((wp * TPB + tl) * NT + nt) * 128 + slot];
                us += part[(size_t)((wp * TPB + tl) * NT + nt) * 128 + slot + 2];
            }
            const float gg = bf16_round(gs), u = bf16_round(us);
            y[(size_t)r * n_out + (blockIdx.x * TPB + tl) * 8 + f] = __float2bfloat16_rn(bf16_round(gg / (1.0f + expf(-gg))) * u);
        }
    } else {
        for (int i = threadIdx.x; i < TPB * NT * 128; i += WARPS * 32) {
            const int f = i & 15, r = (i >> 4) % (8 * NT), tl = i / (128 * NT);
            const int n = (blockIdx.x * TPB + tl) * 16 + f;
            if (r >= rows || n >= N) continue;
            const int nt = r >> 3, rr = r & 7;
            const int slot = (((f & 7) * 4 + (rr >> 1)) * 4) + (f >> 3) * 2 + (rr & 1);
            float a = 0.f;
#pragma unroll
            for (int wp = 0; wp < WARPS; ++wp) a += part[(size_t)((wp * TPB + tl) * NT + nt) * 128 + slot];
            y[(size_t)r * N + n] = __float2bfloat16_rn(a);
        }
    }
}

template <int STEPS, int WARPS, int TPB, int NT, int MODE>
static cudaError_t launch_gemv_mma(const void* x, const void* wf, int rows, int N, void* y, cudaStream_t s, const GemvNorm& nz,
                                   const PfArgs& pf) {
    constexpr int K = STEPS * 32 * WARPS;
    const int tiles = MODE == 1 ? N / 16 : (N + 15) / 16;
    const size_t smem = (size_t)rows * (K * 2 + 64) + (size_t)WARPS * TPB * NT * 512;
    auto kern = linear_gemv_mma_kernel<STEPS, WARPS, TPB, NT, MODE>;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    return t3b_launch(kern, dim3(tiles / TPB), dim3(WARPS * 32), smem, s, (const bf16*)x, (const uint4*)wf, rows, N, (bf16*)y, nz, pf);
}

// shapes of the tensor-core GEMV: K = 1024 (8 warps x 4 steps) or 4096 (16 warps x 8 steps); gate_up needs N / 2 % 16 == 0
bool t3b_gemv_mma_ok(int N, int K, int mode) {
    if (K == 1024) return mode ? (N % 32 == 0) : true;
    return K == 4096 && mode == 0;
}

// W bf16 [N, K] (nn.Linear) -> the fragment order of linear_gemv_mma_kernel: out[tile][ks][half][lane] (16 bytes each) =
// W[row(tile, lane >> 2, half)][32 ks + 8 (lane & 3) .. + 8); rows past N (the 8194-wide head) repeat row N - 1 (never stored).
__global__ void gemv_frag_kernel(const uint4* __restrict__ w, int N, int K, int gate_up, uint4* __restrict__ out, size_t total) {
    const int ksteps = K / 32;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int lane = (int)(i & 31), half = (int)((i >> 5) & 1);
        const size_t rest = i >> 6;
        const int ks = (int)(rest % (size_t)ksteps), tile = (int)(rest / (size_t)ksteps);
        const int g = lane >> 2, t = lane & 3;
        int row = gate_up ? (half ? N / 2 + tile * 8 + g : tile * 8 + g) : tile * 16 + g + 8 * half;
        row = min(row, N - 1);
        out[i] = w[((size_t)row * K + ks * 32 + t * 8) / 8];
    }
}
size_t t3b_gemv_frag_bytes_impl(int N, int K, int gate_up) {
    const size_t tiles = gate_up ? (size_t)N / 16 : ((size_t)N + 15) / 16;
    return tiles * 16 * (size_t)K * 2;
}
int32_t t3b_gemv_frag_launch(const void* w, int N, int K, int gate_up, void* out, cudaStream_t s) {
    T3B_CHECK_SHAPE(t3b_gemv_mma_ok(N, K, gate_up), "gemv repack: N %d K %d gate_up %d is not a shape of the tensor-core GEMV", N, K, gate_up);
    const size_t total = t3b_gemv_frag_bytes_impl(N, K, gate_up) / 16;
    gemv_frag_kernel<<<1184, 256, 0, s>>>((const uint4*)w, N, K, gate_up, (uint4*)out, total);
    T3B_LAUNCH_CHECK();
    return T3B_OK;
}

// mode: 0 = y[rows, N]; 1 = gate_up with fused SiLU*mul -> y[rows, N/2]
int32_t t3b_linear_gemv_launch(const void* x, const void* w, int rows, int N, int K, void* y, int mode, cudaStream_t s,
                               const void* res_in, void* res_out, const void* norm_w, float eps, int has_res, PfArgs pf, const void* wf) {
    GemvNorm nz = {(const bf16*)res_in, (bf16*)res_out, (const bf16*)norm_w, eps, has_res};
    if (norm_w != nullptr) {
        T3B_CHECK_SHAPE(K == 1024, "gemv: the fused norm needs K == 1024 (got %d)", K);
        T3B_CHECK_ARG(res_out != nullptr && (!has_res || res_in != nullptr) && res_in != res_out, "gemv: fused norm needs res_out != res_in");
    }
    T3B_CHECK_SHAPE(mode == 0 || N % 4 == 0, "gemv: gate_up N %d", N);
    cudaError_t e;
    if (wf != nullptr) {        // tensor-core GEMV on the fragment-ordered copy of w
        T3B_CHECK_SHAPE(t3b_gemv_mma_ok(N, K, mode), "gemv (tensor-core): N %d K %d", N, K);
        T3B_CHECK_SHAPE(rows >= 1 && rows <= 16, "gemv (tensor-core): rows %d not in [1,16]", rows);
        if (rows <= 8) {
            if (K == 4096) e = launch_gemv_mma<8, 16, 1, 1, 0>(x, wf, rows, N, y, s, nz, pf);
            else if (mode) e = launch_gemv_mma<4, 8, 2, 1, 1>(x, wf, rows, N, y, s, nz, pf);
            else e = launch_gemv_mma<4, 8, 1, 1, 0>(x, wf, rows, N, y, s, nz, pf);
        } else {             // 9..16 rows: two 8-row B operands per weight fragment (a row's numbers are the same in both shapes)
            if (K == 4096) e = launch_gemv_mma<8, 16, 1, 2, 0>(x, wf, rows, N, y, s, nz, pf);
            else if (mode) e = launch_gemv_mma<4, 8, 2, 2, 1>(x, wf, rows, N, y, s, nz, pf);
            else e = launch_gemv_mma<4, 8, 1, 2, 0>(x, wf, rows, N, y, s, nz, pf);
        }
        T3B_CUDA(e);
        return T3B_OK;
    }
    T3B_CHECK_SHAPE(rows >= 1 && rows <= 4, "gemv: rows %d not in [1,4]", rows);
    T3B_CHECK_SHAPE(K % 1024 == 0 && K <= 8192, "gemv: K %d must be a multiple of 1024 (<= 8192)", K);
    if (rows == 1) e = mode ? dispatch_gemv_r<1, 1>(x, w, rows, N, K, y, s, nz, pf) : dispatch_gemv_r<1, 0>(x, w, rows, N, K, y, s, nz, pf);
    else if (rows == 2) e = mode ? dispatch_gemv_r<2, 1>(x, w, rows, N, K, y, s, nz, pf) : dispatch_gemv_r<2, 0>(x, w, rows, N, K, y, s, nz, pf);
    else e = mode ? dispatch_gemv_r<4, 1>(x, w, rows, N, K, y, s, nz, pf) : dispatch_gemv_r<4, 0>(x, w, rows, N, K, y, s, nz, pf);
    T3B_CUDA(e);
    return T3B_OK;
}
