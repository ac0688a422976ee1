// Measurement helper (include/tdq_b200.h: tdq_ffma_probe): what the FP32 pipe sustains for the
// register patterns of the jet GEMM once operand delivery from shared memory is taken away.
#include <stdio.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/tdq_b200.h"
#include "../../tensordiffeq_b200/csrc/fused_kernel.cuh"
#include "../../tensordiffeq_b200/csrc/tmem.cuh"

namespace {

__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long ra, rb, rc, rd;
    asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rc) : "f"(c.x), "f"(c.y));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    float2 d;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
    return d;
}

// mode 1: acc[c][p][n] += a[c][p] * w[n] with scalar FFMA, every operand in a vector register
__global__ void __launch_bounds__(256) ffma_outer_kernel(const float* __restrict__ in, float* out, int iters) {
    float a[4][4], w[4], acc[4][4][4];
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int p = 0; p < 4; ++p) a[c][p] = in[(c * 4 + p) * 256 + threadIdx.x];
#pragma unroll
    for (int n = 0; n < 4; ++n) w[n] = in[(16 + n) * 256 + threadIdx.x];
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int p = 0; p < 4; ++p)
#pragma unroll
            for (int n = 0; n < 4; ++n) acc[c][p][n] = 0.f;
#pragma unroll 1
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int p = 0; p < 4; ++p)
#pragma unroll
                    for (int n = 0; n < 4; ++n) acc[c][p][n] = fmaf(a[c][p], w[n], acc[c][p][n]);
            float t = w[0]; w[0] = w[1]; w[1] = w[2]; w[2] = w[3]; w[3] = t;   // nothing hoistable
        }
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int p = 0; p < 4; ++p)
#pragma unroll
            for (int n = 0; n < 4; ++n) s += acc[c][p][n];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// mode 3: the same update with packed FFMA2 over point pairs: acc[c][pp][n].xy += a[c][pp].xy * w[n]
__global__ void __launch_bounds__(256) ffma2_outer_kernel(const float* __restrict__ in, float* out, int iters) {
    float2 a[4][2], acc[4][2][4];
    float w[4];
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int p = 0; p < 2; ++p)
            a[c][p] = make_float2(in[(c * 4 + 2 * p) * 256 + threadIdx.x], in[(c * 4 + 2 * p + 1) * 256 + threadIdx.x]);
#pragma unroll
    for (int n = 0; n < 4; ++n) w[n] = in[(16 + n) * 256 + threadIdx.x];
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int p = 0; p < 2; ++p)
#pragma unroll
            for (int n = 0; n < 4; ++n) acc[c][p][n] = make_float2(0.f, 0.f);
#pragma unroll 1
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int n = 0; n < 4; ++n) {
                const float2 ww = make_float2(w[n], w[n]);
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int p = 0; p < 2; ++p) acc[c][p][n] = ffma2(a[c][p], ww, acc[c][p][n]);
            }
            float t = w[0]; w[0] = w[1]; w[1] = w[2]; w[2] = w[3]; w[3] = t;
        }
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int p = 0; p < 2; ++p)
#pragma unroll
            for (int n = 0; n < 4; ++n) s += acc[c][p][n].x + acc[c][p][n].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// modes 10/11: the jet kernel's own GEMM loops (gemm_act_w / gemm_wgrad) run back to back on one CTA per
// SM out of shared memory, nothing else -- the rate those loops sustain in isolation.
template <int WHICH, int MAP>
__global__ void __launch_bounds__(tdq::kThreads, 1) gemm_loop_kernel(float* out, int iters) {
    using S = tdq::Shape<2, 1, 0, 128, 4>;
    extern __shared__ __align__(1024) float smem[];
    float* As = smem;
    float* Zs = smem + S::ACT;
    float* Ws = smem + 2 * S::ACT;
    for (int i = threadIdx.x; i < 2 * S::ACT + S::HP * S::HP; i += blockDim.x) smem[i] = 1e-3f * (float)((i * 2654435761u) >> 22);
    __syncthreads();
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    // MAP 0: lane = 4 * (point / row group) + (neuron / column group); MAP 1: lane = 8 * (neuron group) + (point group)
    const int l8 = MAP == 0 ? (lane >> 2) : (lane & 7), l4 = MAP == 0 ? (lane & 3) : (lane >> 3);
    const int pg = (w % S::WPG) * 8 + l8, ng = (w / S::WPG) * 4 + l4;
    const int kg = (w & 1) * 8 + l8, jg = (w >> 1) * 4 + l4;
    float sum = 0.f;
    if (WHICH == 0) {
        float acc[S::C][S::TP][4];
        for (int c = 0; c < S::C; ++c)
            for (int p = 0; p < S::TP; ++p)
                for (int n = 0; n < 4; ++n) acc[c][p][n] = 0.f;
        for (int it = 0; it < iters; ++it) {
            tdq::gemm_act_w<S>(it & 1 ? Zs : As, Ws, pg, ng * 4, acc);
            __syncthreads();
        }
        for (int c = 0; c < S::C; ++c)
            for (int p = 0; p < S::TP; ++p)
                for (int n = 0; n < 4; ++n) sum += acc[c][p][n];
    } else if (WHICH == 1) {
        constexpr int KT = S::KT, JH = KT / 2;
        for (int it = 0; it < iters; ++it) {
            for (int pass = 0; pass < 2; ++pass) {
                float g[KT][JH];
                for (int i = 0; i < KT; ++i)
                    for (int j = 0; j < JH; ++j) g[i][j] = 0.f;
                tdq::gemm_wgrad<S, JH>(As, Zs, kg * KT, jg * KT + pass * JH, g);
                for (int i = 0; i < KT; ++i)
                    for (int j = 0; j < JH; ++j) sum += g[i][j];
            }
            __syncthreads();
        }
    } else if (WHICH == 2) {
        constexpr int KT = S::KT;
        for (int it = 0; it < iters; ++it) {
            float g[KT][KT];
            for (int i = 0; i < KT; ++i)
                for (int j = 0; j < KT; ++j) g[i][j] = 0.f;
            tdq::gemm_wgrad<S, KT, false>(As, Zs, kg * KT, jg * KT, g);
            for (int i = 0; i < KT; ++i)
                for (int j = 0; j < KT; ++j) sum += g[i][j];
            __syncthreads();
        }
    } else {
        // scalar FFMA, 8x8 outputs, float4 operands along the reduction (the round-1 first version)
        constexpr int KT = S::KT, NQ = S::P / 4;
        for (int it = 0; it < iters; ++it) {
            float g[KT][KT];
            for (int i = 0; i < KT; ++i)
                for (int j = 0; j < KT; ++j) g[i][j] = 0.f;
#pragma unroll 1
            for (int u = 0; u < S::C * NQ; ++u) {
                const int c = u / NQ, q = u - c * NQ;
                float4 av[KT], zv[KT];
#pragma unroll
                for (int i = 0; i < KT; ++i) av[i] = *reinterpret_cast<const float4*>(As + ((c * S::HP + kg * KT + i) * S::PG + (q ^ S::swz(kg * KT + i))) * 4);
#pragma unroll
                for (int j = 0; j < KT; ++j) zv[j] = *reinterpret_cast<const float4*>(Zs + ((c * S::HP + jg * KT + j) * S::PG + (q ^ S::swz(jg * KT + j))) * 4);
#pragma unroll
                for (int i = 0; i < KT; ++i)
#pragma unroll
                    for (int j = 0; j < KT; ++j) {
                        float s2 = g[i][j];
                        s2 = fmaf(av[i].x, zv[j].x, s2);
                        s2 = fmaf(av[i].y, zv[j].y, s2);
                        s2 = fmaf(av[i].z, zv[j].z, s2);
                        s2 = fmaf(av[i].w, zv[j].w, s2);
                        g[i][j] = s2;
                    }
            }
            for (int i = 0; i < KT; ++i)
                for (int j = 0; j < KT; ++j) sum += g[i][j];
            __syncthreads();
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = sum;
}

// modes 40/41: forward/dgrad GEMM loop for the Cahn-Hilliard channel set (C = 6) with 2-point and 4-point thread tiles
template <int TP>
__global__ void __launch_bounds__(tdq::kThreads, 1) gemm_c6_kernel(float* out, int iters) {
    using S = tdq::Shape<4, 1, 0, 128, TP>;
    extern __shared__ __align__(1024) float smem[];
    float* As = smem;
    float* Ws = smem + S::ACT;
    for (int i = threadIdx.x; i < S::ACT + S::HP * S::HP; i += blockDim.x) smem[i] = 1e-3f * (float)((i * 2654435761u) >> 22);
    __syncthreads();
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int pg = (w % S::WPG) * 8 + (lane >> 2), ng = (w / S::WPG) * 4 + (lane & 3);
    float acc[S::C][S::TP][4];
    for (int c = 0; c < S::C; ++c)
        for (int p = 0; p < S::TP; ++p)
            for (int n = 0; n < 4; ++n) acc[c][p][n] = 0.f;
    for (int it = 0; it < iters; ++it) {
        tdq::gemm_act_w<S>(As, Ws, pg, ng * 4, acc);
        __syncthreads();
    }
    float sum = 0.f;
    for (int c = 0; c < S::C; ++c)
        for (int p = 0; p < S::TP; ++p)
            for (int n = 0; n < 4; ++n) sum += acc[c][p][n];
    out[blockIdx.x * blockDim.x + threadIdx.x] = sum;
}

// mode 20: TMEM as per-thread storage -- every thread parks 256 values in its TMEM row, reads them back
// several times; returns 1.0 if every value survived, else 0 (checked on the device)
__global__ void __launch_bounds__(256, 1) tmem_roundtrip_kernel(int* bad, int iters) {
    __shared__ uint32_t slot;
    const int w = threadIdx.x >> 5;
    if (w == 0) tdq::tmem_alloc_512(&slot);
    tdq::tmem_fence_before_sync();
    __syncthreads();
    tdq::tmem_fence_after_sync();
    const uint32_t base = slot;
    int nbad = 0;
    for (int it = 0; it < iters; ++it) {
        for (int blk = 0; blk < 16; ++blk) {
            float v[16];
            for (int i = 0; i < 16; ++i) v[i] = (float)(threadIdx.x * 1000 + blk * 16 + i + it + blockIdx.x * 7);
            tdq::tmem_st16(tdq::tmem_addr(base, w, (w >> 2) * 256 + blk * 16), v);
        }
        tdq::tmem_wait_st();
        for (int blk = 15; blk >= 0; --blk) {
            float v[16];
            tdq::tmem_ld16(tdq::tmem_addr(base, w, (w >> 2) * 256 + blk * 16), v);
            tdq::tmem_wait_ld();
            for (int i = 0; i < 16; ++i) nbad += v[i] != (float)(threadIdx.x * 1000 + blk * 16 + i + it + blockIdx.x * 7);
        }
    }
    if (nbad) atomicAdd(bad, nbad);
    tdq::tmem_fence_before_sync();
    __syncthreads();
    if (w == 0) tdq::tmem_dealloc_512(base);
}

// modes 30/31: tcgen05 3xTF32 GEMM out of shared memory, D[128x128] = X[128xK] * Y[128xK]^T with K = 64,
// X ~ hi + lo (hi = tf32-rounded, lo = remainder): D = Xhi*Yhi + Xlo*Yhi + Xhi*Ylo accumulated in TMEM.
// Operands are written in the no-swizzle K-major canonical layout (8 rows x 16 bytes core matrices).
// mode 30 returns effective TFLOP/s (2*M*N*K per triple of MMAs), mode 31 the max error relative to max |D|
// against a double-precision evaluation of the same product.
namespace t5 {
constexpr int M = 128, N = 128, K = 64;
__device__ int g_round_inputs = 0;
__device__ __forceinline__ float tf32_round(float x);
__device__ __forceinline__ float xval_raw(int m, int k) { return sinf(0.37f * (float)m + 1.3f * (float)k) * 1.7f + 0.01f * (float)(k % 7); }
__device__ __forceinline__ float yval_raw(int n, int k) { return cosf(0.11f * (float)n - 0.9f * (float)k) * 0.8f - 0.02f * (float)(n % 5); }
__device__ __forceinline__ float tf32_round(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}
__device__ __forceinline__ float xval(int m, int k) { float v = xval_raw(m, k); return g_round_inputs == 1 ? tf32_round(v) : (g_round_inputs == 2 ? rintf(v * 16.f) * 0.0625f : v); }
__device__ __forceinline__ float yval(int n, int k) { float v = yval_raw(n, k); return g_round_inputs == 1 ? tf32_round(v) : (g_round_inputs == 2 ? rintf(v * 16.f) * 0.0625f : v); }
// byte offset of element (row, k) in the canonical K-major no-swizzle layout with LBO = 128, SBO = (K/4)*128
__device__ __forceinline__ int koff(int row, int k) { return (row & 7) * 16 + (k >> 2) * 128 + (row >> 3) * ((K / 4) * 128) + (k & 3) * 4; }
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((addr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;                       // descriptor version (Blackwell)
    return d;                                     // layout_type 0 = no swizzle
}
__device__ __forceinline__ void mma_tf32(uint32_t d_tmem, uint64_t a, uint64_t b, uint32_t idesc, int accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a), "l"(b), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
        (uint32_t)__cvta_generic_to_shared(bar)) : "memory");
}
}  // namespace t5

__global__ void __launch_bounds__(256, 1) tf32x3_kernel(float* out, int iters, int check, int variant) {
    using namespace t5;
    extern __shared__ __align__(1024) unsigned char sm[];
    unsigned char* Xh = sm;
    unsigned char* Xl = sm + M * K * 4;
    unsigned char* Yh = sm + 2 * M * K * 4;
    unsigned char* Yl = sm + 3 * M * K * 4;
    __shared__ uint32_t slot;
    __shared__ uint64_t bar;
    const int tid = threadIdx.x, w = tid >> 5;
    for (int i = tid; i < M * K; i += 256) {
        int r = i / K, k = i - r * K;
        float x = xval(r, k), xh = tf32_round(x);
        float y = yval(r, k), yh = tf32_round(y);
        *reinterpret_cast<float*>(Xh + koff(r, k)) = xh;
        *reinterpret_cast<float*>(Xl + koff(r, k)) = x - xh;
        *reinterpret_cast<float*>(Yh + koff(r, k)) = yh;
        *reinterpret_cast<float*>(Yl + koff(r, k)) = y - yh;
    }
    if (tid == 0) tdq::mbar_init(&bar, 1);
    if (w == 0) tdq::tmem_alloc_512(&slot);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy smem writes -> async proxy (UMMA) reads
    tdq::tmem_fence_before_sync();
    __syncthreads();
    tdq::tmem_fence_after_sync();
    const uint32_t tbase = slot;
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    const uint32_t LBO = 128, SBO = (K / 4) * 128;
    uint32_t phase = 0;
    for (int it = 0; it < iters; ++it) {
        if (tid == 0) {
            const uint32_t xh = (uint32_t)__cvta_generic_to_shared(Xh), xl = (uint32_t)__cvta_generic_to_shared(Xl);
            const uint32_t yh = (uint32_t)__cvta_generic_to_shared(Yh), yl = (uint32_t)__cvta_generic_to_shared(Yl);
            int acc = 0;
            if (variant == 3) {            // all correction terms first, the dominant products last
                for (int ks = 0; ks < K / 8; ++ks) {
                    const uint32_t o = ks * 256;
                    mma_tf32(tbase, smem_desc(xl + o, LBO, SBO), smem_desc(yh + o, LBO, SBO), idesc, acc); acc = 1;
                    mma_tf32(tbase, smem_desc(xh + o, LBO, SBO), smem_desc(yl + o, LBO, SBO), idesc, 1);
                }
                for (int ks = 0; ks < K / 8; ++ks) {
                    const uint32_t o = ks * 256;
                    mma_tf32(tbase, smem_desc(xh + o, LBO, SBO), smem_desc(yh + o, LBO, SBO), idesc, 1);
                }
            } else
            for (int ks = 0; ks < K / 8; ++ks) {                      // one MMA = K 8 (32 bytes = 2 core matrices)
                const uint32_t o = ks * 256;
                if (variant == 1) {            // plain TF32
                    mma_tf32(tbase, smem_desc(xh + o, LBO, SBO), smem_desc(yh + o, LBO, SBO), idesc, acc); acc = 1;
                } else if (variant == 2) {     // four terms
                    mma_tf32(tbase, smem_desc(xl + o, LBO, SBO), smem_desc(yl + o, LBO, SBO), idesc, acc); acc = 1;
                    mma_tf32(tbase, smem_desc(xl + o, LBO, SBO), smem_desc(yh + o, LBO, SBO), idesc, 1);
                    mma_tf32(tbase, smem_desc(xh + o, LBO, SBO), smem_desc(yl + o, LBO, SBO), idesc, 1);
                    mma_tf32(tbase, smem_desc(xh + o, LBO, SBO), smem_desc(yh + o, LBO, SBO), idesc, 1);
                } else {
                    mma_tf32(tbase, smem_desc(xl + o, LBO, SBO), smem_desc(yh + o, LBO, SBO), idesc, acc); acc = 1;
                    mma_tf32(tbase, smem_desc(xh + o, LBO, SBO), smem_desc(yl + o, LBO, SBO), idesc, 1);
                    mma_tf32(tbase, smem_desc(xh + o, LBO, SBO), smem_desc(yh + o, LBO, SBO), idesc, 1);
                }
            }
            mma_commit(&bar);
        }
        tdq::mbar_wait(&bar, phase);
        phase ^= 1;
        tdq::tmem_fence_after_sync();
    }
    // read D: warp w (w < 4) lanes 32w..32w+31 = rows; columns 0..127
    float maxerr = 0.f, maxval = 0.f;
    if (w < 4) {
        const int m = w * 32 + (tid & 31);
        for (int cb = 0; cb < N; cb += 16) {
            float v[16];
            tdq::tmem_ld16(tdq::tmem_addr(tbase, w, cb), v);
            tdq::tmem_wait_ld();
            if (check) {
                for (int j = 0; j < 16; ++j) {
                    double ref = 0.0;
                    for (int k = 0; k < K; ++k) ref += (double)xval(m, k) * (double)yval(cb + j, k);
                    float got = v[j];
                    if (variant == 9) { got = 0.f; for (int k = 0; k < K; ++k) got = fmaf(xval(m, k), yval(cb + j, k), got); }
                    maxerr = fmaxf(maxerr, (float)fabs((double)got - ref));
                    maxval = fmaxf(maxval, (float)fabs(ref));
                }
            } else {
                maxval += v[0];
            }
        }
    }
    __shared__ float red_e[256], red_v[256];
    red_e[tid] = maxerr;
    red_v[tid] = maxval;
    tdq::tmem_fence_before_sync();
    __syncthreads();
    if (tid == 0) {
        float e = 0.f, v = 0.f;
        for (int i = 0; i < 256; ++i) { e = fmaxf(e, red_e[i]); v = fmaxf(v, red_v[i]); }
        out[blockIdx.x * 2] = e;
        out[blockIdx.x * 2 + 1] = v;
    }
    if (w == 0) tdq::tmem_dealloc_512(tbase);
}

// modes 42/43: the same 3xTF32 product with the operand layout the jet kernel already uses for its
// activations: planes of [128 rows][32 fp32] (one 128-byte line per row), 16-byte chunks XOR-swizzled
// with (row & 7) = the hardware's SWIZZLE_128B K-major canonical layout.  "hi" is the raw fp32 value (the
// tensor core ignores the low 13 mantissa bits), "lo" = x - trunc_tf32(x).  K = 64 = 2 planes.
namespace t5 {
__device__ __forceinline__ int sw128_off(int plane, int row, int k) {      // byte offset, k in [0,32)
    return plane * (128 * 128) + row * 128 + ((((k >> 2) ^ (row & 7)) & 7) << 4) + (k & 3) * 4;
}
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t addr) {
    uint64_t d = 0;
    d |= (uint64_t)((addr >> 4) & 0x3FFF);
    d |= (uint64_t)1 << 16;                          // leading byte offset (unused for swizzled K-major): 16 B
    d |= (uint64_t)((1024 >> 4) & 0x3FFF) << 32;     // stride between 8-row groups: 1024 B
    d |= (uint64_t)1 << 46;                          // descriptor version
    d |= (uint64_t)2 << 61;                          // SWIZZLE_128B
    return d;
}
}  // namespace t5

__global__ void __launch_bounds__(256, 1) tf32x3_sw128_kernel(float* out, int iters, int check) {
    using namespace t5;
    extern __shared__ __align__(1024) unsigned char sm[];
    unsigned char* Xh = sm;
    unsigned char* Xl = sm + 2 * 128 * 128;
    unsigned char* Yh = sm + 4 * 128 * 128;
    unsigned char* Yl = sm + 6 * 128 * 128;
    __shared__ uint32_t slot;
    __shared__ uint64_t bar;
    const int tid = threadIdx.x, w = tid >> 5;
    for (int i = tid; i < M * K; i += 256) {
        int r = i / K, k = i - r * K;
        float x = xval_raw(r, k), y = yval_raw(r, k);
        float xt = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u), yt = __uint_as_float(__float_as_uint(y) & 0xFFFFE000u);
        *reinterpret_cast<float*>(Xh + sw128_off(k >> 5, r, k & 31)) = x;
        *reinterpret_cast<float*>(Xl + sw128_off(k >> 5, r, k & 31)) = x - xt;
        *reinterpret_cast<float*>(Yh + sw128_off(k >> 5, r, k & 31)) = y;
        *reinterpret_cast<float*>(Yl + sw128_off(k >> 5, r, k & 31)) = y - yt;
    }
    if (tid == 0) tdq::mbar_init(&bar, 1);
    if (w == 0) tdq::tmem_alloc_512(&slot);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tdq::tmem_fence_before_sync();
    __syncthreads();
    tdq::tmem_fence_after_sync();
    const uint32_t tbase = slot;
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    uint32_t phase = 0;
    for (int it = 0; it < iters; ++it) {
        if (tid == 0) {
            const uint32_t xh = (uint32_t)__cvta_generic_to_shared(Xh), xl = (uint32_t)__cvta_generic_to_shared(Xl);
            const uint32_t yh = (uint32_t)__cvta_generic_to_shared(Yh), yl = (uint32_t)__cvta_generic_to_shared(Yl);
            int acc = 0;
            for (int plane = 0; plane < 2; ++plane)
                for (int ks = 0; ks < 4; ++ks) {                  // K = 8 per instruction = 32 bytes along the row
                    const uint32_t o = plane * (128 * 128) + ks * 32;
                    mma_tf32(tbase, smem_desc_sw128(xl + o), smem_desc_sw128(yh + o), idesc, acc); acc = 1;
                    mma_tf32(tbase, smem_desc_sw128(xh + o), smem_desc_sw128(yl + o), idesc, 1);
                    mma_tf32(tbase, smem_desc_sw128(xh + o), smem_desc_sw128(yh + o), idesc, 1);
                }
            mma_commit(&bar);
        }
        tdq::mbar_wait(&bar, phase);
        phase ^= 1;
        tdq::tmem_fence_after_sync();
    }
    float maxerr = 0.f, maxval = 0.f;
    if (w < 4) {
        const int m = w * 32 + (tid & 31);
        for (int cb = 0; cb < N; cb += 16) {
            float v[16];
            tdq::tmem_ld16(tdq::tmem_addr(tbase, w, cb), v);
            tdq::tmem_wait_ld();
            if (check) {
                for (int j = 0; j < 16; ++j) {
                    double ref = 0.0;
                    for (int k = 0; k < K; ++k) ref += (double)xval_raw(m, k) * (double)yval_raw(cb + j, k);
                    maxerr = fmaxf(maxerr, (float)fabs((double)v[j] - ref));
                    maxval = fmaxf(maxval, (float)fabs(ref));
                }
            } else {
                maxval += v[0];
            }
        }
    }
    __shared__ float red_e[256], red_v[256];
    red_e[tid] = maxerr;
    red_v[tid] = maxval;
    tdq::tmem_fence_before_sync();
    __syncthreads();
    if (tid == 0) {
        float e = 0.f, v = 0.f;
        for (int i = 0; i < 256; ++i) { e = fmaxf(e, red_e[i]); v = fmaxf(v, red_v[i]); }
        out[blockIdx.x * 2] = e;
        out[blockIdx.x * 2 + 1] = v;
    }
    if (w == 0) tdq::tmem_dealloc_512(tbase);
}

// modes 44..47: one K = 32 product with the B operand MN-major (N = 4 channel planes x 32 points, K = feature
// rows), A K-major -- the activation GEMM of tc_kernel.cuh.  Inputs on a 1/16 grid so the tf32 product is exact;
// returns max |D - ref| / max |ref|.
//   44: no swizzle, 8-row x 16-byte core matrices, SBO = stride between 4-point chunks, LBO = between row groups
//   45: the same with LBO / SBO exchanged
//   46: SWIZZLE_128B_BASE32B (rows of 128 bytes, 32-byte chunk ^= row & 3), LBO = plane stride, SBO = 512
//   47: the same with LBO / SBO exchanged
__global__ void __launch_bounds__(256, 1) tf32_mn_kernel(float* out, int variant, int iters = 0) {
    using namespace t5;
    extern __shared__ __align__(1024) unsigned char sm[];
    unsigned char* Xa = sm;                        // A: [128 rows m][32 k] K-major SW128
    unsigned char* Yb = sm + 128 * 128;            // B: 64 KB
    __shared__ uint32_t slot;
    __shared__ uint64_t bar;
    const int tid = threadIdx.x, w = tid >> 5;
    auto xv = [](int m, int k) { return rintf(xval_raw(m, k) * 16.f) * 0.0625f; };
    auto yv = [](int n, int k) { return rintf(yval_raw(n, k) * 16.f) * 0.0625f; };
    const bool core = variant < 2;
    for (int i = tid; i < 128 * 32; i += 256) {
        int r = i / 32, k = i - r * 32;
        *reinterpret_cast<float*>(Xa + sw128_off(0, r, k)) = xv(r, k);
        int n = r, c = n >> 5, p = n & 31;          // B element (n, k): channel plane c, point p, feature row k
        int off;
        if (core) off = ((c * 8 + (p >> 2)) * 16 + (k >> 3)) * 128 + (k & 7) * 16 + (p & 3) * 4;      // 16 row groups per chunk column
        else off = c * 16384 + k * 128 + ((((p >> 3) ^ (k & 3)) & 3) << 5) + (p & 7) * 4;
        *reinterpret_cast<float*>(Yb + off) = yv(n, k);
    }
    if (tid == 0) tdq::mbar_init(&bar, 1);
    if (w == 0) tdq::tmem_alloc_512(&slot);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tdq::tmem_fence_before_sync();
    __syncthreads();
    tdq::tmem_fence_after_sync();
    const uint32_t tbase = slot;
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    // modes 48..50 (iters > 0): throughput of 48 back-to-back MMAs per commit; variant 2 = B MN-major (BASE32B),
    // 4 = both K-major, 5 = A from Tensor Memory + B K-major.  out[0] = clock cycles per MMA.
    if (iters > 0) {
        if (tid == 0) {
            const uint32_t xa = (uint32_t)__cvta_generic_to_shared(Xa), yb = (uint32_t)__cvta_generic_to_shared(Yb);
            uint64_t bmn = 0;
            bmn |= (uint64_t)((yb >> 4) & 0x3FFF);
            bmn |= (uint64_t)((16384 >> 4) & 0x3FFF) << 16;
            bmn |= (uint64_t)((512 >> 4) & 0x3FFF) << 32;
            bmn |= (uint64_t)1 << 46;
            bmn |= (uint64_t)1 << 61;
            const uint64_t ak = smem_desc_sw128(xa), bk = smem_desc_sw128(yb);
            const uint32_t idk = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
            uint32_t phase = 0;
            long long t0 = clock64();
            for (int it = 0; it < iters; ++it) {
#pragma unroll 4
                for (int r = 0; r < 12; ++r)
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) {
                        if (variant == 2) mma_tf32(tbase, ak + ks * 2, bmn + ks * 64, idesc, 1);
                        else if (variant == 4) mma_tf32(tbase, ak + ks * 2, bk + ks * 2, idk, 1);
                        else if (variant == 6) mma_tf32(tbase, bmn + ks * 64, bmn + ks * 64, idesc | (1u << 15), 1);   // A and B MN-major
                        else if (variant == 7)
                            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                         "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tbase),
                                         "r"(tbase + 256 + ks * 8), "l"(bmn + ks * 64), "r"(idesc), "r"(1) : "memory");
                        else
                            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                         "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tbase),
                                         "r"(tbase + 256 + ks * 8), "l"(bk + ks * 2), "r"(idk), "r"(1) : "memory");
                    }
                mma_commit(&bar);
                tdq::mbar_wait(&bar, phase);
                phase ^= 1;
            }
            long long t1 = clock64();
            out[blockIdx.x * 2] = (float)(t1 - t0) / (float)(iters * 48);
            out[blockIdx.x * 2 + 1] = 1.f;
        }
        tdq::tmem_fence_before_sync();
        __syncthreads();
        if (w == 0) tdq::tmem_dealloc_512(tbase);
        return;
    }
    if (tid == 0) {
        const uint32_t xa = (uint32_t)__cvta_generic_to_shared(Xa), yb = (uint32_t)__cvta_generic_to_shared(Yb);
        for (int ks = 0; ks < 4; ++ks) {
            uint64_t bd = 0;
            uint32_t addr, lbo, sbo;
            if (core) { addr = yb + ks * 128; sbo = 2048; lbo = 128; }
            else { addr = yb + ks * 1024; lbo = 16384; sbo = 512; }
            if (variant & 1) { uint32_t t = lbo; lbo = sbo; sbo = t; }
            bd |= (uint64_t)((addr >> 4) & 0x3FFF);
            bd |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
            bd |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
            bd |= (uint64_t)1 << 46;
            bd |= (uint64_t)(core ? 0 : 1) << 61;
            mma_tf32(tbase, smem_desc_sw128(xa + ks * 32), bd, idesc, ks != 0);
        }
        mma_commit(&bar);
    }
    tdq::mbar_wait(&bar, 0);
    tdq::tmem_fence_after_sync();
    float maxerr = 0.f, maxval = 0.f;
    if (w < 4) {
        const int m = w * 32 + (tid & 31);
        for (int cb = 0; cb < N; cb += 16) {
            float v[16];
            tdq::tmem_ld16(tdq::tmem_addr(tbase, w, cb), v);
            tdq::tmem_wait_ld();
            for (int j = 0; j < 16; ++j) {
                double ref = 0.0;
                for (int k = 0; k < 32; ++k) ref += (double)xv(m, k) * (double)yv(cb + j, k);
                if (m == 1 && cb == 32 && j < 4) { out[2 + 2 * j] = v[j]; out[3 + 2 * j] = (float)ref; }
                maxerr = fmaxf(maxerr, (float)fabs((double)v[j] - ref));
                maxval = fmaxf(maxval, (float)fabs(ref));
            }
        }
    }
    __shared__ float red_e[256], red_v[256];
    red_e[tid] = maxerr;
    red_v[tid] = maxval;
    tdq::tmem_fence_before_sync();
    __syncthreads();
    if (tid == 0) {
        float e = 0.f, v = 0.f;
        for (int i = 0; i < 256; ++i) { e = fmaxf(e, red_e[i]); v = fmaxf(v, red_v[i]); }
        out[0] = e;
        out[1] = v;
    }
    if (w == 0) tdq::tmem_dealloc_512(tbase);
}

// modes 48..52: tensor-core throughput of one MMA shape (M = N = 128, K = 8, tf32) by operand source, 48 MMAs per
// commit issued from a fully unrolled, branch-free loop (a single thread issues: anything else in the loop is
// measured too).  V: 0 = A K-major / B MN-major (BASE32B), 1 = both K-major, 2 = A in Tensor Memory / B K-major,
// 3 = both MN-major, 4 = A in Tensor Memory / B MN-major.  out[0] = clock cycles per MMA.
template <int V>
__global__ void __launch_bounds__(256, 1) tf32_rate_kernel(float* out, int iters) {
    using namespace t5;
    extern __shared__ __align__(1024) unsigned char sm[];
    __shared__ uint32_t slot;
    __shared__ uint64_t bar;
    const int tid = threadIdx.x, w = tid >> 5;
    for (int i = tid; i < (V >= 5 ? 12 : 5) * 128 * 128 / 4; i += 256) reinterpret_cast<float*>(sm)[i] = 0.001f * (float)(i & 255);
    if (tid == 0) tdq::mbar_init(&bar, 1);
    if (w == 0) tdq::tmem_alloc_512(&slot);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tdq::tmem_fence_before_sync();
    __syncthreads();
    tdq::tmem_fence_after_sync();
    const uint32_t tbase = slot;
    if (tid == 0) {
        const uint32_t xa = (uint32_t)__cvta_generic_to_shared(sm), yb = xa + (V >= 5 ? 4 : 1) * 128 * 128;
        auto mn = [](uint32_t addr) {
            uint64_t d = 0;
            d |= (uint64_t)((addr >> 4) & 0x3FFF);
            d |= (uint64_t)((16384 >> 4) & 0x3FFF) << 16;
            d |= (uint64_t)((512 >> 4) & 0x3FFF) << 32;
            d |= (uint64_t)1 << 46;
            d |= (uint64_t)1 << 61;
            return d;
        };
        const uint64_t ak = smem_desc_sw128(xa), bk = smem_desc_sw128(yb), amn = mn(xa), bmn = mn(yb);
        constexpr uint32_t id0 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        constexpr uint32_t idesc = id0 | ((V == 0 || V == 3 || V == 4 || V >= 5) ? (1u << 16) : 0u) | (V == 3 ? (1u << 15) : 0u);
        uint32_t phase = 0;
        long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int r = 0; r < 12; ++r)
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {
                    if (V >= 5) {
                        // streaming operands as in the jet kernel: A walks 4 K-major planes (r & 3), B walks 128 feature
                        // rows of an MN-major image; V == 6 alternates two B images against the same A (hi.lo, hi.hi)
                        const int q = r & 3;
                        const uint64_t a2 = ak + (uint64_t)(q * 1024 + ks * 2);
                        const uint64_t b2 = bmn + (uint64_t)((q * 32 + ks * 8) * 8) + (uint64_t)((V == 6 && (r & 4)) ? 4096 : 0);
                        asm volatile("tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, 1;" ::"r"(tbase), "l"(a2), "l"(b2), "r"(idesc) : "memory");
                        continue;
                    }
                    const uint64_t a = V == 3 ? amn + ks * 64 : ak + ks * 2;
                    const uint64_t b = (V == 0 || V == 3 || V == 4) ? bmn + ks * 64 : bk + ks * 2;
                    if (V == 2 || V == 4)
                        asm volatile("tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, 1;" ::"r"(tbase), "r"(tbase + 256 + ks * 8),
                                     "l"(b), "r"(idesc) : "memory");
                    else
                        asm volatile("tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, 1;" ::"r"(tbase), "l"(a), "l"(b), "r"(idesc) : "memory");
                }
            mma_commit(&bar);
            tdq::mbar_wait(&bar, phase);
            phase ^= 1;
        }
        long long t1 = clock64();
        out[blockIdx.x] = (float)(t1 - t0) / (float)(iters * 48);
    }
    tdq::tmem_fence_before_sync();
    __syncthreads();
    if (w == 0) tdq::tmem_dealloc_512(tbase);
}

}  // namespace

// returns TFLOP/s (best of 3 after one warm-up) or a negative CUDA error code
static float tdq_probe_run(int mode, int n_sm, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (mode >= 48 && mode <= 54) {
        const size_t smem = (mode >= 53 ? 12 : 5) * 128 * 128;
        void (*kk)(float*, int) = mode == 48 ? tf32_rate_kernel<0> : mode == 49 ? tf32_rate_kernel<1> : mode == 50 ? tf32_rate_kernel<2>
                                  : mode == 51 ? tf32_rate_kernel<3> : mode == 52 ? tf32_rate_kernel<4> : mode == 53 ? tf32_rate_kernel<5> : tf32_rate_kernel<6>;
        cudaFuncSetAttribute(kk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        float* o = nullptr;
        if (cudaMalloc((void**)&o, (size_t)n_sm * 4) != cudaSuccess) return -1.f;
        kk<<<n_sm, 256, smem, st>>>(o, 200);
        kk<<<n_sm, 256, smem, st>>>(o, 2000);
        float h1 = -1.f;
        cudaMemcpyAsync(&h1, o, 4, cudaMemcpyDeviceToHost, st);
        cudaError_t e = cudaStreamSynchronize(st);
        cudaFree(o);
        return e == cudaSuccess ? h1 : -3.f;
    }
    if (mode >= 44 && mode <= 47) {
        const size_t smem = 5 * 128 * 128;
        cudaFuncSetAttribute(tf32_mn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        float* o = nullptr;
        if (cudaMalloc((void**)&o, 64) != cudaSuccess) return -1.f;
        tf32_mn_kernel<<<1, 256, smem, st>>>(o, mode - 44);
        float h2[10] = {-1.f, -1.f};
        cudaMemcpyAsync(h2, o, 40, cudaMemcpyDeviceToHost, st);
        cudaError_t e = cudaStreamSynchronize(st);
        cudaFree(o);
        if (e != cudaSuccess) return -3.f;
        fprintf(stderr, "mn probe %d: err %g max %g | D/ref: %g/%g %g/%g %g/%g %g/%g\n", mode, h2[0], h2[1], h2[2], h2[3], h2[4], h2[5], h2[6], h2[7], h2[8], h2[9]);
        return h2[1] > 0.f ? h2[0] / h2[1] : -4.f;
    }
    if (mode == 42 || mode == 43) {
        const size_t smem = 8 * 128 * 128;
        cudaFuncSetAttribute(tf32x3_sw128_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        float* o = nullptr;
        if (cudaMalloc((void**)&o, (size_t)n_sm * 2 * 4) != cudaSuccess) return -1.f;
        if (mode == 42) {
            tf32x3_sw128_kernel<<<1, 256, smem, st>>>(o, 1, 1);
            float h2[2] = {-1.f, -1.f};
            cudaMemcpyAsync(h2, o, 8, cudaMemcpyDeviceToHost, st);
            cudaError_t e = cudaStreamSynchronize(st);
            cudaFree(o);
            if (e != cudaSuccess) return -3.f;
            return h2[1] > 0.f ? h2[0] / h2[1] : -4.f;
        }
        cudaEvent_t a, b;
        cudaEventCreate(&a);
        cudaEventCreate(&b);
        const int it = 2000;
        float best = 0.f;
        for (int rep = 0; rep < 3; ++rep) {
            cudaEventRecord(a, st);
            tf32x3_sw128_kernel<<<n_sm, 256, smem, st>>>(o, it, 0);
            cudaEventRecord(b, st);
            if (cudaEventSynchronize(b) != cudaSuccess) { cudaFree(o); return -3.f; }
            float ms = 0.f;
            cudaEventElapsedTime(&ms, a, b);
            double fl = 2.0 * t5::M * t5::N * t5::K * (double)it * n_sm;
            float tf = (float)(fl / (ms * 1e-3) / 1e12);
            if (rep > 0 && tf > best) best = tf;
        }
        cudaEventDestroy(a);
        cudaEventDestroy(b);
        cudaFree(o);
        return best;
    }
    if (mode == 40 || mode == 41) {
        const int tp = mode == 40 ? 2 : 4;
        const size_t smem = (size_t)(6 * 128 * (tp * 8) + 128 * 128) * 4;
        void (*kk)(float*, int) = mode == 40 ? gemm_c6_kernel<2> : gemm_c6_kernel<4>;
        cudaFuncSetAttribute(kk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        float* o = nullptr;
        if (cudaMalloc((void**)&o, (size_t)n_sm * 256 * 4) != cudaSuccess) return -1.f;
        cudaEvent_t a, b;
        cudaEventCreate(&a);
        cudaEventCreate(&b);
        const int it = 200;
        float best = 0.f;
        for (int rep = 0; rep < 3; ++rep) {
            cudaEventRecord(a, st);
            kk<<<n_sm, 256, smem, st>>>(o, it);
            cudaEventRecord(b, st);
            if (cudaEventSynchronize(b) != cudaSuccess) { cudaFree(o); return -3.f; }
            float ms = 0.f;
            cudaEventElapsedTime(&ms, a, b);
            double fl = 2.0 * 6.0 * (tp * 8) * 128.0 * 128.0 * it * n_sm;
            float tf = (float)(fl / (ms * 1e-3) / 1e12);
            if (rep > 0 && tf > best) best = tf;
        }
        cudaEventDestroy(a);
        cudaEventDestroy(b);
        cudaFree(o);
        return cudaGetLastError() == cudaSuccess ? best : -2.f;
    }
    if (mode >= 30 && mode <= 39) {
        const size_t smem = 4 * (size_t)t5::M * t5::K * 4;
        cudaFuncSetAttribute(tf32x3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        float* o = nullptr;
        if (cudaMalloc((void**)&o, (size_t)n_sm * 2 * 4) != cudaSuccess) return -1.f;
        if (mode != 30) {
            int variant = mode == 31 ? 0 : (mode == 39 ? 9 : mode - 31);
            int rnd = 0;
            if (mode == 35) { variant = 1; rnd = 1; }
            if (mode == 36) { variant = 1; rnd = 2; }      // inputs on a 1/16 grid: products and sums exact in fp32
            if (mode == 37) { variant = 9; rnd = 1; }      // fp32 FMA on the tf32-rounded inputs      // tf32-exact inputs, single term: isolates product / accumulate precision
            cudaMemcpyToSymbolAsync(t5::g_round_inputs, &rnd, 4, 0, cudaMemcpyHostToDevice, st);
            tf32x3_kernel<<<1, 256, smem, st>>>(o, 1, 1, variant);
            float h2[2] = {-1.f, -1.f};
            cudaMemcpyAsync(h2, o, 8, cudaMemcpyDeviceToHost, st);
            cudaError_t e = cudaStreamSynchronize(st);
            cudaFree(o);
            if (e != cudaSuccess) return -3.f;
            return h2[1] > 0.f ? h2[0] / h2[1] : -4.f;
        }
        cudaEvent_t a, b;
        cudaEventCreate(&a);
        cudaEventCreate(&b);
        const int it = 2000;
        float best = 0.f;
        for (int rep = 0; rep < 3; ++rep) {
            cudaEventRecord(a, st);
            tf32x3_kernel<<<n_sm, 256, smem, st>>>(o, it, 0, 0);
            cudaEventRecord(b, st);
            if (cudaEventSynchronize(b) != cudaSuccess) { cudaFree(o); return -3.f; }
            float ms = 0.f;
            cudaEventElapsedTime(&ms, a, b);
            double fl = 2.0 * t5::M * t5::N * t5::K * (double)it * n_sm;
            float tf = (float)(fl / (ms * 1e-3) / 1e12);
            if (rep > 0 && tf > best) best = tf;
        }
        cudaEventDestroy(a);
        cudaEventDestroy(b);
        cudaFree(o);
        return best;
    }
    if (mode == 20) {
        int* bad = nullptr;
        if (cudaMalloc((void**)&bad, 4) != cudaSuccess) return -1.f;
        cudaMemsetAsync(bad, 0, 4, st);
        tmem_roundtrip_kernel<<<n_sm * 2, 256, 0, st>>>(bad, 8);
        int hb = -1;
        cudaMemcpyAsync(&hb, bad, 4, cudaMemcpyDeviceToHost, st);
        cudaError_t e = cudaStreamSynchronize(st);
        cudaFree(bad);
        if (e != cudaSuccess) return -3.f;
        return hb == 0 ? 1.f : 0.f;
    }
    if (mode >= 10) {
        using S = tdq::Shape<2, 1, 0, 128, 4>;
        const size_t smem = (size_t)(2 * S::ACT + S::HP * S::HP) * 4;
        void (*ks[8])(float*, int) = {gemm_loop_kernel<0, 0>, gemm_loop_kernel<1, 0>, gemm_loop_kernel<2, 0>, gemm_loop_kernel<3, 0>,
                                      gemm_loop_kernel<0, 1>, gemm_loop_kernel<1, 1>, gemm_loop_kernel<2, 1>, gemm_loop_kernel<3, 1>};
        auto kk = ks[mode - 10];
        cudaFuncSetAttribute(kk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        float* o = nullptr;
        if (cudaMalloc((void**)&o, (size_t)n_sm * 256 * 4) != cudaSuccess) return -1.f;
        cudaEvent_t a, b;
        cudaEventCreate(&a);
        cudaEventCreate(&b);
        const int it = 200;
        float best = 0.f;
        for (int rep = 0; rep < 3; ++rep) {
            cudaEventRecord(a, st);
            kk<<<n_sm, 256, smem, st>>>(o, it);
            cudaEventRecord(b, st);
            cudaEventSynchronize(b);
            float ms = 0.f;
            cudaEventElapsedTime(&ms, a, b);
            double fl = 2.0 * 128.0 * 128.0 * 128.0 * it * n_sm;     // one 128^3 GEMM per iteration per CTA
            float tf = (float)(fl / (ms * 1e-3) / 1e12);
            if (rep > 0 && tf > best) best = tf;
        }
        cudaEventDestroy(a);
        cudaEventDestroy(b);
        cudaFree(o);
        return cudaGetLastError() == cudaSuccess ? best : -2.f;
    }
    const int blocks = (mode == 2 || mode == 4) ? 2 * n_sm : n_sm, iters = 2048;
    float *in = nullptr, *out = nullptr;
    if (cudaMalloc((void**)&in, 20 * 256 * 4) != cudaSuccess) return -1.f;
    if (cudaMalloc((void**)&out, (size_t)blocks * 256 * 4) != cudaSuccess) { cudaFree(in); return -1.f; }
    cudaMemsetAsync(in, 0, 20 * 256 * 4, st);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best = 0.f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0, st);
        if (mode <= 2) ffma_outer_kernel<<<blocks, 256, 0, st>>>(in, out, iters);
        else ffma2_outer_kernel<<<blocks, 256, 0, st>>>(in, out, iters);
        cudaEventRecord(e1, st);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        double fl = 2.0 * 64 * 4 * (double)iters * blocks * 256;
        float tf = (float)(fl / (ms * 1e-3) / 1e12);
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(in);
    cudaFree(out);
    return cudaGetLastError() == cudaSuccess ? best : -2.f;
}


// FP32 FFMA throughput of the device, measured: 8 independent FMA chains per thread
__global__ void __launch_bounds__(256) ffma_peak_kernel(float* out, int iters, float a, float b) {
    float x0 = threadIdx.x, x1 = x0 + 1.f, x2 = x0 + 2.f, x3 = x0 + 3.f, x4 = x0 + 4.f, x5 = x0 + 5.f, x6 = x0 + 6.f,
          x7 = x0 + 7.f, x8 = x0 + 8.f, x9 = x0 + 9.f, xa = x0 + 10.f, xb = x0 + 11.f, xc = x0 + 12.f, xd = x0 + 13.f,
          xe = x0 + 14.f, xf = x0 + 15.f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
            x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
            x8 = fmaf(x8, a, b); x9 = fmaf(x9, a, b); xa = fmaf(xa, a, b); xb = fmaf(xb, a, b);
            xc = fmaf(xc, a, b); xd = fmaf(xd, a, b); xe = fmaf(xe, a, b); xf = fmaf(xf, a, b);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7 + x8 + x9 + xa + xb + xc + xd + xe + xf;
}


// ---- C entry points of libtdq_probe.so (tools/probe.py binds them).  Developer micro-benchmarks behind the design
// decisions recorded in DESIGN.md / profiles/r01_probes.txt; not part of the product library.
extern "C" int tdq_probe_ffma_peak(int device, float* tflops) {
    if (!tflops || cudaSetDevice(device) != cudaSuccess) return 1;
    int n_sm = 0;
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, device);
    const int blocks = n_sm * 8, iters = 4096;
    float* out = nullptr;
    if (cudaMalloc((void**)&out, (size_t)blocks * 256 * 4) != cudaSuccess) return 1;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best = 0.f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0, 0);
        ffma_peak_kernel<<<blocks, 256>>>(out, iters, 0.999f, 0.001f);
        cudaEventRecord(e1, 0);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        double fl = 2.0 * 16 * 8 * (double)iters * blocks * 256;
        float tf = (float)(fl / (ms * 1e-3) / 1e12);
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    if (cudaGetLastError() != cudaSuccess) return 1;
    *tflops = best;
    return 0;
}

/* mode:
 *   0        constant-operand FFMA (= tdq_probe_ffma_peak)
 *   1 / 2    the jet GEMM's 4 points x 4 neurons x 4 channels outer-product update as scalar FFMA, operands in
 *            vector registers, one / two 256-thread CTAs per SM
 *   3 / 4    the same update as packed FFMA2 over point pairs
 *   10..17   the jet kernel's own GEMM loops alone, out of shared memory (10 forward/dgrad, 11 wgrad in two
 *            passes, 12 wgrad single pass, 13 wgrad scalar FFMA; +4: alternative lane mapping)
 *   20       Tensor Memory round trip (tcgen05.st / tcgen05.ld as per-thread storage): 1 = intact
 *   30       tcgen05.mma kind::tf32, 3xTF32 split, D[128x128] = X[128x64] Y^T out of shared memory: effective
 *            fp32-equivalent TFLOP/s
 *   31..37   accuracy of that product, max |error| / max |D| against double: 31 3xTF32, 32 plain TF32, 33 four
 *            terms, 34 correction terms first, 35 tf32-rounded inputs single term, 36 inputs on a 1/16 grid
 *            (exact), 37 / 39 scalar fp32 FMA on rounded / raw inputs
 *   40 / 41  forward/dgrad FMA loop for the 6-channel set with 2- / 4-point thread tiles
 *   42 / 43  accuracy / rate of the 3xTF32 product on the jet kernels' K-major SWIZZLE_128B activation planes
 *   44..47   one product with the B operand MN-major: 44 / 45 no swizzle (returns zeros: unsupported for tf32),
 *            46 / 47 SWIZZLE_128B_BASE32B with LBO = plane stride (46 exact) / LBO and SBO exchanged
 *   48..54   clock cycles per tcgen05.mma (M = N = 128, K = 8, tf32) by operand source: 48 A K-major / B MN-major,
 *            49 both K-major, 50 A in Tensor Memory / B K-major, 51 both MN-major, 52 A in Tensor Memory / B
 *            MN-major, 53 / 54 operands streaming through 64 KB each (no reuse) */
extern "C" int tdq_probe(int device, int mode, float* value) {
    if (!value) return 1;
    if (mode == 0) return tdq_probe_ffma_peak(device, value);
    if (mode < 0 || (mode > 4 && (mode < 10 || mode > 17) && mode != 20 && (mode < 30 || mode > 54))) return 2;
    if (cudaSetDevice(device) != cudaSuccess) return 1;
    int n_sm = 0;
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, device);
    float v = tdq_probe_run(mode, n_sm, nullptr);
    if (v < 0.f) return 3;
    *value = v;
    return 0;
}
