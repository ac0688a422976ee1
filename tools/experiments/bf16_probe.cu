// Probe (developer experiment): tcgen05.mma kind::f16 with bf16 inputs, A operand in Tensor Memory (two bf16 per 32-bit
// column), B operand in shared memory K-major / MN-major SWIZZLE_128B -- which packing order and descriptor fields give
// D[m][n] = sum_k A[m][k] B[n][k]?  Build and run on the GPU box:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/bf16_probe tools/experiments/bf16_probe.cu && /tmp/bf16_probe
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "../../tensordiffeq_b200/csrc/fused_kernel.cuh"
#include "../../tensordiffeq_b200/csrc/tmem.cuh"

using namespace tdq;

__host__ __device__ constexpr uint32_t idesc_bf16(int M, int N, int b_mn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)b_mn << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// mode bit 0: A packing (0: even k in the low half), bit 1: B MN-major, bits 2..: variant of the MN descriptor
__global__ void __launch_bounds__(128, 1) probe_kernel(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ D,
                                                       int mode, int KT) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t tmem_slot;
    __shared__ uint64_t bar;
    const int tid = threadIdx.x, w = tid >> 5;
    if (tid == 0) mbar_init(&bar, 1);
    if (w == 0) tmem_alloc_512(&tmem_slot);
    tmem_fence_before_sync();
    __syncthreads();
    tmem_fence_after_sync();
    const uint32_t tb = tmem_slot;
    const uint32_t trow = tb + ((uint32_t)(w * 32) << 16);
    const bool mn = mode & 2;
    // ---- B -> shared memory (bf16)
    for (int i = tid; i < 128 * KT; i += 128) {
        const int n = i / KT, k = i % KT;
        const __nv_bfloat16 v = __float2bfloat16(B[n * KT + k]);
        size_t off;
        if (!mn) {   // K-major: row n = one 128-byte line (64 bf16 along K), 8-row groups 1024 B apart
            off = (size_t)(n >> 3) * 1024 + (n & 7) * 128 + ((((k * 2) >> 4) ^ (n & 7)) << 4) + ((k * 2) & 15);
        } else {     // MN-major: row k = one 128-byte line of 64 n; 8-row groups 1024 B apart; 64-wide n atoms 16 KB apart
            const int an = n >> 6, wn = n & 63;
            off = (size_t)an * 16384 + (size_t)(k >> 3) * 1024 + (k & 7) * 128 + ((((wn * 2) >> 4) ^ (k & 7)) << 4) + ((wn * 2) & 15);
        }
        *reinterpret_cast<__nv_bfloat16*>(smem + off) = v;
    }
    // ---- A -> Tensor Memory, thread m = row m: KT / 2 columns of packed pairs
    {
        const int m = tid;
        for (int j = 0; j < KT / 2; j += 8) {
            float packed[8];
            for (int jj = 0; jj < 8; ++jj) {
                const float a0 = A[m * KT + 2 * (j + jj)], a1 = A[m * KT + 2 * (j + jj) + 1];
                const __nv_bfloat16 b0 = __float2bfloat16(a0), b1 = __float2bfloat16(a1);
                const uint16_t u0 = *reinterpret_cast<const uint16_t*>(&b0), u1 = *reinterpret_cast<const uint16_t*>(&b1);
                const uint32_t word = (mode & 1) ? ((uint32_t)u0 << 16 | u1) : ((uint32_t)u1 << 16 | u0);
                packed[jj] = __uint_as_float(word);
            }
            tmem_st8(trow + 256 + j, packed);
        }
        tmem_wait_st();
    }
    fence_proxy_async_smem();
    tmem_fence_before_sync();
    __syncthreads();
    if (tid == 0) {
        tmem_fence_after_sync();
        const uint32_t a0 = smem_u32(smem) >> 4;
        for (int ks = 0; ks < KT / 16; ++ks) {
            uint64_t bd;
            if (!mn) {
                const uint32_t lo = a0 + ks * 2 + (1u << 16), hi = (1024u >> 4) | (1u << 14) | (2u << 29);     // +32 B per 16 k
                asm("mov.b64 %0, {%1, %2};" : "=l"(bd) : "r"(lo), "r"(hi));
            } else {
                const int var = mode >> 2;
                uint32_t lbo = 16384u >> 4, sbo = 1024u >> 4;
                if (var == 1) { lbo = 1024u >> 4; sbo = 16384u >> 4; }
                const uint32_t lo = a0 + ks * (2048u >> 4) + (lbo << 16), hi = sbo | (1u << 14) | (2u << 29);     // 16 k rows = 2 groups
                asm("mov.b64 %0, {%1, %2};" : "=l"(bd) : "r"(lo), "r"(hi));
            }
            const uint32_t idesc = idesc_bf16(128, 128, mn ? 1 : 0);
            if (ks == 0)
                asm volatile("tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, 0;" ::"r"(tb), "r"(tb + 256 + ks * 8), "l"(bd), "r"(idesc) : "memory");
            else
                asm volatile("tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, 1;" ::"r"(tb), "r"(tb + 256 + ks * 8), "l"(bd), "r"(idesc) : "memory");
        }
        umma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    tmem_fence_after_sync();
    for (int j = 0; j < 128; j += 16) {
        float v[16];
        tmem_ld16(trow + j, v);
        tmem_wait_ld();
        for (int i = 0; i < 16; ++i) D[tid * 128 + j + i] = v[i];
    }
    tmem_fence_before_sync();
    __syncthreads();
    if (w == 0) tmem_dealloc_512(tb);
}

static float bf(float x) { return __bfloat162float(__float2bfloat16(x)); }

int main() {
    const int KT = 64;
    float *A, *B, *D;
    cudaMallocManaged(&A, 128 * KT * 4);
    cudaMallocManaged(&B, 128 * KT * 4);
    cudaMallocManaged(&D, 128 * 128 * 4);
    srand(1);
    for (int i = 0; i < 128 * KT; ++i) { A[i] = bf((rand() % 2001 - 1000) / 500.f); B[i] = bf((rand() % 2001 - 1000) / 500.f); }
    cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    for (int mode = 0; mode < 8; ++mode) {
        if (!(mode & 2) && (mode >> 2)) continue;
        for (int i = 0; i < 128 * 128; ++i) D[i] = -777.f;
        probe_kernel<<<1, 128, 64 * 1024>>>(A, B, D, mode, KT);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("mode %d: %s\n", mode, cudaGetErrorString(e)); return 1; }
        double mx = 0, mr = 0;
        for (int m = 0; m < 128; ++m)
            for (int n = 0; n < 128; ++n) {
                double r = 0;
                for (int k = 0; k < KT; ++k) r += (double)A[m * KT + k] * B[n * KT + k];
                mx = fmax(mx, fabs(r - D[m * 128 + n]));
                mr = fmax(mr, fabs(r));
            }
        printf("mode %d (A packing %d, B %s, variant %d): max |err| %.4g (max |D| %.4g)  D[0][0..3] = %g %g %g %g\n", mode, mode & 1,
               (mode & 2) ? "MN-major" : "K-major", mode >> 2, mx, mr, D[0], D[1], D[2], D[3]);
    }
    return 0;
}
