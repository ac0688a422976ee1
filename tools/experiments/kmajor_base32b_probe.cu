// Probe (developer experiment): can a kind::tf32 MMA read a K-major operand stored with the 32-byte-atom 128-byte swizzle
// (SWIZZLE_128B_BASE32B: 32-byte chunk index XOR (row & 3)), i.e. the layout the jet kernel's MN-major fp32 planes have?
// If so, the weight-gradient GEMM could read zbar straight from the plane the data-gradient GEMM used, and the K-major
// re-store of zbar would go.  D[m][n] = sum_k A[m][k] B[n][k], M = N = 128, K = 32 (one plane); A K-major SWIZZLE_128B.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/kb32 tools/experiments/kmajor_base32b_probe.cu && for m in 0 1 2 3 4; do /tmp/kb32 $m; done   (a faulting mode poisons the context: one process per mode)
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "../../tensordiffeq_b200/csrc/fused_kernel.cuh"
#include "../../tensordiffeq_b200/csrc/tmem.cuh"

using namespace tdq;

// mode: 0 = B K-major SWIZZLE_128B (control); 1.. = B in BASE32B storage, descriptor variants
__global__ void __launch_bounds__(128, 1) probe_kernel(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ D, int mode) {
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
    float* As = reinterpret_cast<float*>(smem);
    float* Bs = reinterpret_cast<float*>(smem + 16384);
    for (int i = tid; i < 128 * 32; i += 128) {
        const int r = i / 32, k = i % 32;
        As[r * 32 + ((((k >> 2) ^ (r & 7)) << 2) | (k & 3))] = A[i];                       // SWIZZLE_128B: 16-byte chunks
        if (mode == 0) Bs[r * 32 + ((((k >> 2) ^ (r & 7)) << 2) | (k & 3))] = B[i];
        else Bs[r * 32 + ((((k >> 3) ^ (r & 3)) << 3) | (k & 7))] = B[i];                  // BASE32B: 32-byte chunks
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (tid == 0) {
        const uint32_t a0 = smem_u32(As) >> 4, b0 = smem_u32(Bs) >> 4;
        for (int ks = 0; ks < 4; ++ks) {
            uint64_t ad, bd;
            {
                const uint32_t lo = a0 + ks * 2 + (1u << 16), hi = (1024u >> 4) | (1u << 14) | (2u << 29);
                asm("mov.b64 %0, {%1, %2};" : "=l"(ad) : "r"(lo), "r"(hi));
            }
            {
                uint32_t sbo = 1024u >> 4, lt = 2u, lbo = 1u;
                if (mode == 1) { lt = 1u; }
                if (mode == 2) { lt = 1u; sbo = 512u >> 4; }
                if (mode == 3) { lt = 1u; lbo = 512u >> 4; }
                if (mode == 4) { lt = 1u; sbo = 512u >> 4; lbo = 1024u >> 4; }
                const uint32_t lo = b0 + ks * 2 + (lbo << 16), hi = sbo | (1u << 14) | (lt << 29);
                asm("mov.b64 %0, {%1, %2};" : "=l"(bd) : "r"(lo), "r"(hi));
            }
            umma_tf32(tb, ad, bd, umma_idesc_tf32(128, 128), ks > 0);
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

static float tf(float x) { uint32_t u; memcpy(&u, &x, 4); u &= 0xFFFFE000u; memcpy(&x, &u, 4); return x; }

int main(int argc, char** argv) {
    const int only = argc > 1 ? atoi(argv[1]) : -1;
    float *A, *B, *D;
    cudaMallocManaged(&A, 128 * 32 * 4);
    cudaMallocManaged(&B, 128 * 32 * 4);
    cudaMallocManaged(&D, 128 * 128 * 4);
    srand(1);
    for (int i = 0; i < 128 * 32; ++i) { A[i] = tf((rand() % 2001 - 1000) / 500.f); B[i] = tf((rand() % 2001 - 1000) / 500.f); }
    cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    for (int mode = 0; mode < 5; ++mode) {
        if (only >= 0 && mode != only) continue;
        for (int i = 0; i < 128 * 128; ++i) D[i] = -777.f;
        probe_kernel<<<1, 128, 64 * 1024>>>(A, B, D, mode);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("mode %d: %s\n", mode, cudaGetErrorString(e)); return 1; }
        double mx = 0, mr = 0;
        for (int m = 0; m < 128; ++m)
            for (int n = 0; n < 128; ++n) {
                double r = 0;
                for (int k = 0; k < 32; ++k) r += (double)A[m * 32 + k] * B[n * 32 + k];
                mx = fmax(mx, fabs(r - D[m * 128 + n]));
                mr = fmax(mr, fabs(r));
            }
        printf("mode %d: max |err| %.4g (max |D| %.4g)  D[0][0..3] = %g %g %g %g\n", mode, mx, mr, D[0], D[1], D[2], D[3]);
    }
    return 0;
}
