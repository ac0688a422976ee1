// Measurement helper (include/tdq_b200.h: tdq_ffma_probe): what the FP32 pipe sustains for the
// register patterns of the jet GEMM once operand delivery from shared memory is taken away.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/tdq_b200.h"

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

}  // namespace

// returns TFLOP/s (best of 3 after one warm-up) or a negative CUDA error code
extern "C" __attribute__((visibility("hidden"))) float tdq_probe_run(int mode, int n_sm, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
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
