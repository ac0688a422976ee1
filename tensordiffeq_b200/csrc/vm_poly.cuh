// Device side of vm_poly.h: forward and reverse sweep over the compiled program, one point per lane, the value
// table V and its adjoints laid out [entry][32 lanes] in shared memory.
// STRIDE = 32 there; the CPU harness (tests/harness/jet_host.cu) runs the same code with STRIDE = 1.
#pragma once
#ifndef TDQ_HD
#define TDQ_HD __host__ __device__ __forceinline__
#endif
#include "vm_poly.h"

namespace tdq {

// V[0 .. kTcVmInputs) must hold the inputs; fills V[kTcVmInputs + i]
template <int STRIDE = 32>
TDQ_HD void tc_vm_forward(const TcOp* __restrict__ ops, int n_ops, float* V) {
    for (int i = 0; i < n_ops; ++i) {
        const TcOp o = ops[i];
        const float X = V[o.a * STRIDE], Y = V[o.b * STRIDE];
        float v;
        if (o.kind == TC_POLY) {
            v = fmaf(o.cx, X, o.c0) + fmaf(o.cxy, X, o.cy) * Y;
        } else {
            const float x = o.cx * X;
            switch (o.kind) {
                case TC_DIV: v = x / (o.cy * Y); break;
                case TC_SIN: v = sinf(x); break;
                case TC_COS: v = cosf(x); break;
                case TC_EXP: v = expf(x); break;
                case TC_TANH: v = tanhf(x); break;
                case TC_SQRT: v = sqrtf(x); break;
                case TC_LOG: v = logf(x); break;
                default: v = powf(x, o.c0); break;
            }
        }
        V[(kTcVmInputs + i) * STRIDE] = v;
    }
}

// A[] (same indexing as V) must be zero except the seed; on return A[0 .. kTcVmInputs) are the input adjoints
template <int STRIDE = 32>
TDQ_HD void tc_vm_reverse(const TcOp* __restrict__ ops, int n_ops, const float* V, float* A) {
    for (int i = n_ops - 1; i >= 0; --i) {
        const TcOp o = ops[i];
        const float g = A[(kTcVmInputs + i) * STRIDE];
        const float X = V[o.a * STRIDE], Y = V[o.b * STRIDE];
        float dx, dy = 0.f;
        if (o.kind == TC_POLY) {
            dx = g * fmaf(o.cxy, Y, o.cx);
            dy = g * fmaf(o.cxy, X, o.cy);
        } else {
            const float v = V[(kTcVmInputs + i) * STRIDE], x = o.cx * X;
            switch (o.kind) {
                case TC_DIV: { const float y = o.cy * Y; dx = g * o.cx / y; dy = -g * v * o.cy / y; } break;
                case TC_SIN: dx = g * cosf(x) * o.cx; break;
                case TC_COS: dx = -g * sinf(x) * o.cx; break;
                case TC_EXP: dx = g * v * o.cx; break;
                case TC_TANH: dx = g * (1.f - v * v) * o.cx; break;
                case TC_SQRT: dx = g * 0.5f / v * o.cx; break;
                case TC_LOG: dx = g / X; break;
                default: dx = g * o.c0 * powf(x, o.c0 - 1.f) * o.cx; break;
            }
        }
        A[o.a * STRIDE] += dx;
        A[o.b * STRIDE] += dy;          // after the first update: a == b (x * x) accumulates both halves
    }
}

}  // namespace tdq
