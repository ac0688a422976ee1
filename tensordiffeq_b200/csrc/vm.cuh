// Per-point evaluator of the traced residual / loss-integrand program and its reverse
// sweep.  Replaces the body of the user's f_model / deriv_model callback (models.py:157,
// boundaries.py:238-239) and the MSE / g_MSE wrappers (utils.py:38-45) once the host
// has traced them into tdq_op form.  ~1e2 scalar ops per point against ~1e6 flops of
// MLP work, so a small interpreter is the right tool: one thread per point, slots in
// local memory, program read uniformly by the warp.
#pragma once
#ifndef TDQ_HD
#define TDQ_HD __host__ __device__ __forceinline__
#endif
#include "../../include/tdq_b200.h"

namespace tdq {

struct VmIn {
    const float* jets;   // [C]
    const float* coord;  // [d]
    float target, lam;
    const float* vars;   // [n_vars] (shared memory copy)
};

TDQ_HD float vm_powi(float x, int n) {
    float r = 1.f;
    int m = n < 0 ? -n : n;
    float b = x;
    while (m) { if (m & 1) r *= b; b *= b; m >>= 1; }
    return n < 0 ? 1.f / r : r;
}

// Every per-point array (slot, adj, jets, coord, jet_bar, var_bar) is indexed [i * STRIDE]: STRIDE = 1 for
// thread-private arrays, 32 for arrays laid out [i][lane] in shared memory (tc_kernel.cuh: thread-private arrays
// of this size live in local memory, and with nearly all of L1 carved out as shared memory they thrash).
// forward: fills slot[0..n_ops)
template <int STRIDE = 1>
TDQ_HD void vm_forward(const tdq_op* __restrict__ ops, int n_ops, const VmIn& in, float* slot) {
    for (int i = 0; i < n_ops; ++i) {
        tdq_op o = ops[i];
        float v;
        switch (o.op) {
            case TDQ_OP_CONST: v = o.imm; break;
            case TDQ_OP_JET: v = in.jets[o.a * STRIDE]; break;
            case TDQ_OP_COORD: v = in.coord[o.a * STRIDE]; break;
            case TDQ_OP_TARGET: v = in.target; break;
            case TDQ_OP_LAMBDA: v = in.lam; break;
            case TDQ_OP_VAR: v = in.vars[o.a]; break;
            case TDQ_OP_ADD: v = slot[o.a * STRIDE] + slot[o.b * STRIDE]; break;
            case TDQ_OP_SUB: v = slot[o.a * STRIDE] - slot[o.b * STRIDE]; break;
            case TDQ_OP_MUL: v = slot[o.a * STRIDE] * slot[o.b * STRIDE]; break;
            case TDQ_OP_DIV: v = slot[o.a * STRIDE] / slot[o.b * STRIDE]; break;
            case TDQ_OP_NEG: v = -slot[o.a * STRIDE]; break;
            case TDQ_OP_SIN: v = sinf(slot[o.a * STRIDE]); break;
            case TDQ_OP_COS: v = cosf(slot[o.a * STRIDE]); break;
            case TDQ_OP_EXP: v = expf(slot[o.a * STRIDE]); break;
            case TDQ_OP_TANH: v = tanhf(slot[o.a * STRIDE]); break;
            case TDQ_OP_SQRT: v = sqrtf(slot[o.a * STRIDE]); break;
            case TDQ_OP_LOG: v = logf(slot[o.a * STRIDE]); break;
            case TDQ_OP_POWI: v = vm_powi(slot[o.a * STRIDE], (int)o.imm); break;
            case TDQ_OP_POWF: v = powf(slot[o.a * STRIDE], o.imm); break;
            default: v = 0.f;
        }
        slot[i * STRIDE] = v;
    }
}

// reverse: adj[] must be zero on entry except adj[out] = seed.  Accumulates into
// jet_bar[C], var_bar[n_vars] and returns d/d lambda.
template <int STRIDE = 1>
TDQ_HD float vm_reverse(const tdq_op* __restrict__ ops, int n_ops, const float* slot, float* adj,
                                            float* jet_bar, float* var_bar) {
    float lam_bar = 0.f;
    for (int i = n_ops - 1; i >= 0; --i) {
        float g = adj[i * STRIDE];
        if (g == 0.f) continue;
        tdq_op o = ops[i];
        switch (o.op) {
            case TDQ_OP_JET: jet_bar[o.a * STRIDE] += g; break;
            case TDQ_OP_LAMBDA: lam_bar += g; break;
            case TDQ_OP_VAR: var_bar[o.a * STRIDE] += g; break;
            case TDQ_OP_ADD: adj[o.a * STRIDE] += g; adj[o.b * STRIDE] += g; break;
            case TDQ_OP_SUB: adj[o.a * STRIDE] += g; adj[o.b * STRIDE] -= g; break;
            case TDQ_OP_MUL: { float x = slot[o.a * STRIDE], y = slot[o.b * STRIDE]; adj[o.a * STRIDE] += g * y; adj[o.b * STRIDE] += g * x; } break;
            case TDQ_OP_DIV: { float y = slot[o.b * STRIDE]; adj[o.a * STRIDE] += g / y; adj[o.b * STRIDE] -= g * slot[i * STRIDE] / y; } break;
            case TDQ_OP_NEG: adj[o.a * STRIDE] -= g; break;
            case TDQ_OP_SIN: adj[o.a * STRIDE] += g * cosf(slot[o.a * STRIDE]); break;
            case TDQ_OP_COS: adj[o.a * STRIDE] -= g * sinf(slot[o.a * STRIDE]); break;
            case TDQ_OP_EXP: adj[o.a * STRIDE] += g * slot[i * STRIDE]; break;
            case TDQ_OP_TANH: adj[o.a * STRIDE] += g * (1.f - slot[i * STRIDE] * slot[i * STRIDE]); break;
            case TDQ_OP_SQRT: adj[o.a * STRIDE] += g * 0.5f / slot[i * STRIDE]; break;
            case TDQ_OP_LOG: adj[o.a * STRIDE] += g / slot[o.a * STRIDE]; break;
            case TDQ_OP_POWI: { int n = (int)o.imm; adj[o.a * STRIDE] += g * (float)n * vm_powi(slot[o.a * STRIDE], n - 1); } break;
            case TDQ_OP_POWF: adj[o.a * STRIDE] += g * o.imm * powf(slot[o.a * STRIDE], o.imm - 1.f); break;
            default: break;
        }
    }
    return lam_bar;
}

}  // namespace tdq
