// The per-point residual program in the form the tensor-core kernel evaluates (tc_kernel.cuh).
//
// The traced program (tdq_op, include/tdq_b200.h) is a chain of tiny ops; interpreted one per step it costs a
// dependent shared-memory round trip and a multi-way branch per op while 16 warps wait for the one warp that
// runs it.  tc_compile() folds it at term-creation time into a shorter list over one value table
//     V[0 .. 8)            jet channels            V[8 .. 11)       coordinates
//     V[11] target         V[12] lambda            V[13 .. 21)      trainable PDE coefficients
//     V[kTcVmInputs + i]   result of compiled op i
// in which input-loading ops are table references, constants and constant factors are folded into
// coefficients, integer powers are multiplications, and every add / subtract / multiply / negate is one
//     POLY:  v = c0 + cx X + cy Y + cxy X Y                    (X = V[a], Y = V[b])
// whose reverse is  adj[a] += g (cx + cxy Y),  adj[b] += g (cy + cxy X).  The remaining ops (divide,
// transcendental) keep a switch.  Same arithmetic as vm.cuh up to the association of the folded constants.
#pragma once
#include <math.h>
#include <stdint.h>

#include <vector>

#include "../../include/tdq_b200.h"

namespace tdq {

enum { TC_POLY = 0, TC_DIV, TC_SIN, TC_COS, TC_EXP, TC_TANH, TC_SQRT, TC_LOG, TC_POWF };

struct alignas(16) TcOp {          // 32 bytes
    int32_t kind, a, b, pad;
    float c0, cx, cy, cxy;
};

constexpr int kTcVmJets = 8, kTcVmCoord = 8, kTcVmTarget = 11, kTcVmLambda = 12, kTcVmVars = 13;
constexpr int kTcVmInputs = kTcVmVars + TDQ_MAX_VARS;     // table entries ahead of the op results
constexpr int kTcVmMaxOps = 220;      // value + adjoint tables of (inputs + ops) x 32 lanes fit one 64 KB region

struct TcProgram {
    std::vector<TcOp> ops;
    int out = 0, f = -1;        // table indices of the loss integrand and of the reported residual
};

// `ops` already carries the variant's jet-channel numbering.  Returns false if the program does not fit.
inline bool tc_compile(const std::vector<tdq_op>& ops, int out_slot, int f_slot, TcProgram* prog) {
    struct Ref { bool is_const; float k; float s; int idx; };      // value = k (const)  or  s * V[idx]
    std::vector<Ref> ref(ops.size());
    prog->ops.clear();
    auto emit = [&](int kind, int a, int b, float c0, float cx, float cy, float cxy) {
        TcOp o{kind, a, b, 0, c0, cx, cy, cxy};
        prog->ops.push_back(o);
        return kTcVmInputs + (int)prog->ops.size() - 1;
    };
    auto materialise = [&](const Ref& r) {          // table index holding exactly the value
        if (!r.is_const && r.s == 1.f) return r.idx;
        if (r.is_const) return emit(TC_POLY, 0, 0, r.k, 0.f, 0.f, 0.f);
        return emit(TC_POLY, r.idx, 0, 0.f, r.s, 0.f, 0.f);
    };
    for (size_t i = 0; i < ops.size(); ++i) {
        const tdq_op& o = ops[i];
        Ref r{false, 0.f, 1.f, 0};
        switch (o.op) {
            case TDQ_OP_CONST: r = Ref{true, o.imm, 1.f, 0}; break;
            case TDQ_OP_JET: r.idx = o.a; break;
            case TDQ_OP_COORD: r.idx = kTcVmCoord + o.a; break;
            case TDQ_OP_TARGET: r.idx = kTcVmTarget; break;
            case TDQ_OP_LAMBDA: r.idx = kTcVmLambda; break;
            case TDQ_OP_VAR: r.idx = kTcVmVars + o.a; break;
            case TDQ_OP_NEG: r = ref[o.a]; if (r.is_const) r.k = -r.k; else r.s = -r.s; break;
            case TDQ_OP_ADD:
            case TDQ_OP_SUB: {
                const Ref x = ref[o.a], y = ref[o.b];
                const float sg = o.op == TDQ_OP_ADD ? 1.f : -1.f;
                if (x.is_const && y.is_const) r = Ref{true, x.k + sg * y.k, 1.f, 0};
                else r.idx = emit(TC_POLY, x.is_const ? 0 : x.idx, y.is_const ? 0 : y.idx, (x.is_const ? x.k : 0.f) + (y.is_const ? sg * y.k : 0.f),
                                  x.is_const ? 0.f : x.s, y.is_const ? 0.f : sg * y.s, 0.f);
            } break;
            case TDQ_OP_MUL: {
                const Ref x = ref[o.a], y = ref[o.b];
                if (x.is_const && y.is_const) r = Ref{true, x.k * y.k, 1.f, 0};
                else if (x.is_const) r = Ref{false, 0.f, x.k * y.s, y.idx};
                else if (y.is_const) r = Ref{false, 0.f, y.k * x.s, x.idx};
                else r.idx = emit(TC_POLY, x.idx, y.idx, 0.f, 0.f, 0.f, x.s * y.s);
            } break;
            case TDQ_OP_POWI: {
                const Ref x = ref[o.a];
                int n = (int)o.imm;
                if (x.is_const) { r = Ref{true, powf(x.k, (float)n), 1.f, 0}; break; }
                if (n == 0) { r = Ref{true, 1.f, 1.f, 0}; break; }
                const bool inv = n < 0;
                if (inv) n = -n;
                // x^n by repeated squaring on the table entry V[x.idx]; the scale s^n is carried in the reference
                int result = -1, base = x.idx;
                for (int m = n; m; m >>= 1) {
                    if (m & 1) result = result < 0 ? base : emit(TC_POLY, result, base, 0.f, 0.f, 0.f, 1.f);
                    if (m >> 1) base = emit(TC_POLY, base, base, 0.f, 0.f, 0.f, 1.f);
                }
                float sc = 1.f;
                for (int k = 0; k < n; ++k) sc *= x.s;
                if (!inv) r = Ref{false, 0.f, sc, result};
                else {
                    int one = emit(TC_POLY, 0, 0, 1.f, 0.f, 0.f, 0.f);
                    r.idx = emit(TC_DIV, one, result, 0.f, 1.f, sc, 0.f);
                }
            } break;
            case TDQ_OP_DIV: {
                const Ref x = ref[o.a], y = ref[o.b];
                if (x.is_const && y.is_const) { r = Ref{true, x.k / y.k, 1.f, 0}; break; }
                if (y.is_const) { r = x.is_const ? Ref{true, x.k / y.k, 1.f, 0} : Ref{false, 0.f, x.s / y.k, x.idx}; break; }
                const int xi = x.is_const ? materialise(x) : x.idx;
                r.idx = emit(TC_DIV, xi, y.idx, 0.f, x.is_const ? 1.f : x.s, y.s, 0.f);
            } break;
            case TDQ_OP_SIN: case TDQ_OP_COS: case TDQ_OP_EXP: case TDQ_OP_TANH: case TDQ_OP_SQRT: case TDQ_OP_LOG: case TDQ_OP_POWF: {
                const Ref x = ref[o.a];
                static const int kind_of[] = {TC_SIN, TC_COS, TC_EXP, TC_TANH, TC_SQRT, TC_LOG};
                const int kind = o.op == TDQ_OP_POWF ? TC_POWF : kind_of[o.op - TDQ_OP_SIN];
                const int xi = x.is_const ? materialise(x) : x.idx;
                r.idx = emit(kind, xi, 0, o.op == TDQ_OP_POWF ? o.imm : 0.f, x.is_const ? 1.f : x.s, 0.f, 0.f);
            } break;
            default: return false;
        }
        ref[i] = r;
    }
    // the integrand and the reported residual must be table entries of their own (the reverse sweep seeds one,
    // the kernel stores the other)
    auto own_entry = [&](int slot) {
        const Ref& r = ref[slot];
        if (!r.is_const && r.s == 1.f && r.idx >= kTcVmInputs) return r.idx;
        if (r.is_const) return emit(TC_POLY, 0, 0, r.k, 0.f, 0.f, 0.f);
        return emit(TC_POLY, r.idx, 0, 0.f, r.s, 0.f, 0.f);
    };
    prog->out = own_entry(out_slot);
    prog->f = f_slot >= 0 ? own_entry(f_slot) : -1;
    return (int)prog->ops.size() <= kTcVmMaxOps;
}

}  // namespace tdq
