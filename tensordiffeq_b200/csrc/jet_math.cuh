// Taylor-mode jets through tanh, forward and reverse.
//
// The reference obtains u_x, u_xx, ... by nesting tf.gradients through the
// keras MLP (e.g. examples/AC-SA.py:36-44); here every neuron carries, next to
// its value, the pure derivatives of its pre-activation along each input
// dimension, and tanh is pushed through with Faa di Bruno's formula.
//   a = tanh(z0), s = 1 - a^2
//   sigma1 = s, sigma2 = -2as, sigma3 = s(6a^2-2), sigma4 = 8as(2-3a^2),
//   sigma5 = 16s^2 - 32a^2 s - 72 a^2 s^2 + 48 a^4 s
#pragma once
#ifndef TDQ_HD
#define TDQ_HD __host__ __device__ __forceinline__
#endif
#include <cuda_runtime.h>

#include "f32x2.cuh"

namespace tdq {

// tanh to ~2e-7 ABSOLUTE error (what the loss / gradient parity needs: tanh values enter sums with O(1) weights),
// branch-free: 1 - 2 / (exp(2x) + 1) on the special-function unit (ex2.approx, rcp.approx); saturates to +-1 for
// large |x| without special cases.  The library tanhf (1-2 ulp relative) costs three times the instructions and a
// divergent branch; TDQ_TANH_LIBM selects it (and the host harness always uses it).
TDQ_HD float tanh_acc(float x) {
#if defined(__CUDA_ARCH__) && !defined(TDQ_TANH_LIBM)
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * 2.8853900817779268f));       // exp(2x)
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(e + 1.0f));
    return fmaf(-2.0f, r, 1.0f);
#else
    return tanhf(x);
#endif
}

// The formulas below are templated on the value type T: float, or F2 (f32x2.cuh) = two points at a time through the
// packed fp32 instructions.  jm_fma marks where a multiply-add is meant to fuse (for float the compiler would contract
// it anyway; a packed add after a packed multiply is never contracted).
TDQ_HD float jm_fma(float a, float b, float c) { return a * b + c; }

TDQ_HD F2 tanh_acc(F2 x) {
#if defined(__CUDA_ARCH__) && !defined(TDQ_TANH_LIBM)
    const F2 t = x * F2(2.8853900817779268f);
    F2 e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.x) : "f"(t.x));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.y) : "f"(t.y));
    const F2 e1 = e + F2(1.0f);
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r.x) : "f"(e1.x));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r.y) : "f"(e1.y));
    return jm_fma(F2(-2.0f), r, F2(1.0f));
#else
    return F2(tanhf(x.x), tanhf(x.y));
#endif
}

template <class T>
struct SigT {
    T a, s1, s2, s3, s4, s5;
};
using Sig = SigT<float>;

// NSIG = highest sigma index needed (forward of order K: K; reverse: K+1)
template <int NSIG, class T>
TDQ_HD SigT<T> make_sig(T a) {
    SigT<T> g;
    g.a = a;
    const T a2 = a * a;
    const T s = T(1.0f) - a2;
    g.s1 = s;
    g.s2 = T(-2.0f) * a * s;
    g.s3 = T(0.f); g.s4 = T(0.f); g.s5 = T(0.f);
    if (NSIG >= 3) g.s3 = s * jm_fma(T(6.0f), a2, T(-2.0f));
    if (NSIG >= 4) g.s4 = T(8.0f) * a * s * jm_fma(T(-3.0f), a2, T(2.0f));
    if (NSIG >= 5) g.s5 = s * (T(16.0f) * s - T(32.0f) * a2 - T(72.0f) * a2 * s + T(48.0f) * a2 * a2);
    return g;
}

// forward along one direction: z[1..K] -> out[1..K]   (arrays indexed 0..K-1 = order 1..K)
template <int K, class T>
TDQ_HD void jet_fwd(const SigT<T>& g, const T* z, T* out) {
    if (K >= 1) out[0] = g.s1 * z[0];
    if (K >= 2) out[1] = jm_fma(g.s2 * z[0], z[0], g.s1 * z[1]);
    if (K >= 3) {
        const T z1s = z[0] * z[0];
        out[2] = jm_fma(g.s3 * z1s, z[0], jm_fma(T(3.0f) * g.s2 * z[0], z[1], g.s1 * z[2]));
    }
    if (K >= 4) {
        const T z1s = z[0] * z[0];
        out[3] = jm_fma(g.s4 * z1s, z1s,
                        jm_fma(T(6.0f) * g.s3 * z1s, z[1],
                               jm_fma(T(3.0f) * g.s2 * z[1], z[1], jm_fma(T(4.0f) * g.s2 * z[0], z[2], g.s1 * z[3]))));
    }
}

// reverse along one direction: given abar[1..K] (cotangents of the outputs) and z[1..K],
// write zbar[1..K] and return the contribution to zbar0.
// d a_k / d z0 is a_k's formula with every sigma_i replaced by sigma_{i+1}.
template <int K, class T>
TDQ_HD T jet_bwd(const SigT<T>& g, const T* z, const T* ab, T* zb) {
    T zb0 = T(0.f);
    if (K == 1) {
        zb[0] = g.s1 * ab[0];
        zb0 = ab[0] * g.s2 * z[0];
    } else if (K == 2) {
        const T t = g.s2 * z[0];
        zb[0] = jm_fma(T(2.0f) * t, ab[1], g.s1 * ab[0]);
        zb[1] = g.s1 * ab[1];
        zb0 = jm_fma(ab[1], jm_fma(g.s3 * z[0], z[0], g.s2 * z[1]), ab[0] * t);
    } else if (K == 3) {
        const T z1s = z[0] * z[0], t = g.s2 * z[0];
        const T d2 = jm_fma(g.s3, z1s, g.s2 * z[1]);                       // d a_2 / d z0
        zb[0] = jm_fma(T(3.0f) * d2, ab[2], jm_fma(T(2.0f) * t, ab[1], g.s1 * ab[0]));
        zb[1] = jm_fma(T(3.0f) * t, ab[2], g.s1 * ab[1]);
        zb[2] = g.s1 * ab[2];
        const T d3 = jm_fma(g.s4 * z1s, z[0], jm_fma(T(3.0f) * g.s3 * z[0], z[1], g.s2 * z[2]));
        zb0 = jm_fma(ab[2], d3, jm_fma(ab[1], d2, ab[0] * t));
    } else if (K == 4) {
        const T z1s = z[0] * z[0], t = g.s2 * z[0];
        const T d2 = jm_fma(g.s3, z1s, g.s2 * z[1]);
        const T d3 = jm_fma(g.s4 * z1s, z[0], jm_fma(T(3.0f) * g.s3 * z[0], z[1], g.s2 * z[2]));
        const T d4 = jm_fma(g.s5 * z1s, z1s,
                            jm_fma(T(6.0f) * g.s4 * z1s, z[1],
                                   jm_fma(T(3.0f) * g.s3 * z[1], z[1], jm_fma(T(4.0f) * g.s3 * z[0], z[2], g.s2 * z[3]))));
        zb[0] = jm_fma(T(4.0f) * d3, ab[3], jm_fma(T(3.0f) * d2, ab[2], jm_fma(T(2.0f) * t, ab[1], g.s1 * ab[0])));
        zb[1] = jm_fma(T(6.0f) * d2, ab[3], jm_fma(T(3.0f) * t, ab[2], g.s1 * ab[1]));
        zb[2] = jm_fma(T(4.0f) * t, ab[3], g.s1 * ab[2]);
        zb[3] = g.s1 * ab[3];
        zb0 = jm_fma(ab[3], d4, jm_fma(ab[2], d3, jm_fma(ab[1], d2, ab[0] * t)));
    }
    return zb0;
}

// Channel layout shared by every kernel: c = 0 value; then dimension 0 orders 1..K0,
// dimension 1 orders 1..K1, dimension 2 orders 1..K2.
template <int K0, int K1, int K2>
struct JetCfg {
    static constexpr int C = 1 + K0 + K1 + K2;
    static constexpr int MAXK = (K0 > K1 ? (K0 > K2 ? K0 : K2) : (K1 > K2 ? K1 : K2));
    static constexpr int OFF0 = 1, OFF1 = 1 + K0, OFF2 = 1 + K0 + K1;

    // v[c] holds (z0, z-derivatives) on entry and (a0, a-derivatives) on exit
    TDQ_HD static void fwd(float* v) {
        float a = tanh_acc(v[0]);
        Sig g = make_sig<MAXK>(a);
        v[0] = a;
        float o[4];
        if (K0 > 0) { jet_fwd<K0>(g, v + OFF0, o); for (int i = 0; i < K0; ++i) v[OFF0 + i] = o[i]; }
        if (K1 > 0) { jet_fwd<K1>(g, v + OFF1, o); for (int i = 0; i < K1; ++i) v[OFF1 + i] = o[i]; }
        if (K2 > 0) { jet_fwd<K2>(g, v + OFF2, o); for (int i = 0; i < K2; ++i) v[OFF2 + i] = o[i]; }
    }
    // st[c] = stash (a0, z-derivatives) -> act[c] = (a0, a-derivatives)
    template <class T>
    TDQ_HD static void recompute(const T* st, T* act) {
        const SigT<T> g = make_sig<MAXK>(st[0]);
        act[0] = st[0];
        if (K0 > 0) jet_fwd<K0>(g, st + OFF0, act + OFF0);
        if (K1 > 0) jet_fwd<K1>(g, st + OFF1, act + OFF1);
        if (K2 > 0) jet_fwd<K2>(g, st + OFF2, act + OFF2);
    }
    // st[c] = stash; ab[c] = cotangent of (a0, a-derivs); zb[c] <- cotangent of (z0, z-derivs)
    template <class T>
    TDQ_HD static void bwd(const T* st, const T* ab, T* zb) {
        const SigT<T> g = make_sig<MAXK + 1>(st[0]);
        T zb0 = ab[0] * g.s1;
        if (K0 > 0) zb0 += jet_bwd<K0>(g, st + OFF0, ab + OFF0, zb + OFF0);
        if (K1 > 0) zb0 += jet_bwd<K1>(g, st + OFF1, ab + OFF1, zb + OFF1);
        if (K2 > 0) zb0 += jet_bwd<K2>(g, st + OFF2, ab + OFF2, zb + OFF2);
        zb[0] = zb0;
    }
};

// Second-order jets in two dimensions WITH the mixed partial: channels (u, u_x, u_xx, u_y, u_yy, u_xy), selected by
// K2 = -1 (a tag, not an order).  tf.gradients(tf.gradients(u, x), y) of examples/testing.py:23-31 needs it.
//   a_xy = sigma2 z_x z_y + sigma1 z_xy
template <>
struct JetCfg<2, 2, -1> {
    static constexpr int C = 6;
    static constexpr int MAXK = 2;
    static constexpr int OFF0 = 1, OFF1 = 3, OFF2 = 5, XY = 5;
    template <class T>
    TDQ_HD static void recompute(const T* st, T* act) {
        const SigT<T> g = make_sig<2>(st[0]);
        act[0] = st[0];
        jet_fwd<2>(g, st + OFF0, act + OFF0);
        jet_fwd<2>(g, st + OFF1, act + OFF1);
        act[XY] = jm_fma(g.s2 * st[OFF0], st[OFF1], g.s1 * st[XY]);
    }
    template <class T>
    TDQ_HD static void bwd(const T* st, const T* ab, T* zb) {
        const SigT<T> g = make_sig<3>(st[0]);
        T zb0 = ab[0] * g.s1;
        zb0 += jet_bwd<2>(g, st + OFF0, ab + OFF0, zb + OFF0);
        zb0 += jet_bwd<2>(g, st + OFF1, ab + OFF1, zb + OFF1);
        const T zx = st[OFF0], zy = st[OFF1], axy = ab[XY];
        zb[XY] = g.s1 * axy;
        zb[OFF0] = jm_fma(g.s2 * zy, axy, zb[OFF0]);
        zb[OFF1] = jm_fma(g.s2 * zx, axy, zb[OFF1]);
        zb0 = jm_fma(axy, jm_fma(g.s3 * zx, zy, g.s2 * st[XY]), zb0);
        zb[0] = zb0;
    }
};

}  // namespace tdq
