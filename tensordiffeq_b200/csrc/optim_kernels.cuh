// Small kernels around the fused jet kernel: deterministic slab reduction, the
// Keras-semantics Adam update fused with the weight re-pack, and the device-side
// L-BFGS iteration.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "tc_image.cuh"

namespace tdq {

constexpr int kMaxTerms = 16;
constexpr int kGvecTail = 16;     // gvec = [P flat grads][8 var grads][loss][pad]

struct SlabList {
    const float* ptr[kMaxTerms];
    int rows[kMaxTerms];
    int perm[kMaxTerms];      // 1: hidden-matrix blocks in the tensor-core kernel's order (tc_slab_off, tc_image.cuh)
    int n;
};

struct AdamDevCfg {
    float alpha, beta1, beta2, eps, sign;
};

// keras.optimizers.Adam.apply_gradients (fit.py:103-107) as TF's ResourceApplyAdam executes it, for ONE flat
// parameter, followed by the re-pack of that parameter into the padded matrices / operand images the jet kernels
// read.  i = packed index, fi = flat index (utils.py:7-29 layout).
__device__ __forceinline__ void adam_pack_one(int i, int fi, float g_in, int off_wh, int HP, float* __restrict__ params,
                                              float* __restrict__ m, float* __restrict__ v, const float* __restrict__ mask,
                                              const AdamDevCfg& cfg, int do_adam, float* __restrict__ wpack,
                                              float* __restrict__ wpack_t, float* __restrict__ wimg) {
    float val = 0.f;
    if (fi >= 0) {
        val = params[fi];
        if (do_adam) {
            float g = cfg.sign * g_in;
            float mm = m[fi], vv = v[fi];
            mm += (g - mm) * (1.f - cfg.beta1);
            vv += (g * g - vv) * (1.f - cfg.beta2);
            m[fi] = mm;
            v[fi] = vv;
            float upd = (mm * cfg.alpha) / (sqrtf(vv) + cfg.eps);
            if (mask) upd *= mask[fi];
            val -= upd;
            params[fi] = val;
        }
    }
    wpack[i] = val;
    if (i >= off_wh) {
        int r = i - off_wh;
        int l = r / (HP * HP);
        int kj = r - l * HP * HP;
        int k = kj / HP, j = kj - k * HP;
        wpack_t[(size_t)l * HP * HP + j * HP + k] = val;
        if (wimg) tc_img_store(wimg, l, k, j, val);       // 128-wide nets: tensor-core operand images
    }
}

// gvec[flat] = sum over the launches' slabs, rows in a fixed order (bit-reproducible).
// Replaces the tail of tape.gradient (models.py:198): the cross-point SUM of per-example parameter gradients, and
// the loss reduce_mean partial sums (utils.py:38-45).  A block owns kReduceCols consecutive positions of a slab
// row (of the first slab's layout, so its reads are contiguous) and splits the rows over kReduceRows thread rows:
// every load of a column is independent of the others, the partial sums meet in shared memory in a fixed order.
// do_adam (single GPU): the thread that holds a parameter's gradient applies the Adam step and re-packs the
// parameter right away -- no second pass over the parameters, no extra launch.
constexpr int kReduceCols = 32, kReduceRows = 8;
__global__ void __launch_bounds__(kReduceCols * kReduceRows)
reduce_slabs_kernel(SlabList sl, int pp, int off_sc, const int* __restrict__ flat_of_packed, float* __restrict__ gvec,
                    int n_flat, int off_wh, int HP, int perm_major, int do_adam, AdamDevCfg cfg,
                    float* __restrict__ params, float* __restrict__ m, float* __restrict__ v,
                    const float* __restrict__ mask, float* __restrict__ wpack, float* __restrict__ wpack_t,
                    float* __restrict__ wimg) {
    __shared__ float part[kReduceRows][kReduceCols + 1];
    const int x = threadIdx.x, y = threadIdx.y;
    const int q = blockIdx.x * kReduceCols + x;    // position inside a slab row of the major layout
    int i = q, ip = q;                             // packed index; position in a permuted (tensor-core kernel) row
    bool live = false, is_tail = false;
    int fi = -1, sc = 0;
    if (q < pp) {
        if (q >= off_wh) {
            const int r = q - off_wh, mt = r / (HP * HP), qq = r - mt * HP * HP;
            if (perm_major) {                      // q counts in tc_slab_off order: invert it
                const int t = qq >> 2, k = t % HP, j = (t / HP) * 4 + (qq & 3);
                i = off_wh + mt * HP * HP + k * HP + j;
            } else {
                ip = off_wh + mt * HP * HP + tc_slab_off(qq / HP, qq % HP, HP);
            }
        }
        fi = flat_of_packed[i];
        sc = i - off_sc;                           // scalar block: [0] d b_out (has fi), [1] loss, [2..9] d vars
        is_tail = (sc >= 1 && sc < 10);
        live = fi >= 0 || is_tail;
    }
    float s = 0.f;
    if (live) {
        for (int t = 0; t < sl.n; ++t) {
            const float* p = sl.ptr[t] + (sl.perm[t] ? ip : i);
            const int rows = sl.rows[t];
            float s0 = 0.f, s1 = 0.f;
            int r = y;
            for (; r + kReduceRows < rows; r += 2 * kReduceRows) {
                s0 += p[(size_t)r * pp];
                s1 += p[(size_t)(r + kReduceRows) * pp];
            }
            if (r < rows) s0 += p[(size_t)r * pp];
            s += s0 + s1;
        }
    }
    part[y][x] = s;
    __syncthreads();
    if (y != 0 || !live) return;
#pragma unroll
    for (int r = 1; r < kReduceRows; ++r) s += part[r][x];
    if (fi >= 0) {
        gvec[fi] = s;
        if (do_adam) adam_pack_one(i, fi, s, off_wh, HP, params, m, v, mask, cfg, 1, wpack, wpack_t, wimg);
    } else if (sc == 1) {
        gvec[n_flat + 8] = s;                      // loss
    } else {
        gvec[n_flat + (sc - 2)] = s;               // d vars
    }
}

// The same update for the whole flat vector from an already reduced (all-reduced) gradient, or the plain re-pack
// (do_adam = 0) after tdq_set_params / an L-BFGS step.  One thread per packed element.
__global__ void adam_pack_kernel(int pp, int off_wh, int HP, const int* __restrict__ flat_of_packed,
                                 const float* __restrict__ gvec, float* __restrict__ params, float* __restrict__ m,
                                 float* __restrict__ v, const float* __restrict__ mask, AdamDevCfg cfg, int do_adam,
                                 float* __restrict__ wpack, float* __restrict__ wpack_t, float* __restrict__ wimg) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= pp) return;
    int fi = flat_of_packed[i];
    adam_pack_one(i, fi, (fi >= 0 && do_adam) ? gvec[fi] : 0.f, off_wh, HP, params, m, v, mask, cfg, do_adam, wpack, wpack_t, wimg);
}

// Adam on the handful of trainable PDE coefficients (DiscoveryModel.tf_optimizer_vars, models.py:341,348)
__global__ void adam_vars_kernel(int n, const float* __restrict__ gvec_vars, float* vars, float* m, float* v,
                                 AdamDevCfg cfg) {
    int i = threadIdx.x;
    if (i >= n) return;
    float g = cfg.sign * gvec_vars[i];
    float mm = m[i], vv = v[i];
    mm += (g - mm) * (1.f - cfg.beta1);
    vv += (g * g - vv) * (1.f - cfg.beta2);
    m[i] = mm;
    v[i] = vv;
    vars[i] -= (mm * cfg.alpha) / (sqrtf(vv) + cfg.eps);
}

// Adam on a lambda vector when it is not fused into the jet kernel's epilogue
__global__ void adam_vec_kernel(long long n, const float* __restrict__ g_in, float* x, float* m, float* v,
                                AdamDevCfg cfg) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float g = cfg.sign * g_in[i];
    float mm = m[i], vv = v[i];
    mm += (g - mm) * (1.f - cfg.beta1);
    vv += (g * g - vv) * (1.f - cfg.beta2);
    m[i] = mm;
    v[i] = vv;
    x[i] -= (mm * cfg.alpha) / (sqrtf(vv) + cfg.eps);
}

// ---------------------------------------------------------------------------
// L-BFGS (optimizers.py:214-402), one CTA, whole iteration on the device.
// ---------------------------------------------------------------------------
constexpr int kLbfgsThreads = 1024;
constexpr int kLbfgsHist = 50;      // nCorrection, optimizers.py:223

struct LbfgsState {
    int n_iter;      // state.nIter
    int k;           // pairs held
    int head;        // physical index of the oldest pair
    int status;      // 0 continue, 1 no progress along direction (gtd > -tolX)
    float Hdiag, t, f, f_old, gtd, sum_abs_g, sum_abs_dt;
    float ro[kLbfgsHist], al[kLbfgsHist];
};

__device__ __forceinline__ float block_sum(float v, float* sh) {
    // fixed-shape tree: shuffle within warps, then warp 0 over the 32 partials
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    float r = (threadIdx.x < (blockDim.x >> 5)) ? sh[threadIdx.x] : 0.f;
    if (w == 0) {
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
        if (lane == 0) sh[32] = r;
    }
    __syncthreads();
    return sh[32];
}

__device__ __forceinline__ float block_dot(const float* a, const float* b, int n, float* sh) {
    float s = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) s = fmaf(a[i], b[i], s);
    return block_sum(s, sh);
}

// One pass of the `for epoch in t_` body up to and including `x += t*d` (optimizers.py:250-339).
// g = gradient at x (from the last evaluation).  S holds s = d*t ("old_dirs"), Y holds y ("old_stps").
__global__ void __launch_bounds__(kLbfgsThreads) lbfgs_iter_kernel(LbfgsState* st, int n, float lr, float* x,
                                                                   const float* __restrict__ g, float* g_old, float* d,
                                                                   float* S, float* Y, float* q, float f_new_valid,
                                                                   const float* __restrict__ loss_dev) {
    __shared__ float sh[40];
    const int tid = threadIdx.x, nt = blockDim.x;
    __shared__ LbfgsState s;
    if (tid == 0) s = *st;
    __syncthreads();
    if (tid == 0 && f_new_valid != 0.f) s.f = *loss_dev;
    __syncthreads();
    // :252-256 -- t for this pass, from the counter BEFORE it moves (quirk Q12)
    float sum_abs_g;
    {
        float a = 0.f;
        for (int i = tid; i < n; i += nt) a += fabsf(g[i]);
        sum_abs_g = block_sum(a, sh);
    }
    float t = (s.n_iter == 1) ? fminf(1.f, 1.f / sum_abs_g) : lr;
    int n_iter = s.n_iter + 1;
    int k = s.k, head = s.head;
    float Hdiag = s.Hdiag;
    if (n_iter == 1) {
        for (int i = tid; i < n; i += nt) d[i] = -g[i];
        k = 0; head = 0; Hdiag = 1.f;
    } else {
        // y = g - g_old ; s = d*t  -> candidate slot
        int slot = (k == kLbfgsHist) ? head : (head + k) % kLbfgsHist;
        float* Sn = S + (size_t)slot * n;
        float* Yn = Y + (size_t)slot * n;
        // do not clobber the oldest pair unless the new one is accepted: stage in q / d is still needed
        float ys_p = 0.f, yy_p = 0.f;
        for (int i = tid; i < n; i += nt) {
            float y = g[i] - g_old[i];
            float sv = d[i] * t;
            ys_p = fmaf(y, sv, ys_p);
            yy_p = fmaf(y, y, yy_p);
        }
        float ys = block_sum(ys_p, sh);
        float yy = block_sum(yy_p, sh);
        if (ys > 1e-10f) {                                    // :277-289
            for (int i = tid; i < n; i += nt) {
                Yn[i] = g[i] - g_old[i];
                Sn[i] = d[i] * t;
            }
            if (k == kLbfgsHist) head = (head + 1) % kLbfgsHist; else k += 1;
            Hdiag = ys / yy;
            __syncthreads();
            if (tid == 0) s.ro[slot] = 1.f / ys;             // ro = 1/dot(y, s), :296-298
        }
        __syncthreads();
        // two-loop recursion :302-313
        for (int i = tid; i < n; i += nt) q[i] = -g[i];
        __syncthreads();
        for (int i = k - 1; i >= 0; --i) {
            int ph = (head + i) % kLbfgsHist;
            float al = block_dot(S + (size_t)ph * n, q, n, sh) * s.ro[ph];
            if (tid == 0) s.al[i] = al;
            const float* Yi = Y + (size_t)ph * n;
            for (int j = tid; j < n; j += nt) q[j] = fmaf(-al, Yi[j], q[j]);
            __syncthreads();
        }
        for (int j = tid; j < n; j += nt) q[j] *= Hdiag;
        __syncthreads();
        for (int i = 0; i < k; ++i) {
            int ph = (head + i) % kLbfgsHist;
            float be = block_dot(Y + (size_t)ph * n, q, n, sh) * s.ro[ph];
            float coef = s.al[i] - be;
            const float* Si = S + (size_t)ph * n;
            for (int j = tid; j < n; j += nt) q[j] = fmaf(coef, Si[j], q[j]);
            __syncthreads();
        }
        for (int j = tid; j < n; j += nt) d[j] = q[j];
    }
    __syncthreads();
    // g_old = g ; f_old = f   :315-316
    for (int i = tid; i < n; i += nt) g_old[i] = g[i];
    float gtd = block_dot(g, d, n, sh);                       // :325
    int status = 0;
    float t2 = t;
    if (gtd > -1e-12f) {                                      // :328-330
        status = 1;
    } else {
        t2 = (n_iter == 1) ? fminf(1.f, 1.f / sum_abs_g) : lr;   // :333-337
        float a = 0.f;
        for (int i = tid; i < n; i += nt) {
            float dt = d[i] * t2;
            x[i] += dt;                                       // :339
            a += fabsf(dt);
        }
        float sdt = block_sum(a, sh);
        if (tid == 0) s.sum_abs_dt = sdt;
    }
    __syncthreads();
    if (tid == 0) {
        s.n_iter = n_iter; s.k = k; s.head = head; s.Hdiag = Hdiag; s.t = t2; s.f_old = s.f;
        s.gtd = gtd; s.status = status; s.sum_abs_g = sum_abs_g;
        *st = s;
    }
}

// sum |g| at the new point, for the optimality test (optimizers.py:367-371)
__global__ void __launch_bounds__(kLbfgsThreads) lbfgs_check_kernel(LbfgsState* st, int n, const float* __restrict__ g,
                                                                    const float* __restrict__ loss_dev) {
    __shared__ float sh[40];
    float a = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) a += fabsf(g[i]);
    float s = block_sum(a, sh);
    if (threadIdx.x == 0) { st->sum_abs_g = s; st->f = *loss_dev; }
}

}  // namespace tdq
