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

// ---------------------------------------------------------------------------
// Multi-GPU: slab reduction + all-reduce over NVLink peer memory + Adam update in ONE kernel (replaces
// reduce_slabs_kernel -> ncclAllReduce -> adam_pack_kernel; fit.py:149-153's implicit NcclAllReduce).
// Every rank runs the same grid.  A block owns a fixed set of 32-column chunks of the packed gradient:
//   1. it sums its columns over the local slab rows (as reduce_slabs_kernel does) into the rank's own exchange
//      buffer xbuf[parity][packed index];
//   2. it publishes "my chunks of step s are written" by storing s into flags[my rank][block] on EVERY peer
//      (plain NVLink stores into the peers' memory, after a system-scope fence);
//   3. it waits until the flags of all ranks for this block hold s, then every thread adds up the W ranks'
//      values of its column -- loads from the peers' xbuf over NVLink -- in rank order, so all ranks get the same
//      bits, writes gvec and applies the Adam step / re-pack right there.
// No grid-wide barrier: block b only ever waits for block b of the peers, and the grid is sized to be fully
// resident, so every block makes progress.  Two exchange buffers alternate with the step parity: a rank can only
// be one exchange ahead of its slowest peer.  A wait that exceeds ~4 s (a dead peer) raises *status instead of
// hanging the GPU.
// ---------------------------------------------------------------------------
constexpr int kMaxPeers = 8;
struct PeerTable {
    float* xbuf[kMaxPeers];        // each rank's exchange buffer [2][pp_pad] (own entry = local pointer)
    unsigned* flags[kMaxPeers];    // each rank's flag array [kMaxPeers][max blocks]
    int rank, world, max_blocks, pp_pad;
};

__global__ void __launch_bounds__(kReduceCols * kReduceRows)
reduce_exchange_adam_kernel(SlabList sl, PeerTable pt, unsigned step, int pp, int off_sc, const int* __restrict__ flat_of_packed,
                            float* __restrict__ gvec, int n_flat, int off_wh, int HP, int perm_major, int do_adam, AdamDevCfg cfg,
                            float* __restrict__ params, float* __restrict__ m, float* __restrict__ v,
                            const float* __restrict__ mask, float* __restrict__ wpack, float* __restrict__ wpack_t,
                            float* __restrict__ wimg, int* status) {
    __shared__ float part[kReduceRows][kReduceCols + 1];
    const int x = threadIdx.x, y = threadIdx.y;
    const int n_chunks = (pp + kReduceCols - 1) / kReduceCols;
    float* mine = pt.xbuf[pt.rank] + (size_t)(step & 1u) * pt.pp_pad;
    // ---- 1. local sums of this block's chunks -> own exchange buffer
    for (int ch = blockIdx.x; ch < n_chunks; ch += gridDim.x) {
        const int q = ch * kReduceCols + x;
        int i = q, ip = q;
        if (q < pp && q >= off_wh) {
            const int r = q - off_wh, mt = r / (HP * HP), qq = r - mt * HP * HP;
            if (perm_major) {
                const int t = qq >> 2, k = t % HP, j = (t / HP) * 4 + (qq & 3);
                i = off_wh + mt * HP * HP + k * HP + j;
            } else {
                ip = off_wh + mt * HP * HP + tc_slab_off(qq / HP, qq % HP, HP);
            }
        }
        float s = 0.f;
        if (q < pp) {
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
        __syncthreads();
        part[y][x] = s;
        __syncthreads();
        if (y == 0 && q < pp) {
#pragma unroll
            for (int r = 1; r < kReduceRows; ++r) s += part[r][x];
            mine[i] = s;
        }
    }
    // ---- 2. publish: thread x of row 0 tells rank x (the eight stores travel side by side, each behind its own fence;
    //         the barrier makes the other threads' buffer writes precede the fences)
    __syncthreads();
    if (y == 0 && x < pt.world) {
        __threadfence_system();
        unsigned* f = pt.flags[x] + (size_t)pt.rank * pt.max_blocks + blockIdx.x;
        asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(f), "r"(step) : "memory");
    }
    // ---- 3. wait for this block's chunks on every rank
    if (y == 0 && x < pt.world) {
        const unsigned* f = pt.flags[pt.rank] + (size_t)x * pt.max_blocks + blockIdx.x;
        const long long t0 = clock64();
        unsigned got;
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(got) : "l"(f) : "memory");
            // (a peer that is already one exchange ahead has overwritten the flag with step + 1)
            if ((int)(got - step) < 0 && clock64() - t0 > 8000000000LL) { atomicExch(status, 1); break; }
        } while ((int)(got - step) < 0);
    }
    __syncthreads();
    // ---- sum over ranks in rank order, Adam, re-pack.  Thread row y fetches rank y's value of column x (all rows'
    // NVLink loads of all of the block's chunks are in flight together), row 0 adds them up in rank order.
    constexpr int kMaxChunks = 4;                 // chunks per block: host sizes the grid accordingly
    float val[kMaxChunks];
    int idx[kMaxChunks];
#pragma unroll
    for (int c = 0; c < kMaxChunks; ++c) {
        const int ch = blockIdx.x + c * gridDim.x;
        const int q = ch * kReduceCols + x;
        val[c] = 0.f;
        idx[c] = -1;
        if (ch < n_chunks && q < pp) {
            int i = q;
            if (perm_major && q >= off_wh) {
                const int r = q - off_wh, mt = r / (HP * HP), qq = r - mt * HP * HP;
                const int t = qq >> 2, k = t % HP, j = (t / HP) * 4 + (qq & 3);
                i = off_wh + mt * HP * HP + k * HP + j;
            }
            idx[c] = i;
            if (y < pt.world) {
                const float* src = pt.xbuf[y] + (size_t)(step & 1u) * pt.pp_pad + i;
                asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(val[c]) : "l"(src));
            }
        }
    }
#pragma unroll
    for (int c = 0; c < kMaxChunks; ++c) {
        __syncthreads();
        part[y][x] = val[c];
        __syncthreads();
        const int i = idx[c];
        if (y != 0 || i < 0) continue;
        const int fi = flat_of_packed[i];
        const int sc = i - off_sc;
        const bool is_tail = (sc >= 1 && sc < 10);
        if (fi < 0 && !is_tail) continue;
        float s = 0.f;
        for (int r = 0; r < pt.world; ++r) s += part[r][x];
        if (fi >= 0) {
            gvec[fi] = s;
            if (do_adam) adam_pack_one(i, fi, s, off_wh, HP, params, m, v, mask, cfg, 1, wpack, wpack_t, wimg);
        } else if (sc == 1) {
            gvec[n_flat + 8] = s;
        } else {
            gvec[n_flat + (sc - 2)] = s;
        }
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
// L-BFGS (optimizers.py:214-402): the whole iteration and its stop tests on the device.
//
// One cooperative launch per iteration, a few dozen CTAs: the two-loop recursion is a chain of 2 k dependent
// (dot, axpy) pairs over the flat parameter vector; every axpy pass also accumulates the NEXT dot, so a pair
// costs one sweep of two vectors and one grid barrier.  Dots are reduced in two fixed stages (thread-strided
// partials -> block tree -> the CTAs' partials summed in CTA order by every CTA), hence bit-reproducible and
// identical on all CTAs.  The host enqueues iterations back to back and looks at the state only every few
// iterations: once a stop test fires, `done` is set and the remaining enqueued kernels return at once.
// ---------------------------------------------------------------------------
constexpr int kLbfgsThreads = 1024;
constexpr int kLbfgsHist = 50;      // nCorrection, optimizers.py:223
constexpr int kLbfgsMaxCtas = 160;
constexpr int kLbfgsRounds = 4;     // rotating partial-sum buffers (a CTA may lag one barrier behind)

struct LbfgsState {
    int n_iter;      // state.nIter
    int k;           // pairs held
    int head;        // physical index of the oldest pair
    int status;      // 0 continue, 1 no progress along direction (gtd > -tolX)
    float Hdiag, t, f, f_old, gtd, sum_abs_g, sum_abs_dt;
    float ro[kLbfgsHist], al[kLbfgsHist];
    // loop control (optimizers.py:341-380), evaluated on the device
    int done;        // 0 running; 1 gtd > -tolX; 2 maxIter; 3 maxEval; 4 optimality; 5 step size; 6 |f| < tolX
    int n_hist;      // entries written to the f history
    int n_eval;      // function evaluations so far
    int max_iter;
    float max_eval;
};

__device__ __forceinline__ float block_sum(float v, float* sh) {
    // fixed-shape tree: shuffle within warps, then warp 0 over the warps' partials
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

// grid-wide barrier for a cooperative launch (all CTAs resident): generation counter in global memory
__device__ __forceinline__ void grid_barrier(unsigned* bar, unsigned& gen) {
    __syncthreads();
    if (threadIdx.x == 0) {
        gen += 1;
        __threadfence();
        unsigned arrived = atomicAdd(bar, 1u) + 1u;
        const unsigned target = gen * gridDim.x;
        if (arrived != target) {
            while (*reinterpret_cast<volatile unsigned*>(bar) < target) { }
        }
        __threadfence();
    }
    __syncthreads();
}

// sums NV values over the whole grid; every CTA returns the same bits.  part: [kLbfgsRounds][NV][kLbfgsMaxCtas]
template <int NV>
__device__ __forceinline__ void grid_sum(float (&v)[NV], float* part, unsigned* bar, unsigned& gen, unsigned& round, float* sh) {
    if (gridDim.x == 1) {                          // small nets: the block tree is the whole reduction
#pragma unroll
        for (int q = 0; q < NV; ++q) v[q] = block_sum(v[q], sh);
        return;
    }
    float* mine = part + (size_t)(round % kLbfgsRounds) * NV * kLbfgsMaxCtas;
    round += 1;
#pragma unroll
    for (int q = 0; q < NV; ++q) {
        float b = block_sum(v[q], sh);
        if (threadIdx.x == 0) mine[q * kLbfgsMaxCtas + blockIdx.x] = b;
    }
    grid_barrier(bar, gen);
#pragma unroll
    for (int q = 0; q < NV; ++q) {
        float a = 0.f;
        if (threadIdx.x < 32) {
            const volatile float* src = mine + q * kLbfgsMaxCtas;
            for (int c = threadIdx.x; c < (int)gridDim.x; c += 32) a += src[c];     // CTA order within a lane is fixed
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
            if (threadIdx.x == 0) sh[33 + q] = a;
        }
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < NV; ++q) v[q] = sh[33 + q];
    __syncthreads();
}

// One pass of the `for epoch in t_` body up to and including `x += t*d` (optimizers.py:250-339).
// g = gradient at x (from the last evaluation).  S holds s = d*t ("old_dirs"), Y holds y ("old_stps").
// Cooperative launch, gridDim.x <= kLbfgsMaxCtas.  `bar` is a zero-initialised counter that only this kernel touches.
__global__ void __launch_bounds__(kLbfgsThreads) lbfgs_iter_kernel(LbfgsState* st, int n, float lr, float* x,
                                                                   const float* __restrict__ g, float* g_old, float* d,
                                                                   float* S, float* Y, float* q, float* part,
                                                                   unsigned* bar) {
    __shared__ float sh[40];
    __shared__ float al_s[kLbfgsHist];
    __shared__ LbfgsState s;
    const int tid = threadIdx.x;
    const int gtid = blockIdx.x * blockDim.x + tid, gstride = gridDim.x * blockDim.x;
    if (tid == 0) s = *st;
    __syncthreads();
    if (s.done) return;                                       // a stop test fired earlier: nothing moves any more
    // the barrier counter keeps counting across launches: this launch starts at the generation it finds
    unsigned gen = 0, round = 0;
    if (tid == 0) gen = *reinterpret_cast<volatile unsigned*>(bar) / gridDim.x;
    // (every CTA reads bar before any CTA of THIS launch can have added to it a full generation: the first
    //  barrier below cannot complete without this CTA's own arrival)
    // :252-256 -- t for this pass, from the counter BEFORE it moves (quirk Q12)
    float r1[1] = {0.f};
    for (int i = gtid; i < n; i += gstride) r1[0] += fabsf(g[i]);
    grid_sum<1>(r1, part, bar, gen, round, sh);
    const float sum_abs_g = r1[0];
    float t = (s.n_iter == 1) ? fminf(1.f, 1.f / sum_abs_g) : lr;
    int n_iter = s.n_iter + 1;
    int k = s.k, head = s.head;
    float Hdiag = s.Hdiag;
    float gtd;
    if (n_iter == 1) {
        float r[1] = {0.f};
        for (int i = gtid; i < n; i += gstride) {
            float gi = g[i];
            d[i] = -gi;
            g_old[i] = gi;                                    // :315
            r[0] = fmaf(gi, -gi, r[0]);                       // gtd = g . d   :325
        }
        k = 0; head = 0; Hdiag = 1.f;
        grid_sum<1>(r, part, bar, gen, round, sh);
        gtd = r[0];
    } else {
        // y = g - g_old ; s = d*t  -> candidate slot  :273-275
        const int slot = (k == kLbfgsHist) ? head : (head + k) % kLbfgsHist;
        float* Sn = S + (size_t)slot * n;
        float* Yn = Y + (size_t)slot * n;
        float r2[2] = {0.f, 0.f};
        for (int i = gtid; i < n; i += gstride) {
            float y = g[i] - g_old[i];
            float sv = d[i] * t;
            r2[0] = fmaf(y, sv, r2[0]);
            r2[1] = fmaf(y, y, r2[1]);
        }
        grid_sum<2>(r2, part, bar, gen, round, sh);
        const float ys = r2[0], yy = r2[1];
        float ro_new = 0.f;
        const bool accept = ys > 1e-10f;                      // :277-289
        if (accept) {
            for (int i = gtid; i < n; i += gstride) {
                Yn[i] = g[i] - g_old[i];
                Sn[i] = d[i] * t;
            }
            if (k == kLbfgsHist) head = (head + 1) % kLbfgsHist; else k += 1;
            Hdiag = ys / yy;
            ro_new = 1.f / ys;                                // ro = 1/dot(y, s), :296-298
        }
        auto ro_of = [&](int ph) { return (accept && ph == slot) ? ro_new : s.ro[ph]; };
        // two-loop recursion :302-313.  Each thread owns the same elements of q in every pass, so only the dots
        // need a barrier; the axpy of step i and the dot of step i -+ 1 share one sweep.
        float r[1] = {0.f};
        {
            const float* Sk = S + (size_t)((head + k - 1) % kLbfgsHist) * n;
            for (int j = gtid; j < n; j += gstride) {
                float qj = -g[j];
                q[j] = qj;
                if (k > 0) r[0] = fmaf(Sk[j], qj, r[0]);
            }
        }
        for (int i = k - 1; i >= 0; --i) {
            const int ph = (head + i) % kLbfgsHist;
            grid_sum<1>(r, part, bar, gen, round, sh);
            const float al = r[0] * ro_of(ph);
            if (tid == 0) al_s[i] = al;
            const float* Yi = Y + (size_t)ph * n;
            const float* Sm = i > 0 ? S + (size_t)((head + i - 1) % kLbfgsHist) * n : nullptr;
            r[0] = 0.f;
            for (int j = gtid; j < n; j += gstride) {
                float qj = fmaf(-al, Yi[j], q[j]);
                q[j] = qj;
                if (Sm) r[0] = fmaf(Sm[j], qj, r[0]);
            }
        }
        __syncthreads();
        {
            const float* Y0 = k > 0 ? Y + (size_t)(head % kLbfgsHist) * n : nullptr;
            r[0] = 0.f;
            for (int j = gtid; j < n; j += gstride) {
                float qj = q[j] * Hdiag;                      // :310
                q[j] = qj;
                if (Y0) r[0] = fmaf(Y0[j], qj, r[0]);
            }
        }
        for (int i = 0; i < k; ++i) {
            const int ph = (head + i) % kLbfgsHist;
            grid_sum<1>(r, part, bar, gen, round, sh);
            const float be = r[0] * ro_of(ph);
            const float coef = al_s[i] - be;
            const float* Si = S + (size_t)ph * n;
            const float* Yn1 = i + 1 < k ? Y + (size_t)((head + i + 1) % kLbfgsHist) * n : nullptr;
            r[0] = 0.f;
            for (int j = gtid; j < n; j += gstride) {
                float qj = fmaf(coef, Si[j], q[j]);
                q[j] = qj;
                if (Yn1) r[0] = fmaf(Yn1[j], qj, r[0]);
            }
        }
        // d = r ; g_old = g ; gtd = g . d   :313-325
        r[0] = 0.f;
        for (int j = gtid; j < n; j += gstride) {
            float dj = q[j], gj = g[j];
            d[j] = dj;
            g_old[j] = gj;
            r[0] = fmaf(gj, dj, r[0]);
        }
        grid_sum<1>(r, part, bar, gen, round, sh);
        gtd = r[0];
        if (tid == 0 && blockIdx.x == 0 && accept) st->ro[slot] = ro_new;
    }
    int status = 0;
    float t2 = t, sdt = s.sum_abs_dt;
    if (gtd > -1e-12f) {                                      // :328-330
        status = 1;
    } else {
        t2 = (n_iter == 1) ? fminf(1.f, 1.f / sum_abs_g) : lr;   // :333-337
        float r[1] = {0.f};
        for (int i = gtid; i < n; i += gstride) {
            float dt = d[i] * t2;
            x[i] += dt;                                       // :339
            r[0] += fabsf(dt);
        }
        grid_sum<1>(r, part, bar, gen, round, sh);
        sdt = r[0];
    }
    if (tid == 0 && blockIdx.x == 0) {
        st->n_iter = n_iter; st->k = k; st->head = head; st->Hdiag = Hdiag; st->t = t2; st->f_old = s.f;
        st->gtd = gtd; st->status = status; st->sum_abs_g = sum_abs_g; st->sum_abs_dt = sdt;
        if (status == 1) st->done = 1;
    }
}

// After the evaluation at the new point (evaluated != 0) or, on the last allowed iteration, without one:
// record f, count the evaluation and run the stop tests of optimizers.py:357-380 in their order.
// first != 0: the initial evaluation before the loop (:233-244).
__global__ void __launch_bounds__(kLbfgsThreads) lbfgs_check_kernel(LbfgsState* st, int n, const float* __restrict__ g,
                                                                    const float* __restrict__ loss_dev, float* f_hist,
                                                                    int evaluated, int first) {
    __shared__ float sh[40];
    if (st->done) return;
    float sag = st->sum_abs_g, f = st->f;
    if (evaluated) {
        float a = 0.f;
        for (int i = threadIdx.x; i < n; i += blockDim.x) a += fabsf(g[i]);
        sag = block_sum(a, sh);
        f = *loss_dev;
    }
    if (threadIdx.x != 0) return;
    st->sum_abs_g = sag;
    st->f = f;
    f_hist[st->n_hist] = f;
    st->n_hist += 1;
    st->n_eval += 1;
    const float tol_fun = 1e-12f, tol_x = 1e-12f;             // :221-222
    int done = 0;
    if (first) {
        if (sag <= tol_fun) done = 4;                         // :241-244
    } else {
        if (st->n_iter == st->max_iter) done = 2;             // :357-358
        else if ((float)st->n_eval >= st->max_eval) done = 3; // :360-363
        else if (sag <= tol_fun) done = 4;                    // :365-369
        else if (st->sum_abs_dt <= tol_x) done = 5;           // :371-375
        else if (fabsf(f) < tol_x) done = 6;                  // :377 (tf.abs(f, f_old): quirk Q8)
    }
    st->done = done;
}

// ---------------------------------------------------------------------------
// Graph-mode Newton phase (fit.py:65-80 -> tfp.optimizer.lbfgs_minimize): L-BFGS with 10 correction pairs and the
// Hager-Zhang line search.  The line search's control flow (bracket / secant^2 / bisect: a handful of scalar
// comparisons per trial step) runs on the host in api.cu; everything that touches a vector is one of these kernels.
// ---------------------------------------------------------------------------
constexpr int kHzHist = 10;       // tfp default num_correction_pairs
struct HzScalars {
    float f, df, gd, gmax, ys, yy;
};
struct HzRho {
    float rho[kHzHist];
};

// d = -H g by the two-loop recursion over the k stored pairs (ring starting at `head`), H0 = (s.y / y.y) I of the
// newest pair; out->gd = g . d, out->gmax = max |g_i|.  One CTA.
__global__ void __launch_bounds__(1024) hz_direction_kernel(int n, const float* __restrict__ g, const float* __restrict__ S,
                                                            const float* __restrict__ Y, int k, int head, HzRho rho, float gamma,
                                                            float* q, float* d, HzScalars* out) {
    __shared__ float sh[40];
    __shared__ float al[kHzHist];
    const int tid = threadIdx.x, nt = blockDim.x;
    float mx = 0.f;
    for (int j = tid; j < n; j += nt) { float gj = g[j]; q[j] = gj; mx = fmaxf(mx, fabsf(gj)); }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    __shared__ float shm[32];
    if ((tid & 31) == 0) shm[tid >> 5] = mx;
    __syncthreads();
    if (tid == 0) { float m = 0.f; for (int w = 0; w < (nt >> 5); ++w) m = fmaxf(m, shm[w]); out->gmax = m; }
    for (int i = k - 1; i >= 0; --i) {
        const int ph = (head + i) % kHzHist;
        const float* Si = S + (size_t)ph * n;
        const float* Yi = Y + (size_t)ph * n;
        float p = 0.f;
        for (int j = tid; j < n; j += nt) p = fmaf(Si[j], q[j], p);
        const float a = rho.rho[ph] * block_sum(p, sh);
        if (tid == 0) al[i] = a;
        for (int j = tid; j < n; j += nt) q[j] = fmaf(-a, Yi[j], q[j]);
        __syncthreads();
    }
    if (k > 0)
        for (int j = tid; j < n; j += nt) q[j] *= gamma;
    __syncthreads();
    for (int i = 0; i < k; ++i) {
        const int ph = (head + i) % kHzHist;
        const float* Si = S + (size_t)ph * n;
        const float* Yi = Y + (size_t)ph * n;
        float p = 0.f;
        for (int j = tid; j < n; j += nt) p = fmaf(Yi[j], q[j], p);
        const float be = rho.rho[ph] * block_sum(p, sh);
        const float coef = al[i] - be;
        for (int j = tid; j < n; j += nt) q[j] = fmaf(coef, Si[j], q[j]);
        __syncthreads();
    }
    float p = 0.f;
    for (int j = tid; j < n; j += nt) { float dj = -q[j]; d[j] = dj; p = fmaf(g[j], dj, p); }
    p = block_sum(p, sh);
    if (tid == 0) out->gd = p;
}

// trial point of the line search: params = x0 + c d
__global__ void hz_trial_kernel(int n, const float* __restrict__ x0, const float* __restrict__ d, float c, float* __restrict__ params) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) params[i] = fmaf(c, d[i], x0[i]);
}

// after the evaluation at the trial point: out->f = loss, out->df = g . d
__global__ void __launch_bounds__(1024) hz_dot_kernel(int n, const float* __restrict__ g, const float* __restrict__ d,
                                                      const float* __restrict__ loss_dev, HzScalars* out) {
    __shared__ float sh[40];
    float p = 0.f;
    for (int j = threadIdx.x; j < n; j += blockDim.x) p = fmaf(g[j], d[j], p);
    p = block_sum(p, sh);
    if (threadIdx.x == 0) { out->df = p; out->f = *loss_dev; }
}

// accept the trial step c: s = c d, y = g - g_old into the pair slot when s . y > 0; g_old = g; x0 = params
__global__ void __launch_bounds__(1024) hz_accept_kernel(int n, float c, const float* __restrict__ d, const float* __restrict__ g,
                                                         float* g_old, const float* __restrict__ params, float* x0, float* Sn,
                                                         float* Yn, HzScalars* out) {
    __shared__ float sh[40];
    float ys = 0.f, yy = 0.f;
    for (int j = threadIdx.x; j < n; j += blockDim.x) {
        float y = g[j] - g_old[j], sv = c * d[j];
        ys = fmaf(y, sv, ys);
        yy = fmaf(y, y, yy);
    }
    ys = block_sum(ys, sh);
    yy = block_sum(yy, sh);
    for (int j = threadIdx.x; j < n; j += blockDim.x) {
        float gj = g[j];
        if (ys > 0.f) { Yn[j] = gj - g_old[j]; Sn[j] = c * d[j]; }
        g_old[j] = gj;
        x0[j] = params[j];
    }
    if (threadIdx.x == 0) { out->ys = ys; out->yy = yy; }
}

}  // namespace tdq
