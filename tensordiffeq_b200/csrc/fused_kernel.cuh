// The hot path: one persistent kernel per loss term that, for a tile of collocation
// points, runs the tanh-MLP forward carrying Taylor jets, evaluates the traced
// residual / loss integrand, and sweeps back through the jets accumulating parameter
// gradients.  Replaces, for one term, everything between
//   CollocationSolverND.update_loss()  (tensordiffeq/models.py:80-192)
//   tape.gradient(loss, variables)     (tensordiffeq/models.py:195-199)
// i.e. the keras forward (networks.py:10-20), the nested tf.gradients of f_model
// (e.g. examples/AC-SA.py:36-44) and TF's reverse graph.
//
// Shape of the computation (HP = padded hidden width, C = jet channels, P points/tile):
//   * hidden->hidden layers are C*P x HP x HP fp32 GEMMs on packed FFMA2, register tile
//     TP points x 4 neurons x C channels per thread, operands in shared memory;
//   * the layer's HP x HP weights (64 KB at HP=128) arrive by one bulk TMA copy
//     (cp.async.bulk + mbarrier) issued a phase ahead, so the copy hides behind the
//     previous GEMM / element-wise phase;
//   * per-layer (tanh value, pre-activation derivatives) needed by the reverse sweep are
//     parked per thread in Tensor Memory (tcgen05.st / tcgen05.ld, tmem.cuh); layers that do
//     not fit use a per-CTA global scratch slab that is rewritten every tile and lives in L2;
//   * weight gradients of the hidden layers accumulate in a per-CTA slab with vector
//     reductions (red.global.add.v4.f32, L2 resident); small gradients (biases, first / last
//     layer) accumulate in shared memory.  A second kernel sums the slabs in a fixed order,
//     so results are bit-reproducible run to run.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "jet_math.cuh"
#include "tmem.cuh"
#include "vm.cuh"

namespace tdq {

constexpr int kThreads = 256;
constexpr int kMaxHidden = 12;      // tanh layers
constexpr int kScalars = 16;        // [0] d b_out, [1] loss, [2..9] d vars

// Packed (padded) layout shared by the weight buffer the kernel reads and the
// per-CTA gradient slab it writes.  L = number of tanh layers.
struct PackLayout {
    int HP, L;
    __host__ __device__ int off_w0() const { return 0; }                     // [3][HP]
    __host__ __device__ int off_b() const { return 3 * HP; }                 // [L][HP]
    __host__ __device__ int off_wout() const { return (3 + L) * HP; }        // [HP]
    __host__ __device__ int off_sc() const { return (4 + L) * HP; }          // [kScalars]
    __host__ __device__ int off_wh() const { return (4 + L) * HP + kScalars; }  // [L-1][HP][HP]
    __host__ __device__ int small() const { return off_wh(); }
    __host__ __device__ int total() const { return off_wh() + (L - 1) * HP * HP; }
};

struct AdamDev {
    float alpha, beta1, beta2, eps, sign;   // alpha = lr*sqrt(1-b2^t)/(1-b1^t)
};

struct alignas(16) TcOp;       // vm_poly.h

// One launch evaluates up to kMaxSeg loss terms ("segments"): the big residual term and the small BC / IC terms
// of update_loss() (models.py:91-188) share one persistent kernel, one gradient slab and one tile numbering
// (segment s owns tiles tile0 .. tile0 + ceil(n / P) - 1), instead of one launch each.  Every segment is
// evaluated with the launch's jet configuration (the smallest compiled one that covers all of them); a segment
// whose program reads fewer channels seeds the others with exact zeros.
constexpr int kMaxSeg = 8;
struct Seg {
    const float* X;          // [d][x_stride] SoA, device (first point of the evaluated range)
    long long x_stride;      // distance between coordinate columns
    const float* target;     // [n] or null
    float* lam;              // [n] or null
    float* lam_m;            // Adam moments of lambda (lam_mode 2)
    float* lam_v;
    float* lam_grad;         // [n] or null (lam_mode >= 1)
    float* f_out;            // [n] or null : slot[f_slot]
    float* u_out;            // [n] or null : network output
    long long n;             // points
    long long tile0;         // first tile of the segment
    float inv_n;
    int lam_mode;            // 0 none, 1 write gradient, 2 gradient + fused Adam ascent/descent
    int pair;
    int ops_off, n_ops, out_slot, f_slot;      // the segment's program inside KParams::ops (slots relative to ops_off)
    int tc_off, tc_n_ops, tc_out, tc_f;        // ... inside KParams::tc_ops (tensor-core kernel, vm_poly.h)
};

struct KParams {
    Seg seg[kMaxSeg];
    int n_seg;
    long long n_tiles;       // of all segments
    AdamDev lam_adam;
    // programs of all segments, back to back
    const tdq_op* ops;
    int n_ops;
    const float* vars;
    int n_vars;
    // network
    int d, L;
    const float* wpack;      // PackLayout, [in][out] hidden matrices
    const float* wpack_t;    // hidden matrices transposed, [L-1][HP][HP] ([out][in])
    const float* wimg;       // tensor-core operand images of the hidden matrices (tc_kernel.cuh), or null
    const TcOp* tc_ops;      // the programs compiled for the tensor-core kernel (vm_poly.h), or null
    int tc_n_ops;
    // scratch
    float* stash;            // [grid][L][C*4][256][TP]
    float* slab;             // [grid][PackLayout::total]
    int dbg;                 // developer aid (TDQ_DBG, read by tdq_create).  FMA kernel: 1 skip fwd/dgrad GEMM loops,
                             // 2 skip FMA wgrad loops, 4 keep the whole stash in global memory (no TMEM), 8 no
                             // tensor-core wgrad inside the FMA kernel.  api.cu: 16 never launch the tensor-core
                             // kernel (24 = pure FP32 FMA everywhere, the baseline of tests/test_gpu_tc*.py)
};

// segment of a tile (tiles are numbered segment after segment)
__device__ __forceinline__ int seg_of_tile(const KParams& prm, long long tile) {
    int si = 0;
    while (si + 1 < prm.n_seg && tile >= prm.seg[si + 1].tile0) ++si;
    return si;
}

// ---- PTX helpers: mbarrier + 1-D bulk TMA -----------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_1d(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

template <int TP>
struct VecT;
template <>
struct VecT<4> { using type = float4; };
template <>
struct VecT<2> { using type = float2; };
template <>
struct VecT<1> { using type = float; };

template <int TP>
__device__ __forceinline__ void vec_to(const typename VecT<TP>::type& v, float* o);
template <>
__device__ __forceinline__ void vec_to<4>(const float4& v, float* o) { o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; }
template <>
__device__ __forceinline__ void vec_to<2>(const float2& v, float* o) { o[0] = v.x; o[1] = v.y; }
template <int TP>
__device__ __forceinline__ typename VecT<TP>::type vec_from(const float* o);
template <>
__device__ __forceinline__ float4 vec_from<4>(const float* o) { return make_float4(o[0], o[1], o[2], o[3]); }
template <>
__device__ __forceinline__ float2 vec_from<2>(const float* o) { return make_float2(o[0], o[1]); }

// Static shape of one instantiation.
template <int K0, int K1, int K2, int HP_, int TP_>
struct Shape {
    using J = JetCfg<K0, K1, K2>;
    static constexpr int C = J::C;
    static constexpr int HP = HP_;
    static constexpr int TP = TP_;
    static constexpr int NG = HP / 4;            // neuron groups (4 neurons each)
    static constexpr int PG = kThreads / NG;     // point groups (TP points each)
    static constexpr int P = TP * PG;            // points per tile
    static constexpr int WPG = PG / 8;           // warps along the point-group axis
    static constexpr int KT = HP / 16;           // wgrad thread tile KT x KT
    static constexpr int ACT = C * HP * P;       // floats in one activation buffer
    static_assert(PG >= 8 && PG % 8 == 0, "tile shape");
    static_assert(TP == 4 || TP == 2, "TP");
    // Swizzle of the point-granule index inside a feature row.
    // HW128 shapes (4-point granules, 32 points per tile: one feature row = one 128-byte line of 8 sixteen-byte
    // chunks): chunk ^= row & 7, which is exactly the tensor core's SWIZZLE_128B K-major canonical layout, so a
    // channel plane [128 rows][32 points] of an activation buffer can be handed to tcgen05.mma as it is (the
    // wgrad GEMM runs there).  It also keeps the FFMA2 loops' operand loads and the element-wise stores spread
    // over the banks.
    // Other TP == 4 shapes: rows 8 apart (the SIMT wgrad operand walk) land in 8 different 16-byte slots of a
    // 128-byte line, and so do the 8 lanes of a quarter-warp in the element-wise stores; within an aligned block
    // of 8 rows the value is s for rows 0..3 and s ^ 2 for rows 4..7.
    static constexpr bool HW128 = (TP == 4 && HP == 128);
    __device__ __forceinline__ static int swz(int row) {
        return HW128 ? (row & 7) : (TP == 4 ? (((row >> 3) & 7) ^ (((row >> 2) & 3) << 1)) : (((row >> 3) & 3) << 1));
    }
    static constexpr int SWZ_HALF = TP == 4 ? 2 : 0;   // non-HW128: swz(row + 4) ^ swz(row) inside an aligned block of 8
    // float offset of (channel c, feature row, point granule g) in an activation buffer
    __device__ __forceinline__ static int aoff(int c, int row, int g) {
        return ((c * HP + row) * PG + (g ^ swz(row))) * TP;
    }
    // shared memory carve-up (floats)
    static constexpr int SM_AS = 0;
    static constexpr int SM_ZS = ACT;
    static constexpr int SM_WS = 2 * ACT;                     // HP*HP
    static constexpr int SM_END_BIG = 2 * ACT + HP * HP;
};

__host__ __device__ inline int small_floats(int HP, int L) { return (4 + L) * HP + kScalars; }

// dynamic shared memory (bytes) an instantiation needs for a net of L tanh layers
template <class S>
__host__ __device__ inline size_t smem_bytes(int L, int n_ops) {
    size_t f = S::SM_END_BIG;
    f += 2 * (size_t)small_floats(S::HP, L);   // small params + their gradient accumulators
    f += (size_t)kThreads * S::C;              // red: output-layer partial sums
    f += (size_t)S::WPG * 4 * S::HP;           // redB: cross-warp reduction over points (up to 4 quantities)
    f += 3 * S::P;                             // Xs
    f += (size_t)S::C * S::P;                  // Ub
    f += 64;                                   // warp scratch (loss / var reductions)
    size_t b = f * 4 + (size_t)n_ops * sizeof(tdq_op) + 64;
    return (b + 127) & ~(size_t)127;
}


// ---- packed FP32 FMA (Blackwell FFMA2: two fp32 FMAs per lane per instruction on an aligned register
//      pair).  A scalar register broadcast to both halves costs nothing (ptxas folds make_float2(b,b)
//      into the instruction's .F32 operand form).  Measured with tdq_ffma_probe on B200: the jet GEMM's
//      outer-product update sustains 58 TFLOP/s as scalar FFMA (register-bank conflicts on three vector
//      operands) and 65 TFLOP/s as FFMA2, and FFMA2 halves the issue slots the update needs, which
//      leaves room for the shared-memory loads.
// Packed values live in 64-bit registers for the whole GEMM loop (packing / unpacking fp32 pairs around
// every instruction makes ptxas shuffle registers); accumulation is in place ("+l").
typedef unsigned long long u64;
__device__ __forceinline__ u64 pack2(float lo, float hi) {
    u64 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(u64 v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
// acc.xy += a.xy * b.xy
__device__ __forceinline__ void ffma2_acc(u64& acc, u64 a, u64 b) {
    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(a), "l"(b));
}
// acc.xy += a.xy * w   (w broadcast to both halves)
__device__ __forceinline__ void ffma2_acc_bcast(u64& acc, u64 a, float w) {
    asm("{\n\t.reg .b64 wb;\n\tmov.b64 wb, {%2, %2};\n\tfma.rn.f32x2 %0, %1, wb, %0;\n\t}" : "+l"(acc) : "l"(a), "f"(w));
}

template <int TP>
struct PackedVec;                    // TP fp32 values as TP/2 64-bit registers, loaded in one instruction
template <>
struct PackedVec<4> {
    ulonglong2 v;
    __device__ __forceinline__ void load(const float* p) { v = *reinterpret_cast<const ulonglong2*>(p); }
    __device__ __forceinline__ u64 pair(int i) const { return i == 0 ? v.x : v.y; }
};
template <>
struct PackedVec<2> {
    u64 v;
    __device__ __forceinline__ void load(const float* p) { v = *reinterpret_cast<const u64*>(p); }
    __device__ __forceinline__ u64 pair(int) const { return v; }
};

// ---- the two GEMM shapes of the kernel.  Both are software-pipelined by hand: the operands of step
//      k+1 are requested from shared memory before the FMAs of step k issue.
//
// acc[c][p][n] += sum_k A[c][k][p] * W[k][n]   (A: activation buffer, W: [HP][HP] row-major);
// FFMA2 runs over pairs of points (adjacent in A's rows), W[k][n] is the broadcast operand.
template <class S>
__device__ __forceinline__ void gemm_act_w(const float* __restrict__ A, const float* __restrict__ W, int pg, int n0,
                                           float (&acc)[S::C][S::TP][4]) {
    constexpr int C = S::C, TP = S::TP, HP = S::HP, PG = S::PG, P = S::P, NP = TP / 2;
    u64 acc2[C][NP][4];
#pragma unroll
    for (int c = 0; c < C; ++c)
#pragma unroll
        for (int pp = 0; pp < NP; ++pp)
#pragma unroll
            for (int ni = 0; ni < 4; ++ni) acc2[c][pp][ni] = pack2(acc[c][2 * pp][ni], acc[c][2 * pp + 1][ni]);
    float4 wq[2];
    PackedVec<TP> aq[2][C];
    // float offset, inside an aligned block of 8 feature rows, of this thread's granule in row kk
    int go[8];
#pragma unroll
    for (int kk = 0; kk < 8; ++kk) go[kk] = kk * P + (pg ^ S::swz(kk)) * TP;     // exact for HW128 (swz = row & 7)
    {
        const float* ap = A + (pg ^ S::swz(0)) * TP;
        wq[0] = *reinterpret_cast<const float4*>(W + n0);
#pragma unroll
        for (int c = 0; c < C; ++c) aq[0][c].load(ap + c * HP * P);
    }
#pragma unroll 2
    for (int k0 = 0; k0 < HP; k0 += 8) {
        const int g0 = pg ^ S::swz(k0);
        const float* ap0 = A + (k0 * PG + g0) * TP;                           // rows k0 .. k0+3
        const float* ap1 = A + (k0 * PG + (g0 ^ S::SWZ_HALF)) * TP;           // rows k0+4 .. k0+7
        const float* apn = A + ((k0 + 8) * PG + (pg ^ S::swz(k0 + 8))) * TP;
        const float* ab = A + k0 * P;                                         // HW128: block base
        const float* wp = W + k0 * HP + n0;
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
            const int cur = kk & 1, nxt = cur ^ 1;
            if (kk < 7) {
                wq[nxt] = *reinterpret_cast<const float4*>(wp + (kk + 1) * HP);
                if (S::HW128) {
#pragma unroll
                    for (int c = 0; c < C; ++c) aq[nxt][c].load(ab + go[kk + 1] + c * HP * P);
                } else {
                    const float* ap = (kk + 1 < 4) ? ap0 : ap1;
#pragma unroll
                    for (int c = 0; c < C; ++c) aq[nxt][c].load(ap + (c * HP + kk + 1) * P);
                }
            } else if (k0 + 8 < HP) {
                wq[nxt] = *reinterpret_cast<const float4*>(wp + 8 * HP);
#pragma unroll
                for (int c = 0; c < C; ++c) aq[nxt][c].load(apn + c * HP * P);
            }
            const float wv[4] = {wq[cur].x, wq[cur].y, wq[cur].z, wq[cur].w};
#pragma unroll
            for (int c = 0; c < C; ++c)
#pragma unroll
                for (int pp = 0; pp < NP; ++pp)
#pragma unroll
                    for (int ni = 0; ni < 4; ++ni) ffma2_acc_bcast(acc2[c][pp][ni], aq[cur][c].pair(pp), wv[ni]);
        }
    }
#pragma unroll
    for (int c = 0; c < C; ++c)
#pragma unroll
        for (int pp = 0; pp < NP; ++pp)
#pragma unroll
            for (int ni = 0; ni < 4; ++ni) unpack2(acc2[c][pp][ni], acc[c][2 * pp][ni], acc[c][2 * pp + 1][ni]);
}

// g[i][j] += sum_{c, p} A[c][row_a + i][p] * Z[c][row_z + j][p]   (both activation buffers), a KT x JH block.
// The reduction index p is the packed direction: two partial sums (even / odd points) per output,
// folded at the end.  JH = KT/2 keeps the packed accumulators at 64 registers; the kernel makes two
// passes over the operands for the two column halves.
template <class S, int JH, bool DB = true>
__device__ __forceinline__ void gemm_wgrad(const float* __restrict__ A, const float* __restrict__ Z, int row_a, int row_z,
                                           float (&g)[S::KT][JH]) {
    constexpr int C = S::C, TP = S::TP, HP = S::HP, PG = S::PG, P = S::P, KT = S::KT;
    constexpr int NQ = P / 4, U = C * NQ;     // units: 4 points of one channel
    static_assert(U % 2 == 0, "unit count");
    u64 g2[KT][JH];
#pragma unroll
    for (int i = 0; i < KT; ++i)
#pragma unroll
        for (int j = 0; j < JH; ++j) g2[i][j] = pack2(g[i][j], 0.f);
    ulonglong2 av[DB ? 2 : 1][KT], zv[DB ? 2 : 1][JH];
    auto load = [&](int u, int buf) {
        const int c = u / NQ, q = u - c * NQ;
#pragma unroll
        for (int i = 0; i < KT; ++i) {
            const int row = row_a + i;
            const int gq4 = TP == 4 ? (q ^ S::swz(row)) : ((2 * q) ^ S::swz(row));
            av[buf][i] = *reinterpret_cast<const ulonglong2*>(A + ((c * HP + row) * PG + gq4) * TP);
        }
#pragma unroll
        for (int j = 0; j < JH; ++j) {
            const int row = row_z + j;
            const int gq4 = TP == 4 ? (q ^ S::swz(row)) : ((2 * q) ^ S::swz(row));
            zv[buf][j] = *reinterpret_cast<const ulonglong2*>(Z + ((c * HP + row) * PG + gq4) * TP);
        }
    };
    auto fma_unit = [&](int buf) {
#pragma unroll
        for (int i = 0; i < KT; ++i)
#pragma unroll
            for (int j = 0; j < JH; ++j) {
                ffma2_acc(g2[i][j], av[buf][i].x, zv[buf][j].x);
                ffma2_acc(g2[i][j], av[buf][i].y, zv[buf][j].y);
            }
    };
    if (DB) {
        load(0, 0);
#pragma unroll 1
        for (int u = 0; u < U; u += 2) {
            load(u + 1, 1);
            fma_unit(0);
            if (u + 2 < U) load(u + 2, 0);
            fma_unit(DB ? 1 : 0);
        }
    } else {
#pragma unroll 1
        for (int u = 0; u < U; ++u) {
            load(u, 0);
            fma_unit(0);
        }
    }
#pragma unroll
    for (int i = 0; i < KT; ++i)
#pragma unroll
        for (int j = 0; j < JH; ++j) {
            float lo, hi;
            unpack2(g2[i][j], lo, hi);
            g[i][j] = lo + hi;
        }
}

template <int K0, int K1, int K2, int HP, int TP, bool BWD>
__global__ void __launch_bounds__(kThreads, 1) fused_jet_kernel(const __grid_constant__ KParams prm) {
    using S = Shape<K0, K1, K2, HP, TP>;
    using J = typename S::J;
    using Vec = typename VecT<TP>::type;
    constexpr int C = S::C, NG = S::NG, PG = S::PG, P = S::P, WPG = S::WPG, KT = S::KT;

    extern __shared__ __align__(1024) float smem[];
    float* As = smem + S::SM_AS;
    float* Zs = smem + S::SM_ZS;
    float* Ws = smem + S::SM_WS;
    const int L = prm.L;
    const PackLayout lay{HP, L};
    const int nsmall = lay.small();
    float* sp = smem + S::SM_END_BIG;           // small params (PackLayout order)
    float* sg = sp + nsmall;                    // their gradient accumulators
    float* red = sg + nsmall;                   // [WNG][C][P] = kThreads*C
    float* redB = red + kThreads * C;           // [WPG][4][HP]
    float* Xs = redB + WPG * 4 * HP;            // [3][P]
    float* Ub = Xs + 3 * P;                     // [C][P]
    float* wsc = Ub + C * P;                    // 64 floats of warp scratch
    tdq_op* ops_s = reinterpret_cast<tdq_op*>(wsc + 64);
    __shared__ uint64_t wbar;

    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    // lane -> (point group, neuron group) inside the warp's 8 x 4 patch.  HW128: lane bits (b4 b3 b2 b1 b0) =
    // (ng1 pg2 pg1 pg0 ng0), so a quarter-warp covers 4 point groups x 2 neuron groups whose element-wise stores
    // (rows 4 ng + i, chunk pg ^ (row & 7)) fall into 8 different 16-byte slots.
    constexpr int NGX1 = 1, NGX2 = S::HW128 ? 16 : 2;             // shuffle distances across neuron groups
    constexpr int PGX1 = S::HW128 ? 2 : 4, PGX2 = S::HW128 ? 4 : 8, PGX3 = S::HW128 ? 8 : 16;   // across point groups
    const int pg_l = S::HW128 ? ((lane >> 1) & 7) : (lane >> 2);
    const int ng_l = S::HW128 ? ((lane & 1) | ((lane >> 4) << 1)) : (lane & 3);
    const int wpg = w % WPG, wng = w / WPG;
    const int pg = wpg * 8 + pg_l, ng = wng * 4 + ng_l;
    const int n0 = ng * 4;                      // first of this thread's 4 neurons
    const int p0 = pg * TP;                     // first of this thread's TP points (within the tile)
    // wgrad mapping: 16 x 16 thread tiles of KT x KT
    const int kg = (w & 1) * 8 + (lane >> 2), jg = (w >> 1) * 4 + (lane & 3);

    const int d = prm.d;
    const long long n_tiles = prm.n_tiles;
    float* slab = prm.slab + (size_t)blockIdx.x * lay.total();
    float* stash = BWD ? prm.stash + (size_t)blockIdx.x * L * (C * 4) * kThreads * TP : nullptr;
    const float* whid = prm.wpack + lay.off_wh();
    const float* whid_t = prm.wpack_t;
    constexpr uint32_t kWBytes = HP * HP * 4;

    // ---- prologue: small params, program, barrier
    for (int i = tid; i < nsmall; i += kThreads) { sp[i] = prm.wpack[i]; sg[i] = 0.f; }
    for (int i = tid; i < prm.n_ops; i += kThreads) ops_s[i] = prm.ops[i];
    if (tid == 0) mbar_init(&wbar, 1);
    // Tensor Memory as stash: each thread parks its (tanh value, pre-activation derivatives) of the first
    // NT layers in its own TMEM row instead of the global scratch slab; the two warps sharing a lane
    // quarter take one half of the 512 columns each.  Layers beyond NT (deep nets) use the slab.
    constexpr int NV = C * TP * 4;                 // stash values per thread per layer
    __shared__ uint32_t tmem_slot;
    uint32_t tmem_base = 0;
    int NT = 0;
    // wgrad on the tensor cores (HW128 shapes): dW_l of the current tile and layer accumulates in TMEM columns
    // kDwCol .. kDwCol+127 (lane = row k of W_l, column = j), which leaves 192 stash columns per thread.
    constexpr bool TCW = BWD && S::HW128;
    constexpr int kDwCol = 384;
    const bool tcw = TCW && L >= 2 && !(prm.dbg & 8);
    __shared__ uint64_t mbar_mma[2];               // one per half of the lo-image staging area (Ws)
    uint32_t mph0 = 0, mph1 = 0;                   // their phase parities
    if (TCW && tid == 0) { mbar_init(&mbar_mma[0], 1); mbar_init(&mbar_mma[1], 1); }
    if (BWD) {
        if (w == 0) tmem_alloc_512(&tmem_slot);
        tmem_fence_before_sync();
    }
    __syncthreads();
    const int tcols = tcw ? kDwCol / 2 : 256;      // stash columns of this warp's half
    if (BWD) {
        tmem_fence_after_sync();
        tmem_base = tmem_slot;
        NT = (prm.dbg & 4) ? 0 : (tcols / NV < L ? tcols / NV : L);
    }
    const int tcol0 = (w >> 2) * tcols;            // this warp's half of the stash columns
    uint32_t wphase = 0;
    const bool has_tiles = (long long)blockIdx.x < n_tiles;
    if (L >= 2 && has_tiles && tid == 0) {
        mbar_expect_tx(&wbar, kWBytes);
        tma_load_1d(Ws, whid, kWBytes, &wbar);             // W_2 for the first forward GEMM
    }
    const float* b_s = sp + lay.off_b();
    const float* wout_s = sp + lay.off_wout();
    float loss_acc = 0.f;        // thread 0 only
    bool first_tile = true;

    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const Seg& sq = prm.seg[seg_of_tile(prm, tile)];       // the term this tile belongs to
        const long long base = (tile - sq.tile0) * P;
        const long long n = sq.n;
        const tdq_op* ops_t = ops_s + sq.ops_off;
        // ---- coordinates of this tile
        for (int i = tid; i < d * P; i += kThreads) {
            int dim = i / P, p = i - dim * P;
            long long gi = base + p;
            Xs[dim * P + p] = gi < n ? sq.X[(size_t)dim * sq.x_stride + gi] : 0.f;
        }
        __syncthreads();

        float acc[C][TP][4];
        // ================= forward =================
        // layer 1: z0 = x.W0 + b, dz/dx_i = W0[i], higher orders 0
#pragma unroll
        for (int c = 0; c < C; ++c)
#pragma unroll
            for (int pi = 0; pi < TP; ++pi)
#pragma unroll
                for (int ni = 0; ni < 4; ++ni) acc[c][pi][ni] = 0.f;
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) {
            float bz = b_s[n0 + ni];
#pragma unroll
            for (int pi = 0; pi < TP; ++pi) acc[0][pi][ni] = bz;
        }
        for (int dim = 0; dim < d; ++dim) {
#pragma unroll
            for (int ni = 0; ni < 4; ++ni) {
                float wv = sp[lay.off_w0() + dim * HP + n0 + ni];
#pragma unroll
                for (int pi = 0; pi < TP; ++pi) acc[0][pi][ni] += Xs[dim * P + p0 + pi] * wv;
                if (K0 > 0 && dim == 0) {
#pragma unroll
                    for (int pi = 0; pi < TP; ++pi) acc[J::OFF0][pi][ni] = wv;
                }
                if (K1 > 0 && dim == 1) {
#pragma unroll
                    for (int pi = 0; pi < TP; ++pi) acc[K1 > 0 ? J::OFF1 : 0][pi][ni] = wv;
                }
                if (K2 > 0 && dim == 2) {
#pragma unroll
                    for (int pi = 0; pi < TP; ++pi) acc[K2 > 0 ? J::OFF2 : 0][pi][ni] = wv;
                }
            }
        }

        for (int l = 1; l <= L; ++l) {
            if (l >= 2) {
                // ---- GEMM: z^l[c][p][n] = sum_k a^{l-1}[c][k][p] W_l[k][n]
                mbar_wait(&wbar, wphase);
                wphase ^= 1;
#pragma unroll
                for (int c = 0; c < C; ++c)
#pragma unroll
                    for (int pi = 0; pi < TP; ++pi)
#pragma unroll
                        for (int ni = 0; ni < 4; ++ni) acc[c][pi][ni] = 0.f;
#pragma unroll
                for (int ni = 0; ni < 4; ++ni) {
                    float bz = b_s[(l - 1) * HP + n0 + ni];
#pragma unroll
                    for (int pi = 0; pi < TP; ++pi) acc[0][pi][ni] = bz;
                }
                if (!(prm.dbg & 1)) gemm_act_w<S>(As, Ws, pg, n0, acc);
                __syncthreads();   // everyone is done with As and Ws
                if (tid == 0) {
                    if (l < L) {
                        mbar_expect_tx(&wbar, kWBytes);
                        tma_load_1d(Ws, whid + (size_t)(l - 1) * HP * HP, kWBytes, &wbar);       // W_{l+1}
                    } else if (BWD) {
                        mbar_expect_tx(&wbar, kWBytes);
                        tma_load_1d(Ws, whid_t + (size_t)(L - 2) * HP * HP, kWBytes, &wbar);    // W_L^T
                    }
                }
            }
            // ---- element-wise: tanh jets.  acc <- stash values (tanh value, pre-activation
            //      derivatives); the activations become the next layer's operand in As.
#pragma unroll
            for (int ni = 0; ni < 4; ++ni) {
                float actv[C][TP];
#pragma unroll
                for (int pi = 0; pi < TP; ++pi) {
                    float st[C], act[C];
                    st[0] = tanh_acc(acc[0][pi][ni]);
#pragma unroll
                    for (int c = 1; c < C; ++c) st[c] = acc[c][pi][ni];
                    J::recompute(st, act);
#pragma unroll
                    for (int c = 0; c < C; ++c) { acc[c][pi][ni] = st[c]; actv[c][pi] = act[c]; }
                }
#pragma unroll
                for (int c = 0; c < C; ++c)
                    *reinterpret_cast<Vec*>(As + S::aoff(c, n0 + ni, pg)) = vec_from<TP>(actv[c]);
            }
            if (BWD) {
                if (l - 1 < NT) {
#pragma unroll
                    for (int c = 0; c < C; ++c)
#pragma unroll
                        for (int h = 0; h < TP / 2; ++h) {
                            float v[8];
#pragma unroll
                            for (int i = 0; i < 8; ++i) v[i] = acc[c][2 * h + (i >> 2)][i & 3];
                            tmem_st8(tmem_addr(tmem_base, w, tcol0 + (l - 1) * NV + (c * (TP / 2) + h) * 8), v);
                        }
                } else {
#pragma unroll
                    for (int c = 0; c < C; ++c)
#pragma unroll
                        for (int ni = 0; ni < 4; ++ni) {
                            float v[TP];
#pragma unroll
                            for (int pi = 0; pi < TP; ++pi) v[pi] = acc[c][pi][ni];
                            *reinterpret_cast<Vec*>(stash + ((size_t)((l - 1) * C * 4 + c * 4 + ni) * kThreads + tid) * TP) =
                                vec_from<TP>(v);
                        }
                }
            }
            __syncthreads();   // As holds a^l
        }

        // ================= output layer + integrand =================
        // u_c[p] = sum_n a^L[c][n][p] w_out[n] (+ b_out on c = 0); thread-partial over its 4 neurons,
        // read back from As (its own writes)
        {
            float part[C][TP];
#pragma unroll
            for (int c = 0; c < C; ++c)
#pragma unroll
                for (int pi = 0; pi < TP; ++pi) part[c][pi] = 0.f;
#pragma unroll
            for (int ni = 0; ni < 4; ++ni) {
                float wo = wout_s[n0 + ni];
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    Vec av = *reinterpret_cast<const Vec*>(As + S::aoff(c, n0 + ni, pg));
                    float a[TP];
                    vec_to<TP>(av, a);
#pragma unroll
                    for (int pi = 0; pi < TP; ++pi) part[c][pi] = fmaf(a[pi], wo, part[c][pi]);
                }
            }
#pragma unroll
            for (int c = 0; c < C; ++c)
#pragma unroll
                for (int pi = 0; pi < TP; ++pi) {
                    float v = part[c][pi];
                    v += __shfl_xor_sync(0xffffffffu, v, NGX1);     // over the 4 neuron groups of the warp
                    v += __shfl_xor_sync(0xffffffffu, v, NGX2);
                    if (ng_l == 0) red[(wng * C + c) * P + p0 + pi] = v;
                }
        }
        __syncthreads();

        if (tid < ((P + 31) & ~31)) {   // whole warps: the reductions below use full-mask shuffles
            const int p = tid;
            const bool inrange = p < P;
            const long long gi = base + p;
            const bool valid = inrange && gi < n;
            float jets[C];
#pragma unroll
            for (int c = 0; c < C; ++c) {
                float v = 0.f;
                if (inrange)
                    for (int q = 0; q < NG / 4; ++q) v += red[(q * C + c) * P + p];
                jets[c] = v;
            }
            jets[0] += sp[lay.off_sc()];   // b_out
            float coord[3] = {0.f, 0.f, 0.f};
            for (int dim = 0; dim < d; ++dim) coord[dim] = inrange ? Xs[dim * P + p] : 0.f;
            VmIn in;
            in.jets = jets;
            in.coord = coord;
            in.target = (valid && sq.target) ? sq.target[gi] : 0.f;
            in.lam = (valid && sq.lam) ? sq.lam[gi] : 0.f;
            in.vars = prm.vars;
            float slot[TDQ_MAX_OPS];
            vm_forward(ops_t, sq.n_ops, in, slot);
            float outv = slot[sq.out_slot];
            if (valid && sq.f_out && sq.f_slot >= 0) sq.f_out[gi] = slot[sq.f_slot];
            if (valid && sq.u_out) sq.u_out[gi] = jets[0];
            float lsum = 0.f;
            if (BWD) {
                float seed;
                if (sq.pair) {
                    float other = __shfl_xor_sync(0xffffffffu, outv, 1);
                    float r = (p & 1) ? (other - outv) : (outv - other);
                    seed = ((p & 1) ? -2.f : 2.f) * r * sq.inv_n;
                    lsum = (p & 1) ? 0.f : r * r * sq.inv_n;
                } else {
                    seed = sq.inv_n;
                    lsum = outv * sq.inv_n;
                }
                if (!valid) { seed = 0.f; lsum = 0.f; }
                float adj[TDQ_MAX_OPS];
                for (int i = 0; i < sq.n_ops; ++i) adj[i] = 0.f;
                adj[sq.out_slot] = seed;
                float jb[C];
#pragma unroll
                for (int c = 0; c < C; ++c) jb[c] = 0.f;
                float vb[TDQ_MAX_VARS];
#pragma unroll
                for (int v = 0; v < TDQ_MAX_VARS; ++v) vb[v] = 0.f;
                float lam_bar = vm_reverse(ops_t, sq.n_ops, slot, adj, jb, vb);
                if (inrange) {
#pragma unroll
                    for (int c = 0; c < C; ++c) Ub[c * P + p] = jb[c];
                }
                if (valid && sq.lam_mode >= 1) {
                    if (sq.lam_grad) sq.lam_grad[gi] = lam_bar;
                    if (sq.lam_mode == 2) {
                        // keras Adam on this point's lambda, gradient sign-flipped for ascent (fit.py:101,104)
                        float g = prm.lam_adam.sign * lam_bar;
                        float m = sq.lam_m[gi], v = sq.lam_v[gi];
                        m += (g - m) * (1.f - prm.lam_adam.beta1);
                        v += (g * g - v) * (1.f - prm.lam_adam.beta2);
                        sq.lam_m[gi] = m;
                        sq.lam_v[gi] = v;
                        sq.lam[gi] = in.lam - (m * prm.lam_adam.alpha) / (sqrtf(v) + prm.lam_adam.eps);
                    }
                }
                // reductions over the tile: loss, d b_out (= sum of cotangents of u), d vars
                float q0 = lsum, q1 = jb[0];
#pragma unroll
                for (int o = 16; o >= 1; o >>= 1) {
                    q0 += __shfl_xor_sync(0xffffffffu, q0, o);
                    q1 += __shfl_xor_sync(0xffffffffu, q1, o);
                }
                for (int v = 0; v < prm.n_vars; ++v) {
                    float qv = vb[v];
#pragma unroll
                    for (int o = 16; o >= 1; o >>= 1) qv += __shfl_xor_sync(0xffffffffu, qv, o);
                    if (lane == 0) wsc[16 + w * TDQ_MAX_VARS + v] = qv;
                }
                if (lane == 0) { wsc[w * 2] = q0; wsc[w * 2 + 1] = q1; }
            }
        }
        if (!BWD) {
            __syncthreads();   // red / Xs reuse by the next tile
            if (L >= 2 && tile + gridDim.x < n_tiles && tid == 0) {
                mbar_expect_tx(&wbar, kWBytes);
                tma_load_1d(Ws, whid, kWBytes, &wbar);
            }
            first_tile = false;
            continue;
        }
        __syncthreads();
        if (tid == 0) {
            constexpr int NW = (P + 31) / 32;
            float l0 = 0.f, l1 = 0.f;
            for (int i = 0; i < NW; ++i) { l0 += wsc[i * 2]; l1 += wsc[i * 2 + 1]; }
            loss_acc += l0;
            sg[lay.off_sc()] += l1;
            for (int v = 0; v < prm.n_vars; ++v) {
                float s = 0.f;
                for (int i = 0; i < NW; ++i) s += wsc[16 + i * TDQ_MAX_VARS + v];
                sg[lay.off_sc() + 2 + v] += s;
            }
        }

        // ================= backward =================
        // cotangent of a^L: ab[c][p][n] = ubar[c][p] * w_out[n];  d w_out[n] += sum_{c,p} a^L[c][n][p] ubar[c][p]
        {
            float ub[C][TP];
#pragma unroll
            for (int c = 0; c < C; ++c) {
                Vec uv = *reinterpret_cast<const Vec*>(Ub + c * P + p0);
                vec_to<TP>(uv, ub[c]);
            }
            float gw[4];
#pragma unroll
            for (int ni = 0; ni < 4; ++ni) {
                float wo = wout_s[n0 + ni];
                float s = 0.f;
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    Vec av = *reinterpret_cast<const Vec*>(As + S::aoff(c, n0 + ni, pg));
                    float a[TP];
                    vec_to<TP>(av, a);
#pragma unroll
                    for (int pi = 0; pi < TP; ++pi) {
                        s = fmaf(a[pi], ub[c][pi], s);
                        acc[c][pi][ni] = ub[c][pi] * wo;
                    }
                }
                gw[ni] = s;
            }
            // reduce over the point groups
#pragma unroll
            for (int ni = 0; ni < 4; ++ni) {
                float v = gw[ni];
                v += __shfl_xor_sync(0xffffffffu, v, PGX1);     // over the 8 point groups of the warp
                v += __shfl_xor_sync(0xffffffffu, v, PGX2);
                v += __shfl_xor_sync(0xffffffffu, v, PGX3);
                gw[ni] = v;
            }
            if (WPG == 1) {
                if (pg_l == 0) {
#pragma unroll
                    for (int ni = 0; ni < 4; ++ni) sg[lay.off_wout() + n0 + ni] += gw[ni];
                }
            } else {
                if (pg_l == 0) {
#pragma unroll
                    for (int ni = 0; ni < 4; ++ni) redB[wpg * 4 * HP + n0 + ni] = gw[ni];
                }
                __syncthreads();
                for (int i = tid; i < HP; i += kThreads) {
                    float s = 0.f;
                    for (int q = 0; q < WPG; ++q) s += redB[q * 4 * HP + i];
                    sg[lay.off_wout() + i] += s;
                }
                __syncthreads();
            }
        }

        // stash of `layer` -> registers, from TMEM (first NT layers) or the global slab
        auto load_stash = [&](int layer, float (&dst)[C][TP][4]) {
            if (layer - 1 < NT) {
#pragma unroll
                for (int c = 0; c < C; ++c)
#pragma unroll
                    for (int h = 0; h < TP / 2; ++h) {
                        float v[8];
                        tmem_ld8(tmem_addr(tmem_base, w, tcol0 + (layer - 1) * NV + (c * (TP / 2) + h) * 8), v);
                        tmem_wait_ld();
#pragma unroll
                        for (int i = 0; i < 8; ++i) dst[c][2 * h + (i >> 2)][i & 3] = v[i];
                    }
            } else {
#pragma unroll
                for (int c = 0; c < C; ++c)
#pragma unroll
                    for (int ni = 0; ni < 4; ++ni) {
                        Vec sv = *reinterpret_cast<const Vec*>(
                            stash + ((size_t)((layer - 1) * C * 4 + c * 4 + ni) * kThreads + tid) * TP);
                        float v[TP];
                        vec_to<TP>(sv, v);
#pragma unroll
                        for (int pi = 0; pi < TP; ++pi) dst[c][pi][ni] = v[pi];
                    }
            }
        };
        tmem_wait_st();      // this thread's forward stash stores have landed
        bool tcw_pending = false;

        for (int l = L; l >= 1; --l) {
            // ---- element-wise reverse through tanh jets: acc (= cotangent of a^l) -> cotangent of z^l
            float gb[4] = {0.f, 0.f, 0.f, 0.f};
            {
                float stv[C][TP][4];
                load_stash(l, stv);
#pragma unroll
                for (int pi = 0; pi < TP; ++pi)
#pragma unroll
                    for (int ni = 0; ni < 4; ++ni) {
                        float st[C], ab[C], zb[C];
#pragma unroll
                        for (int c = 0; c < C; ++c) { st[c] = stv[c][pi][ni]; ab[c] = acc[c][pi][ni]; }
                        J::bwd(st, ab, zb);
#pragma unroll
                        for (int c = 0; c < C; ++c) acc[c][pi][ni] = zb[c];
                        gb[ni] += zb[0];
                    }
            }
            // ---- bias gradient (and first-layer weights when l == 1): sums over the tile's points
            float gq[4][4];   // [quantity][ni] : 0 = bias, 1.. = W0 rows (l == 1)
            int nq = 1;
#pragma unroll
            for (int ni = 0; ni < 4; ++ni) gq[0][ni] = gb[ni];
            if (l == 1) {
                nq = 1 + d;
#pragma unroll
                for (int dim = 0; dim < 3; ++dim) {
                    if (dim < d) {
#pragma unroll
                        for (int ni = 0; ni < 4; ++ni) {
                            float s = 0.f;
#pragma unroll
                            for (int pi = 0; pi < TP; ++pi) {
                                s = fmaf(Xs[dim * P + p0 + pi], acc[0][pi][ni], s);
                                if (dim == 0 && K0 > 0) s += acc[J::OFF0][pi][ni];
                                if (dim == 1 && K1 > 0) s += acc[K1 > 0 ? J::OFF1 : 0][pi][ni];
                                if (dim == 2 && K2 > 0) s += acc[K2 > 0 ? J::OFF2 : 0][pi][ni];
                            }
                            gq[1 + dim][ni] = s;
                        }
                    }
                }
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (q < nq) {
#pragma unroll
                    for (int ni = 0; ni < 4; ++ni) {
                        float v = gq[q][ni];
                        v += __shfl_xor_sync(0xffffffffu, v, PGX1);
                        v += __shfl_xor_sync(0xffffffffu, v, PGX2);
                        v += __shfl_xor_sync(0xffffffffu, v, PGX3);
                        gq[q][ni] = v;
                    }
                }
            }
            if (WPG == 1) {
                if (pg_l == 0) {
#pragma unroll
                    for (int ni = 0; ni < 4; ++ni) {
                        sg[lay.off_b() + (l - 1) * HP + n0 + ni] += gq[0][ni];
                        if (l == 1) {
#pragma unroll
                            for (int dim = 0; dim < 3; ++dim)
                                if (dim < d) sg[lay.off_w0() + dim * HP + n0 + ni] += gq[1 + dim][ni];
                        }
                    }
                }
            } else {
                if (pg_l == 0) {
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        if (q < nq) {
#pragma unroll
                            for (int ni = 0; ni < 4; ++ni) redB[(wpg * 4 + q) * HP + n0 + ni] = gq[q][ni];
                        }
                }
                __syncthreads();
                for (int i = tid; i < nq * HP; i += kThreads) {
                    int q = i / HP, nn = i - q * HP;
                    float s = 0.f;
                    for (int r = 0; r < WPG; ++r) s += redB[(r * 4 + q) * HP + nn];
                    if (q == 0) sg[lay.off_b() + (l - 1) * HP + nn] += s;
                    else sg[lay.off_w0() + (q - 1) * HP + nn] += s;
                }
                __syncthreads();
            }
            if (TCW && tcw_pending) {
                // dW_{l+1} of this tile is complete in TMEM once the tensor core has drained: Ws (which held the
                // lo images) is free for the next weight matrix, As / Zs may be overwritten, and the sums go
                // to the slab.
                if (C >= 2) {
                    if ((C - 2) & 1) { mbar_wait(&mbar_mma[1], mph1); mph1 ^= 1; } else { mbar_wait(&mbar_mma[0], mph0); mph0 ^= 1; }
                }
                if ((C - 1) & 1) { mbar_wait(&mbar_mma[1], mph1); mph1 ^= 1; } else { mbar_wait(&mbar_mma[0], mph0); mph0 ^= 1; }
                tmem_fence_after_sync();
                if (tid == 0) {
                    if (l >= 2) {
                        mbar_expect_tx(&wbar, kWBytes);
                        tma_load_1d(Ws, whid_t + (size_t)(l - 2) * HP * HP, kWBytes, &wbar);     // W_l^T
                    } else if (tile + gridDim.x < n_tiles) {
                        mbar_expect_tx(&wbar, kWBytes);
                        tma_load_1d(Ws, whid, kWBytes, &wbar);                                   // W_2 for the next tile
                    }
                }
                // thread (w, lane) reads row k = 32 (w % 4) + lane, columns 64 (w / 4) .. + 63
                const int kr = (w & 3) * 32 + lane, jc = (w >> 2) * 64;
                float* gdst = slab + lay.off_wh() + (size_t)(l - 1) * HP * HP + (size_t)kr * HP + jc;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    float v[16];
                    tmem_ld16(tmem_addr(tmem_base, w, kDwCol + jc + q * 16), v);
                    tmem_wait_ld();
#pragma unroll
                    for (int j = 0; j < 16; j += 4) {
                        if (first_tile)
                            *reinterpret_cast<float4*>(gdst + q * 16 + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                        else
                            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(gdst + q * 16 + j), "f"(v[j]),
                                         "f"(v[j + 1]), "f"(v[j + 2]), "f"(v[j + 3])
                                         : "memory");
                    }
                }
                tmem_fence_before_sync();      // the next layer's first MMA overwrites these columns (after a barrier)
                tcw_pending = false;
            }
            if (l == 1) break;

            // ---- operands of this layer's two GEMMs: Zs <- cotangent of z^l ; As <- a^{l-1} (recomputed)
#pragma unroll
            for (int c = 0; c < C; ++c)
#pragma unroll
                for (int ni = 0; ni < 4; ++ni) {
                    float v[TP];
#pragma unroll
                    for (int pi = 0; pi < TP; ++pi) v[pi] = acc[c][pi][ni];
                    *reinterpret_cast<Vec*>(Zs + S::aoff(c, n0 + ni, pg)) = vec_from<TP>(v);
                }
            {
                float stv[C][TP][4];
                load_stash(l - 1, stv);
#pragma unroll
                for (int ni = 0; ni < 4; ++ni) {
                    float actv[C][TP];
#pragma unroll
                    for (int pi = 0; pi < TP; ++pi) {
                        float st[C], act[C];
#pragma unroll
                        for (int c = 0; c < C; ++c) st[c] = stv[c][pi][ni];
                        J::recompute(st, act);
#pragma unroll
                        for (int c = 0; c < C; ++c) actv[c][pi] = act[c];
                    }
#pragma unroll
                    for (int c = 0; c < C; ++c)
                        *reinterpret_cast<Vec*>(As + S::aoff(c, n0 + ni, pg)) = vec_from<TP>(actv[c]);
                }
            }
            if (TCW && tcw) fence_proxy_async_smem();    // As / Zs are tensor-core operands below
            __syncthreads();

            // ---- wgrad: dW_l[k][j] += sum_{c,p} a^{l-1}[c][k][p] zbar^l[c][j][p]
            if (!(TCW && tcw)) {
                constexpr int VW = KT >= 4 ? 4 : 2;           // slab vector width
                float g[KT][KT];
#pragma unroll
                for (int i = 0; i < KT; ++i)
#pragma unroll
                    for (int j = 0; j < KT; ++j) g[i][j] = 0.f;
                if (!(prm.dbg & 2)) gemm_wgrad<S, KT, false>(As, Zs, kg * KT, jg * KT, g);
                float* gdst = slab + lay.off_wh() + (size_t)(l - 2) * HP * HP + (size_t)(kg * KT) * HP + jg * KT;
                if (first_tile) {
#pragma unroll
                    for (int i = 0; i < KT; ++i)
#pragma unroll
                        for (int j = 0; j < KT; j += VW) {
                            if (VW == 4)
                                *reinterpret_cast<float4*>(gdst + i * HP + j) = make_float4(g[i][j], g[i][j + 1], g[i][j + 2], g[i][j + 3]);
                            else
                                *reinterpret_cast<float2*>(gdst + i * HP + j) = make_float2(g[i][j], g[i][j + 1]);
                        }
                } else {
                    // Add into the running sums with vector reductions (red.global.add.v4.f32): the values
                    // travel to L2 and are added there, nothing comes back, so no round trip is exposed.
                    // Every slab word is touched by exactly one thread per tile and tiles are sequential, so
                    // the order of additions -- hence the result -- is the same run to run.
#pragma unroll
                    for (int i = 0; i < KT; ++i)
#pragma unroll
                        for (int j = 0; j < KT; j += VW) {
                            if (VW == 4)
                                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(gdst + i * HP + j), "f"(g[i][j]),
                                             "f"(g[i][j + 1]), "f"(g[i][j + 2]), "f"(g[i][j + 3])
                                             : "memory");
                            else
                                asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(gdst + i * HP + j), "f"(g[i][j]), "f"(g[i][j + 1])
                                             : "memory");
                        }
                }
            }

            // ---- dgrad: cotangent of a^{l-1}[c][p][k] = sum_j zbar^l[c][j][p] W_l[k][j]   (Ws = W_l^T)
            mbar_wait(&wbar, wphase);
            wphase ^= 1;
#pragma unroll
            for (int c = 0; c < C; ++c)
#pragma unroll
                for (int pi = 0; pi < TP; ++pi)
#pragma unroll
                    for (int ni = 0; ni < 4; ++ni) acc[c][pi][ni] = 0.f;
            if (!(prm.dbg & 1)) gemm_act_w<S>(Zs, Ws, pg, n0, acc);
            __syncthreads();   // As, Zs, Ws free
            if (TCW && tcw) {
                // ---- wgrad on the tensor cores, 3xTF32: with x = hi + lo (hi = x truncated to tf32, which is what
                // the tensor core reads out of the raw fp32 words; lo = x - hi, exact in fp32),
                //     a . z  ~=  a_lo . z_hi  +  a_hi . z_lo  +  a_hi . z_hi          (lo . lo ~ 2^-22 dropped)
                // The hi operands are the channel planes of As / Zs where they lie; the lo images of one channel
                // (2 x 16 KB) are staged in one half of Ws (free now that dgrad has read W_l^T), the two halves
                // alternating between channels so the threads stage channel c + 1 while the tensor core works on c.
                constexpr uint32_t kPlane = HP * P * 4;                        // bytes of one channel plane (16 KB)
                constexpr uint32_t idesc = umma_idesc_tf32(HP, HP);
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    const int hf = c & 1;
                    if (c >= 2) {        // the MMAs of channel c - 2 must have read this half
                        if (hf) { mbar_wait(&mbar_mma[1], mph1); mph1 ^= 1; } else { mbar_wait(&mbar_mma[0], mph0); mph0 ^= 1; }
                    }
                    float* Alo = Ws + hf * (2 * HP * P);
                    float* Zlo = Alo + HP * P;
#pragma unroll
                    for (int ni = 0; ni < 4; ++ni) {
                        const int o = S::aoff(0, n0 + ni, pg);
                        float4 a = *reinterpret_cast<const float4*>(As + c * HP * P + o);
                        float4 z = *reinterpret_cast<const float4*>(Zs + c * HP * P + o);
                        a.x -= __uint_as_float(__float_as_uint(a.x) & 0xFFFFE000u);
                        a.y -= __uint_as_float(__float_as_uint(a.y) & 0xFFFFE000u);
                        a.z -= __uint_as_float(__float_as_uint(a.z) & 0xFFFFE000u);
                        a.w -= __uint_as_float(__float_as_uint(a.w) & 0xFFFFE000u);
                        z.x -= __uint_as_float(__float_as_uint(z.x) & 0xFFFFE000u);
                        z.y -= __uint_as_float(__float_as_uint(z.y) & 0xFFFFE000u);
                        z.z -= __uint_as_float(__float_as_uint(z.z) & 0xFFFFE000u);
                        z.w -= __uint_as_float(__float_as_uint(z.w) & 0xFFFFE000u);
                        *reinterpret_cast<float4*>(Alo + o) = a;
                        *reinterpret_cast<float4*>(Zlo + o) = z;
                    }
                    fence_proxy_async_smem();
                    tmem_fence_before_sync();
                    __syncthreads();
                    if (tid == 0) {
                        tmem_fence_after_sync();
                        const uint32_t a_hi = smem_u32(As) + c * kPlane, z_hi = smem_u32(Zs) + c * kPlane;
                        const uint32_t a_lo = smem_u32(Alo), z_lo = smem_u32(Zlo);
                        const uint32_t dcol = tmem_base + kDwCol;
#pragma unroll
                        for (int ks = 0; ks < P / 8; ++ks) {                  // K = 8 points per instruction = 32 bytes
                            const uint32_t o = ks * 32;
                            umma_tf32(dcol, umma_desc_sw128(a_lo + o), umma_desc_sw128(z_hi + o), idesc, (c | ks) != 0);
                            umma_tf32(dcol, umma_desc_sw128(a_hi + o), umma_desc_sw128(z_lo + o), idesc, 1);
                            umma_tf32(dcol, umma_desc_sw128(a_hi + o), umma_desc_sw128(z_hi + o), idesc, 1);
                        }
                        umma_commit(&mbar_mma[hf]);
                    }
                }
                tcw_pending = true;
            } else
            if (tid == 0) {
                if (l - 1 >= 2) {
                    mbar_expect_tx(&wbar, kWBytes);
                    tma_load_1d(Ws, whid_t + (size_t)(l - 3) * HP * HP, kWBytes, &wbar);   // W_{l-1}^T
                } else if (tile + gridDim.x < n_tiles) {
                    mbar_expect_tx(&wbar, kWBytes);
                    tma_load_1d(Ws, whid, kWBytes, &wbar);                                   // W_2 for the next tile
                }
            }
        }
        first_tile = false;
        __syncthreads();   // Xs / sg / red reuse
    }

    // ---- epilogue: publish this CTA's slab (small part; hidden matrices were accumulated in place)
    if (BWD) {
        tmem_fence_before_sync();
        __syncthreads();
        if (w == 0) tmem_dealloc_512(tmem_base);
        if (tid == 0) sg[lay.off_sc() + 1] = loss_acc;
        __syncthreads();
        for (int i = tid; i < nsmall; i += kThreads) slab[i] = sg[i];
        if (!has_tiles) {
            for (int i = tid; i < (L - 1) * HP * HP; i += kThreads) slab[lay.off_wh() + i] = 0.f;
        }
    }
}

}  // namespace tdq
