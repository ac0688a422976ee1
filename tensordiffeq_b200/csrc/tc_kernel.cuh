// Tensor-core version of the fused jet kernel for 128-wide nets with at most 4 jet channels (the
// north-star Allen-Cahn shape): every hidden-layer GEMM of the step -- forward, data gradient and weight
// gradient -- runs on tcgen05.mma with fp32 accumulation in Tensor Memory, as the split
//     x . y  ~=  x_hi . y_hi  +  x_hi . y_lo  +  x_lo . y_hi,      x_hi = x truncated to tf32, x_lo = x - x_hi:
//   * x_hi . y_hi is one kind::tf32 MMA per K = 8 on the RAW fp32 words (the tensor core reads their tf32 part);
//   * the two correction terms are kind::f16 MMAs on bf16 images (K = 16 per instruction, i.e. half the tensor time of a
//     tf32 MMA per product): x_hi there is x rounded to bf16, x_lo the bf16 image of x - trunc_tf32(x).  Their A operand
//     always comes from Tensor Memory (two bf16 per column), their B operand from a bf16 image in shared memory.
// Error per product ~2^-19 (the corrections carry 8 bits of a 2^-11-sized term); measured against the FP32 FMA kernel
// the loss agrees to ~4e-6 and every gradient block to ~2e-6 (tests/test_gpu_tcw.py, test_gpu_tc_shapes.py), against
// the reference goldens as before (tolerance 1e-4).  Round 1 ran all three terms as kind::tf32 (48 instead of 32
// MMA-equivalents per GEMM, fp32 lo images: 13.6 vs 13.2 ms per launch; tc16_kernel.cuh still does).  Everything
// else (tanh jets and their reverse, the traced residual program, the small-parameter gradients) stays in
// FP32 on the CUDA cores, and is the same math as fused_kernel.cuh (models.py:80-199 of the reference).
//
// One persistent CTA per SM, 32-point tiles, 16 + 1 warps:
//   * 16 worker warps.  Thread (w, lane) owns neuron n = 32 (w % 4) + lane -- which is also its Tensor Memory
//     lane -- and the 8 points 8 (w / 4) .. + 7 of the tile, all C channels.  GEMM results arrive in its own
//     TMEM row (D^T[n][(c, p)]); it applies the bias, tanh jets, writes the next operand into shared memory
//     and parks the stash (tanh value, pre-activation derivatives) back into the same TMEM columns.
//   * 1 control warp: runs warp-uniform (so descriptors live in uniform registers), one lane issues the
//     weight-image TMA loads and all tcgen05.mma instructions; it talks to the workers through mbarriers only.
//
// Shared memory (bytes):  R0 64K | R1 64K | R2 64K | small
//   forward : R0 = a^{l-1} fp32, R1 = its bf16 images, R2 = weight ring (4 x 16 KB fp32 planes)
//   backward: R1 = zbar^l fp32, R2 = its bf16 images; R0 = weight ring during dgrad, then a^{l-1} fp32 for wgrad;
//             the bf16 images of a^{l-1} go to Tensor Memory (A operand of wgrad's correction MMAs), and so do the
//             bf16 images of the weight operand of every activation GEMM
// fp32 activation buffers are channel-planar [c][feature row 0..127][32 points], one row = one 128-byte line, in one
// of two swizzles, because the tensor core reads fp32 operands transposed (MN-major: N = (c, p) along the line,
// K = feature rows; planes LBO = 16 KB apart) only in the SWIZZLE_128B_BASE32B layout (32-byte chunk index XOR
// (row & 3)), while the weight-gradient GEMM reads the same kind of plane K-major (M or N = feature rows,
// K = points) in SWIZZLE_128B (16-byte chunk index XOR (row & 7)); probe modes 42..47 (tools/probe) pin both down.
// a^{l-1} is MN-major in the forward sweep and K-major in the backward sweep; zbar^l is written ONCE, MN-major for dgrad:
// wgrad reads the same fp32 planes K-major through a BASE32B descriptor (SBO = 512, LBO = 1024;
// tools/experiments/kmajor_base32b_probe.cu), and the same bf16 images as the K-major tiles they also are.
// bf16 images (layouts pinned by tools/experiments/bf16_probe.cu), plain SWIZZLE_128B on [128 rows][64 bf16] planes:
//   MN-major (activation GEMMs): a 64-wide N atom = 2 channels x 32 points; region = [hi16 atom 0 | lo16 atom 0 | hi16
//   atom 1 | lo16 atom 1], 16 feature rows per MMA.  The same rows read K-major (wgrad) are 64 K elements = the 32 points
//   of channel 2 j, then of channel 2 j + 1: a correction MMA takes 16 points of one channel from zbar's hi16 (lo16) image
//   against [a_lo16 (a_hi16) of those 16 points] = 8 Tensor Memory columns, written IN PLACE of the dgrad result by the
//   two threads that own the points (each also writes 4 of its partner's columns: a barrier of the warp pair orders it).
// Weight images (tdq::tc_img_*; written by the pack kernels): per hidden matrix and direction, 4 K-blocks of
// [128 x 32] fp32 in K-major SWIZZLE_128B order; one 16 KB bulk copy brings one plane = one ring slot.
//
// Tensor Memory (512 columns): block l-2 of NB = 32 C columns holds, for hidden layer l = 2..L, first the GEMM
// result z^l, then in place the stash of layer l; in the backward sweep the data gradient w.r.t. a^{l-1} and then
// the bf16 images of a^{l-1}.  Columns 384..511 hold dW_l of the current tile.  The bf16 images of the weight operand
// of an activation GEMM (64 + 64 columns: W rounded to bf16, its lo parts) go where nothing is live at that moment:
// the dW columns in the forward sweep and for the first dgrad of a tile, the columns above the live stash blocks
// otherwise.  Layer 1 needs no stash (it is recomputed from x).
#pragma once
#include <stdio.h>

#include <type_traits>

#include <cuda_bf16.h>

#include "fused_kernel.cuh"
#include "tc_image.cuh"
#include "vm_poly.cuh"

namespace tdq {

// MN-major fp32 operand = SWIZZLE_128B_BASE32B: 32 elements along N per 128-byte line, 32-byte chunk index XOR (row & 3),
// groups of 4 K rows 512 bytes apart (SBO), 32-element groups along N one channel plane apart (LBO); the
// descriptors themselves are built in the control warp (dk / dmn).
// float offset of (row, point) in an MN-major plane
__device__ __forceinline__ int tc_mn32_off(int row, int p) { return row * 32 + ((((p >> 3) ^ (row & 3)) << 3) | (p & 7)); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// Developer build (-DTDQ_TC_PROFILE): warps 0 and 5 of CTA 0 accumulate clock() deltas per phase and print them.
#ifdef TDQ_TC_PROFILE
#define TCP(i) do { unsigned _t = clock(); tcp_acc[i] += (_t - tcp_t0) >> 4; tcp_t0 = _t; } while (0)
#else
#define TCP(i) do { } while (0)
#endif

template <int N>
__device__ __forceinline__ void worker_sync() { asm volatile("bar.sync 1, %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ float tf32_lo(float x) { return x - __uint_as_float(__float_as_uint(x) & 0xFFFFE000u); }
__device__ __forceinline__ float4 tf32_lo4(float4 v) { return make_float4(tf32_lo(v.x), tf32_lo(v.y), tf32_lo(v.z), tf32_lo(v.w)); }

// two floats -> one 32-bit word of two bf16 (round to nearest), the FIRST in the low half (= the even K element of a
// Tensor Memory A operand, the lower address in shared memory)
__device__ __forceinline__ uint32_t pack_bf16(float first, float second) {
    __nv_bfloat162 t = __floats2bfloat162_rn(first, second);
    return *reinterpret_cast<uint32_t*>(&t);
}
// 8 values -> their bf16 image (h) and the bf16 image of their lo parts x - trunc_tf32(x) (l), 16 bytes each
__device__ __forceinline__ uint32_t lo16_pair(float a, float b) {      // one packed subtraction for the two lo parts
    const F2 lo = F2(a, b) - F2(__uint_as_float(__float_as_uint(a) & 0xFFFFE000u), __uint_as_float(__float_as_uint(b) & 0xFFFFE000u));
    return pack_bf16(lo.x, lo.y);
}
__device__ __forceinline__ void split16(const float* x, uint4& h, uint4& l) {
    h.x = pack_bf16(x[0], x[1]); h.y = pack_bf16(x[2], x[3]); h.z = pack_bf16(x[4], x[5]); h.w = pack_bf16(x[6], x[7]);
    l.x = lo16_pair(x[0], x[1]); l.y = lo16_pair(x[2], x[3]); l.z = lo16_pair(x[4], x[5]); l.w = lo16_pair(x[6], x[7]);
}

// N x 32-bit columns of this thread's TMEM lane <-> registers, N = 8 or 16
template <int N>
__device__ __forceinline__ void tmem_ldn(uint32_t taddr, float* v) {
    if (N == 16) tmem_ld16(taddr, v); else tmem_ld8(taddr, v);
}
template <int N>
__device__ __forceinline__ void tmem_stn(uint32_t taddr, const float* v) {
    if (N == 16) tmem_st16(taddr, v); else tmem_st8(taddr, v);
}

// NW = number of worker warps (8 or 16): NH = NW / 4 warps share a TMEM lane quarter and split the tile's 32
// points into NH groups of PT = 32 / NH
// FWD_ONLY: forward sweep + residual program only (tdq_predict): no stash, no reverse sweep, no slab (a compile-time
// switch: as a run-time flag its branches cost the training instantiation 3 %)
template <int K0, int K1, int K2, int NW, bool FWD_ONLY = false>
__global__ void __launch_bounds__((NW + 1) * 32, 1) fused_jet_tc_kernel(const __grid_constant__ KParams prm) {
    using J = JetCfg<K0, K1, K2>;
    constexpr int C = J::C;
    static_assert(C <= 4, "tensor-core jet kernel: at most 4 channels");
    static_assert(NW == 16, "16 worker warps (8 points per thread: one 16-byte bf16 piece per channel)");
    constexpr int HP = 128, P = 32, NB = C * 32;
    constexpr int NH = NW / 4, PT = P / NH, NJ = PT / 4, kWorkers = NW * 32, kThreadsAll = (NW + 1) * 32;

    extern __shared__ __align__(1024) float smem[];
    float* R0 = smem;
    float* R1 = R0 + 4 * kTcPlaneF;
    float* R2 = R1 + 4 * kTcPlaneF;
    const int L = prm.L;
    const PackLayout lay{HP, L};
    const int nsmall = lay.small();
    float* sp = R2 + 4 * kTcPlaneF;             // small params (PackLayout order)
    float* sgb = sp + nsmall;                   // [NH][L][HP] bias-gradient accumulators, one copy per point group
    float* sgs = sgb + NH * L * HP;             // [kScalars] d b_out, loss, d vars (warp 0)
    float* Xs = sgs + kScalars;                 // [3][P]
    float* Ub = Xs + 3 * P;                     // [4][P]
    float* red = Ub + 4 * P;                    // [4][4][P]
    float* wsc = red + 4 * 4 * P;               // 64
    TcOp* ops_s = reinterpret_cast<TcOp*>(wsc + 64);        // 32-byte records: wsc + 64 floats keeps them 16-byte aligned
    __shared__ uint64_t bar_full[4], bar_empty[4], bar_wlo, bar_opnd, bar_done, bar_lo_ready[4], bar_ch[4], bar_wg, bar_r0;
    __shared__ uint32_t tmem_slot;

    const int tid = threadIdx.x, lane = tid & 31;
    const int w = __shfl_sync(0xffffffffu, tid >> 5, 0);      // warp-uniform for the compiler: role branches do not diverge
    const int d = prm.d;
    const long long n_tiles = prm.n_tiles;
    float* slab = prm.slab + (size_t)blockIdx.x * lay.total();
    const bool has_tiles = (long long)blockIdx.x < n_tiles;

    // ---- prologue
    for (int i = tid; i < nsmall; i += kThreadsAll) sp[i] = prm.wpack[i];
    for (int i = tid; i < NH * L * HP + kScalars; i += kThreadsAll) sgb[i] = 0.f;
    for (int i = tid; i < prm.tc_n_ops; i += kThreadsAll) ops_s[i] = prm.tc_ops[i];
    if (tid == 0) {
        for (int i = 0; i < 4; ++i) { mbar_init(&bar_full[i], 1); mbar_init(&bar_empty[i], 1); }
        mbar_init(&bar_wlo, NW);
        mbar_init(&bar_opnd, NW); mbar_init(&bar_done, 1);
        for (int i = 0; i < 4; ++i) { mbar_init(&bar_lo_ready[i], NW); mbar_init(&bar_ch[i], 1); }
        mbar_init(&bar_wg, 1); mbar_init(&bar_r0, NW);
    }
    if (w == NW) tmem_alloc_512(&tmem_slot);
    tmem_fence_before_sync();
    __syncthreads();
    tmem_fence_after_sync();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_slot, 0);

    if (w == NW) {
        // =========================== control warp ===========================
        // The whole warp runs the control loop so that addresses and descriptors stay warp-uniform (uniform
        // registers feed tcgen05.mma directly); one lane issues.
        if (has_tiles) {
            const int leader = lane == 0;
            uint32_t pf = 0, pe = 0, p_wlo = 0, p_opnd = 0, p_lo = 0, p_wg = 0, p_r0 = 0, p_ch = 0;
#ifdef TDQ_TC_PROFILE
            unsigned tcp_acc[16];
            for (int i = 0; i < 16; ++i) tcp_acc[i] = 0;
            unsigned tcp_t0 = clock();
#endif      // pf / pe: one parity bit per ring slot
            constexpr uint32_t idesc_mn = umma_idesc_tf32(128, NB) | (1u << 16);   // B operand MN-major
            constexpr uint32_t idesc_k = umma_idesc_tf32(128, 128);
            constexpr uint32_t idesc16_mn = umma_idesc_bf16(128, NB) | (1u << 16); // bf16 correction terms
            constexpr uint32_t idesc16_k = umma_idesc_bf16(128, 128);
            const uint32_t r0 = smem_u32(R0), r1 = smem_u32(R1), r2 = smem_u32(R2);
            // The weight ring has 4 slots of 16 KB (one [128][32] plane).  Only the hi planes of a GEMM's 4 K-blocks are
            // streamed from L2, while the workers are busy with the previous phase.  The lo parts never pass through
            // shared memory: each worker thread derives the lo values of its own row from the resident hi planes and
            // parks them in Tensor Memory (128 columns that are dead at that moment), and the last third of the GEMM's
            // MMAs reads its A operand from there.  No load is issued during a GEMM, so none can be late, and the
            // weight stream from L2 is 64 KB per GEMM instead of 128.
            auto wait_full = [&](int s) { mbar_wait(&bar_full[s], (pf >> s) & 1); pf ^= 1u << s; };
            auto wait_empty = [&](int s) { mbar_wait(&bar_empty[s], (pe >> s) & 1); pe ^= 1u << s; };
            auto tma_plane = [&](float* ring, int s, int m, int dir) {
                if (!leader) return;
                mbar_expect_tx(&bar_full[s], kTcPlaneB);
                // (no proxy fence here: whoever wrote this region with ordinary stores fenced before signalling)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                                 smem_u32(ring + s * kTcPlaneF)),
                             "l"(prm.wimg + tc_img_off(m, dir, s)), "r"(kTcPlaneB), "r"(smem_u32(&bar_full[s]))
                             : "memory");
            };
            // after a GEMM (its hi planes have all been consumed, and read by the workers): load the next one's
            auto prefetch_next = [&](float* ring, int m, int dir, bool load) {
#pragma unroll
                for (int q = 0; q < 4; ++q) { wait_empty(q); if (load) tma_plane(ring, q, m, dir); }
            };
            auto prefetch_first = [&](float* ring, int m, int dir) {
#pragma unroll
                for (int q = 0; q < 4; ++q) tma_plane(ring, q, m, dir);
            };
            // Descriptors are built once; an operand further along is the same descriptor plus (bytes >> 4) in the
            // start-address field (shared-memory addresses stay below 2^18, so the 14-bit field never carries).
            // Every operand descriptor is {low word = (base >> 4) + compile-time offset + LBO field, high word = a
            // constant of the layout}, built from ONE uniform base: cheap to rematerialise, so the compiler does not
            // park dozens of precomputed 64-bit descriptors in uniform registers and spill them.
            const uint32_t a0 = r0 >> 4;
            constexpr uint32_t kReg = (4 * kTcPlaneB) >> 4;          // R0 -> R1 -> R2 in descriptor address units
            auto dk = [&](uint32_t off16) {                          // K-major SWIZZLE_128B plane at byte offset 16 * off16
                const uint32_t lo = a0 + off16 + (1u << 16), hi = (1024u >> 4) | (1u << 14) | (2u << 29);
                uint64_t d;
                asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(hi));
                return d;
            };
            auto dmn = [&](uint32_t off16) {                         // MN-major SWIZZLE_128B_BASE32B, planes 16 KB apart
                const uint32_t lo = a0 + off16 + ((kTcPlaneB >> 4) << 16), hi = (512u >> 4) | (1u << 14) | (1u << 29);
                uint64_t d;
                asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(hi));
                return d;
            };
            // the SAME fp32 plane read K-major (N = feature rows, K = the 32 points of a line): layout type BASE32B with
            // SBO = 512 (4-row swizzle period), LBO = 1024 (tools/experiments/kmajor_base32b_probe.cu, mode 4) -- the weight
            // gradient reads zbar straight from the plane the data gradient used, no K-major copy of it is ever stored
            auto dkb = [&](uint32_t off16) {
                const uint32_t lo = a0 + off16 + ((1024u >> 4) << 16), hi = (512u >> 4) | (1u << 14) | (1u << 29);
                uint64_t d;
                asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(hi));
                return d;
            };
            auto dmn16 = [&](uint32_t off16) {                       // bf16 MN-major SWIZZLE_128B: 64-wide N atoms two planes apart
                const uint32_t lo = a0 + off16 + (((2 * kTcPlaneB) >> 4) << 16), hi = (1024u >> 4) | (1u << 14) | (2u << 29);
                uint64_t d;
                asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(hi));
                return d;
            };
            // D^T[n][(c,p)] = sum_k A[n][k] B[(c,p)][k]:  A = weight image streamed through `ring` (descriptor of
            // slot 0: a_desc), B = activation image (hi / lo descriptors); result in TMEM columns dcol .. dcol + NB.
            // On return every slot's last MMAs are committed to bar_empty (the caller waits before reloading).
            auto act_gemm = [&](auto fwd_tag, int m, uint32_t dcol, uint32_t wlo) {
                constexpr bool FWD = decltype(fwd_tag)::value;
#ifdef TDQ_TC_PROFILE
                unsigned g0 = (unsigned)clock();
#endif
                float* ring = FWD ? R2 : R0;
                const int dir = FWD ? 0 : 1;
                constexpr uint32_t oA = FWD ? 2 * kReg : 0, oBh = FWD ? 0 : kReg, oBl = FWD ? kReg : 2 * kReg;     // ring, B hi, B lo regions
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    wait_full(q);
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) {          // hi x hi: kind::tf32 on the raw fp32 words
                        const uint32_t ao = oA + q * (kTcPlaneB >> 4) + ks * 2, bo = (q * 32 + ks * 8) * 128 >> 4;
                        if (leader) { if (q == 0 && ks == 0) umma_tf32_i<false>(dcol, dk(ao), dmn(oBh + bo), idesc_mn); else umma_tf32_i<true>(dcol, dk(ao), dmn(oBh + bo), idesc_mn); }
                    }
                    if (leader) umma_commit(&bar_empty[q]);
                }
#ifdef TDQ_TC_PROFILE
                unsigned g1 = (unsigned)clock();
                tcp_acc[12] += (g1 - g0) >> 4;
#endif
                mbar_wait(&bar_wlo, p_wlo); p_wlo ^= 1;      // the bf16 images of the weight operand are in TMEM columns wlo ..
                tmem_fence_after_sync();
#ifdef TDQ_TC_PROFILE
                unsigned g2 = (unsigned)clock();
                tcp_acc[13] += (g2 - g1) >> 4;
#endif
                // correction terms W_hi16 x a_lo16 + W_lo16 x a_hi16 as kind::f16 (bf16) MMAs, K = 16 each; the B region holds
                // the bf16 images [hi16 atom 0 | lo16 atom 0 | hi16 atom 1 | lo16 atom 1] (one plane each)
#pragma unroll
                for (int k16 = 0; k16 < 8; ++k16) {
                    const uint32_t bo = k16 * 2048 >> 4;
                    if (leader) umma_bf16_ts_i<true>(dcol, wlo + k16 * 8, dmn16(oBl + (kTcPlaneB >> 4) + bo), idesc16_mn);          // W_hi16 x a_lo16
                    if (leader) umma_bf16_ts_i<true>(dcol, wlo + 64 + k16 * 8, dmn16(oBl + bo), idesc16_mn);                       // W_lo16 x a_hi16
                }
                if (leader) umma_commit(&bar_done);
#ifdef TDQ_TC_PROFILE
                tcp_acc[14] += ((unsigned)clock() - g2) >> 4;
#endif
            };
            prefetch_first(R2, 0, 0);
            for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                // ---- forward GEMMs
                for (int l = 2; l <= L; ++l) {
                    TCP(0);
                    mbar_wait(&bar_opnd, p_opnd); p_opnd ^= 1;
                    tmem_fence_after_sync();
                    TCP(1);                                   // waited for the forward operand
                    act_gemm(std::true_type{}, l - 2, tmem_base + (l - 2) * NB, tmem_base + kTcDwCol);
                    TCP(2);                                   // issue (incl. waits for weight planes)
                    prefetch_next(R2, l - 1, 0, l < L);
                    TCP(3);                                   // until the GEMM has drained
                }
                // ---- backward
                TCP(4);
                mbar_wait(&bar_r0, p_r0); p_r0 ^= 1;              // workers no longer read a^L in R0
                if (FWD_ONLY) {                               // prediction: next tile's first weight planes, no reverse sweep
                    if (tile + gridDim.x < n_tiles) prefetch_first(R2, 0, 0);
                    continue;
                }
                prefetch_first(R0, L - 2, 1);
                TCP(5);
                for (int l = L; l >= 2; --l) {
                    mbar_wait(&bar_opnd, p_opnd); p_opnd ^= 1;
                    tmem_fence_after_sync();
                    TCP(6);                                   // waited for zbar
                    act_gemm(std::false_type{}, l - 2, tmem_base + (l - 2) * NB, tmem_base + (l == L ? kTcDwCol : (l - 1) * NB));      // dgrad
                    TCP(7);
                    prefetch_next(R0, 0, 1, false);      // (only consumes the slots' completion phases)
                    TCP(8);
                    // wgrad: dW_l[k][j] = sum_{c,p} a^{l-1}[c][k][p] zbar^l[c][j][p].  a_hi: R0 (K-major), a_lo: Tensor
                    // Memory (each thread parked the lo parts of its own row there: the A operand may come from
                    // TMEM), zbar hi / lo: R1 / R2 (K-major).
                    const uint32_t dw = tmem_base + kTcDwCol;
                    const uint32_t alo = tmem_base + (l - 2) * NB;      // where dgrad's result was, read back by now
#pragma unroll
                    for (int c = 0; c < C; ++c) {
                        mbar_wait(&bar_lo_ready[c], p_lo);    // channel c of a^{l-1} (hi plane, lo columns) is in place
                        tmem_fence_after_sync();
                        if (c == 0) TCP(9);                   // waited for the wgrad operands
                        const uint32_t po = c * (kTcPlaneB >> 4);
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {      // a_hi x zbar_hi: kind::tf32 on the K-major fp32 planes
                            if (leader) { if (c == 0 && ks == 0) umma_tf32_i<false>(dw, dk(po + ks * 2), dkb(kReg + po + ks * 2), idesc_k); else umma_tf32_i<true>(dw, dk(po + ks * 2), dkb(kReg + po + ks * 2), idesc_k); }
                        }
#pragma unroll
                        for (int t = 0; t < 2; ++t) {        // 16 points per MMA and term: a_lo16 x zbar_hi16, a_hi16 x zbar_lo16; zbar's images
                            // are dgrad's own (R2: atom c / 2 = planes 2 (c / 2) [hi16], + 1 [lo16]; channel c % 2 = 64 bytes into a row)
                            const uint32_t zo = 2 * kReg + (2 * (c >> 1)) * (kTcPlaneB >> 4) + (c & 1) * 4 + t * 2;
                            if (leader) umma_bf16_ts_i<true>(dw, alo + c * 32 + t * 16, dk(zo), idesc16_k);
                            if (leader) umma_bf16_ts_i<true>(dw, alo + c * 32 + t * 16 + 8, dk(zo + (kTcPlaneB >> 4)), idesc16_k);
                        }
                        if (leader) umma_commit(&bar_ch[c]);
                    }
                    p_lo ^= 1;
                    if (leader) umma_commit(&bar_wg);
                    TCP(10);                                  // wgrad issue
                    // The next GEMM's weight planes go where this wgrad's operand planes lie (R0 before a dgrad,
                    // R2 before the next tile's first forward GEMM): plane c is dead once channel c has drained.
                    const bool more = l > 2 || tile + gridDim.x < n_tiles;
                    float* ring = l > 2 ? R0 : R2;
                    const int nm = l > 2 ? l - 3 : 0, ndir = l > 2 ? 1 : 0;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        // R0 plane q (a^{l-1}[q]) is read by channel q alone; R2 planes 2 j, 2 j + 1 (zbar's bf16 images of the
                        // channel pair j) by both channels of the pair
                        const int wq = l > 2 ? q : ((q | 1) < C ? (q | 1) : C - 1);
                        if (l == 2 || q < C) { mbar_wait(&bar_ch[wq], p_ch); }
                        if (more) tma_plane(ring, q, nm, ndir);
                    }
                    p_ch ^= 1;
                    mbar_wait(&bar_wg, p_wg); p_wg ^= 1;
                    TCP(11);                                  // until wgrad has drained
                }
            }
#ifdef TDQ_TC_PROFILE
            if (blockIdx.x == 0 && n_tiles > 1000 && leader)
                printf("tcp ctl: %u %u %u %u | %u %u | %u %u %u | %u %u %u | act gemm: hi issue %u refill waits %u lo issue %u\n", tcp_acc[0], tcp_acc[1], tcp_acc[2], tcp_acc[3], tcp_acc[4],
                       tcp_acc[5], tcp_acc[6], tcp_acc[7], tcp_acc[8], tcp_acc[9], tcp_acc[10], tcp_acc[11], tcp_acc[12], tcp_acc[13], tcp_acc[14]);
#endif
        }
    } else {
        // =========================== worker warps ===========================
        const int q4 = w & 3, hf = w >> 2;
        const int n = q4 * 32 + lane;                       // this thread's neuron = its TMEM lane
        const int pb = hf * PT;                             // first of its PT points
        const uint32_t trow = tmem_base + ((uint32_t)(q4 * 32) << 16);
        float* sg_b = sgb + hf * L * HP;                    // this point group's bias accumulators (entry n is private)
        const float* b_s = sp + lay.off_b();
        const float* w0_s = sp + lay.off_w0();
        const float wout_n = sp[lay.off_wout() + n];
        uint32_t p_done = 0, p_wg = 0, p_chw = 0;
        float loss_acc = 0.f;                               // tid 0 only
        F2 g_wout2(0.f);                                    // d w_out[n] of this thread's points, all tiles (even / odd points)
        float g_w0[3] = {0.f, 0.f, 0.f};                    // d W0[:, n] likewise
        bool first_tile = true;
#ifdef TDQ_TC_PROFILE
        unsigned tcp_acc[16];
        for (int i = 0; i < 16; ++i) tcp_acc[i] = 0;
        unsigned tcp_t0 = clock();
#endif
        // float offset of this thread's 4-point chunk jj of row n inside a plane, K-major / MN-major swizzle.
        // Rows n and n ^ 4 of a quarter-warp map a chunk pair to the same 32 bytes of their (different) lines, i.e.
        // the same banks: rows with bit 2 set therefore go through each pair in the opposite order.
        const bool swp = (n >> 2) & 1;
        int coffK[NJ], coffM[NJ];
#pragma unroll
        for (int jj = 0; jj < NJ; ++jj) {
            coffK[jj] = n * 32 + (((hf * NJ + jj) ^ (n & 7)) << 2);
            coffM[jj] = tc_mn32_off(n, pb + jj * 4);
        }
        // 16-byte-unit offsets of this thread's bf16 pieces (8 points of one channel = 16 bytes):
        //   MN-major images (B operand of the activation GEMMs' correction terms): [128 feature rows][64 = 2 channels x 32
        //   points] bf16 per atom, SWIZZLE_128B; region = [hi16 atom 0 | lo16 atom 0 | hi16 atom 1 | lo16 atom 1], atom = c / 2
        //   (wgrad's correction terms read the same images K-major)
        int o16M[C];
#pragma unroll
        for (int c = 0; c < C; ++c) {
            o16M[c] = (c >> 1) * 2048 + n * 8 + ((((c & 1) * 4 + hf)) ^ (n & 7));
        }
        // the 8 points x[0..7] of channel c -> fp32 MN-major plane c of `hi` (and the bf16 images in `b16` unless null)
        auto store_mn = [&](float* hi, float* b16, int c, const float* x) {
            const float4 x0 = make_float4(x[0], x[1], x[2], x[3]), x1 = make_float4(x[4], x[5], x[6], x[7]);
            float* hp = hi + c * kTcPlaneF;
            const int oa = swp ? coffM[1] : coffM[0], ob = swp ? coffM[0] : coffM[1];
            *reinterpret_cast<float4*>(hp + oa) = swp ? x1 : x0;
            *reinterpret_cast<float4*>(hp + ob) = swp ? x0 : x1;
            if (b16) {
                uint4 h, l;
                split16(x, h, l);
                uint4* B = reinterpret_cast<uint4*>(b16);
                B[o16M[c]] = h;
                B[1024 + o16M[c]] = l;
            }
        };
        // K-major: fp32 plane c of `hi`; 8 rows of a quarter-warp land in 8 different 16-byte slots, no reordering needed
        auto store_k = [&](float* hi, int c, const float* x) {
            float* hp = hi + c * kTcPlaneF;
            *reinterpret_cast<float4*>(hp + coffK[0]) = make_float4(x[0], x[1], x[2], x[3]);
            *reinterpret_cast<float4*>(hp + coffK[1]) = make_float4(x[4], x[5], x[6], x[7]);
        };

        // pre-activation jet of layer 1 at the thread's points: v[c][p]
        auto layer1_pre = [&](float (&v)[C][PT]) {
#pragma unroll
            for (int c = 0; c < C; ++c)
#pragma unroll
                for (int p = 0; p < PT; ++p) v[c][p] = 0.f;
            const float bz = b_s[n];
#pragma unroll
            for (int p = 0; p < PT; ++p) v[0][p] = bz;
            for (int dim = 0; dim < d; ++dim) {
                const float wv = w0_s[dim * HP + n];
#pragma unroll
                for (int p = 0; p < PT; p += 2) {
                    const F2 z = jm_fma(F2(Xs[dim * P + pb + p], Xs[dim * P + pb + p + 1]), F2(wv), F2(v[0][p], v[0][p + 1]));
                    v[0][p] = z.x; v[0][p + 1] = z.y;
                }
                if (K0 > 0 && dim == 0) {
#pragma unroll
                    for (int p = 0; p < PT; ++p) v[J::OFF0][p] = wv;
                }
                if (K1 > 0 && dim == 1) {
#pragma unroll
                    for (int p = 0; p < PT; ++p) v[K1 > 0 ? J::OFF1 : 0][p] = wv;
                }
                if (K2 > 0 && dim == 2) {
#pragma unroll
                    for (int p = 0; p < PT; ++p) v[K2 > 0 ? J::OFF2 : 0][p] = wv;
                }
            }
        };
        // v: pre-activation jet (tanh not yet applied to channel 0 when apply_tanh) -> on exit v = stash
        // (tanh value, pre-activation derivatives); activations go to `hi` (and their lo image to `lo` if non-null)
        auto act_store = [&](float (&v)[C][PT], bool apply_tanh, float* hi, float* b16) {
            float a[C][8];
#pragma unroll
            for (int p = 0; p < 8; p += 2) {               // two points at a time through the packed fp32 instructions
                F2 st[C], act[C];
#pragma unroll
                for (int c = 0; c < C; ++c) st[c] = F2(v[c][p], v[c][p + 1]);
                if (apply_tanh) { st[0] = tanh_acc(st[0]); v[0][p] = st[0].x; v[0][p + 1] = st[0].y; }
                J::recompute(st, act);
#pragma unroll
                for (int c = 0; c < C; ++c) { a[c][p] = act[c].x; a[c][p + 1] = act[c].y; }
            }
#pragma unroll
            for (int c = 0; c < C; ++c) store_mn(hi, b16, c, a[c]);
        };
        // bf16 images of the running GEMM's weight operand for the correction terms: row n of the 4 resident fp32 planes ->
        // this thread's TMEM row, columns [0, 64): W rounded to bf16, [64, 128): its lo parts (see the ring comment in the
        // control warp); thread (n, hf) covers K columns 8 hf .. + 7 of each plane = 4 packed columns of either image
        uint32_t wf = 0;                         // this thread's phase parities of bar_full[]
        auto make_wlo = [&](const float* ring, uint32_t tcol) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                mbar_wait(&bar_full[q], (wf >> q) & 1); wf ^= 1u << q;           // plane q has landed
                const float4 h0 = *reinterpret_cast<const float4*>(ring + q * kTcPlaneF + coffK[0]);
                const float4 h1 = *reinterpret_cast<const float4*>(ring + q * kTcPlaneF + coffK[1]);
                const float x[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
                uint4 h, l;
                split16(x, h, l);
                const float hw[4] = {__uint_as_float(h.x), __uint_as_float(h.y), __uint_as_float(h.z), __uint_as_float(h.w)};
                const float lw[4] = {__uint_as_float(l.x), __uint_as_float(l.y), __uint_as_float(l.z), __uint_as_float(l.w)};
                tmem_st4(tcol + q * 16 + hf * 4, hw);
                tmem_st4(tcol + 64 + q * 16 + hf * 4, lw);
            }
            tmem_wait_st();
            tmem_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar_wlo);
        };
        auto signal = [&](uint64_t* bar) {       // this warp's shared-memory / TMEM work is published
            fence_proxy_async_smem();
            tmem_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar);
        };
        // this tile's dW of hidden matrix m: TMEM -> slab (row n, columns (128 / NH) hf .. )
        auto drain_dw = [&](int m) {
            constexpr int NC = 128 / NH;
            // slab order tc_slab_off: element (k, j) at ((j / 4) * 128 + k) * 4 + j % 4 -- a warp instruction below
            // covers 32 consecutive rows of one 4-column group = 512 contiguous bytes
            float* gdst = slab + lay.off_wh() + (size_t)m * HP * HP + ((size_t)(hf * NC / 4) * HP + n) * 4;
#pragma unroll
            for (int qq = 0; qq < NC / 16; ++qq) {
                float v[16];
                tmem_ld16(trow + kTcDwCol + hf * NC + qq * 16, v);
                tmem_wait_ld();
#pragma unroll
                for (int j = 0; j < 16; j += 4) {
                    float* g4 = gdst + (size_t)((qq * 16 + j) / 4) * HP * 4;
                    if (first_tile)
                        *reinterpret_cast<float4*>(g4) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                    else
                        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(g4), "f"(v[j]),
                                     "f"(v[j + 1]), "f"(v[j + 2]), "f"(v[j + 3])
                                     : "memory");
                }
            }
        };

        // coordinates of a tile -> Xs, then layer 1 at the thread's points: v <- activations (a^1 jet)
        auto layer1_act = [&](long long tile_, float (&v)[C][PT]) {
            worker_sync<kWorkers>();                         // the previous readers of Xs are done
            const Seg& sq1 = prm.seg[seg_of_tile(prm, tile_)];
            for (int i = tid; i < d * P; i += kWorkers) {
                int dim = i / P, p = i - dim * P;
                long long gi = (tile_ - sq1.tile0) * P + p;
                Xs[dim * P + p] = gi < sq1.n ? sq1.X[(size_t)dim * sq1.x_stride + gi] : 0.f;
            }
            worker_sync<kWorkers>();
            layer1_pre(v);
#pragma unroll
            for (int p = 0; p < PT; p += 2) {
                F2 st[C], act[C];
#pragma unroll
                for (int c = 0; c < C; ++c) st[c] = F2(v[c][p], v[c][p + 1]);
                st[0] = tanh_acc(st[0]);
                J::recompute(st, act);
#pragma unroll
                for (int c = 0; c < C; ++c) { v[c][p] = act[c].x; v[c][p + 1] = act[c].y; }
            }
        };

        float v[C][PT];
        bool l1_ready = false;           // v already holds this tile's layer-1 activations (computed at the end of the previous tile)
        for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            const Seg& sq = prm.seg[seg_of_tile(prm, tile)];      // the term this tile belongs to
            const long long base = (tile - sq.tile0) * P;
            // ================= forward =================
            if (!l1_ready) layer1_act(tile, v);
            TCP(0);
#pragma unroll
            for (int c = 0; c < C; ++c) store_mn(R0, R1, c, v[c]);
            signal(&bar_opnd);
            make_wlo(R2, trow + kTcDwCol);
            TCP(1);
            for (int l = 2; l <= L; ++l) {
                mbar_wait(&bar_done, p_done); p_done ^= 1;
                tmem_fence_after_sync();
                TCP(2);
                const uint32_t tcol = trow + (l - 2) * NB + pb;
#pragma unroll
                for (int c = 0; c < C; ++c) tmem_ldn<PT>(tcol + c * 32, v[c]);
                tmem_wait_ld();
                const float bz = b_s[(l - 1) * HP + n];
#pragma unroll
                for (int p = 0; p < PT; p += 2) {
                    const F2 z = F2(v[0][p], v[0][p + 1]) + F2(bz);
                    v[0][p] = z.x; v[0][p + 1] = z.y;
                }
                act_store(v, true, R0, l < L ? R1 : nullptr);
                if (l < L) signal(&bar_opnd);
                if (!FWD_ONLY) {
#pragma unroll
                    for (int c = 0; c < C; ++c) tmem_stn<PT>(tcol + c * 32, v[c]);  // stash, in place of z^l
                }
                if (l < L) make_wlo(R2, trow + kTcDwCol);
                TCP(3);
            }
            worker_sync<kWorkers>();                         // R0 holds a^L
            TCP(4);

            // ================= output layer + integrand =================
            {
                constexpr int G = kWorkers / 128;            // neuron groups per (channel, point)
                const int g = tid >> 7, cp = tid & 127, c = cp >> 5, p = cp & 31;
                if (c < C) {
                    const float* ap = R0 + c * kTcPlaneF;
                    const float* wo = sp + lay.off_wout();
                    float s = 0.f;
#pragma unroll 8
                    for (int k = g * (128 / G); k < (g + 1) * (128 / G); ++k) s = fmaf(ap[tc_mn32_off(k, p)], wo[k], s);
                    red[(g * 4 + c) * P + p] = s;
                }
            }
            worker_sync<kWorkers>();
            TCP(5);
            if (w == 0) {
                // The residual program (compiled form, vm_poly.h) runs on one warp, one point per lane.  Its value
                // and adjoint tables live in R2 (idle between the last forward GEMM and the first zbar store),
                // laid out [entry][lane].
                constexpr int G = kWorkers / 128;
                const int p = lane;
                const long long gi = base + p;
                const bool valid = gi < sq.n;
                const TcOp* ops_t = ops_s + sq.tc_off;
                const int n_ops = sq.tc_n_ops, n_ent = kTcVmInputs + n_ops;
                float* V = R2 + lane;
                float* A = V + n_ent * 32;
                float u0 = 0.f;
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    float sj = c == 0 ? sp[lay.off_sc()] : 0.f;   // b_out
#pragma unroll
                    for (int g = 0; g < G; ++g) sj += red[(g * 4 + c) * P + p];
                    V[c * 32] = sj;
                    if (c == 0) u0 = sj;
                }
#pragma unroll
                for (int dim = 0; dim < 3; ++dim) V[(kTcVmCoord + dim) * 32] = dim < d ? Xs[dim * P + p] : 0.f;
                const float lam_in = (valid && sq.lam) ? sq.lam[gi] : 0.f;
                V[kTcVmTarget * 32] = (valid && sq.target) ? sq.target[gi] : 0.f;
                V[kTcVmLambda * 32] = lam_in;
                for (int vi = 0; vi < prm.n_vars; ++vi) V[(kTcVmVars + vi) * 32] = prm.vars[vi];
                for (int i = 0; i < n_ent; ++i) A[i * 32] = 0.f;
                tc_vm_forward(ops_t, n_ops, V);
                const float outv = V[sq.tc_out * 32];
                if (valid && sq.f_out && sq.tc_f >= 0) sq.f_out[gi] = V[sq.tc_f * 32];
                if (valid && sq.u_out) sq.u_out[gi] = u0;
                if (!FWD_ONLY) {
                float seed, lsum;
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
                A[sq.tc_out * 32] = seed;
                tc_vm_reverse(ops_t, n_ops, V, A);
#pragma unroll
                for (int c = 0; c < C; ++c) Ub[c * P + p] = A[c * 32];
                const float jb0 = A[0], lam_bar = A[kTcVmLambda * 32];
                if (valid && sq.lam_mode >= 1) {
                    if (sq.lam_grad) sq.lam_grad[gi] = lam_bar;
                    if (sq.lam_mode == 2) {
                        // keras Adam on this point's lambda, gradient sign-flipped for ascent (fit.py:101,104)
                        float g = prm.lam_adam.sign * lam_bar;
                        float m = sq.lam_m[gi], vv = sq.lam_v[gi];
                        m += (g - m) * (1.f - prm.lam_adam.beta1);
                        vv += (g * g - vv) * (1.f - prm.lam_adam.beta2);
                        sq.lam_m[gi] = m;
                        sq.lam_v[gi] = vv;
                        sq.lam[gi] = lam_in - (m * prm.lam_adam.alpha) / (sqrtf(vv) + prm.lam_adam.eps);
                    }
                }
                float q0 = lsum, q1 = jb0;
#pragma unroll
                for (int o = 16; o >= 1; o >>= 1) {
                    q0 += __shfl_xor_sync(0xffffffffu, q0, o);
                    q1 += __shfl_xor_sync(0xffffffffu, q1, o);
                }
                if (lane == 0) { loss_acc += q0; sgs[0] += q1; }
                for (int vi = 0; vi < prm.n_vars; ++vi) {
                    float qv = A[(kTcVmVars + vi) * 32];
#pragma unroll
                    for (int o = 16; o >= 1; o >>= 1) qv += __shfl_xor_sync(0xffffffffu, qv, o);
                    if (lane == 0) sgs[2 + vi] += qv;
                }
                }
            }
            worker_sync<kWorkers>();                         // Ub is complete
            TCP(6);
            if (FWD_ONLY) {                              // prediction ends here: R0 and the tables in R2 are free again
                __syncwarp();
                if (lane == 0) mbar_arrive(&bar_r0);
                continue;
            }

            // ================= backward =================
            // cotangent of a^L: v[c][p] = ubar[c][p] w_out[n];  d w_out[n] += sum_{c,p} a^L[c][n][p] ubar[c][p]
#pragma unroll
            for (int c = 0; c < C; ++c)
#pragma unroll
                for (int jj = 0; jj < NJ; ++jj) {
                    const float4 ub = *reinterpret_cast<const float4*>(Ub + c * P + pb + jj * 4);
                    const float4 av = *reinterpret_cast<const float4*>(R0 + c * kTcPlaneF + coffM[jj]);
                    g_wout2 = jm_fma(F2(av.x, av.y), F2(ub.x, ub.y), g_wout2);
                    g_wout2 = jm_fma(F2(av.z, av.w), F2(ub.z, ub.w), g_wout2);
                    const F2 v01 = F2(ub.x, ub.y) * F2(wout_n), v23 = F2(ub.z, ub.w) * F2(wout_n);
                    v[c][jj * 4 + 0] = v01.x; v[c][jj * 4 + 1] = v01.y;
                    v[c][jj * 4 + 2] = v23.x; v[c][jj * 4 + 3] = v23.y;
                }
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar_r0);             // R0 may become the weight ring
            tmem_wait_st();                                  // this thread's stash stores have landed
            TCP(7);

            for (int l = L; l >= 1; --l) {
                // ---- reverse through the tanh jets of layer l: v (cotangent of a^l) -> cotangent of z^l
                {
                    float s[C][PT];
                    if (l >= 2) {
                        const uint32_t tcol = trow + (l - 2) * NB + pb;
#pragma unroll
                        for (int c = 0; c < C; ++c) tmem_ldn<PT>(tcol + c * 32, s[c]);
                        tmem_wait_ld();
                    } else {
                        layer1_pre(s);
#pragma unroll
                        for (int p = 0; p < PT; p += 2) {
                            const F2 t = tanh_acc(F2(s[0][p], s[0][p + 1]));
                            s[0][p] = t.x; s[0][p + 1] = t.y;
                        }
                    }
                    F2 gb(0.f);
#pragma unroll
                    for (int p = 0; p < PT; p += 2) {
                        F2 st[C], ab[C], zb[C];
#pragma unroll
                        for (int c = 0; c < C; ++c) { st[c] = F2(s[c][p], s[c][p + 1]); ab[c] = F2(v[c][p], v[c][p + 1]); }
                        J::bwd(st, ab, zb);
#pragma unroll
                        for (int c = 0; c < C; ++c) { v[c][p] = zb[c].x; v[c][p + 1] = zb[c].y; }
                        gb += zb[0];
                    }
                    sg_b[(l - 1) * HP + n] += gb.x + gb.y;
                }
                TCP(8);
                if (l == 1) {
#pragma unroll
                    for (int dim = 0; dim < 3; ++dim) {
                        if (dim < d) {
                            float s0 = 0.f;
#pragma unroll
                            for (int p = 0; p < PT; ++p) {
                                s0 = fmaf(Xs[dim * P + pb + p], v[0][p], s0);
                                if (dim == 0 && K0 > 0) s0 += v[J::OFF0][p];
                                if (dim == 1 && K1 > 0) s0 += v[K1 > 0 ? J::OFF1 : 0][p];
                                if (dim == 2 && K2 > 0) s0 += v[K2 > 0 ? J::OFF2 : 0][p];
                            }
                            g_w0[dim] += s0;
                        }
                    }
                    // the last wgrad of the tile is still running: do the next tile's layer 1 in its shadow
                    l1_ready = tile + gridDim.x < n_tiles;
                    if (l1_ready) layer1_act(tile + gridDim.x, v);
                    TCP(8);
                    if (L >= 2) {
                        mbar_wait(&bar_wg, p_wg); p_wg ^= 1;     // dW_2 of this tile is complete; R0..R2 are free again
                        p_chw ^= 1;                               // (its per-channel completions are not waited for)
                        tmem_fence_after_sync();
                        drain_dw(0);
                        // Every warp has read its share of the dW columns before any warp's next make_wlo (top of the tile
                        // loop) writes W_2's bf16 images across ALL of them: nothing else orders those accesses (measured
                        // alternatives: images of W_2 in the dead stash columns, or a drain that reads exactly the columns the
                        // thread later writes -- both slower than this barrier, which costs 0.45 %).
                        tmem_fence_before_sync();
                        worker_sync<kWorkers>();
                        tmem_fence_after_sync();
                    }
                    TCP(9);
                    break;
                }
                TCP(9);
                // ---- zbar^l -> R1 (and its lo image -> R2), MN-major: operand of dgrad.  Plane c of R1 / R2 still holds
                // the K-major zbar^{l+1} the running wgrad reads: it is free once that wgrad's channel c has drained.
                // (the bf16 images of channels 2 j, 2 j + 1 share the atom planes 2 j, 2 j + 1 of R2: both wgrad channels must
                //  have drained before either is stored)
#pragma unroll
                for (int c = 0; c < C; c += 2) {
                    if (l < L) { mbar_wait(&bar_ch[c], p_chw); if (c + 1 < C) mbar_wait(&bar_ch[c + 1], p_chw); }
                    store_mn(R1, R2, c, v[c]);
                    if (c + 1 < C) store_mn(R1, R2, c + 1, v[c + 1 < C ? c + 1 : c]);
                }
                if (l < L) p_chw ^= 1;
                signal(&bar_opnd);
                // lo parts of W_l: into the dW columns for the first dgrad of the tile (no wgrad has run yet), else into
                // the columns above the live stash blocks (stash / a_lo of layer l + 1: dead, that wgrad has drained)
                make_wlo(R0, trow + (l == L ? kTcDwCol : (l - 1) * NB));
                TCP(10);
                if (l < L) {                     // in the shadow of dgrad (the next wgrad comes after it)
                    mbar_wait(&bar_wg, p_wg); p_wg ^= 1;     // (all channels have drained by now)
                    tmem_fence_after_sync();
                    drain_dw(l - 1);
                }
                TCP(11);
                // ---- a^{l-1}, recomputed while the tensor core runs dgrad
                float a[C][PT];
                {
                    float s[C][PT];
                    if (l - 1 >= 2) {
                        const uint32_t tcol = trow + (l - 3) * NB + pb;
#pragma unroll
                        for (int c = 0; c < C; ++c) tmem_ldn<PT>(tcol + c * 32, s[c]);
                        tmem_wait_ld();
                    } else {
                        layer1_pre(s);
#pragma unroll
                        for (int p = 0; p < PT; p += 2) {
                            const F2 t = tanh_acc(F2(s[0][p], s[0][p + 1]));
                            s[0][p] = t.x; s[0][p + 1] = t.y;
                        }
                    }
#pragma unroll
                    for (int p = 0; p < PT; p += 2) {
                        F2 st[C], act[C];
#pragma unroll
                        for (int c = 0; c < C; ++c) st[c] = F2(s[c][p], s[c][p + 1]);
                        J::recompute(st, act);
#pragma unroll
                        for (int c = 0; c < C; ++c) { a[c][p] = act[c].x; a[c][p + 1] = act[c].y; }
                    }
                }
                TCP(12);
                mbar_wait(&bar_done, p_done); p_done ^= 1;     // dgrad done: result in TMEM; R0, R1, R2 free
                tmem_fence_after_sync();
                TCP(13);
                // Channel by channel, so that the tensor core starts on channel c while c + 1 is being stored:
                //   zbar^l[c] again, K-major this time (wgrad contracts over the points) -> R1 / R2;
                //   cotangent of a^{l-1}[c] (dgrad result) -> v[c];
                //   a^{l-1}[c]: hi K-major -> R0 (B operand of wgrad), lo -> the TMEM columns just read (A operand)
                const uint32_t tcol_l = trow + (l - 2) * NB + pb;
                // (zbar^l needs nothing here: the tensor core reads its MN-major fp32 planes in R1 K-major, and its bf16 images
                //  in R2 -- SWIZZLE_128B rows of 2 channels x 32 points -- ARE K-major tiles with K = the points of a channel)
#pragma unroll
                for (int c = 0; c < C; ++c) tmem_ldn<PT>(tcol_l + c * 32, v[c]);       // cotangent of a^{l-1} (dgrad result)
                tmem_wait_ld();
                // The bf16 images of a^{l-1}[c] go where that result was, 16 columns per PAIR of point groups:
                // [a_lo16 of 16 points | a_hi16 of 16 points] -- one K = 16 correction MMA per term and 16 points, against the
                // hi16 / lo16 image of zbar.  A thread therefore writes 4 columns its partner (the other point group of the
                // pair, another warp) has just read: the two warps meet at a barrier of their own first.
                tmem_fence_before_sync();
                asm volatile("bar.sync %0, 64;" ::"r"(2 + q4 * 2 + (hf >> 1)) : "memory");
                tmem_fence_after_sync();
                const uint32_t tpair = trow + (l - 2) * NB + (hf >> 1) * 16 + (hf & 1) * 4;
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    store_k(R0, c, a[c]);                                     // a^{l-1}[c], K-major fp32 (A operand of hi x hi)
                    uint4 ah, al;
                    split16(a[c], ah, al);
                    const float alw[4] = {__uint_as_float(al.x), __uint_as_float(al.y), __uint_as_float(al.z), __uint_as_float(al.w)};
                    const float ahw[4] = {__uint_as_float(ah.x), __uint_as_float(ah.y), __uint_as_float(ah.z), __uint_as_float(ah.w)};
                    tmem_st4(tpair + c * 32, alw);
                    tmem_st4(tpair + c * 32 + 8, ahw);
                    tmem_wait_st();
                    signal(&bar_lo_ready[c]);
                }
                TCP(14);
                TCP(15);
            }
            first_tile = false;
        }
#ifdef TDQ_TC_PROFILE
        if (blockIdx.x == 0 && lane == 0 && (w == 0 || w == 5) && n_tiles > 1000) {
            printf("tcp w%d a: %u %u %u %u %u %u %u %u\n", w, tcp_acc[0], tcp_acc[1], tcp_acc[2], tcp_acc[3], tcp_acc[4], tcp_acc[5], tcp_acc[6], tcp_acc[7]);
            printf("tcp w%d b: %u %u %u %u %u %u %u %u\n", w, tcp_acc[8], tcp_acc[9], tcp_acc[10], tcp_acc[11], tcp_acc[12], tcp_acc[13], tcp_acc[14], tcp_acc[15]);
        }
#endif
        // ---- this thread's register accumulators -> shared memory (R0 is free: [NH][4][HP])
        worker_sync<kWorkers>();
        R0[(hf * 4 + 0) * HP + n] = g_wout2.x + g_wout2.y;
#pragma unroll
        for (int dim = 0; dim < 3; ++dim) R0[(hf * 4 + 1 + dim) * HP + n] = g_w0[dim];
        if (tid == 0) sgs[1] = loss_acc;
    }

    // ---- epilogue: publish this CTA's small gradients (hidden matrices were accumulated in place)
    tmem_fence_before_sync();
    __syncthreads();
    if (w == NW) tmem_dealloc_512(tmem_base);
    if (FWD_ONLY) return;
    for (int i = tid; i < nsmall; i += kThreadsAll) {
        float g = 0.f;
        if (i >= lay.off_sc()) {
            g = sgs[i - lay.off_sc()];
        } else if (has_tiles) {
            const int row = i / HP, nn = i - row * HP;           // PackLayout: rows 0..2 W0, 3..2+L biases, 3+L w_out
#pragma unroll
            for (int h = 0; h < NH; ++h) {
                if (row < 3) g += R0[(h * 4 + 1 + row) * HP + nn];
                else if (row < 3 + L) g += sgb[(h * L + (row - 3)) * HP + nn];
                else g += R0[(h * 4 + 0) * HP + nn];
            }
        }
        slab[i] = g;
    }
    if (!has_tiles) {
        for (int i = tid; i < (L - 1) * HP * HP; i += kThreadsAll) slab[lay.off_wh() + i] = 0.f;
    }
}

}  // namespace tdq
