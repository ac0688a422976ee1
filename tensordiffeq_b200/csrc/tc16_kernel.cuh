// The tensor-core jet kernel for 5 .. 8 jet channels (Poisson-type C = 5, Cahn-Hilliard C = 6, three-dimensional
// second-order C = 7): the sibling of tc_kernel.cuh with 16-point tiles.  Same pipeline -- every hidden-layer GEMM
// (forward, data gradient, weight gradient) as 3xTF32 tcgen05.mma with fp32 accumulators in Tensor Memory, 16 worker
// warps + 1 control warp, weight planes streamed by bulk TMA, lo parts of the weight operand in TMEM -- so read
// tc_kernel.cuh's header first; only the tile geometry differs:
//
//   * a 128-byte operand line holds TWO channels x 16 points, interleaved in groups of 4 points (element
//     8 (p / 4) + 4 (c % 2) + p % 4 is point p of channel c = 2 pl + c % 2), so the C channels fill NP = ceil(C / 2)
//     planes of [128 feature rows][32] and a Tensor Memory block is NB = 32 NP <= 128 columns; an odd C leaves one
//     all-zero dummy channel (it multiplies into nothing);
//   * the same line is at once an MN-major atom of the activation GEMMs (N = (plane, channel parity, point)) and a
//     32-element K slice of the weight-gradient GEMM (which contracts over channels AND points anyway), so the
//     descriptors, swizzles and instruction shapes of tc_kernel.cuh carry over with N = NB;
//   * worker thread (w, lane) owns neuron n = 32 (w % 4) + lane and the 4 points 4 (w / 4) .. + 3 of the tile, all
//     channels: per plane it moves two 16-byte pieces (one per channel of the pair) that are the two halves of one
//     32-byte swizzle chunk -- exactly the "pair" the 32-point kernel moves per channel, so the bank-conflict
//     reasoning (store_pair) is unchanged -- and its 8 values of a plane are 8 consecutive Tensor Memory columns;
//   * zbar is stored once per layer (hi and lo planes, MN-major, for dgrad); wgrad reads the same planes K-major through
//     the BASE32B descriptor (dkb), as in tc_kernel.cuh.
// The math (tanh jets, their reverse, residual program, small-parameter gradients) is fused_kernel.cuh's, i.e.
// models.py:80-199 of the reference.
#pragma once
#include "tc_kernel.cuh"

namespace tdq {

// FWD_ONLY: forward sweep + residual program only (tdq_predict): no stash, no reverse sweep, no slab (a compile-time
// switch: as a run-time flag its branches cost the training instantiation 3 %)
template <int K0, int K1, int K2, int NW, bool FWD_ONLY = false>
__global__ void __launch_bounds__((NW + 1) * 32, 1) fused_jet_tc16_kernel(const __grid_constant__ KParams prm) {
    using J = JetCfg<K0, K1, K2>;
    constexpr int C = J::C;
    static_assert(C >= 5 && C <= 8, "16-point tensor-core jet kernel: 5 .. 8 channels");
    static_assert(NW == 16, "16 worker warps");
    constexpr int HP = 128, P = 16, NP = (C + 1) / 2, CP = 2 * NP, NB = NP * 32;
    constexpr int NH = NW / 4, PT = P / NH, kWorkers = NW * 32, kThreadsAll = (NW + 1) * 32;
    static_assert(PT == 4, "4 points per thread");
    constexpr int WPT = 32 / NH, WNJ = WPT / 4;         // K columns of a weight plane per thread (lo parts), float4 pieces

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
    float* Xs = sgs + kScalars;                 // [3][32] (16 used)
    float* Ub = Xs + 3 * 32;                    // [CP][P] = 128
    float* red = Ub + 4 * 32;                   // [4][CP][P] = 512
    float* wsc = red + 4 * 4 * 32;              // 64
    TcOp* ops_s = reinterpret_cast<TcOp*>(wsc + 64);
    __shared__ uint64_t bar_full[4], bar_empty[4], bar_wlo, bar_opnd, bar_done, bar_lo_ready[4], bar_ch[4], bar_wg, bar_r0;
    __shared__ uint32_t tmem_slot;

    const int tid = threadIdx.x, lane = tid & 31;
    const int w = __shfl_sync(0xffffffffu, tid >> 5, 0);      // warp-uniform for the compiler
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
        // =========================== control warp (tc_kernel.cuh's, with planes in place of channels) ===========
        if (has_tiles) {
            const int leader = lane == 0;
            uint32_t pf = 0, pe = 0, p_wlo = 0, p_opnd = 0, p_lo = 0, p_wg = 0, p_r0 = 0, p_ch = 0;
            constexpr uint32_t idesc_mn = umma_idesc_tf32(128, NB) | (1u << 16);   // B operand MN-major
            constexpr uint32_t idesc_k = umma_idesc_tf32(128, 128);
            const uint32_t r0 = smem_u32(R0);
            auto wait_full = [&](int s) { mbar_wait(&bar_full[s], (pf >> s) & 1); pf ^= 1u << s; };
            auto wait_empty = [&](int s) { mbar_wait(&bar_empty[s], (pe >> s) & 1); pe ^= 1u << s; };
            auto tma_plane = [&](float* ring, int s, int m, int dir) {
                if (!leader) return;
                mbar_expect_tx(&bar_full[s], kTcPlaneB);
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                                 smem_u32(ring + s * kTcPlaneF)),
                             "l"(prm.wimg + tc_img_off(m, dir, s)), "r"(kTcPlaneB), "r"(smem_u32(&bar_full[s]))
                             : "memory");
            };
            auto prefetch_next = [&](float* ring, int m, int dir, bool load) {
#pragma unroll
                for (int q = 0; q < 4; ++q) { wait_empty(q); if (load) tma_plane(ring, q, m, dir); }
            };
            auto prefetch_first = [&](float* ring, int m, int dir) {
#pragma unroll
                for (int q = 0; q < 4; ++q) tma_plane(ring, q, m, dir);
            };
            const uint32_t a0 = r0 >> 4;
            constexpr uint32_t kReg = (4 * kTcPlaneB) >> 4;          // R0 -> R1 -> R2 in descriptor address units
            auto dk = [&](uint32_t off16) {                          // K-major SWIZZLE_128B plane at byte offset 16 * off16
                const uint32_t lo = a0 + off16 + (1u << 16), hi = (1024u >> 4) | (1u << 14) | (2u << 29);
                uint64_t dd;
                asm("mov.b64 %0, {%1, %2};" : "=l"(dd) : "r"(lo), "r"(hi));
                return dd;
            };
            auto dmn = [&](uint32_t off16) {                         // MN-major SWIZZLE_128B_BASE32B, planes 16 KB apart
                const uint32_t lo = a0 + off16 + ((kTcPlaneB >> 4) << 16), hi = (512u >> 4) | (1u << 14) | (1u << 29);
                uint64_t dd;
                asm("mov.b64 %0, {%1, %2};" : "=l"(dd) : "r"(lo), "r"(hi));
                return dd;
            };
            // an MN-major plane read K-major (see tc_kernel.cuh: BASE32B, SBO = 512, LBO = 1024): wgrad takes zbar's hi and lo
            // planes as dgrad left them
            auto dkb = [&](uint32_t off16) {
                const uint32_t lo = a0 + off16 + ((1024u >> 4) << 16), hi = (512u >> 4) | (1u << 14) | (1u << 29);
                uint64_t dd;
                asm("mov.b64 %0, {%1, %2};" : "=l"(dd) : "r"(lo), "r"(hi));
                return dd;
            };
            auto act_gemm = [&](auto fwd_tag, int m, uint32_t dcol, uint32_t wlo) {
                constexpr bool FWD = decltype(fwd_tag)::value;
                constexpr uint32_t oA = FWD ? 2 * kReg : 0, oBh = FWD ? 0 : kReg, oBl = FWD ? kReg : 2 * kReg;     // ring, B hi, B lo regions
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    wait_full(q);
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) {
                        const uint32_t ao = oA + q * (kTcPlaneB >> 4) + ks * 2, bo = (q * 32 + ks * 8) * 128 >> 4;
                        if (leader) { if (q == 0 && ks == 0) umma_tf32_keep_i<false>(dcol, dk(ao), dmn(oBl + bo), idesc_mn); else umma_tf32_keep_i<true>(dcol, dk(ao), dmn(oBl + bo), idesc_mn); }
                        if (leader) umma_tf32_reuse_i(dcol, dk(ao), dmn(oBh + bo), idesc_mn);      // same weight operand: from the collector
                    }
                    if (leader) umma_commit(&bar_empty[q]);
                }
                mbar_wait(&bar_wlo, p_wlo); p_wlo ^= 1;      // the lo parts of the weight operand are in TMEM columns wlo ..
                tmem_fence_after_sync();
#pragma unroll
                for (int q = 0; q < 4; ++q) {
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) {
                        const uint32_t bo = (q * 32 + ks * 8) * 128 >> 4;
                        if (leader) umma_tf32_ts_i<true>(dcol, wlo + q * 32 + ks * 8, dmn(oBh + bo), idesc_mn);
                    }
                }
                if (leader) umma_commit(&bar_done);
            };
            prefetch_first(R2, 0, 0);
            for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                for (int l = 2; l <= L; ++l) {
                    mbar_wait(&bar_opnd, p_opnd); p_opnd ^= 1;
                    tmem_fence_after_sync();
                    act_gemm(std::true_type{}, l - 2, tmem_base + (l - 2) * NB, tmem_base + kTcDwCol);
                    prefetch_next(R2, l - 1, 0, l < L);
                }
                mbar_wait(&bar_r0, p_r0); p_r0 ^= 1;              // workers no longer read a^L in R0
                if (FWD_ONLY) {                               // prediction: next tile's first weight planes, no reverse sweep
                    if (tile + gridDim.x < n_tiles) prefetch_first(R2, 0, 0);
                    continue;
                }
                prefetch_first(R0, L - 2, 1);
                for (int l = L; l >= 2; --l) {
                    mbar_wait(&bar_opnd, p_opnd); p_opnd ^= 1;
                    tmem_fence_after_sync();
                    act_gemm(std::false_type{}, l - 2, tmem_base + (l - 2) * NB, tmem_base + (l == L ? kTcDwCol : (l - 1) * NB));      // dgrad
                    prefetch_next(R0, 0, 1, false);
                    // wgrad: dW_l[k][j] = sum over (channel, point) of a^{l-1} zbar^l, one plane = 32 of them at a time
                    const uint32_t dw = tmem_base + kTcDwCol;
                    const uint32_t alo = tmem_base + (l - 2) * NB;
#pragma unroll
                    for (int pl = 0; pl < NP; ++pl) {
                        mbar_wait(&bar_lo_ready[pl], p_lo);
                        tmem_fence_after_sync();
                        const uint32_t po = pl * (kTcPlaneB >> 4);
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            if (leader) { if (pl == 0 && ks == 0) umma_tf32_ts_i<false>(dw, alo + pl * 32 + ks * 8, dkb(kReg + po + ks * 2), idesc_k); else umma_tf32_ts_i<true>(dw, alo + pl * 32 + ks * 8, dkb(kReg + po + ks * 2), idesc_k); }
                            if (leader) umma_tf32_keep_i<true>(dw, dk(po + ks * 2), dkb(2 * kReg + po + ks * 2), idesc_k);
                            if (leader) umma_tf32_reuse_i(dw, dk(po + ks * 2), dkb(kReg + po + ks * 2), idesc_k);   // same a_hi operand: from the collector
                        }
                        if (leader) umma_commit(&bar_ch[pl]);
                    }
                    p_lo ^= 1;
                    if (leader) umma_commit(&bar_wg);
                    const bool more = l > 2 || tile + gridDim.x < n_tiles;
                    float* ring = l > 2 ? R0 : R2;
                    const int nm = l > 2 ? l - 3 : 0, ndir = l > 2 ? 1 : 0;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        if (q < NP) { mbar_wait(&bar_ch[q], p_ch); }      // (a plane beyond NP holds no operand: free already)
                        if (more) tma_plane(ring, q, nm, ndir);
                    }
                    p_ch ^= 1;
                    mbar_wait(&bar_wg, p_wg); p_wg ^= 1;
                }
            }
        }
    } else {
        // =========================== worker warps ===========================
        const int q4 = w & 3, hf = w >> 2;
        const int n = q4 * 32 + lane;                       // this thread's neuron = its TMEM lane
        const int pb = hf * PT;                             // first of its 4 points
        const uint32_t trow = tmem_base + ((uint32_t)(q4 * 32) << 16);
        float* sg_b = sgb + hf * L * HP;
        const float* b_s = sp + lay.off_b();
        const float* w0_s = sp + lay.off_w0();
        const float wout_n = sp[lay.off_wout() + n];
        uint32_t p_done = 0, p_wg = 0, p_chw = 0;
        float loss_acc = 0.f;                               // tid 0 only
        float g_wout = 0.f, g_w0[3] = {0.f, 0.f, 0.f};
        bool first_tile = true;
        // float offsets of this thread's 16-byte piece of row n for the even (s = 0) / odd (s = 1) channel of a plane:
        // line position 8 hf + 4 s
        const bool swp = (n >> 2) & 1;
        int coffK[2], coffM[2], wcoff[WNJ];
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            coffK[s] = n * 32 + (((hf * 2 + s) ^ (n & 7)) << 2);
            coffM[s] = tc_mn32_off(n, hf * 8 + s * 4);
        }
#pragma unroll
        for (int jj = 0; jj < WNJ; ++jj) wcoff[jj] = n * 32 + (((hf * WNJ + jj) ^ (n & 7)) << 2);
        auto store_pair = [&](float* hi, float* lo, int off0, int off1, float4 x0, float4 x1) {
            const int oa = swp ? off1 : off0, ob = swp ? off0 : off1;
            const float4 xa = swp ? x1 : x0, xb = swp ? x0 : x1;
            *reinterpret_cast<float4*>(hi + oa) = xa;
            *reinterpret_cast<float4*>(hi + ob) = xb;
            if (lo) {
                *reinterpret_cast<float4*>(lo + oa) = tf32_lo4(xa);
                *reinterpret_cast<float4*>(lo + ob) = tf32_lo4(xb);
            }
        };
        auto store_pair_k = [&](float* hi, float* lo, int off0, int off1, float4 x0, float4 x1) {
            *reinterpret_cast<float4*>(hi + off0) = x0;
            *reinterpret_cast<float4*>(hi + off1) = x1;
            if (lo) {
                *reinterpret_cast<float4*>(lo + off0) = tf32_lo4(x0);
                *reinterpret_cast<float4*>(lo + off1) = tf32_lo4(x1);
            }
        };
        auto f4 = [](const float* x) { return make_float4(x[0], x[1], x[2], x[3]); };
        // the values of plane pl (channels 2 pl, 2 pl + 1; v has CP rows, the dummy one zero) -> MN-major planes
        auto store_plane_mn = [&](float (&v)[CP][PT], int pl, float* hi, float* lo) {
            store_pair(hi + pl * kTcPlaneF, lo ? lo + pl * kTcPlaneF : nullptr, coffM[0], coffM[1], f4(v[2 * pl]), f4(v[2 * pl + 1]));
        };
        auto store_plane_k = [&](float (&v)[CP][PT], int pl, float* hi, float* lo) {
            store_pair_k(hi + pl * kTcPlaneF, lo ? lo + pl * kTcPlaneF : nullptr, coffK[0], coffK[1], f4(v[2 * pl]), f4(v[2 * pl + 1]));
        };
        // this thread's columns of a Tensor Memory block: plane pl at pl * 32 + 8 hf, 4 points of channel 2 pl then 4 of 2 pl + 1
        // (v[2 pl] and v[2 pl + 1] are adjacent in memory: one 8-column access per plane)
        auto tm_ld = [&](uint32_t tblock, float (&v)[CP][PT]) {
#pragma unroll
            for (int pl = 0; pl < NP; ++pl) tmem_ld8(tblock + pl * 32 + hf * 8, &v[2 * pl][0]);
        };
        auto tm_st = [&](uint32_t tblock, float (&v)[CP][PT]) {
#pragma unroll
            for (int pl = 0; pl < NP; ++pl) tmem_st8(tblock + pl * 32 + hf * 8, &v[2 * pl][0]);
        };

        auto layer1_pre = [&](float (&v)[CP][PT]) {
#pragma unroll
            for (int c = 0; c < CP; ++c)
#pragma unroll
                for (int p = 0; p < PT; ++p) v[c][p] = 0.f;
            const float bz = b_s[n];
#pragma unroll
            for (int p = 0; p < PT; ++p) v[0][p] = bz;
            for (int dim = 0; dim < d; ++dim) {
                const float wv = w0_s[dim * HP + n];
#pragma unroll
                for (int p = 0; p < PT; ++p) v[0][p] = fmaf(Xs[dim * 32 + pb + p], wv, v[0][p]);
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
        // stash (tanh value, pre-activation derivatives) in v -> activations in a (dummy channel 0)
        auto recompute_all = [&](const float (&v)[CP][PT], float (&a)[CP][PT]) {
#pragma unroll
            for (int p = 0; p < PT; p += 2) {              // two points at a time through the packed fp32 instructions (f32x2.cuh)
                F2 st[C], act[C];
#pragma unroll
                for (int c = 0; c < C; ++c) st[c] = F2(v[c][p], v[c][p + 1]);
                J::recompute(st, act);
#pragma unroll
                for (int c = 0; c < C; ++c) { a[c][p] = act[c].x; a[c][p + 1] = act[c].y; }
                if (CP > C) { a[CP - 1][p] = 0.f; a[CP - 1][p + 1] = 0.f; }
            }
        };
        uint32_t wf = 0;                         // this thread's phase parities of bar_full[]
        auto make_wlo = [&](const float* ring, uint32_t tcol) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                mbar_wait(&bar_full[q], (wf >> q) & 1); wf ^= 1u << q;           // hi plane q has landed
                float lo[WPT];
#pragma unroll
                for (int jj = 0; jj < WNJ; ++jj) {
                    const float4 h4 = *reinterpret_cast<const float4*>(ring + q * kTcPlaneF + wcoff[jj]);
                    lo[jj * 4 + 0] = tf32_lo(h4.x); lo[jj * 4 + 1] = tf32_lo(h4.y);
                    lo[jj * 4 + 2] = tf32_lo(h4.z); lo[jj * 4 + 3] = tf32_lo(h4.w);
                }
                tmem_stn<WPT>(tcol + q * 32 + hf * WPT, lo);
            }
            tmem_wait_st();
            tmem_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar_wlo);
        };
        auto signal = [&](uint64_t* bar) {
            fence_proxy_async_smem();
            tmem_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar);
        };
        auto drain_dw = [&](int m) {
            constexpr int NC = 128 / NH;
            float* gdst = slab + lay.off_wh() + (size_t)m * HP * HP + ((size_t)(hf * NC / 4) * HP + n) * 4;
#pragma unroll
            for (int qq = 0; qq < NC / 16; ++qq) {
                float vv[16];
                tmem_ld16(trow + kTcDwCol + hf * NC + qq * 16, vv);
                tmem_wait_ld();
#pragma unroll
                for (int j = 0; j < 16; j += 4) {
                    float* g4 = gdst + (size_t)((qq * 16 + j) / 4) * HP * 4;
                    if (first_tile)
                        *reinterpret_cast<float4*>(g4) = make_float4(vv[j], vv[j + 1], vv[j + 2], vv[j + 3]);
                    else
                        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(g4), "f"(vv[j]),
                                     "f"(vv[j + 1]), "f"(vv[j + 2]), "f"(vv[j + 3])
                                     : "memory");
                }
            }
        };
        // coordinates of a tile -> Xs, then layer 1 at the thread's points: v <- activations (a^1 jet)
        auto layer1_act = [&](long long tile_, float (&v)[CP][PT]) {
            worker_sync<kWorkers>();                         // the previous readers of Xs are done
            const Seg& sq1 = prm.seg[seg_of_tile(prm, tile_)];
            for (int i = tid; i < d * P; i += kWorkers) {
                int dim = i / P, p = i - dim * P;
                long long gi = (tile_ - sq1.tile0) * P + p;
                Xs[dim * 32 + p] = gi < sq1.n ? sq1.X[(size_t)dim * sq1.x_stride + gi] : 0.f;
            }
            worker_sync<kWorkers>();
            float s[CP][PT];
            layer1_pre(s);
#pragma unroll
            for (int p = 0; p < PT; ++p) s[0][p] = tanh_acc(s[0][p]);
            recompute_all(s, v);
        };

        float v[CP][PT];
        bool l1_ready = false;
        for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            const Seg& sq = prm.seg[seg_of_tile(prm, tile)];
            const long long base = (tile - sq.tile0) * P;
            // ================= forward =================
            if (!l1_ready) layer1_act(tile, v);
#pragma unroll
            for (int pl = 0; pl < NP; ++pl) store_plane_mn(v, pl, R0, R1);
            signal(&bar_opnd);
            make_wlo(R2, trow + kTcDwCol);
            for (int l = 2; l <= L; ++l) {
                mbar_wait(&bar_done, p_done); p_done ^= 1;
                tmem_fence_after_sync();
                const uint32_t tblock = trow + (l - 2) * NB;
                tm_ld(tblock, v);
                tmem_wait_ld();
                const float bz = b_s[(l - 1) * HP + n];
                float a[CP][PT];
#pragma unroll
                for (int p = 0; p < PT; ++p) v[0][p] = tanh_acc(v[0][p] + bz);
                if (CP > C) {
#pragma unroll
                    for (int p = 0; p < PT; ++p) v[CP - 1][p] = 0.f;     // the dummy channel's GEMM result is 0 anyway
                }
                recompute_all(v, a);
#pragma unroll
                for (int pl = 0; pl < NP; ++pl) store_plane_mn(a, pl, R0, l < L ? R1 : nullptr);
                if (l < L) signal(&bar_opnd);
                if (!FWD_ONLY) tm_st(tblock, v);                     // stash, in place of z^l
                if (l < L) make_wlo(R2, trow + kTcDwCol);
            }
            worker_sync<kWorkers>();                         // R0 holds a^L

            // ================= output layer + integrand =================
            {
                constexpr int G = kWorkers / 128;            // neuron groups per (channel, point)
                const int g = tid >> 7, cp = tid & 127, c = cp >> 4, p = cp & 15;
                if (c < C) {
                    const float* ap = R0 + (c >> 1) * kTcPlaneF;
                    const float* wo = sp + lay.off_wout();
                    const int lp = (p >> 2) * 8 + (c & 1) * 4 + (p & 3);
                    float s = 0.f;
#pragma unroll 8
                    for (int k = g * (128 / G); k < (g + 1) * (128 / G); ++k) s = fmaf(ap[tc_mn32_off(k, lp)], wo[k], s);
                    red[(g * CP + c) * P + p] = s;
                }
            }
            worker_sync<kWorkers>();
            if (w == 0) {
                constexpr int G = kWorkers / 128;
                const int p = lane & (P - 1);               // lanes 16..31 shadow lanes 0..15 and contribute nothing
                const long long gi = base + p;
                const bool valid = lane < P && gi < sq.n;
                const TcOp* ops_t = ops_s + sq.tc_off;
                const int n_ops = sq.tc_n_ops, n_ent = kTcVmInputs + n_ops;
                float* V = R2 + lane;
                float* A = V + n_ent * 32;
                float u0 = 0.f;
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    float sj = c == 0 ? sp[lay.off_sc()] : 0.f;   // b_out
#pragma unroll
                    for (int g = 0; g < G; ++g) sj += red[(g * CP + c) * P + p];
                    V[c * 32] = sj;
                    if (c == 0) u0 = sj;
                }
#pragma unroll
                for (int dim = 0; dim < 3; ++dim) V[(kTcVmCoord + dim) * 32] = dim < d ? Xs[dim * 32 + p] : 0.f;
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
                if (lane < P) {
#pragma unroll
                    for (int c = 0; c < C; ++c) Ub[c * P + p] = A[c * 32];
                    if (CP > C) Ub[(CP - 1) * P + p] = 0.f;
                }
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
                float q0 = lsum, q1 = valid ? jb0 : 0.f;
#pragma unroll
                for (int o = 16; o >= 1; o >>= 1) {
                    q0 += __shfl_xor_sync(0xffffffffu, q0, o);
                    q1 += __shfl_xor_sync(0xffffffffu, q1, o);
                }
                if (lane == 0) { loss_acc += q0; sgs[0] += q1; }
                for (int vi = 0; vi < prm.n_vars; ++vi) {
                    float qv = valid ? A[(kTcVmVars + vi) * 32] : 0.f;
#pragma unroll
                    for (int o = 16; o >= 1; o >>= 1) qv += __shfl_xor_sync(0xffffffffu, qv, o);
                    if (lane == 0) sgs[2 + vi] += qv;
                }
                }
            }
            worker_sync<kWorkers>();                         // Ub is complete
            if (FWD_ONLY) {                              // prediction ends here: R0 and the tables in R2 are free again
                __syncwarp();
                if (lane == 0) mbar_arrive(&bar_r0);
                continue;
            }

            // ================= backward =================
            // cotangent of a^L: v[c][p] = ubar[c][p] w_out[n];  d w_out[n] += sum_{c,p} a^L[c][n][p] ubar[c][p]
#pragma unroll
            for (int c = 0; c < CP; ++c) {
                const float4 ub = *reinterpret_cast<const float4*>(Ub + c * P + pb);
                const float4 av = *reinterpret_cast<const float4*>(R0 + (c >> 1) * kTcPlaneF + coffM[c & 1]);
                g_wout = fmaf(av.x, ub.x, g_wout); g_wout = fmaf(av.y, ub.y, g_wout);
                g_wout = fmaf(av.z, ub.z, g_wout); g_wout = fmaf(av.w, ub.w, g_wout);
                v[c][0] = ub.x * wout_n; v[c][1] = ub.y * wout_n; v[c][2] = ub.z * wout_n; v[c][3] = ub.w * wout_n;
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar_r0);             // R0 may become the weight ring
            tmem_wait_st();                                  // this thread's stash stores have landed

            for (int l = L; l >= 1; --l) {
                // ---- reverse through the tanh jets of layer l: v (cotangent of a^l) -> cotangent of z^l
                {
                    float s[CP][PT];
                    if (l >= 2) {
                        tm_ld(trow + (l - 2) * NB, s);
                        tmem_wait_ld();
                    } else {
                        layer1_pre(s);
#pragma unroll
                        for (int p = 0; p < PT; ++p) s[0][p] = tanh_acc(s[0][p]);
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
                        if (CP > C) { v[CP - 1][p] = 0.f; v[CP - 1][p + 1] = 0.f; }
                        gb += zb[0];
                    }
                    sg_b[(l - 1) * HP + n] += gb.x + gb.y;
                }
                if (l == 1) {
#pragma unroll
                    for (int dim = 0; dim < 3; ++dim) {
                        if (dim < d) {
                            float s0 = 0.f;
#pragma unroll
                            for (int p = 0; p < PT; ++p) {
                                s0 = fmaf(Xs[dim * 32 + pb + p], v[0][p], s0);
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
                    if (L >= 2) {
                        mbar_wait(&bar_wg, p_wg); p_wg ^= 1;     // dW_2 of this tile is complete; R0..R2 are free again
                        p_chw ^= 1;
                        tmem_fence_after_sync();
                        drain_dw(0);
                        // every warp has read its share of the dW columns before any warp's next make_wlo writes the weight
                        // operand's lo parts across all of them (see tc_kernel.cuh)
                        tmem_fence_before_sync();
                        worker_sync<kWorkers>();
                        tmem_fence_after_sync();
                    }
                    break;
                }
                // ---- zbar^l -> R1 (and its lo image -> R2), MN-major: operand of dgrad.  Plane pl of R1 / R2 still holds
                // the K-major zbar^{l+1} the running wgrad reads: it is free once that wgrad's plane pl has drained.
#pragma unroll
                for (int pl = 0; pl < NP; ++pl) {
                    if (l < L) mbar_wait(&bar_ch[pl], p_chw);
                    store_plane_mn(v, pl, R1, R2);
                }
                if (l < L) p_chw ^= 1;
                signal(&bar_opnd);
                make_wlo(R0, trow + (l == L ? kTcDwCol : (l - 1) * NB));
                if (l < L) {                     // in the shadow of dgrad (the next wgrad comes after it)
                    mbar_wait(&bar_wg, p_wg); p_wg ^= 1;
                    tmem_fence_after_sync();
                    drain_dw(l - 1);
                }
                // ---- a^{l-1}, recomputed while the tensor core runs dgrad
                float a[CP][PT];
                {
                    float s[CP][PT];
                    if (l - 1 >= 2) {
                        tm_ld(trow + (l - 3) * NB, s);
                        tmem_wait_ld();
                    } else {
                        layer1_pre(s);
#pragma unroll
                        for (int p = 0; p < PT; ++p) s[0][p] = tanh_acc(s[0][p]);
                    }
                    recompute_all(s, a);
                }
                mbar_wait(&bar_done, p_done); p_done ^= 1;     // dgrad done: result in TMEM; R0, R1, R2 free
                tmem_fence_after_sync();
                // plane by plane, so that the tensor core starts on plane pl while pl + 1 is being stored
                const uint32_t tblock = trow + (l - 2) * NB;
#pragma unroll
                for (int pl = 0; pl < NP; ++pl) {
                    // (zbar^l needs nothing: the tensor core reads its MN-major hi / lo planes in R1 / R2 K-major)
                    const uint32_t tc0 = tblock + pl * 32 + hf * 8;
                    tmem_ld8(tc0, &v[2 * pl][0]);                          // cotangent of a^{l-1} (dgrad result), both channels of the plane
                    store_plane_k(a, pl, R0, nullptr);                     // a^{l-1} hi, K-major (B operand of wgrad)
#pragma unroll
                    for (int p = 0; p < PT; ++p) { a[2 * pl][p] = tf32_lo(a[2 * pl][p]); a[2 * pl + 1][p] = tf32_lo(a[2 * pl + 1][p]); }
                    tmem_wait_ld();                      // v has arrived: its columns may be overwritten
                    tmem_st8(tc0, &a[2 * pl][0]);                          // lo of a^{l-1}: A operand of wgrad, from TMEM
                    tmem_wait_st();
                    signal(&bar_lo_ready[pl]);
                }
                if (CP > C) {
#pragma unroll
                    for (int p = 0; p < PT; ++p) v[CP - 1][p] = 0.f;
                }
            }
            first_tile = false;
        }
        // ---- this thread's register accumulators -> shared memory (R0 is free: [NH][4][HP])
        worker_sync<kWorkers>();
        R0[(hf * 4 + 0) * HP + n] = g_wout;
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
            const int row = i / HP, nn = i - row * HP;
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
