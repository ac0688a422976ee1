// Two half-tiles in flight: the forward-only tensor-core jet kernel behind tdq_predict for 128-wide nets with at most 4
// jet channels.  A 32-point tile is split into two 16-point halves A and B, each owned by its own group of 8 worker
// warps; ONE control warp issues the tensor-core work of both, alternately (GEMM A, GEMM B, GEMM A, ...), so that a
// group's element-wise phase (tanh jets, operand stores) runs while the tensor core works on the OTHER half -- in
// tc_kernel.cuh the two take turns.  Both halves use the same weight planes (streamed once per layer through the 4-slot
// ring) and the same lo parts of the weight operand in Tensor Memory (derived by the leading group A alone, so that A's
// GEMM never waits for B's element-wise phase).
//
// Geometry (the 16-point layout of tc16_kernel.cuh): a 128-byte operand line of half h holds two channels x 16 points,
// element 16 hf + 8 (c % 2) + i = point 8 hf + i of channel c (hf = which 8 points, i < 8); half h owns planes 2 h and
// 2 h + 1 of a 64 KB region and the 64 Tensor Memory columns (2 (l - 2) + h) * 64 .. of layer l; a worker thread owns one
// neuron x 8 points x all channels = 16 consecutive columns per plane.  MMAs are M = 128, N = 64, K = 8 (3xTF32 as in
// tc_kernel.cuh).  Read tc_kernel.cuh's header for the operand layouts and the ring.
//
// EXPERIMENT, NOT BUILT INTO THE LIBRARY (round 2).  Forward sweep only; it was wired behind tdq_predict, passed the predict
// goldens, and measured 5.42 ms for 2 M points against 5.48 ms for tc_kernel.cuh's forward-only mode: the overlap of
// element-wise work with the other half's GEMM is bought with N = 64 MMAs that read the 4 KB weight operand once per HALF,
// i.e. 640 KB instead of 512 KB through the shared-memory data pipe per layer and 32-point tile -- and that pipe, already
// 74 % busy in tc_kernel.cuh (profiles/r02_ncu_tc_kernel.txt), is the ceiling.  Kept as the record of the measurement
// (DESIGN.md section 9).  To build it: include it from csrc/inst_tc.cu and add a TcVariant whose launcher starts
// fused_jet_tc2_fwd_kernel with 17 x 32 threads.
#pragma once
#include "../../tensordiffeq_b200/csrc/tc_kernel.cuh"

namespace tdq {

template <int N>
__device__ __forceinline__ void group_sync(int id) { asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(N) : "memory"); }

template <int K0, int K1, int K2>
__global__ void __launch_bounds__(17 * 32, 1) fused_jet_tc2_fwd_kernel(const __grid_constant__ KParams prm) {
    using J = JetCfg<K0, K1, K2>;
    constexpr int C = J::C;
    static_assert(C == 3 || C == 4, "half-tile kernel: 3 or 4 channels (two planes per half)");
    constexpr int HP = 128, P = 32, PH = 16, CP = 4, NBH = 64, NW = 16, PT = 8;
    constexpr int kGroup = 8 * 32, kThreadsAll = 17 * 32;

    extern __shared__ __align__(1024) float smem[];
    float* R0 = smem;
    float* R1 = R0 + 4 * kTcPlaneF;
    float* R2 = R1 + 4 * kTcPlaneF;
    const int L = prm.L;
    const PackLayout lay{HP, L};
    const int nsmall = lay.small();
    float* sp = R2 + 4 * kTcPlaneF;             // small params (PackLayout order)
    float* Xs = sp + nsmall;                    // [2 halves][3][16]
    float* red = Xs + 2 * 3 * PH;               // [2 halves][4 neuron groups][CP][16]
    TcOp* ops_s = reinterpret_cast<TcOp*>(red + 2 * 4 * CP * PH);     // (nsmall, 96 and 512 are multiples of 4 floats: 16-byte aligned)
    __shared__ uint64_t bar_full[4], bar_empty[4], bar_wlo, bar_wlo_free, bar_opnd[2], bar_done[2];
    __shared__ uint32_t tmem_slot;

    const int tid = threadIdx.x, lane = tid & 31;
    const int w = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int d = prm.d;
    const long long n_tiles = prm.n_tiles;
    const bool has_tiles = (long long)blockIdx.x < n_tiles;

    for (int i = tid; i < nsmall; i += kThreadsAll) sp[i] = prm.wpack[i];
    for (int i = tid; i < prm.tc_n_ops; i += kThreadsAll) ops_s[i] = prm.tc_ops[i];
    if (tid == 0) {
        for (int i = 0; i < 4; ++i) { mbar_init(&bar_full[i], 1); mbar_init(&bar_empty[i], 1); }
        mbar_init(&bar_wlo, 8);                 // the leading group's 8 warps
        mbar_init(&bar_wlo_free, 1);
        for (int h = 0; h < 2; ++h) { mbar_init(&bar_opnd[h], 8); mbar_init(&bar_done[h], 1); }
    }
    if (w == NW) tmem_alloc_512(&tmem_slot);
    tmem_fence_before_sync();
    __syncthreads();
    tmem_fence_after_sync();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_slot, 0);

    if (w == NW) {
        // =========================== control warp ===========================
        if (has_tiles) {
            const int leader = lane == 0;
            uint32_t pf = 0, pe = 0, p_wlo = 0, p_opnd[2] = {0, 0};
            constexpr uint32_t idesc_mn = umma_idesc_tf32(128, NBH) | (1u << 16);   // B operand MN-major, N = 64
            const uint32_t a0 = smem_u32(R0) >> 4;
            constexpr uint32_t kReg = (4 * kTcPlaneB) >> 4;
            auto wait_full = [&](int s) { mbar_wait(&bar_full[s], (pf >> s) & 1); pf ^= 1u << s; };
            auto wait_empty = [&](int s) { mbar_wait(&bar_empty[s], (pe >> s) & 1); pe ^= 1u << s; };
            auto tma_plane = [&](int s, int m) {
                if (!leader) return;
                mbar_expect_tx(&bar_full[s], kTcPlaneB);
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                                 smem_u32(R2 + s * kTcPlaneF)),
                             "l"(prm.wimg + tc_img_off(m, 0, s)), "r"(kTcPlaneB), "r"(smem_u32(&bar_full[s]))
                             : "memory");
            };
            auto dk = [&](uint32_t off16) {
                const uint32_t lo = a0 + off16 + (1u << 16), hi = (1024u >> 4) | (1u << 14) | (2u << 29);
                uint64_t dd;
                asm("mov.b64 %0, {%1, %2};" : "=l"(dd) : "r"(lo), "r"(hi));
                return dd;
            };
            auto dmn = [&](uint32_t off16) {
                const uint32_t lo = a0 + off16 + ((kTcPlaneB >> 4) << 16), hi = (512u >> 4) | (1u << 14) | (1u << 29);
                uint64_t dd;
                asm("mov.b64 %0, {%1, %2};" : "=l"(dd) : "r"(lo), "r"(hi));
                return dd;
            };
            // forward GEMM of half h: D^T[n][(plane, line position)] = sum_k W^T[n][k] a_h[(..)][k]; B operand = the half's
            // two planes (2 h, 2 h + 1) of R0 (hi) / R1 (lo image); A = ring in R2, lo parts of W in TMEM at wlo
            auto gemm_half = [&](auto h_tag, uint32_t dcol, uint32_t wlo) {
                constexpr int h = decltype(h_tag)::value;
                constexpr uint32_t oA = 2 * kReg, oBh = 0 + h * 2 * (kTcPlaneB >> 4), oBl = kReg + h * 2 * (kTcPlaneB >> 4);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    if (h == 0) wait_full(q);
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) {
                        const uint32_t ao = oA + q * (kTcPlaneB >> 4) + ks * 2, bo = (q * 32 + ks * 8) * 128 >> 4;
                        if (leader) { if (q == 0 && ks == 0) umma_tf32_i<false>(dcol, dk(ao), dmn(oBl + bo), idesc_mn); else umma_tf32_i<true>(dcol, dk(ao), dmn(oBl + bo), idesc_mn); }
                        if (leader) umma_tf32_i<true>(dcol, dk(ao), dmn(oBh + bo), idesc_mn);
                    }
                    if (h == 1 && leader) umma_commit(&bar_empty[q]);      // the slot is free once BOTH halves have used it
                }
                if (h == 0) { mbar_wait(&bar_wlo, p_wlo); p_wlo ^= 1; tmem_fence_after_sync(); }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) {
                        const uint32_t bo = (q * 32 + ks * 8) * 128 >> 4;
                        if (leader) umma_tf32_ts_i<true>(dcol, wlo + q * 32 + ks * 8, dmn(oBh + bo), idesc_mn);
                    }
                }
                if (leader) umma_commit(&bar_done[h]);
                if (h == 1 && leader) umma_commit(&bar_wlo_free);           // the lo parts of this layer's weights are no longer read
            };
#pragma unroll
            for (int q = 0; q < 4; ++q) tma_plane(q, 0);
            for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                for (int l = 2; l <= L; ++l) {
                    mbar_wait(&bar_opnd[0], p_opnd[0]); p_opnd[0] ^= 1;
                    tmem_fence_after_sync();
                    gemm_half(std::integral_constant<int, 0>{}, tmem_base + ((l - 2) * 2 + 0) * NBH, tmem_base + kTcDwCol);
                    mbar_wait(&bar_opnd[1], p_opnd[1]); p_opnd[1] ^= 1;
                    tmem_fence_after_sync();
                    gemm_half(std::integral_constant<int, 1>{}, tmem_base + ((l - 2) * 2 + 1) * NBH, tmem_base + kTcDwCol);
                    // next layer's planes (next tile's first layer after the last GEMM of this one, once both groups are
                    // done with the tables they keep in R1 -- R2 itself is only the ring here)
                    const bool more = l < L || tile + gridDim.x < n_tiles;
#pragma unroll
                    for (int q = 0; q < 4; ++q) { wait_empty(q); if (more) tma_plane(q, l < L ? l - 1 : 0); }
                }
                // (a group stores the next tile's operands into its own planes itself, after its program phase: no cross-group
                //  hazard on R0 / R1)
            }
        }
    } else {
        // =========================== worker warps: group g owns half g ===========================
        const int g = w >> 3, ww = w & 7, q4 = ww & 3, hf = ww >> 2;
        const int n = q4 * 32 + lane;
        const int gtid = tid & (kGroup - 1);                 // thread index inside the group
        const uint32_t trow = tmem_base + ((uint32_t)(q4 * 32) << 16);
        const float* b_s = sp + lay.off_b();
        const float* w0_s = sp + lay.off_w0();
        float* Xg = Xs + g * 3 * PH;
        float* redg = red + g * 4 * CP * PH;
        float* R0h = R0 + g * 2 * kTcPlaneF;                 // this half's planes
        float* R1h = R1 + g * 2 * kTcPlaneF;
        uint32_t p_done = 0, p_wfree = 1;      // the lo-part columns start out free (waiting on parity 1 of a fresh barrier passes)
        const bool swp = (n >> 2) & 1;
        int coffM[2][2];                                     // [channel parity][16-byte half of the chunk]
#pragma unroll
        for (int s = 0; s < 2; ++s)
#pragma unroll
            for (int j = 0; j < 2; ++j) coffM[s][j] = tc_mn32_off(n, hf * 16 + s * 8 + j * 4);
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
        // activations a[c][8 points] -> the half's MN-major planes (channel c: plane c / 2, chunk 2 hf + c % 2)
        auto store_act = [&](float (&a)[CP][PT], float* hi, float* lo) {
#pragma unroll
            for (int c = 0; c < CP; ++c)
                store_pair(hi + (c >> 1) * kTcPlaneF, lo ? lo + (c >> 1) * kTcPlaneF : nullptr, coffM[c & 1][0], coffM[c & 1][1],
                           make_float4(a[c][0], a[c][1], a[c][2], a[c][3]), make_float4(a[c][4], a[c][5], a[c][6], a[c][7]));
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
                for (int p = 0; p < PT; ++p) v[0][p] = fmaf(Xg[dim * PH + hf * 8 + p], wv, v[0][p]);
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
        auto recompute_all = [&](const float (&v)[CP][PT], float (&a)[CP][PT]) {
#pragma unroll
            for (int p = 0; p < PT; ++p) {
                float st[C], act[C];
#pragma unroll
                for (int c = 0; c < C; ++c) st[c] = v[c][p];
                J::recompute(st, act);
#pragma unroll
                for (int c = 0; c < C; ++c) a[c][p] = act[c];
                if (CP > C) a[CP - 1][p] = 0.f;
            }
        };
        auto signal = [&](uint64_t* bar) {
            fence_proxy_async_smem();
            tmem_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar);
        };
        // lo parts of the weight operand of the coming GEMM pair: the leading group alone, each thread its row n and 16 K
        // columns of every K-block (the ring's hi planes are resident; the previous pair no longer reads the TMEM columns)
        uint32_t wf = 0;
        auto make_wlo = [&](uint32_t tcol) {
            mbar_wait(&bar_wlo_free, p_wfree); p_wfree ^= 1;
            tmem_fence_after_sync();
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                mbar_wait(&bar_full[q], (wf >> q) & 1); wf ^= 1u << q;
                float lo[16];
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) {
                    const float4 h4 = *reinterpret_cast<const float4*>(R2 + q * kTcPlaneF + n * 32 + (((hf * 4 + jj) ^ (n & 7)) << 2));
                    lo[jj * 4 + 0] = tf32_lo(h4.x); lo[jj * 4 + 1] = tf32_lo(h4.y);
                    lo[jj * 4 + 2] = tf32_lo(h4.z); lo[jj * 4 + 3] = tf32_lo(h4.w);
                }
                tmem_st16(tcol + q * 32 + hf * 16, lo);
            }
            tmem_wait_st();
            tmem_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar_wlo);
        };

        for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            const Seg& sq = prm.seg[seg_of_tile(prm, tile)];
            const long long base = (tile - sq.tile0) * P + g * PH;          // first point of this half
            group_sync<kGroup>(1 + g);                                       // previous readers of Xg / redg are done
            for (int i = gtid; i < d * PH; i += kGroup) {
                int dim = i / PH, p = i - dim * PH;
                long long gi = base + p;
                Xg[dim * PH + p] = gi < sq.n ? sq.X[(size_t)dim * sq.x_stride + gi] : 0.f;
            }
            group_sync<kGroup>(1 + g);
            float v[CP][PT], a[CP][PT];
            layer1_pre(v);
#pragma unroll
            for (int p = 0; p < PT; ++p) v[0][p] = tanh_acc(v[0][p]);
            recompute_all(v, a);
            store_act(a, R0h, R1h);
            signal(&bar_opnd[g]);
            if (g == 0) make_wlo(trow + kTcDwCol);
            for (int l = 2; l <= L; ++l) {
                mbar_wait(&bar_done[g], p_done); p_done ^= 1;
                tmem_fence_after_sync();
                const uint32_t tblock = trow + ((l - 2) * 2 + g) * NBH;
                tmem_ld16(tblock + 0 * 32 + hf * 16, &v[0][0]);              // channels 0, 1 (8 points each)
                tmem_ld16(tblock + 1 * 32 + hf * 16, &v[2][0]);              // channels 2, 3
                tmem_wait_ld();
                const float bz = b_s[(l - 1) * HP + n];
#pragma unroll
                for (int p = 0; p < PT; ++p) v[0][p] = tanh_acc(v[0][p] + bz);
                recompute_all(v, a);
                store_act(a, R0h, l < L ? R1h : nullptr);
                if (l < L) {
                    signal(&bar_opnd[g]);
                    if (g == 0) make_wlo(trow + kTcDwCol);
                }
            }
            group_sync<kGroup>(1 + g);                                       // the half's planes of R0 hold a^L

            // ---- output layer: u_c[p] = sum_n a^L[c][n][p] w_out[n]; 256 threads = 4 neuron groups x (4 channels x 16 points)
            {
                const int ng = gtid >> 6, cp = gtid & 63, c = cp >> 4, p = cp & 15;
                if (c < C) {
                    const float* ap = R0h + (c >> 1) * kTcPlaneF;
                    const float* wo = sp + lay.off_wout();
                    const int lp = (p >> 3) * 16 + (c & 1) * 8 + (p & 7);
                    float s = 0.f;
#pragma unroll 8
                    for (int k = ng * 32; k < (ng + 1) * 32; ++k) s = fmaf(ap[tc_mn32_off(k, lp)], wo[k], s);
                    redg[(ng * CP + c) * PH + p] = s;
                }
            }
            group_sync<kGroup>(1 + g);
            if (ww == 0) {
                // residual program for the half's 16 points (lanes 16..31 shadow lanes 0..15); tables in the half's R1 planes
                const int p = lane & (PH - 1);
                const long long gi = base + p;
                const bool valid = lane < PH && gi < sq.n;
                const TcOp* ops_t = ops_s + sq.tc_off;
                const int n_ops = sq.tc_n_ops, n_ent = kTcVmInputs + n_ops;
                float* V = R1h + lane;
                float u0 = 0.f;
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    float sj = c == 0 ? sp[lay.off_sc()] : 0.f;
#pragma unroll
                    for (int q = 0; q < 4; ++q) sj += redg[(q * CP + c) * PH + p];
                    V[c * 32] = sj;
                    if (c == 0) u0 = sj;
                }
#pragma unroll
                for (int dim = 0; dim < 3; ++dim) V[(kTcVmCoord + dim) * 32] = dim < d ? Xg[dim * PH + p] : 0.f;
                V[kTcVmTarget * 32] = (valid && sq.target) ? sq.target[gi] : 0.f;
                V[kTcVmLambda * 32] = (valid && sq.lam) ? sq.lam[gi] : 0.f;
                for (int vi = 0; vi < prm.n_vars; ++vi) V[(kTcVmVars + vi) * 32] = prm.vars[vi];
                (void)n_ent;
                tc_vm_forward(ops_t, n_ops, V);
                if (valid && sq.f_out && sq.tc_f >= 0) sq.f_out[gi] = V[sq.tc_f * 32];
                if (valid && sq.u_out) sq.u_out[gi] = u0;
            }
            // (the next iteration's first group_sync orders the program's reads of R1h / redg before they are overwritten)
        }
    }
    tmem_fence_before_sync();
    __syncthreads();
    if (w == NW) tmem_dealloc_512(tmem_base);
}

}  // namespace tdq
