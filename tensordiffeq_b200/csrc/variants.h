// Registry of compiled instantiations of the fused jet kernel.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

#include <mutex>

#include "fused_kernel.cuh"
#include "tc_image.cuh"

namespace tdq {

struct Variant {
    int k0, k1, k2, hp, tp, bwd;
    int C, P;
    size_t (*smem)(int L, int n_ops);
    cudaError_t (*launch)(const KParams& p, int grid, size_t smem, cudaStream_t s);
    const void* func;
};

// smallest compiled jet configuration that dominates (k0,k1,k2) at padded width hp
const Variant* find_variant(int k0, int k1, int k2, int hp, int bwd, int mixed = 0);
int n_variants();
const Variant* variant_at(int i);

// Opt a kernel into `sm` bytes of dynamic shared memory on the current device, once per (kernel, device) and
// growing only.  The attribute lives in the device's context, so the cache is per device ordinal; it is guarded by
// a mutex because handles on different host threads and tdq_loss_grad_static reach the same launcher, and devices
// beyond the cache always set the attribute.  (A cudaDeviceReset() invalidates the cache: not supported while the
// library is loaded.)
struct SmemOptIn {
    static constexpr int kDevs = 64;
    std::mutex mu;
    size_t configured[kDevs] = {0};
    cudaError_t ensure(const void* func, size_t sm) {
        int dev = -1;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return e;
        std::lock_guard<std::mutex> lk(mu);
        if (dev >= 0 && dev < kDevs && sm <= configured[dev]) return cudaSuccess;
        e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        if (e == cudaSuccess && dev >= 0 && dev < kDevs) configured[dev] = sm;
        return e;
    }
};

template <int K0, int K1, int K2, int HP, int TP, bool BWD>
struct VariantImpl {
    using S = Shape<K0, K1, K2, HP, TP>;
    static size_t smem(int L, int n_ops) { return smem_bytes<S>(L, n_ops); }
    static cudaError_t launch(const KParams& p, int grid, size_t sm, cudaStream_t s) {
        auto k = fused_jet_kernel<K0, K1, K2, HP, TP, BWD>;
        static SmemOptIn opt;
        cudaError_t e = opt.ensure((const void*)k, sm);
        if (e != cudaSuccess) return e;
        k<<<grid, kThreads, sm, s>>>(p);
        return cudaGetLastError();
    }
    static Variant make() {
        return Variant{K0, K1, K2, HP, TP, BWD ? 1 : 0, S::C, S::P, &smem, &launch,
                       (const void*)fused_jet_kernel<K0, K1, K2, HP, TP, BWD>};
    }
};

#define TDQ_TP(K0, K1, K2) (JetCfg<K0, K1, K2>::C <= 4 ? 4 : 2)
#define TDQ_VARIANTS_FOR(K0, K1, K2, HP)                                   \
    VariantImpl<K0, K1, K2, HP, TDQ_TP(K0, K1, K2), false>::make(),       \
    VariantImpl<K0, K1, K2, HP, TDQ_TP(K0, K1, K2), true>::make()

// per-translation-unit tables (one .cu per jet configuration, compiled in parallel)
extern const Variant kVariants_000[];
extern const Variant kVariants_110[];
extern const Variant kVariants_210[];
extern const Variant kVariants_220[];
extern const Variant kVariants_410[];
extern const Variant kVariants_222[];
extern const Variant kVariants_22m[];   // (2, 2, -1): second order in two dimensions with the mixed partial u_xy
constexpr int kVariantsPerCfg = 6;   // 3 widths x {fwd, fwd+bwd}

// tensor-core loss+gradient kernels (tc_kernel.cuh): width 128, at most 4 channels
constexpr int kTcWorkerWarps = 16;
__host__ __device__ inline size_t tc_smem_bytes(int L, int n_tc_ops) {
    size_t f = 3 * 4 * (size_t)kTcPlaneF;                  // R0..R2
    f += (size_t)small_floats(128, L);                    // small params
    f += (size_t)(kTcWorkerWarps / 4) * L * 128 + kScalars;   // bias-gradient accumulators per point group, scalars
    f += 3 * 32 + 4 * 32 + 4 * 4 * 32 + 64;               // Xs, Ub, red, scratch
    size_t b = f * 4 + (size_t)n_tc_ops * 32 + 64;          // + the compiled program (TcOp, vm_poly.h)
    return (b + 127) & ~(size_t)127;
}
// can a tensor-core kernel take this net?  (C channels, L tanh layers)
__host__ __device__ inline bool tc_supported(int C, int HP, int L) {
    // Tensor Memory columns per stash block: 32 per channel with 32-point tiles (C <= 4, tc_kernel.cuh), 32 per channel
    // PAIR with 16-point tiles (5 <= C <= 8, tc16_kernel.cuh).  The blocks of layers 2..L lie below the dW block, and
    // 128 columns for the lo parts of W_l must be free above the blocks still live during dgrad(l), l < L.
    const int NB = C <= 4 ? C * 32 : ((C + 1) / 2) * 32;
    return HP == 128 && C <= 8 && L >= 2 && L <= kTcMaxL && (L - 1) * NB <= kTcDwCol &&
           (L < 3 || (L - 2) * NB + 128 <= kTcDwCol);
}

struct TcVariant {
    int k0, k1, k2, C;
    int P;                   // points per tile
    cudaError_t (*launch)(const KParams& p, int grid, size_t smem, cudaStream_t s);
    const void* func;
    cudaError_t (*launch_fwd)(const KParams& p, int grid, size_t smem, cudaStream_t s);   // forward-only instantiation (tdq_predict)
};
const TcVariant* find_tc_variant(int k0, int k1, int k2);

}  // namespace tdq
