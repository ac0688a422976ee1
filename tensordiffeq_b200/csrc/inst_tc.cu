// Instantiations of the tensor-core jet kernel (128-wide nets, derivative orders with at most 4 channels).
#include "tc_kernel.cuh"
#include "variants.h"

namespace tdq {
namespace {
template <int K0, int K1, int K2, bool FWD>
cudaError_t launch_tc(const KParams& p, int grid, size_t sm, cudaStream_t s) {
    auto k = fused_jet_tc_kernel<K0, K1, K2, kTcWorkerWarps, FWD>;
    static SmemOptIn opt;
    cudaError_t e = opt.ensure((const void*)k, sm);
    if (e != cudaSuccess) return e;
    k<<<grid, (kTcWorkerWarps + 1) * 32, sm, s>>>(p);
    return cudaGetLastError();
}
template <int K0, int K1, int K2>
TcVariant make_tc() {
    return TcVariant{K0, K1, K2, 1 + K0 + K1 + K2, 32, &launch_tc<K0, K1, K2, false>, (const void*)fused_jet_tc_kernel<K0, K1, K2, kTcWorkerWarps>,
                     &launch_tc<K0, K1, K2, true>};
}
const TcVariant kTc[] = {make_tc<0, 0, 0>(), make_tc<1, 1, 0>(), make_tc<2, 1, 0>()};
}  // namespace

const TcVariant* find_tc16_variant(int k0, int k1, int k2);      // inst_tc16.cu

const TcVariant* find_tc_variant(int k0, int k1, int k2) {
    for (const TcVariant& v : kTc)
        if (v.k0 == k0 && v.k1 == k1 && v.k2 == k2) return &v;
    return find_tc16_variant(k0, k1, k2);
}
}  // namespace tdq
