// Instantiations of the 16-point-tile tensor-core jet kernel (128-wide nets, derivative orders with 5 .. 8 channels).
#include "tc16_kernel.cuh"
#include "variants.h"

namespace tdq {
namespace {
template <int K0, int K1, int K2, bool FWD>
cudaError_t launch_tc16(const KParams& p, int grid, size_t sm, cudaStream_t s) {
    auto k = fused_jet_tc16_kernel<K0, K1, K2, kTcWorkerWarps, FWD>;
    static SmemOptIn opt;
    cudaError_t e = opt.ensure((const void*)k, sm);
    if (e != cudaSuccess) return e;
    k<<<grid, (kTcWorkerWarps + 1) * 32, sm, s>>>(p);
    return cudaGetLastError();
}
template <int K0, int K1, int K2>
TcVariant make_tc16() {
    return TcVariant{K0, K1, K2, 1 + K0 + K1 + K2, 16, &launch_tc16<K0, K1, K2, false>, (const void*)fused_jet_tc16_kernel<K0, K1, K2, kTcWorkerWarps>,
                     &launch_tc16<K0, K1, K2, true>};
}
const TcVariant kTc16[] = {make_tc16<2, 2, 0>(), make_tc16<4, 1, 0>(), make_tc16<2, 2, 2>(), make_tc16<2, 2, -1>()};
}  // namespace

const TcVariant* find_tc16_variant(int k0, int k1, int k2) {
    for (const TcVariant& v : kTc16)
        if (v.k0 == k0 && v.k1 == k1 && v.k2 == k2) return &v;
    return nullptr;
}
}  // namespace tdq
