// Instantiations of the fused jet kernel for derivative orders (1,1,0) per input dimension.
#include "variants.h"

namespace tdq {
const Variant kVariants_110[kVariantsPerCfg] = {
    TDQ_VARIANTS_FOR(1, 1, 0, 32),
    TDQ_VARIANTS_FOR(1, 1, 0, 64),
    TDQ_VARIANTS_FOR(1, 1, 0, 128),
};
}  // namespace tdq
