// Instantiations of the fused jet kernel for derivative orders (4,1,0) per input dimension.
#include "variants.h"

namespace tdq {
const Variant kVariants_410[kVariantsPerCfg] = {
    TDQ_VARIANTS_FOR(4, 1, 0, 32),
    TDQ_VARIANTS_FOR(4, 1, 0, 64),
    TDQ_VARIANTS_FOR(4, 1, 0, 128),
};
}  // namespace tdq
