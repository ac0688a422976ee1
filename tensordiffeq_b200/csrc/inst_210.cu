// Instantiations of the fused jet kernel for derivative orders (2,1,0) per input dimension.
#include "variants.h"

namespace tdq {
const Variant kVariants_210[kVariantsPerCfg] = {
    TDQ_VARIANTS_FOR(2, 1, 0, 32),
    TDQ_VARIANTS_FOR(2, 1, 0, 64),
    TDQ_VARIANTS_FOR(2, 1, 0, 128),
};
}  // namespace tdq
