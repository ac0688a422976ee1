// Instantiations of the fused jet kernel for derivative orders (2,2,2) per input dimension.
#include "variants.h"

namespace tdq {
const Variant kVariants_222[kVariantsPerCfg] = {
    TDQ_VARIANTS_FOR(2, 2, 2, 32),
    TDQ_VARIANTS_FOR(2, 2, 2, 64),
    TDQ_VARIANTS_FOR(2, 2, 2, 128),
};
}  // namespace tdq
