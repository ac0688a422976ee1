// Instantiations of the fused jet kernel for second-order jets in two dimensions with the mixed partial (K2 = -1 tag).
#include "variants.h"

namespace tdq {
const Variant kVariants_22m[kVariantsPerCfg] = {
    TDQ_VARIANTS_FOR(2, 2, -1, 32),
    TDQ_VARIANTS_FOR(2, 2, -1, 64),
    TDQ_VARIANTS_FOR(2, 2, -1, 128),
};
}  // namespace tdq
