// Instantiations of the fused jet kernel for derivative orders (0,0,0) per input dimension.
#include "variants.h"

namespace tdq {
const Variant kVariants_000[kVariantsPerCfg] = {
    TDQ_VARIANTS_FOR(0, 0, 0, 32),
    TDQ_VARIANTS_FOR(0, 0, 0, 64),
    TDQ_VARIANTS_FOR(0, 0, 0, 128),
};
}  // namespace tdq
