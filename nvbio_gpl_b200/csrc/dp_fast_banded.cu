// dp_fast_banded.cu -- packed s16x2 DPX banded kernels (placeholder until the kernel lands)
#include "dp_fast.h"
bool dpf_banded_score_supported(const nvb_aligner&, uint32_t, const StrSet&, const StrSet&, uint32_t, uint32_t) { return false; }
int  dpf_banded_score(nvb_ctx*, const nvb_aligner&, uint32_t, StrSet, StrSet, nvb_best_sink*, uint32_t, uint32_t, uint32_t*, cudaStream_t) { return NVB_ERR_UNSUPPORTED; }
