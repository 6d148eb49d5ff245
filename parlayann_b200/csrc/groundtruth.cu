#include "pann_internal.h"
extern "C" int pann_groundtruth(pann_index*, const void*, uint64_t, uint64_t, uint32_t, int64_t, int32_t*, float*) { PANN_FAIL("pann_groundtruth: not built yet"); }
extern "C" int pann_groundtruth_resident(pann_index*, uint32_t, int64_t) { PANN_FAIL("not built yet"); }
extern "C" int pann_groundtruth_fetch(pann_index*, int32_t*, float*) { PANN_FAIL("not built yet"); }
extern "C" int pann_groundtruth_merge(int, uint32_t, uint64_t, uint32_t, const int32_t*, const float*, int32_t*, float*) { PANN_FAIL("not built yet"); }
