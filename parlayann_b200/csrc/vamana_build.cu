#include "pann_internal.h"
extern "C" int pann_vamana_build(pann_index*, uint32_t, uint32_t, double, int, int, uint32_t*, uint32_t*, double*) { PANN_FAIL("pann_vamana_build: not built yet"); }
