// beam_search_inst.cuh — instantiates the K1 kernels of one element type (PANN_INST_DT) so the three
// element types compile in parallel translation units.
#include "beam_search.cuh"

#define PANN_CAT_(a, b) a##b
#define PANN_CAT(a, b) PANN_CAT_(a, b)

template <int METRIC, int CPL, bool GS>
static const void* kernel_of() { return (const void*)&pann_beam_search_kernel<PANN_INST_DT, METRIC, CPL, GS>; }

template <int METRIC>
static const void* pick_cpl(int cpl, bool gs) {
  switch (cpl) {
    case 1: return gs ? kernel_of<METRIC, 1, true>() : kernel_of<METRIC, 1, false>();
    case 2: return gs ? kernel_of<METRIC, 2, true>() : kernel_of<METRIC, 2, false>();
    case 3: return gs ? kernel_of<METRIC, 3, true>() : kernel_of<METRIC, 3, false>();
    default: return gs ? kernel_of<METRIC, 0, true>() : kernel_of<METRIC, 0, false>();
  }
}

const void* PANN_CAT(pann_beam_kernel_dt, PANN_INST_DT)(int metric, int cpl, bool gs) {
  return metric == PANN_M_L2 ? pick_cpl<PANN_M_L2>(cpl, gs) : pick_cpl<PANN_M_MIPS>(cpl, gs);
}
