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

// K3 / K6 of the same element type
#include "vamana_prune.cuh"
template <int METRIC, int CPL, bool GS>
static const void* prune_of() { return (const void*)&pann_prune_kernel<PANN_INST_DT, METRIC, CPL, GS>; }
template <int METRIC>
static const void* pick_prune(int cpl, bool gs) {
  switch (cpl) {
    case 1: return gs ? prune_of<METRIC, 1, true>() : prune_of<METRIC, 1, false>();
    case 2: return gs ? prune_of<METRIC, 2, true>() : prune_of<METRIC, 2, false>();
    case 3: return gs ? prune_of<METRIC, 3, true>() : prune_of<METRIC, 3, false>();
    default: return gs ? prune_of<METRIC, 0, true>() : prune_of<METRIC, 0, false>();
  }
}
const void* PANN_CAT(pann_prune_kernel_dt, PANN_INST_DT)(int metric, int cpl, bool gs) {
  return metric == PANN_M_L2 ? pick_prune<PANN_M_L2>(cpl, gs) : pick_prune<PANN_M_MIPS>(cpl, gs);
}
template <int METRIC>
static const void* pick_edge(int cpl) {
  switch (cpl) {
    case 1: return (const void*)&pann_edge_dist_kernel<PANN_INST_DT, METRIC, 1>;
    case 2: return (const void*)&pann_edge_dist_kernel<PANN_INST_DT, METRIC, 2>;
    case 3: return (const void*)&pann_edge_dist_kernel<PANN_INST_DT, METRIC, 3>;
    default: return (const void*)&pann_edge_dist_kernel<PANN_INST_DT, METRIC, 0>;
  }
}
const void* PANN_CAT(pann_edge_kernel_dt, PANN_INST_DT)(int metric, int cpl) {
  return metric == PANN_M_L2 ? pick_edge<PANN_M_L2>(cpl) : pick_edge<PANN_M_MIPS>(cpl);
}
