// beam_search_inst.cuh — instantiates K1 / K3 / K6 of one element type (PANN_INST_DT) so the three
// element types compile in parallel translation units.  QR = PANN_QR (128-byte rows, query in registers) or 0.
#include "beam_search.cuh"
#include "beam_search_small.cuh"
#include "vamana_prune.cuh"

#define PANN_CAT_(a, b) a##b
#define PANN_CAT(a, b) PANN_CAT_(a, b)

template <int METRIC>
static const void* pick_beam(int qr, bool gs) {
  if (qr == PANN_QR) return gs ? (const void*)&pann_beam_search_kernel<PANN_INST_DT, METRIC, PANN_QR, true>
                               : (const void*)&pann_beam_search_kernel<PANN_INST_DT, METRIC, PANN_QR, false>;
  return gs ? (const void*)&pann_beam_search_kernel<PANN_INST_DT, METRIC, 0, true>
            : (const void*)&pann_beam_search_kernel<PANN_INST_DT, METRIC, 0, false>;
}
const void* PANN_CAT(pann_beam_kernel_dt, PANN_INST_DT)(int metric, int qr, bool gs) {
  return metric == PANN_M_L2 ? pick_beam<PANN_M_L2>(qr, gs) : pick_beam<PANN_M_MIPS>(qr, gs);
}

template <int METRIC>
static const void* pick_beam_small(int qr, bool ext) {
  if (qr == PANN_QR) return ext ? (const void*)&pann_beam_search_small_kernel<PANN_INST_DT, METRIC, PANN_QR, true>
                                : (const void*)&pann_beam_search_small_kernel<PANN_INST_DT, METRIC, PANN_QR, false>;
  return ext ? (const void*)&pann_beam_search_small_kernel<PANN_INST_DT, METRIC, 0, true>
             : (const void*)&pann_beam_search_small_kernel<PANN_INST_DT, METRIC, 0, false>;
}
const void* PANN_CAT(pann_beam_small_kernel_dt, PANN_INST_DT)(int metric, int qr, bool ext) {
  return metric == PANN_M_L2 ? pick_beam_small<PANN_M_L2>(qr, ext) : pick_beam_small<PANN_M_MIPS>(qr, ext);
}

// the WIDE variant (small batches): 128-byte integer rows only
const void* PANN_CAT(pann_beam_small_wide_kernel_dt, PANN_INST_DT)(int metric) {
  if constexpr (pann_quad_rows<PANN_INST_DT, PANN_QR>())
    return metric == PANN_M_L2 ? (const void*)&pann_beam_search_small_kernel<PANN_INST_DT, PANN_M_L2, PANN_QR, false, true>
                               : (const void*)&pann_beam_search_small_kernel<PANN_INST_DT, PANN_M_MIPS, PANN_QR, false, true>;
  else return nullptr;
}

template <int METRIC>
static const void* pick_prune(int qr, bool gs) {
  if (qr == PANN_QR) return gs ? (const void*)&pann_prune_kernel<PANN_INST_DT, METRIC, PANN_QR, true>
                               : (const void*)&pann_prune_kernel<PANN_INST_DT, METRIC, PANN_QR, false>;
  return gs ? (const void*)&pann_prune_kernel<PANN_INST_DT, METRIC, 0, true>
            : (const void*)&pann_prune_kernel<PANN_INST_DT, METRIC, 0, false>;
}
const void* PANN_CAT(pann_prune_kernel_dt, PANN_INST_DT)(int metric, int qr, bool gs) {
  return metric == PANN_M_L2 ? pick_prune<PANN_M_L2>(qr, gs) : pick_prune<PANN_M_MIPS>(qr, gs);
}

template <int METRIC>
static const void* pick_edge(int qr) {
  if (qr == PANN_QR) return (const void*)&pann_edge_dist_kernel<PANN_INST_DT, METRIC, PANN_QR>;
  return (const void*)&pann_edge_dist_kernel<PANN_INST_DT, METRIC, 0>;
}
const void* PANN_CAT(pann_edge_kernel_dt, PANN_INST_DT)(int metric, int qr) {
  return metric == PANN_M_L2 ? pick_edge<PANN_M_L2>(qr) : pick_edge<PANN_M_MIPS>(qr);
}

const void* PANN_CAT(pann_rerank_kernel_dt, PANN_INST_DT)(int metric) {
  return metric == PANN_M_L2 ? (const void*)&pann_rerank_kernel<PANN_INST_DT, PANN_M_L2> : (const void*)&pann_rerank_kernel<PANN_INST_DT, PANN_M_MIPS>;
}
