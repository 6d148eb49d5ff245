// oracle/b200_shim.h — TEST INFRASTRUCTURE: the call-site switch of INTEGRATION.md applied to the reference's own command-line
// programs WITHOUT editing them.  Force-included (-include) ahead of the reference's translation unit:
//   1. the reference headers that DEFINE the hot-path entry points are included first, unchanged (their include guards then
//      keep later #includes of them inert);
//   2. include/parlayann_b200/dropin.h adds parlayANN::b200::{qsearchAll, knn_index, compute_groundtruth};
//   3. forwarders pick the b200:: version for the type combinations the GPU path covers and the reference's own otherwise
//      (the CLI instantiates every quantisation mode, bench/neighborsTime.C:157-164, vamana/neighbors.h:100-190);
//   4. two macros rename the call sites that come AFTER this header (check_nn_recall.h:46-60 calls qsearchAll, vamana/neighbors.h:56
//      declares knn_index<...>) — the textual equivalent of a maintainer writing `b200::` there.
#pragma once
#include "utils/beamSearch.h"
#include "utils/euclidian_point.h"
#include "utils/mips_point.h"
#include "utils/point_range.h"
#include "vamana/index.h"
#include "parlayann_b200/dropin.h"

namespace parlayANN {
template <class PR, class QPR, class QQPR>
struct b200_covers_search
    : std::integral_constant<bool, std::is_same<QPR, QQPR>::value && b200::point_traits<typename PR::Point>::supported &&
                                       b200::point_traits<typename QPR::Point>::supported> {};
template <class QPR, class QQPR>
struct b200_covers_build : std::integral_constant<bool, std::is_same<QPR, QQPR>::value && b200::point_traits<typename QPR::Point>::supported> {};

template <typename PointRange, typename QPointRange, typename QQPointRange, typename indexType>
parlay::sequence<parlay::sequence<indexType>>
b200_qsearchAll(const PointRange& Query_Points, const QPointRange& Q_Query_Points, const QQPointRange& QQ_Query_Points, const Graph<indexType>& G,
                const PointRange& Base_Points, const QPointRange& Q_Base_Points, const QQPointRange& QQ_Base_Points, stats<indexType>& QueryStats,
                const indexType starting_point, const QueryParams& QP) {
  if constexpr (b200_covers_search<PointRange, QPointRange, QQPointRange>::value)
    return b200::qsearchAll<PointRange, QPointRange, QQPointRange, indexType>(Query_Points, Q_Query_Points, QQ_Query_Points, G, Base_Points,
                                                                              Q_Base_Points, QQ_Base_Points, QueryStats, starting_point, QP);
  else
    return qsearchAll<PointRange, QPointRange, QQPointRange, indexType>(Query_Points, Q_Query_Points, QQ_Query_Points, G, Base_Points,
                                                                        Q_Base_Points, QQ_Base_Points, QueryStats, starting_point, QP);
}
template <typename QPointRange, typename QQPointRange, typename indexType>
using b200_knn_index = typename std::conditional<b200_covers_build<QPointRange, QQPointRange>::value, b200::knn_index<QPointRange, QQPointRange, indexType>,
                                                 knn_index<QPointRange, QQPointRange, indexType>>::type;
}  // namespace parlayANN

#define qsearchAll b200_qsearchAll
#define knn_index b200_knn_index
