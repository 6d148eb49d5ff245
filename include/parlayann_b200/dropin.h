// parlayann_b200/dropin.h — the reference-side C++ binding of libpann_cuda.so (include/pann.h).
//
// Header-only; compile it INSIDE the reference tree (it uses the reference's own types, nothing is re-declared):
//     #include "utils/beamSearch.h"   #include "vamana/index.h"   #include "parlayann_b200/dropin.h"
//     g++ ... -I<reference>/algorithms -I<reference>/parlaylib/include -I<this repo>/include \
//         -L<this repo>/parlayann_b200 -lpann_cuda -Wl,-rpath,<this repo>/parlayann_b200
// It gives the batch-level entry points of the hot path the same names, template parameters, argument order, return
// types and error behaviour (message on std::cout, then abort()) as the reference, in namespace parlayANN::b200:
//     qsearchAll   (algorithms/utils/beamSearch.h:537-565)      searchAll (:353-387)
//     knn_index<PR,QPR,idx>::build_index (algorithms/vamana/index.h:150-186)
//     compute_groundtruth (data_tools/compute_groundtruth.cpp:22-60)
// A maintainer switches a call site by adding `b200::`.  The index (points + graph) is kept
// resident on the GPU in a DeviceIndex; the reference-signature overloads look it up in a small cache keyed by the
// PointRange's buffer address, so repeated calls (the search sweep of check_nn_recall.h:214-257) upload once.
#pragma once
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <type_traits>
#include <utility>
#include <vector>

#include "../pann.h"

namespace parlayANN {
namespace b200 {

// element type / metric of the reference's point classes (euclidian_point.h:92-242, mips_point.h:67-139)
template <typename Point> struct point_traits { static constexpr bool supported = false; };
#define PANN_B200_TRAIT(P, T, DT, M)                                                     \
  template <> struct point_traits<P<T>> {                                                \
    static constexpr bool supported = true;                                              \
    static constexpr int dtype = DT, metric = M;                                         \
    using elt = T;                                                                       \
  };
PANN_B200_TRAIT(Euclidian_Point, float, PANN_F32, PANN_L2)
PANN_B200_TRAIT(Euclidian_Point, uint8_t, PANN_U8, PANN_L2)
PANN_B200_TRAIT(Euclidian_Point, int8_t, PANN_I8, PANN_L2)
PANN_B200_TRAIT(Mips_Point, float, PANN_F32, PANN_MIPS)
PANN_B200_TRAIT(Mips_Point, uint8_t, PANN_U8, PANN_MIPS)
PANN_B200_TRAIT(Mips_Point, int8_t, PANN_I8, PANN_MIPS)
#undef PANN_B200_TRAIT
// Quantized_Mips_Point<8, trim> (mips_point.h:288-462): rows of int8, distance = -(int8 inner product) (:339-346) — the
// int8 MIPS kernels compute exactly that (values below 2^24, so the reference's int64 and the kernels' float agree).
template <bool trim, int range> struct point_traits<Quantized_Mips_Point<8, trim, range>> {
  static constexpr bool supported = true;
  static constexpr int dtype = PANN_I8, metric = PANN_MIPS;
  using elt = int8_t;
};

inline void check(int rc) {  // the reference's error convention: message, then abort()
  if (rc != 0) { std::cout << pann_last_error() << std::endl; abort(); }
}

template <typename PointRange> inline size_t row_stride(const PointRange& P) {
  // PointRange::aligned_bytes is private (point_range.h:139); a range of one row has no second row to measure against
  using T = typename point_traits<typename PointRange::Point>::elt;
  return P.size() > 1 ? (size_t)(P.location(1) - P.location(0)) : (size_t)P.dimension() * sizeof(T);
}
template <typename indexType> inline indexType* graph_base(const Graph<indexType>& G) {
  return G[0].begin() - 1;  // flat n x (maxDeg+1) array, slot 0 = degree (graph.h:134-141, :234-242)
}

// PointRange + Graph resident on one GPU (RAII around pann_index)
template <typename PointRange, typename indexType = unsigned int>
class DeviceIndex {
  static_assert(point_traits<typename PointRange::Point>::supported,
                "parlayann_b200 covers Euclidian_Point / Mips_Point over float, uint8_t, int8_t and Quantized_Mips_Point<8> (SURVEY.md §8)");
  static_assert(sizeof(indexType) == 4, "graph ids are 32-bit");
  using T = point_traits<typename PointRange::Point>;

 public:
  DeviceIndex(const PointRange& Points, long max_degree, const Graph<indexType>* G = nullptr, int device = 0) {
    check(pann_index_create(&h_, device, T::dtype, T::metric, Points.size(), Points.dimension(), row_stride(Points),
                            Points.location(0), (uint32_t)max_degree, G ? (const uint32_t*)graph_base(*G) : nullptr));
  }
  DeviceIndex(const PointRange& Points, const Graph<indexType>& G, int device = 0)
      : DeviceIndex(Points, G.max_degree(), &G, device) {}
  DeviceIndex(const DeviceIndex&) = delete;
  DeviceIndex& operator=(const DeviceIndex&) = delete;
  ~DeviceIndex() { pann_index_destroy(h_); }
  void upload_graph(const Graph<indexType>& G) { check(pann_index_set_graph(h_, (const uint32_t*)graph_base(G))); }
  void download_graph(Graph<indexType>& G) { check(pann_index_get_graph(h_, (uint32_t*)graph_base(G))); }
  pann_index* handle() const { return h_; }

 private:
  pann_index* h_ = nullptr;
};

inline pann_query_params to_pann(const QueryParams& QP) {  // types.h:218-231
  pann_query_params q;
  q.k = QP.k; q.beam = QP.beamSize; q.cut = QP.cut;
  // the ABI reads a negative limit as "all"; the reference visits / reads nothing then (beamSearch.h:107, :130)
  q.limit = QP.limit < 0 ? 0 : QP.limit; q.degree_limit = QP.degree_limit < 0 ? 0 : QP.degree_limit;
  return q;
}

// ---- resident-index cache for the reference-signature overloads ----------------------------------------------------
// Keyed by the PointRange's buffer address AND validated on every call: n, dimension, max degree, and a fingerprint of
// the contents (every 61st point row's first bytes; every graph row's degree word and every 61st row in full — about 1 %
// of the bytes, ~10 ms at 10M points).  A buffer that was freed and re-used by another PointRange / Graph, or a graph that
// was rebuilt in place, therefore gets a fresh device copy.  An in-place edit that keeps all degrees and misses every
// sampled row is the caller's to announce: b200::invalidate(Points) drops the entry, b200::clear_cache() drops all
// (they also return the GPU memory).
inline uint64_t fp_mix(uint64_t h, uint64_t x) { h ^= x + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2); return h; }
template <typename PointRange> inline uint64_t points_fingerprint(const PointRange& P) {
  uint64_t h = fp_mix(P.size(), P.dimension());
  const size_t n = P.size(), rb = std::min<size_t>(16, (size_t)P.dimension());
  for (size_t i = 0; i < n; i += 61) { uint64_t w[2] = {0, 0}; memcpy(w, P.location(i), rb); h = fp_mix(fp_mix(h, w[0]), w[1]); }
  return h;
}
template <typename indexType> inline uint64_t graph_fingerprint(const Graph<indexType>& G) {
  uint64_t h = fp_mix(G.size(), G.max_degree());
  const size_t n = G.size(), w = (size_t)G.max_degree() + 1;
  const indexType* g = graph_base(G);
  for (size_t i = 0; i < n; i++) h = fp_mix(h, g[i * w]);
  for (size_t i = 0; i < n; i += 61) for (size_t j = 1; j <= g[i * w] && j < w; j++) h = fp_mix(h, g[i * w + j]);
  return h;
}
struct CacheBase {   // type-erased registry so that invalidate / clear_cache reach every instantiation's map
  virtual void drop(const void* key) = 0;
  virtual void clear() = 0;
  static std::vector<CacheBase*>& all() { static std::vector<CacheBase*> v; return v; }
  static std::mutex& mu() { static std::mutex m; return m; }
};
template <typename PointRange, typename indexType>
struct IndexCache : CacheBase {
  using DI = DeviceIndex<PointRange, indexType>;
  struct Entry { std::unique_ptr<DI> di; size_t n; unsigned dims; long maxdeg; uint64_t pfp, gfp; bool has_graph; };
  std::map<const void*, Entry> map;
  IndexCache() { CacheBase::all().push_back(this); }
  void drop(const void* key) override { map.erase(key); }
  void clear() override { map.clear(); }
  static IndexCache& get() { static IndexCache c; return c; }
};
template <typename PointRange> inline void invalidate(const PointRange& Points) {
  std::lock_guard<std::mutex> lock(CacheBase::mu());
  if (Points.size() == 0) return;
  for (CacheBase* c : CacheBase::all()) c->drop(Points.location(0));
}
inline void clear_cache() {
  std::lock_guard<std::mutex> lock(CacheBase::mu());
  for (CacheBase* c : CacheBase::all()) c->clear();
}
template <typename PointRange, typename indexType>
inline DeviceIndex<PointRange, indexType>& cached_index(const PointRange& Points, const Graph<indexType>& G, bool graph_dirty) {
  using IC = IndexCache<PointRange, indexType>;
  std::lock_guard<std::mutex> lock(CacheBase::mu());
  IC& C = IC::get();
  const void* key = Points.location(0);
  const uint64_t pfp = points_fingerprint(Points), gfp = graph_fingerprint(G);
  auto it = C.map.find(key);
  if (it == C.map.end() || it->second.n != Points.size() || it->second.dims != Points.dimension() || it->second.maxdeg != G.max_degree() ||
      it->second.pfp != pfp || !it->second.has_graph) {
    C.map.erase(key);
    C.map[key] = typename IC::Entry{std::make_unique<typename IC::DI>(Points, G), Points.size(), (unsigned)Points.dimension(), G.max_degree(), pfp, gfp, true};
    return *C.map[key].di;
  }
  if (graph_dirty || it->second.gfp != gfp) { it->second.di->upload_graph(G); it->second.gfp = gfp; }
  return *it->second.di;
}

// ---- searchAll (beamSearch.h:353-387) on a resident index -------------------------------------------------------
template <typename PointRange, typename indexType>
parlay::sequence<parlay::sequence<indexType>>
searchAll(DeviceIndex<PointRange, indexType>& D, const PointRange& Query_Points, stats<indexType>& QueryStats,
          const parlay::sequence<indexType>& starting_points, const QueryParams& QP) {
  if (QP.k > QP.beamSize) {  // beamSearch.h:368-372 / :549-553
    std::cout << "Error: beam search parameter Q = " << QP.beamSize << " same size or smaller than k = " << QP.k << std::endl;
    abort();
  }
  const size_t nq = Query_Points.size();
  const size_t k = (size_t)QP.k;
  std::vector<uint32_t> ids(nq * k), visited(nq), cmps(nq);
  pann_query_params q = to_pann(QP);
  if (nq)
    check(pann_search_batch(D.handle(), Query_Points.location(0), nq, row_stride(Query_Points), (const uint32_t*)starting_points.begin(),
                            (uint32_t)starting_points.size(), &q, ids.data(), nullptr, visited.data(), cmps.data()));
  parlay::sequence<parlay::sequence<indexType>> all_neighbors(nq);
  parlay::parallel_for(0, nq, [&](size_t i) {
    all_neighbors[i] = parlay::tabulate(k, [&](size_t j) { return (indexType)ids[i * k + j]; });
    QueryStats.increment_visited(i, visited[i]);  // stats.h:70-73
    QueryStats.increment_dist(i, cmps[i]);
  });
  return all_neighbors;
}

// reference signature (beamSearch.h:362-367)
template <typename PointRange, typename indexType>
parlay::sequence<parlay::sequence<indexType>>
searchAll(PointRange& Query_Points, Graph<indexType>& G, PointRange& Base_Points, stats<indexType>& QueryStats,
          parlay::sequence<indexType> starting_points, QueryParams& QP) {
  return searchAll(cached_index(Base_Points, G, false), Query_Points, QueryStats, starting_points, QP);
}
template <typename PointRange, typename indexType>
parlay::sequence<parlay::sequence<indexType>>
searchAll(PointRange& Query_Points, Graph<indexType>& G, PointRange& Base_Points, stats<indexType>& QueryStats,
          indexType starting_point, QueryParams& QP) {
  parlay::sequence<indexType> start_points = {starting_point};
  return searchAll(cached_index(Base_Points, G, false), Query_Points, QueryStats, start_points, QP);
}

// exact points only (the re-rank side of a quantised search): no graph on the device; same validation as cached_index
template <typename PointRange, typename indexType>
inline DeviceIndex<PointRange, indexType>& cached_points(const PointRange& Points) {
  using DI = DeviceIndex<PointRange, indexType>;
  struct PC : CacheBase {
    struct Entry { std::unique_ptr<DI> di; size_t n; unsigned dims; uint64_t pfp; };
    std::map<const void*, Entry> map;
    PC() { CacheBase::all().push_back(this); }
    void drop(const void* key) override { map.erase(key); }
    void clear() override { map.clear(); }
  };
  static PC C;
  std::lock_guard<std::mutex> lock(CacheBase::mu());
  const void* key = Points.location(0);
  const uint64_t pfp = points_fingerprint(Points);
  auto it = C.map.find(key);
  if (it == C.map.end() || it->second.n != Points.size() || it->second.dims != Points.dimension() || it->second.pfp != pfp) {
    C.map.erase(key);
    C.map[key] = typename PC::Entry{std::make_unique<DI>(Points, 1), Points.size(), (unsigned)Points.dimension(), pfp};
    return *C.map[key].di;
  }
  return *it->second.di;
}

// ---- qsearchAll (beamSearch.h:537-565), reference signature.
//  * quantised ranges of the same type as the exact ones: beam_search_rerank reduces to filtered_beam_search
//    (:409-410, :445-452) -> pann_search_batch;
//  * a 1-byte quantised copy for the search and its first-level filter (Q == QQ, the -quantize_mode 1 configuration of
//    vamana/neighbors.h:117-127): search on the quantised index, exact re-rank of the first k * rerank_factor frontier
//    entries (:426-444) -> pann_search_rerank_batch.  A second, coarser level (QQ != Q, use_filtering) is not built.
template <typename PointRange, typename QPointRange, typename QQPointRange, typename indexType>
parlay::sequence<parlay::sequence<indexType>>
qsearchAll(const PointRange& Query_Points, const QPointRange& Q_Query_Points, const QQPointRange&, const Graph<indexType>& G,
           const PointRange& Base_Points, const QPointRange& Q_Base_Points, const QQPointRange&, stats<indexType>& QueryStats,
           const indexType starting_point, const QueryParams& QP) {
  static_assert(std::is_same<QPointRange, QQPointRange>::value,
                "parlayann_b200: two-level quantised search (QQ != Q, use_filtering) is not on the GPU path (SURVEY.md §8f.2)");
  parlay::sequence<indexType> starting_points = {starting_point};
  if constexpr (std::is_same<PointRange, QPointRange>::value) {
    return searchAll(cached_index(Base_Points, G, false), Query_Points, QueryStats, starting_points, QP);
  } else {
    if (QP.k > QP.beamSize) {  // beamSearch.h:549-553
      std::cout << "Error: beam search parameter Q = " << QP.beamSize << " same size or smaller than k = " << QP.k << std::endl;
      abort();
    }
    auto& DQ = cached_index(Q_Base_Points, G, false);
    auto& DB = cached_points<PointRange, indexType>(Base_Points);
    const size_t nq = Query_Points.size();
    const size_t k = (size_t)QP.k;
    std::vector<uint32_t> ids(nq * k), visited(nq), cmps(nq);
    pann_query_params q = to_pann(QP);
    if (nq)
      check(pann_search_rerank_batch(DQ.handle(), DB.handle(), Q_Query_Points.location(0), row_stride(Q_Query_Points),
                                     Query_Points.location(0), row_stride(Query_Points), nq, (const uint32_t*)starting_points.begin(), 1, &q,
                                     QP.rerank_factor, ids.data(), nullptr, visited.data(), cmps.data()));
    parlay::sequence<parlay::sequence<indexType>> all_neighbors(nq);
    parlay::parallel_for(0, nq, [&](size_t i) {
      all_neighbors[i] = parlay::tabulate(k, [&](size_t j) { return (indexType)ids[i * k + j]; });
      QueryStats.increment_visited(i, visited[i]);  // beamSearch.h:421-424
      QueryStats.increment_dist(i, cmps[i]);
    });
    return all_neighbors;
  }
}

// ---- knn_index::build_index (vamana/index.h:150-186) -------------------------------------------------------------
// Same class shape as the reference's knn_index: construct with BuildParams, call build_index(G, Points, QPoints,
// BuildStats, sort_neighbors).  The graph is built on the GPU and copied into G; BuildStats receives the same
// per-point counters the reference accumulates (index.h:260-265, :298).
template <typename PointRange, typename QPointRange, typename indexType>
struct knn_index {
  using GraphI = Graph<indexType>;
  BuildParams BP;
  indexType start_point = 0;
  double phase_seconds[3] = {0, 0, 0};  // beam search + prune / bidirect / reverse-edge prune (index.h:313-315)
  knn_index(BuildParams& BP) : BP(BP) {}
  indexType get_start() { return start_point; }
  void set_start() { start_point = 0; }

  // batch_insert (vamana/index.h:188-192): the listed vertices into the existing graph G (uploaded, extended on the GPU,
  // copied back); same argument order and defaults.  `base` must be 2 (the only value the reference's callers use).
  void batch_insert(parlay::sequence<indexType>& inserts, GraphI& G, PointRange& Points, QPointRange&, stats<indexType>& BuildStats,
                    double alpha, bool random_order = false, double base = 2, double max_fraction = .02, bool print = true) {
    static_assert(std::is_same<PointRange, QPointRange>::value, "parlayann_b200: batch_insert takes the same range for Points and QPoints");
    if (base != 2) { std::cout << "parlayann_b200: batch_insert supports base = 2 only" << std::endl; abort(); }
    (void)print;
    const size_t n = Points.size();
    DeviceIndex<PointRange, indexType> D(Points, G);
    std::vector<uint32_t> vis(n), cmp(n);
    check(pann_vamana_batch_insert(D.handle(), (const uint32_t*)inserts.begin(), inserts.size(), (uint32_t)BP.R, (uint32_t)BP.L, alpha,
                                   random_order ? 1 : 0, max_fraction, vis.data(), cmp.data(), phase_seconds));
    D.download_graph(G);
    invalidate(Points);   // any resident copy of this graph is stale now
    parlay::parallel_for(0, n, [&](size_t i) { BuildStats.increment_visited(i, vis[i]); BuildStats.increment_dist(i, cmp[i]); });
  }

  void build_index(GraphI& G, PointRange& Points, QPointRange&, stats<indexType>& BuildStats, bool sort_neighbors = true) {
    // a quantised build passes the quantised range twice (bench/neighborsTime.C:157-164, python/builder.cpp:60-87)
    static_assert(std::is_same<PointRange, QPointRange>::value,
                  "parlayann_b200: build_index takes the same range for Points and QPoints");
    if (BP.single_batch != 0) { std::cout << "parlayann_b200: single_batch builds are not supported" << std::endl; abort(); }
    std::cout << "Building graph..." << std::endl;
    set_start();
    const size_t n = Points.size();
    DeviceIndex<PointRange, indexType> D(Points, G.max_degree());
    std::vector<uint32_t> vis(n), cmp(n);
    std::cout << "number of passes = " << BP.num_passes << std::endl;
    check(pann_vamana_build(D.handle(), (uint32_t)BP.R, (uint32_t)BP.L, BP.alpha, BP.num_passes, sort_neighbors ? 1 : 0,
                            vis.data(), cmp.data(), phase_seconds));
    D.download_graph(G);
    invalidate(Points);   // any resident copy of an earlier graph over these points is stale now
    parlay::parallel_for(0, n, [&](size_t i) { BuildStats.increment_visited(i, vis[i]); BuildStats.increment_dist(i, cmp[i]); });
    // the three phase timers as parlay::internal::timer::total() prints them (index.h:313-315; get_time.h:23-32), including
    // its side effect: the stream's precision stays at 4 afterwards, which the recall figures of the sweep are printed with
    auto report = [](const char* name, double seconds) {
      std::ios::fmtflags cout_settings = std::cout.flags();
      std::cout.precision(4);
      std::cout << std::fixed;
      std::cout << name << ": total: " << seconds << std::endl;
      std::cout.flags(cout_settings);
    };
    report("beam search time", phase_seconds[0]);
    report("bidirect time", phase_seconds[1]);
    report("prune time", phase_seconds[2]);
  }
};

// ---- compute_groundtruth (data_tools/compute_groundtruth.cpp:22-60) -----------------------------------------------
// Returns the same structure (per query k pairs (int id, float dist)), already sorted by (distance, id), so the
// reference's write_ibin (:64-100) can be applied to it unchanged.
template <typename PointRange>
parlay::sequence<parlay::sequence<std::pair<int, float>>> compute_groundtruth(PointRange& B, PointRange& Q, int k) {
  const size_t nq = Q.size();
  DeviceIndex<PointRange, unsigned int> D(B, 1);
  std::vector<int32_t> ids(nq * (size_t)k);
  std::vector<float> ds(nq * (size_t)k);
  if (nq) check(pann_groundtruth(D.handle(), Q.location(0), nq, row_stride(Q), (uint32_t)k, 0, ids.data(), ds.data()));
  auto answers = parlay::tabulate(nq, [&](size_t i) {
    return parlay::tabulate((size_t)k, [&](size_t j) { return std::make_pair((int)ids[i * k + j], ds[i * k + j]); });
  });
  std::cout << "Done computing groundtruth" << std::endl;
  return answers;
}

}  // namespace b200
}  // namespace parlayANN
