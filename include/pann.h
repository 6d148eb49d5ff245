/* pann.h — C ABI of libpann_cuda.so: the B200 (sm_100a) implementation of ParlayANN's data-parallel hot path.
 *
 * The reference (landrumb/ParlayANN) is a header-only C++ template library with no FFI layer; this
 * boundary sits directly below its batch-level functions.  Each entry point names the reference
 * interface it replaces (paths relative to the reference root).  INTEGRATION.md shows the binding a
 * maintainer adds on the reference side (C++ header bodies and the pybind/ctypes stub).
 *
 * Conventions: every call returns 0 on success, non-zero on failure; pann_last_error() returns the
 * message of the calling thread's last failure (the reference prints a message and abort()s, the
 * header layer above this ABI keeps that behaviour).  "host" pointers are borrowed for the call;
 * device memory is owned by the pann_index.  One host thread per pann_index; calls are synchronous
 * unless the name ends in _async.  There is no CPU fallback: without a CUDA device every compute
 * call fails with an error.
 */
#ifndef PANN_H_
#define PANN_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum { PANN_F32 = 0, PANN_U8 = 1, PANN_I8 = 2 } pann_dtype;   /* Euclidian_Point<T> / Mips_Point<T> element types */
typedef enum { PANN_L2 = 0, PANN_MIPS = 1 } pann_metric;               /* euclidian_point.h:92-242 / mips_point.h:67-139 */
typedef enum { PANN_GEN_KIND_U8 = 0, PANN_GEN_KIND_I8 = 1, PANN_GEN_KIND_F32 = 2, PANN_GEN_KIND_F32_NORM = 3 } pann_gen_kind_t;

/* = parlayANN::QueryParams (algorithms/utils/types.h:218-231).  limit < 0 means "number of points";
 * degree_limit < 0 means "max degree of the graph". */
typedef struct {
  int64_t k;
  int64_t beam;
  double cut;
  int64_t limit;
  int64_t degree_limit;
} pann_query_params;

typedef struct pann_index pann_index;

/* library / device */
const char* pann_last_error(void);
int pann_device_count(int* count);
const char* pann_version(void);

/* PointRange<Point> + Graph<unsigned> resident on one GPU.
 * host_points = PointRange::location(0) with n rows of row_stride_bytes (point_range.h:42-141);
 * host_graph  = Graph's flat n x (max_degree+1) uint32 array, slot 0 = degree (graph.h:125-250), may be NULL
 * (an empty graph of that max degree is created: Graph(maxDeg, n), graph.h:134-145). */
int pann_index_create(pann_index** out, int device, int dtype, int metric, uint64_t n, uint32_t dims,
                      uint64_t row_stride_bytes, const void* host_points, uint32_t max_degree,
                      const uint32_t* host_graph);
/* same, but the points are synthesised on the device (SURVEY.md §8d generator; no host copy at all) */
int pann_index_create_synthetic(pann_index** out, int device, int dtype, int metric, uint64_t n, uint32_t dims,
                                uint32_t max_degree, int gen_kind, uint64_t seed, uint64_t stream, uint32_t latent_r,
                                uint32_t n_clusters);
/* rows first_row .. first_row+n of the same (counter-based) stream: one base shard per GPU without any host traffic */
int pann_index_create_synthetic_shard(pann_index** out, int device, int dtype, int metric, uint64_t first_row, uint64_t n,
                                      uint32_t dims, uint32_t max_degree, int gen_kind, uint64_t seed, uint64_t stream,
                                      uint32_t latent_r, uint32_t n_clusters);
void pann_index_destroy(pann_index* idx);
int pann_index_set_graph(pann_index* idx, const uint32_t* host_graph);        /* Graph(char* gFile) after load */
int pann_index_get_graph(pann_index* idx, uint32_t* host_graph_out);          /* for Graph::save */
int pann_index_get_points(pann_index* idx, uint64_t row0, uint64_t nrows, void* host_out, uint64_t out_stride_bytes);
int pann_index_info(pann_index* idx, uint64_t* n, uint32_t* dims, uint32_t* max_degree, int* dtype, int* metric);
/* device time (ms, CUDA events on the index's stream) and kernel-launch count of the last compute call */
int pann_index_last_timing(pann_index* idx, float* kernel_ms, int* launches);
/* algorithmic bytes of the last beam-search call (per query cmps*row_bytes + 4*(ids read from graph rows
 * + rows visited) + row_bytes + 4k), summed on the device — the numerator of the HBM-gather roofline */
int pann_index_last_bytes(pann_index* idx, uint64_t* algorithmic_bytes);

/* qsearchAll / searchAll (algorithms/utils/beamSearch.h:537-565, :353-387): one beam search per query
 * (filtered_beam_search, beamSearch.h:22-214) from the given start points; writes the first k ids of each
 * frontier (and their distances), and the per-query counters of stats<indexType> (stats.h:57-94):
 * visited = |visited list|, cmps = full_dist_cmps.  HOST buffers; copies are part of the call.
 * Fails if k > beam (beamSearch.h:549-553) or a query ends with fewer than k results (:416-419). */
int pann_search_batch(pann_index* idx, const void* host_queries, uint64_t nq, uint64_t q_stride_bytes,
                      const uint32_t* starts, uint32_t nstarts, const pann_query_params* qp,
                      uint32_t* out_ids, float* out_dists, uint32_t* out_visited, uint32_t* out_cmps);

/* The same with ONE START LIST PER QUERY: query i starts from starts[i * starts_stride .. + c_i), c_i = nstarts_q[i]
 * (or starts_stride for every query when nstarts_q is NULL).  Replaces beamSearchRandom (beamSearch.h:309-351: one random
 * start per query, see pann_random_starts) and is what HNSW's layer-0 search needs after its per-query descent
 * (HNSW/HNSW.hpp:1570-1613).  A list longer than the beam is allowed when the cut rule cannot fire (k == 0, k == beam, or
 * MIPS): the frontier is then longer than the beam until the first merge truncates it (beamSearch.h:185), which is what
 * the reference does once its unvisited_frontier buffer (:74-76) is made large enough. */
int pann_search_batch_starts(pann_index* idx, const void* host_queries, uint64_t nq, uint64_t q_stride_bytes,
                             const uint32_t* starts, uint32_t starts_stride, const uint32_t* nstarts_q,
                             const pann_query_params* qp, uint32_t* out_ids, float* out_dists, uint32_t* out_visited,
                             uint32_t* out_cmps);
/* beamSearchRandom's start points (beamSearch.h:327-332): parlay::random_generator()[i] through
 * std::uniform_int_distribution<long>(0, n-1), i = 0..nq-1 */
int pann_random_starts(uint64_t nq, uint64_t n, uint32_t* out);

/* Device-resident variant: queries are uploaded once, results stay on the device until fetched.
 * Used to time the search with inputs already in HBM. */
int pann_queries_upload(pann_index* idx, const void* host_queries, uint64_t nq, uint64_t q_stride_bytes);
int pann_queries_synthetic(pann_index* idx, uint64_t nq, int gen_kind, uint64_t seed, uint64_t stream,
                           uint32_t latent_r, uint32_t n_clusters);
int pann_queries_download(pann_index* idx, void* host_out, uint64_t out_stride_bytes);
int pann_search_resident(pann_index* idx, const uint32_t* starts, uint32_t nstarts, const pann_query_params* qp);
int pann_search_fetch(pann_index* idx, uint32_t* out_ids, float* out_dists, uint32_t* out_visited, uint32_t* out_cmps);

/* device pointers to the results of the last pann_search_resident (valid until the next search on this index): the
 * multi-GPU path all-gathers them over NCCL/NVLink without a trip through the host (queries sharded) */
int pann_search_device_result(pann_index* idx, const uint32_t** d_ids, const float** d_dists, const uint32_t** d_visited,
                              const uint32_t** d_cmps, uint64_t* nq, uint32_t* k);

/* beam_search (beamSearch.h:217-223) with both result lists, for parity tests and the build:
 * per query the whole frontier (<= beam entries) and the whole visited list (sorted by (dist,id)).
 * self_ids (may be NULL): query i IS base row self_ids[i] (beam_search_rerank__ in the build,
 * vamana/index.h:249-259) — host_queries is then ignored and same_as() skips that row. */
int pann_search_full(pann_index* idx, const void* host_queries, uint64_t nq, uint64_t q_stride_bytes,
                     const uint32_t* self_ids, const uint32_t* starts, uint32_t nstarts, const pann_query_params* qp,
                     uint32_t fcap, uint32_t* fr_ids, float* fr_dists, uint32_t* n_frontier,
                     uint32_t vcap, uint32_t* vis_ids, float* vis_dists, uint32_t* n_visited, uint32_t* out_cmps);

/* pann_search_full with one start list per query (see pann_search_batch_starts); fcap >= max(beam, starts_stride) */
int pann_search_full_starts(pann_index* idx, const void* host_queries, uint64_t nq, uint64_t q_stride_bytes,
                            const uint32_t* self_ids, const uint32_t* starts, uint32_t starts_stride, const uint32_t* nstarts_q,
                            const pann_query_params* qp, uint32_t fcap, uint32_t* fr_ids, float* fr_dists, uint32_t* n_frontier,
                            uint32_t vcap, uint32_t* vis_ids, float* vis_dists, uint32_t* n_visited, uint32_t* out_cmps);

/* beam_search_rerank when the searched copy is quantised (beamSearch.h:390-454 with differing byte widths; the
 * -quantize_mode 1 path of vamana/neighbors.h:117-127 and Python's quant=True, graph_index.cpp:141-186): q_idx holds the
 * 1-byte points and the graph, base_idx the exact points (same device, same n / dims / metric).  The search runs on
 * q_idx with the quantised queries; the first min(k * rerank_factor, |frontier|) frontier entries are then re-ranked by
 * exact distance between host_queries and base_idx's rows, sorted by (distance, id), first k returned.  Counters are the
 * search's.  HOST buffers.  QueryParams::rerank_factor defaults to 100 in the reference (types.h:224). */
int pann_search_rerank_batch(pann_index* q_idx, pann_index* base_idx, const void* host_qqueries, uint64_t qq_stride_bytes,
                             const void* host_queries, uint64_t q_stride_bytes, uint64_t nq, const uint32_t* starts,
                             uint32_t nstarts, const pann_query_params* qp, int64_t rerank_factor, uint32_t* out_ids,
                             float* out_dists, uint32_t* out_visited, uint32_t* out_cmps);

/* knn_index::build_index (algorithms/vamana/index.h:150-186): num_passes x batch_insert (:188-316) of all n
 * points in parlay::random_permutation order, start point 0, robustPrune (:63-120) with alpha on the last
 * pass and 1.0 before; optional final per-row sort by distance.  The graph of the index is replaced.
 * stats (may be NULL): n u32 each, as stats<indexType> accumulates them during the build.
 * phase_seconds (may be NULL): beam search+prune / bidirect / reverse-edge prune, as index.h:313-315. */
int pann_vamana_build(pann_index* idx, uint32_t R, uint32_t L, double alpha, int num_passes, int sort_neighbors,
                      uint32_t* out_visited, uint32_t* out_cmps, double phase_seconds[3]);

/* knn_index::batch_insert (algorithms/vamana/index.h:188-316): inserts the listed vertices into the index's EXISTING graph —
 * in the given order or in parlay::random_permutation order (random_order, :211-218), batches per :206-209 / :223-234 with
 * this max_fraction, start point 0, robustPrune with `alpha`.  No final sort.  stats (may be NULL): n u32 each, the
 * increments this call adds to stats<indexType>.  Fails like the reference on an id >= n (:193-199). */
int pann_vamana_batch_insert(pann_index* idx, const uint32_t* inserts, uint64_t m, uint32_t R, uint32_t L, double alpha,
                             int random_order, double max_fraction, uint32_t* out_visited, uint32_t* out_cmps,
                             double phase_seconds[3]);

/* accounting of the last pann_vamana_build, counted on the device: algorithmic_bytes = B_ins summed over
 * all insertions = bytes of the build-mode beam searches (as pann_index_last_bytes counts them) + per robustPrune call
 * (vamana/index.h:63-120) C * row bytes + 4 (R+1) with C the candidates after unique + 2 * 4 (R+1) per inserted point;
 * pair_distances = the reference's own distance_comps of those calls */
int pann_index_build_stats(pann_index* idx, uint64_t* algorithmic_bytes, uint64_t* search_bytes, uint64_t* prune_bytes,
                           uint64_t* prune_calls, uint64_t* pair_distances);

/* parlay::random_permutation<int>(n) with the default seed (parlay/random.h:154-159) — the insertion order. */
int pann_random_permutation(int64_t n, int32_t* out);
/* batch boundaries of batch_insert (vamana/index.h:206-209, :223-234); returns the number of batches */
int64_t pann_batch_schedule(uint64_t n, uint64_t m, int64_t* lo, int64_t* hi, int64_t cap);

/* compute_groundtruth (data_tools/compute_groundtruth.cpp:22-60) + write_ibin's per-row sort (:68-73):
 * exact k nearest base rows of every query; rows sorted by (distance, id).  The base rows are those of the
 * index; id_offset is added to the ids (base sharding across GPUs).  HOST query/result buffers. */
int pann_groundtruth(pann_index* idx, const void* host_queries, uint64_t nq, uint64_t q_stride_bytes, uint32_t k,
                     int64_t id_offset, int32_t* out_ids, float* out_dists);
/* same on the resident query set; results stay on the device (pann_groundtruth_fetch) */
int pann_groundtruth_resident(pann_index* idx, uint32_t k, int64_t id_offset);
int pann_groundtruth_fetch(pann_index* idx, int32_t* out_ids, float* out_dists);
/* One leg of a sharded job (the reference's compute_groundtruth has no sharded form; this is SURVEY.md §8e): base rows
 * [row_lo, row_hi) only (row_hi = 0: to the end; row_lo a multiple of 256), keep_state != 0 continues from the top-k the
 * previous leg left on the device, d_tau_cap (device, nq floats, or NULL) = per-query upper bounds of the final k-th distance
 * — e.g. the k-th distance of the shards' merged samples — that tighten the filter, seen_rows = rows that bound was taken
 * over.  With a bound a result row may hold fewer than k entries; the rest read id -1, distance +inf and sort last in
 * pann_groundtruth_merge*.  Results as pann_groundtruth_resident. */
int pann_groundtruth_resident_range(pann_index* idx, uint32_t k, int64_t id_offset, uint64_t row_lo, uint64_t row_hi,
                                    int keep_state, const float* d_tau_cap, uint64_t seen_rows);
/* which path the last ground-truth call took: launches of the tcgen05 kernel (0 = CUDA-core path only) and, cumulative over
 * the index's life, chunks retaken on the CUDA cores after a candidate-list overflow */
int pann_index_groundtruth_stats(pann_index* idx, int* mma_launches, int* fallback_chunks);
/* No counterpart in the reference; added for callers that may not allocate
 * device memory (CUDA graph capture, serving loops).  Exact k-NN of DEVICE-resident query rows (nq rows of the index's
 * element type, q_stride_bytes apart) against the index's base rows into DEVICE buffers d_out_ids / d_out_dists
 * ([nq,k], rows sorted by (distance, id), as pann_groundtruth).  Calls no cudaMalloc / cudaFree and does not
 * synchronise: it only enqueues kernels on `stream` (a cudaStream_t; NULL = the legacy default stream).  All working
 * memory is ONE file-scope device array of 16 MiB per GPU (pann_groundtruth_static_limit), held from library load to
 * exit: a call whose nq * (8 k + padded row bytes + 256) exceeds it fails with a clear message.  Every call on a
 * device shares that array, so calls that overlap on the GPU race; successive calls on one stream are safe. */
int pann_groundtruth_static(pann_index* idx, const void* d_queries, uint64_t nq, uint64_t q_stride_bytes, uint32_t k,
                            int64_t id_offset, int32_t* d_out_ids, float* d_out_dists, void* stream);
int pann_groundtruth_static_limit(uint64_t* scratch_bytes);
/* k-way merge of per-shard results (each [nq, k], sorted) into the global top-k under the (dist,id) order */
int pann_groundtruth_merge(int device, uint32_t nshards, uint64_t nq, uint32_t k, const int32_t* ids, const float* dists,
                           int32_t* out_ids, float* out_dists);

/* device-pointer variants for the multi-GPU path (base rows sharded over ranks, one NCCL all-gather of the local
 * top-k, then the same merge on the device): nothing touches the host.  stream = cudaStream_t (NULL = default). */
int pann_groundtruth_device_result(pann_index* idx, const int32_t** d_ids, const float** d_dists, uint64_t* nq, uint32_t* k);
int pann_groundtruth_merge_device(int device, uint32_t nshards, uint64_t nq, uint32_t k, const int32_t* d_ids, const float* d_dists,
                                  int32_t* d_out_ids, float* d_out_dists, void* stream);

/* ---- the evaluation harness on the device: search_and_parse's loop (algorithms/utils/check_nn_recall.h:181-268).  For every
 * QueryParams: the search over the RESIDENT query set (pann_queries_upload / _synthetic), tie-aware recall against the resident
 * ground truth (checkRecall, :83-109: the first k ids plus every later id at the k-th distance; pann_groundtruth_resident
 * leaves one, pann_groundtruth_upload sets one from a .ibin), and stats<indexType>::statistics of both counters (stats.h:84-92).
 * One copy brings all results back.  status != 0 marks a setting the reference would abort on (k > beam, fewer than k found). */
typedef struct {
  double recall;
  uint32_t avg_cmps, tail_cmps, avg_visited, tail_visited;
  float device_ms;   /* kernel time of the setting's search (CUDA events) */
  int status;
} pann_sweep_result;
int pann_groundtruth_upload(pann_index* idx, const int32_t* gt_ids, const float* gt_dists, uint64_t nq, uint32_t gt_k);
int pann_search_sweep(pann_index* idx, const uint32_t* starts, uint32_t nstarts, const pann_query_params* qps, uint32_t nset,
                      pann_sweep_result* out);

/* ---- HNSW (algorithms/HNSW/HNSW.hpp).  Node i = point i of the index; the index must have max_degree = 2 m: its graph IS
 * level 0 of the model (degree bound 2m, HNSW.hpp:527-531); the upper levels (bound m) live in the pann_hnsw object as
 * compacted flat arrays.  The model moves in and out in the order of the model file v3 (HNSW.hpp:1655-1685, :690-740):
 * level[n]; for every node, for l = 0..level[v], a count (counts[]) and that many node ids (edges[]); the entrances. */
typedef struct pann_hnsw pann_hnsw;
int pann_hnsw_create(pann_hnsw** out, pann_index* idx, uint32_t m, uint32_t ef_construction, float m_l, float alpha);
void pann_hnsw_destroy(pann_hnsw* h);
/* get_level_random (HNSW.hpp:505-516) as one worker draws it: mt19937{0}, uniform_real_distribution<>(FLT_MIN, 1) */
int pann_hnsw_levels(uint64_t n, float m_l, uint32_t* out);
/* HNSW(begin, end, dim, m_l, m, ef_construction, alpha) (HNSW.hpp:746-824) = batched insert() (:828-1014) of nodes 1..n-1.
 * levels (may be NULL = pann_hnsw_levels): the level of every node.  seconds (may be NULL): total, layer searches,
 * prune_heuristic + forward edges, reverse edges. */
int pann_hnsw_build(pann_hnsw* h, const uint32_t* levels, double seconds[4]);
int pann_hnsw_set_model(pann_hnsw* h, const uint32_t* level, const uint64_t* counts, const uint32_t* edges, const uint32_t* entrance,
                        uint64_t n_entrances);
/* sizes = {#counts, #edges, #entrances}; call with NULL arrays first to size the buffers */
int pann_hnsw_get_model(pann_hnsw* h, uint32_t* level, uint64_t* counts, uint32_t* edges, uint32_t* entrance, uint64_t sizes[3]);
/* HNSW::search (HNSW.hpp:1570-1613) per query: ef = 1 descent from the entrances (search_layer_to, :1545-1568), search_layer(ef)
 * on level 0 with visit limit `limit` (<= 0: none), the answer cut to k plus the entries tied with the k-th distance.
 * HOST buffers; out_ids / out_dists: ocap entries per query (k <= ocap <= ef), out_cnt: entries the reference returns,
 * out_cmps: distance comparisons summed over the layers (search_control::count_cmps). */
int pann_hnsw_search(pann_hnsw* h, const void* host_queries, uint64_t nq, uint64_t q_stride_bytes, uint32_t k, uint32_t ef,
                     int64_t limit, uint32_t ocap, uint32_t* out_ids, float* out_dists, uint32_t* out_cnt, uint32_t* out_cmps);
int pann_hnsw_info(pann_hnsw* h, uint32_t* top_level, uint64_t* upper_rows, uint32_t* n_entrances, double seconds[4]);

/* pinned host memory for callers that want asynchronous copies */
int pann_host_alloc(void** p, size_t bytes);
int pann_host_free(void* p);

/* ---- point conversions of the reference's Python builder / index loader, on the device (host arrays in and out) ----
 * pann_rows_normalize: Mips_Point<float>::normalize over every row, in place (utils/mips_point.h:115-124; called by
 *   python/builder.cpp:45-49 and python/graph_index.cpp:88-90).
 * pann_rows_quant_params_u8 / pann_rows_quantize_u8: Euclidian_Point<uint8_t>::generate_parameters + parameters(min, max, dims)
 *   and translate_point (utils/euclidian_point.h:211-237, :104-108, :183-209): the 1-byte copy of float Euclidean data.
 * pann_rows_quant_params_mips8 / pann_rows_quantize_mips8: Quantized_Mips_Point<8,trim>::generate_parameters and
 *   translate_point (utils/mips_point.h:437-462, :421-434): the int8 copy of normalised float MIPS data.
 * rows: n rows of dims floats, stride_bytes apart; outputs: n rows of dims bytes, out_stride_bytes apart. */
int pann_rows_normalize(int device, float* rows, uint64_t n, uint32_t dims, uint64_t stride_bytes);
int pann_rows_quant_params_u8(int device, const float* rows, uint64_t n, uint32_t dims, uint64_t stride_bytes, float* slope, int64_t* offset);
int pann_rows_quantize_u8(int device, const float* rows, uint64_t n, uint32_t dims, uint64_t stride_bytes, float slope, int64_t offset,
                          uint8_t* out, uint64_t out_stride_bytes);
int pann_rows_quant_params_mips8(int device, const float* rows, uint64_t n, uint32_t dims, uint64_t stride_bytes, int trim, float* max_val);
int pann_rows_quantize_mips8(int device, const float* rows, uint64_t n, uint32_t dims, uint64_t stride_bytes, float max_val,
                             int8_t* out, uint64_t out_stride_bytes);

#ifdef __cplusplus
}
#endif
#endif /* PANN_H_ */
