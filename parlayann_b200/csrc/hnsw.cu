// hnsw.cu — HNSW on one B200: layered graph held as flat fixed-degree arrays, search = ef-1 descent + layer-0 beam search,
// build = the reference's batched insert (algorithms/HNSW/HNSW.hpp:746-1014) with every per-node step a batched kernel.
//
// Layout: level 0 is the pann_index's own graph, n x (2m + 1) words (slot 0 = degree), so every layer-0 search is the plain
// K1 / K1s launch.  The upper layers are compacted: only vertices with level >= 1 own a row; rowmap[v] is that row (the
// same for every upper layer), layer l (1-based) is rows x (m + 1) words.  The reference keeps a heap-allocated
// sequence per (node, level) (HNSW.hpp:86-91); the model file (v3) is converted on the way in and out.
//
// Layer search = filtered_beam_search with QP(ef, ef, 1.35, limit, cap(l)) (HNSW.hpp:1088-1109) = K1 / K1s with per-query
// start lists and the row map (beam_search.cuh).  prune_heuristic (:430-486) and the reverse-edge trim (:963-987) are K3
// with rule 1 / 2 (vamana_prune.cuh); reverse edges are grouped by the same radix sort as the Vamana build.
#include <math.h>
#include <stdio.h>

#include <algorithm>
#include <chrono>
#include <limits>
#include <random>
#include <vector>

#include "pann_internal.h"
#include "vamana_build.h"

struct pann_hnsw {
  pann_index* idx = nullptr;   // points + level-0 graph (max degree 2m); borrowed
  uint32_t m = 0, efc = 0;
  float m_l = 0, alpha = 0;
  std::vector<uint32_t> level;     // host copy, n entries
  std::vector<uint32_t> entrance;  // host
  uint32_t top = 0;                // highest level with a graph allocated
  uint64_t rows = 0;               // vertices with level >= 1
  DevBuf rowmap, up;               // rowmap: n words; up: top x rows x (m+1) words
  DevBuf work[12];
  double last_seconds[4] = {0, 0, 0, 0};
  uint32_t* layer(uint32_t l) const { return reinterpret_cast<uint32_t*>(up.p) + (uint64_t)(l - 1) * rows * (m + 1); }
};

static inline uint32_t al16u(size_t x) { return (uint32_t)((x + 15) / 16 * 16); }

// get_level_random (HNSW.hpp:505-516) as ONE worker draws it: thread_local mt19937{worker_id = 0} through
// uniform_real_distribution<>(FLT_MIN, 1.0), level = uint32(-log(x) * m_l), in node order.  (With several workers the
// reference's assignment depends on the scheduler; this is its single-thread sequence, so a GPU build reproduces the
// model file the reference writes with PARLAY_NUM_THREADS=1.)
extern "C" int pann_hnsw_levels(uint64_t n, float m_l, uint32_t* out) {
  if (!out) PANN_FAIL("pann_hnsw_levels: null argument");
  std::mt19937 gen{0};
  std::uniform_real_distribution<> dis(std::numeric_limits<float>::min(), 1.0);
  for (uint64_t i = 0; i < n; i++) out[i] = (uint32_t)(-log(dis(gen)) * m_l);
  return 0;
}

extern "C" int pann_hnsw_create(pann_hnsw** out, pann_index* idx, uint32_t m, uint32_t efc, float m_l, float alpha) {
  if (!out || !idx) PANN_FAIL("pann_hnsw_create: null argument");
  if (m == 0 || idx->maxdeg != 2 * m) PANN_FAIL("pann_hnsw_create: the index must be created with max_degree = 2 m (level 0 of the model, HNSW.hpp:527-531)");
  if (efc == 0) PANN_FAIL("pann_hnsw_create: ef_construction must be positive");
  pann_hnsw* h = new pann_hnsw();
  h->idx = idx; h->m = m; h->efc = efc; h->m_l = m_l; h->alpha = alpha;
  *out = h;
  return 0;
}

extern "C" void pann_hnsw_destroy(pann_hnsw* h) {
  if (!h) return;
  cudaSetDevice(h->idx->device);
  cudaStreamSynchronize(h->idx->stream);
  h->rowmap.release(); h->up.release();
  for (auto& b : h->work) b.release();
  delete h;
}

// levels known -> row map and empty upper layers
static int hnsw_alloc_layers(pann_hnsw* h) {
  pann_index* idx = h->idx;
  const uint64_t n = idx->n;
  std::vector<uint32_t> rm(n, 0xffffffffu);
  uint64_t rows = 0;
  uint32_t top = 0;
  for (uint64_t v = 0; v < n; v++) {
    if (h->level[v] >= 1) rm[v] = (uint32_t)rows++;
    top = std::max(top, h->level[v]);
  }
  h->rows = rows; h->top = top;
  if (h->rowmap.reserve(4 * n)) return 1;
  PANN_CUDA(cudaMemcpyAsync(h->rowmap.p, rm.data(), 4 * n, cudaMemcpyHostToDevice, idx->stream));
  const size_t ub = (size_t)top * rows * (h->m + 1) * 4;
  if (h->up.reserve(ub + 16)) return 1;
  if (ub) PANN_CUDA(cudaMemsetAsync(h->up.p, 0, ub, idx->stream));
  PANN_CUDA(cudaStreamSynchronize(idx->stream));
  return 0;
}

// ---- model in the file's order (HNSW.hpp:1655-1685): level[n]; for every node, for l = 0..level, a count and the ids
extern "C" int pann_hnsw_set_model(pann_hnsw* h, const uint32_t* level, const uint64_t* counts, const uint32_t* edges,
                                   const uint32_t* entrance, uint64_t nent) {
  if (!h || !level || !counts || (!entrance && nent)) PANN_FAIL("pann_hnsw_set_model: null argument");
  pann_index* idx = h->idx;
  PANN_CUDA(cudaSetDevice(idx->device));
  const uint64_t n = idx->n;
  if (nent == 0) PANN_FAIL("pann_hnsw_set_model: the model has no entrance");
  for (uint64_t i = 0; i < nent; i++) if (entrance[i] >= n) PANN_FAIL("pann_hnsw_set_model: entrance out of range");
  h->level.assign(level, level + n);
  h->entrance.assign(entrance, entrance + nent);
  if (hnsw_alloc_layers(h)) return 1;
  std::vector<uint32_t> rm(n);
  PANN_CUDA(cudaMemcpy(rm.data(), h->rowmap.p, 4 * n, cudaMemcpyDeviceToHost));
  const uint32_t m = h->m, w0 = 2 * m + 1, w1 = m + 1;
  std::vector<uint32_t> g0((size_t)n * w0, 0u), gu((size_t)h->top * h->rows * w1, 0u);
  uint64_t ci = 0, ei = 0;
  for (uint64_t v = 0; v < n; v++)
    for (uint32_t l = 0; l <= level[v]; l++) {
      const uint64_t c = counts[ci++];
      const uint32_t cap = l == 0 ? 2 * m : m;
      if (c > cap) PANN_FAIL("pann_hnsw_set_model: a neighbour list is longer than the level's degree bound (2m at level 0, m above)");
      uint32_t* row = l == 0 ? &g0[v * w0] : &gu[((uint64_t)(l - 1) * h->rows + rm[v]) * w1];
      row[0] = (uint32_t)c;
      for (uint64_t j = 0; j < c; j++) {
        const uint32_t x = edges[ei++];
        if (x >= n || h->level[x] < l) PANN_FAIL("pann_hnsw_set_model: an edge points to a vertex that does not exist on that level");
        row[1 + j] = x;
      }
    }
  PANN_CUDA(cudaMemcpy(idx->graph, g0.data(), g0.size() * 4, cudaMemcpyHostToDevice));
  if (!gu.empty()) PANN_CUDA(cudaMemcpy(h->up.p, gu.data(), gu.size() * 4, cudaMemcpyHostToDevice));
  return 0;
}

extern "C" int pann_hnsw_get_model(pann_hnsw* h, uint32_t* level, uint64_t* counts, uint32_t* edges, uint32_t* entrance, uint64_t sizes[3]) {
  if (!h || !sizes) PANN_FAIL("pann_hnsw_get_model: null argument");
  pann_index* idx = h->idx;
  PANN_CUDA(cudaSetDevice(idx->device));
  const uint64_t n = idx->n;
  if (h->level.size() != n) PANN_FAIL("pann_hnsw_get_model: no model (build or set one first)");
  const uint32_t m = h->m, w0 = 2 * m + 1, w1 = m + 1;
  std::vector<uint32_t> g0((size_t)n * w0), gu((size_t)h->top * h->rows * w1), rm(n);
  PANN_CUDA(cudaMemcpy(g0.data(), idx->graph, g0.size() * 4, cudaMemcpyDeviceToHost));
  if (!gu.empty()) PANN_CUDA(cudaMemcpy(gu.data(), h->up.p, gu.size() * 4, cudaMemcpyDeviceToHost));
  PANN_CUDA(cudaMemcpy(rm.data(), h->rowmap.p, 4 * n, cudaMemcpyDeviceToHost));
  uint64_t ci = 0, ei = 0;
  for (uint64_t v = 0; v < n; v++) {
    if (level) level[v] = h->level[v];
    for (uint32_t l = 0; l <= h->level[v]; l++) {
      const uint32_t* row = l == 0 ? &g0[v * w0] : &gu[((uint64_t)(l - 1) * h->rows + rm[v]) * w1];
      if (counts) counts[ci] = row[0];
      ci++;
      if (edges) for (uint32_t j = 0; j < row[0]; j++) edges[ei + j] = row[1 + j];
      ei += row[0];
    }
  }
  if (entrance) for (size_t i = 0; i < h->entrance.size(); i++) entrance[i] = h->entrance[i];
  sizes[0] = ci; sizes[1] = ei; sizes[2] = h->entrance.size();
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
// small kernels of the build / search plumbing
__global__ void pann_hnsw_fill_eps_kernel(uint32_t* eps, uint32_t* neps, uint32_t S, uint32_t bs, const uint32_t* entrance, uint32_t nent) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (uint64_t)bs * nent) return;
  const uint32_t i = (uint32_t)(t / nent), j = (uint32_t)(t % nent);
  eps[(uint64_t)i * S + j] = entrance[j];
  if (j == 0) neps[i] = nent;
}
// the active nodes' start lists, compacted: task j = batch member act[j]
__global__ void pann_hnsw_gather_kernel(const uint32_t* eps, const uint32_t* neps, uint32_t S, const uint32_t* act, uint32_t na, uint32_t b0,
                                        uint32_t* c_st, uint32_t* c_ns, uint32_t* c_self) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (uint64_t)na * S) return;
  const uint32_t j = (uint32_t)(t / S), k = (uint32_t)(t % S);
  const uint32_t i = act[j];
  if (k < neps[i]) c_st[(uint64_t)j * S + k] = eps[(uint64_t)i * S + k];
  if (k == 0) { c_ns[j] = neps[i]; c_self[j] = b0 + i; }
}
// search results back into the batch's start lists: the whole frontier (insert levels, HNSW.hpp:929-931) or its first entry
// (descent, :877-878)
__global__ void pann_hnsw_scatter_kernel(uint32_t* eps, uint32_t* neps, uint32_t S, const uint32_t* act, uint32_t na,
                                         const uint32_t* c_fr, const uint32_t* c_nfr, int first_only) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (uint64_t)na * S) return;
  const uint32_t j = (uint32_t)(t / S), k = (uint32_t)(t % S);
  const uint32_t i = act[j];
  const uint32_t c = first_only ? 1u : c_nfr[j];
  if (k < c) eps[(uint64_t)i * S + k] = c_fr[(uint64_t)j * S + k];
  if (k == 0) neps[i] = c;
}
__global__ void pann_hnsw_add_kernel(uint32_t* acc, const uint32_t* x, uint64_t n) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) acc[t] += x[t];
}
static inline unsigned nblk(uint64_t n) { return (unsigned)((n + 255) / 256); }

// one layer search over `nq` tasks: queries are base rows self (build) or the resident query set (search)
static int hnsw_layer_search(pann_hnsw* h, uint32_t l, uint64_t nq, const uint8_t* d_queries, const uint32_t* d_self, uint32_t ef, int64_t limit,
                             const uint32_t* h_starts, uint32_t nstarts, const uint32_t* d_starts, uint32_t sstride, const uint32_t* d_ns,
                             uint32_t* fr_ids, float* fr_dists, uint32_t* nfr, uint32_t fcap, uint32_t* cmps, int* status) {
  pann_index* idx = h->idx;
  pann_query_params qp;
  qp.k = ef; qp.beam = ef; qp.cut = 1.35; qp.limit = limit; qp.degree_limit = l == 0 ? 2 * h->m : h->m;   // HNSW.hpp:1096
  SearchOutputs o;
  o.fr_ids = fr_ids; o.fr_dists = fr_dists; o.n_frontier = nfr; o.fcap = fcap; o.cmps = cmps; o.status = status;
  if (d_starts) { o.d_starts = d_starts; o.starts_stride = sstride; o.d_nstarts_q = d_ns; }
  if (l > 0) { o.g_graph = h->layer(l); o.g_maxdeg = h->m; o.g_rowmap = h->rowmap.as<uint32_t>(); }
  int flag = 0;
  if (pann_run_search(idx, d_queries, nq, idx->geom.stride, d_self, h_starts, nstarts, &qp, o, &flag)) return 1;
  if (flag & ~(1 << 2)) PANN_FAIL("HNSW layer search failed (status " + std::to_string(flag) + ")");
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
// HNSW::HNSW(begin, end, dim, m_l, m, ef_construction, alpha) (HNSW.hpp:746-824): node i = point i; node 0 is the first
// entrance; batches [b, min(n, ceil(2b)+1, b + 0.02n)); insert() per batch.
extern "C" int pann_hnsw_build(pann_hnsw* h, const uint32_t* levels, double seconds[4]) {
  if (!h) PANN_FAIL("pann_hnsw_build: null argument");
  pann_index* idx = h->idx;
  PANN_CUDA(cudaSetDevice(idx->device));
  cudaStream_t s = idx->stream;
  const uint64_t n = idx->n;
  const uint32_t m = h->m, efc = h->efc;
  const auto t_begin = std::chrono::steady_clock::now();
  h->level.resize(n);
  if (levels) std::copy(levels, levels + n, h->level.begin());
  else pann_hnsw_levels(n, h->m_l, h->level.data());
  h->entrance.assign(1, 0u);
  if (hnsw_alloc_layers(h)) return 1;
  PANN_CUDA(cudaMemsetAsync(idx->graph, 0, (size_t)n * (2 * m + 1) * 4, s));
  idx->last_launches = 0;

  const uint32_t size_limit = (uint32_t)((double)(uint32_t)n * 0.02);   // HNSW.hpp:771
  const uint64_t max_bs = std::max<uint64_t>(size_limit, 2);
  const uint32_t S = std::max<uint32_t>(efc, 32);   // start-list stride: a level's whole frontier, or the entrance list
  const uint32_t cap0 = 2 * m;
  enum { W_EPS, W_NEPS, W_ACT, W_CST, W_CNS, W_CSELF, W_FR, W_FRD, W_CNT, W_NEW, W_KEYS, W_MISC };
  DevBuf* W = h->work;
  DevBuf d_keys2, d_gstart, d_sort_tmp, d_big, d_gscr, d_ent, d_acct;
  auto release_all = [&]() { d_keys2.release(); d_gstart.release(); d_sort_tmp.release(); d_big.release(); d_gscr.release(); d_ent.release(); d_acct.release(); };
  struct Guard { decltype(release_all)& f; ~Guard() { f(); } } guard{release_all};
  if (W[W_EPS].reserve(4 * max_bs * S) || W[W_NEPS].reserve(4 * max_bs) || W[W_ACT].reserve(4 * max_bs * 2) || W[W_CST].reserve(4 * max_bs * S) ||
      W[W_CNS].reserve(4 * max_bs) || W[W_CSELF].reserve(4 * max_bs) || W[W_FR].reserve(4 * max_bs * S) || W[W_FRD].reserve(4 * max_bs * S) ||
      W[W_CNT].reserve(4 * (4 * max_bs + 8)) || W[W_NEW].reserve(4 * max_bs * cap0) || W[W_KEYS].reserve(8 * max_bs * cap0) ||
      d_keys2.reserve(8 * max_bs * cap0) || d_gstart.reserve(8 * max_bs * cap0) || d_ent.reserve(4 * 4096) || d_acct.reserve(32))
    return 1;
  PANN_CUDA(cudaMemsetAsync(d_acct.p, 0, 32, s));
  int* gs_overflow = reinterpret_cast<int*>(d_acct.as<u64>() + 3);
  int key_bits = 33; while (key_bits < 64 && (n >> (key_bits - 32)) != 0) key_bits++;
  size_t sort_tmp_bytes = 0;
  pann_sort_keys(nullptr, sort_tmp_bytes, W[W_KEYS].as<u64>(), d_keys2.as<u64>(), (int64_t)(max_bs * cap0), key_bits, s);
  if (d_sort_tmp.reserve(sort_tmp_bytes + 16)) return 1;
  uint32_t* eps = W[W_EPS].as<uint32_t>(); uint32_t* neps = W[W_NEPS].as<uint32_t>();
  uint32_t* c_st = W[W_CST].as<uint32_t>(); uint32_t* c_ns = W[W_CNS].as<uint32_t>(); uint32_t* c_self = W[W_CSELF].as<uint32_t>();
  uint32_t* c_fr = W[W_FR].as<uint32_t>(); float* c_frd = W[W_FRD].as<float>();
  uint32_t* c_nfr = W[W_CNT].as<uint32_t>(); uint32_t* c_cmps = c_nfr + max_bs; int* c_status = (int*)(c_cmps + max_bs);
  uint32_t* c_newdeg = (uint32_t*)(c_status + max_bs); uint32_t* c_ngroups = c_newdeg + max_bs;
  double t_search = 0, t_prune = 0, t_rev = 0;
  std::vector<uint32_t> act;
  cudaEvent_t ev[4];
  for (auto& e : ev) PANN_CUDA(cudaEventCreate(&e));
  int rc = 0;

  uint32_t batch_begin = 0, batch_end = 1;
  while (batch_end < n && !rc) {
    batch_begin = batch_end;
    batch_end = std::min({(uint32_t)n, (uint32_t)std::ceil(batch_begin * 2.0f) + 1, batch_begin + size_limit});   // HNSW.hpp:777-783
    if (batch_end <= batch_begin) { pann_set_error("pann_hnsw_build: empty batch (n too small for the 2 % batch limit)"); rc = 1; break; }
    const uint32_t b0 = batch_begin, bs = batch_end - batch_begin;
    const uint32_t level_ep = h->level[h->entrance[0]];
    const uint32_t nent = (uint32_t)h->entrance.size();
    if (nent > S || nent > 4096) { pann_set_error("pann_hnsw_build: more entrances than the start-list stride"); rc = 1; break; }
    PANN_CUDA(cudaMemcpyAsync(d_ent.p, h->entrance.data(), 4 * nent, cudaMemcpyHostToDevice, s));
    pann_hnsw_fill_eps_kernel<<<nblk((uint64_t)bs * nent), 256, 0, s>>>(eps, neps, S, bs, d_ent.as<uint32_t>(), nent);   // eps_u = entrance (:867)
    // two passes over the levels, top down: the ef = 1 descent of the nodes whose own level is below (:869-881), then the
    // insertion of the nodes that exist on the level (:885-1000)
    for (int pass = 0; pass < 2 && !rc; pass++)
      for (int l = (int)level_ep; l >= (pass == 0 ? 1 : 0) && !rc; l--) {
        act.clear();
        for (uint32_t i = 0; i < bs; i++) {
          const uint32_t lu = h->level[b0 + i];
          if (pass == 0 ? (lu < (uint32_t)l) : (lu >= (uint32_t)l)) act.push_back(i);
        }
        const uint32_t na = (uint32_t)act.size();
        if (!na) continue;
        uint32_t* d_act = W[W_ACT].as<uint32_t>();
        PANN_CUDA(cudaMemcpyAsync(d_act, act.data(), 4ull * na, cudaMemcpyHostToDevice, s));
        PANN_CUDA(cudaStreamSynchronize(s));   // `act` is reused by the next level
        pann_hnsw_gather_kernel<<<nblk((uint64_t)na * S), 256, 0, s>>>(eps, neps, S, d_act, na, b0, c_st, c_ns, c_self);
        cudaEventRecord(ev[0], s);
        const uint32_t ef = pass == 0 ? 1u : efc;
        // longest start list of the launch: the entrance list, one descent result, or (insert levels) a whole frontier
        const uint32_t ns_bound = pass == 0 ? nent : S;
        if (hnsw_layer_search(h, (uint32_t)l, na, nullptr, c_self, ef, (int64_t)n, nullptr, ns_bound, c_st, S, c_ns, c_fr, c_frd, c_nfr, S, c_cmps, c_status)) { rc = 1; break; }
        cudaEventRecord(ev[1], s);
        pann_hnsw_scatter_kernel<<<nblk((uint64_t)na * S), 256, 0, s>>>(eps, neps, S, d_act, na, c_fr, c_nfr, pass == 0 ? 1 : 0);
        idx->last_launches += 3;
        if (pass == 0) { PANN_CUDA(cudaEventSynchronize(ev[1])); float ms = 0; cudaEventElapsedTime(&ms, ev[0], ev[1]); t_search += ms * 1e-3; continue; }
        // ---- select_neighbors = prune_heuristic over the frontier (:488-503, :430-486)
        const uint32_t cap = l == 0 ? cap0 : m;
        uint32_t* lgraph = l == 0 ? idx->graph : h->layer((uint32_t)l);
        const uint32_t* rowmap = l == 0 ? nullptr : h->rowmap.as<uint32_t>();
        PruneArgs A;
        memset(&A, 0, sizeof(A));
        A.points = idx->points; A.stride = idx->geom.stride; A.nch = idx->geom.nch;
        A.graph = lgraph; A.maxdeg = cap; A.R = cap; A.n = n; A.alpha = (double)h->alpha; A.alpha_f = h->alpha; A.rule = 1; A.rowmap = rowmap;
        A.perm_lo = c_self;
        A.mode = 0; A.ntasks = na;
        A.vis_ids = c_fr; A.vis_dists = c_frd; A.nvis = c_nfr; A.search_cmps = c_cmps; A.vout = S;
        A.new_out = W[W_NEW].as<uint32_t>(); A.new_deg = c_newdeg;
        if (pann_launch_prune(idx, A, na, d_big, d_gscr, gs_overflow, 0)) { rc = 1; break; }
        // ---- forward edges (:938-942) + reverse-edge keys, grouped by target with values in batch order (:944-947)
        const uint64_t nk = (uint64_t)na * cap;
        pann_write_rows_kernel<<<nblk(nk), 256, 0, s>>>(lgraph, cap, cap, n, c_self, W[W_NEW].as<uint32_t>(), c_newdeg, na, W[W_KEYS].as<u64>(), rowmap);
        cudaEventRecord(ev[2], s);
        size_t tb = sort_tmp_bytes;
        if (pann_sort_keys(d_sort_tmp.p, tb, W[W_KEYS].as<u64>(), d_keys2.as<u64>(), (int64_t)nk, key_bits, s)) { pann_set_error("pann_hnsw_build: radix sort failed"); rc = 1; break; }
        PANN_CUDA(cudaMemsetAsync(c_ngroups, 0, 4, s));
        pann_find_heads_kernel<<<nblk(nk), 256, 0, s>>>(d_keys2.as<u64>(), (uint32_t)nk, n, d_gstart.as<uint32_t>(), c_ngroups);
        // ---- per target: append, or keep the cap closest of old + new (:949-999)
        A.rule = 2; A.mode = 1; A.ntasks = (uint32_t)nk;
        A.sorted = d_keys2.as<u64>(); A.nsorted = (uint32_t)nk; A.gstart = d_gstart.as<uint32_t>(); A.ngroups = c_ngroups;
        // a hub can receive an edge from every node of the batch: the slab holds its old list plus all of them
        if (pann_launch_prune(idx, A, (uint32_t)std::min<uint64_t>(nk, n), d_big, d_gscr, gs_overflow, na + cap)) { rc = 1; break; }
        cudaEventRecord(ev[3], s);
        idx->last_launches += 4 + (key_bits + 7) / 8;
        PANN_CUDA(cudaEventSynchronize(ev[3]));
        float m0 = 0, m1 = 0, m2 = 0;
        cudaEventElapsedTime(&m0, ev[0], ev[1]); cudaEventElapsedTime(&m1, ev[1], ev[2]); cudaEventElapsedTime(&m2, ev[2], ev[3]);
        t_search += m0 * 1e-3; t_prune += m1 * 1e-3; t_rev += m2 * 1e-3;
      }
    if (rc) break;
    // ---- entrance update (:1002-1013): the first node of the batch with the highest level
    uint32_t hi = b0;
    for (uint32_t u = b0; u < batch_end; u++) if (h->level[u] > h->level[hi]) hi = u;
    if (h->level[hi] > level_ep) h->entrance.assign(1, hi);
    else if (h->level[hi] == level_ep) h->entrance.push_back(hi);
  }
  for (auto& e : ev) cudaEventDestroy(e);
  if (rc) return 1;
  int ovf = 0;
  PANN_CUDA(cudaMemcpyAsync(&ovf, gs_overflow, 4, cudaMemcpyDeviceToHost, s));
  PANN_CUDA(cudaStreamSynchronize(s));
  if (ovf) PANN_FAIL("pann_hnsw_build: a candidate list of " + std::to_string(ovf) + " entries exceeds the long-list slab (set PANN_PRUNE_SLAB_CAP)");
  PANN_CUDA(cudaGetLastError());
  h->last_seconds[0] = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
  h->last_seconds[1] = t_search; h->last_seconds[2] = t_prune; h->last_seconds[3] = t_rev;
  if (seconds) for (int i = 0; i < 4; i++) seconds[i] = h->last_seconds[i];
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
// HNSW::search (HNSW.hpp:1570-1613) for a batch: search_layer_to (ef = 1, top level .. 1, :1545-1568), search_layer(ef) on level
// 0, the answer cut to k plus whatever ties the k-th distance (:1594-1600).  HOST buffers: out_ids / out_dists hold `ocap`
// entries per query (ocap >= k), out_cnt how many the reference returns, out_cmps the summed distance comparisons.
extern "C" int pann_hnsw_search(pann_hnsw* h, const void* host_queries, uint64_t nq, uint64_t q_stride_bytes, uint32_t k, uint32_t ef,
                                int64_t limit, uint32_t ocap, uint32_t* out_ids, float* out_dists, uint32_t* out_cnt, uint32_t* out_cmps) {
  if (!h || (!host_queries && nq) || !out_ids) PANN_FAIL("pann_hnsw_search: null argument");
  pann_index* idx = h->idx;
  PANN_CUDA(cudaSetDevice(idx->device));
  if (h->level.size() != idx->n || h->entrance.empty()) PANN_FAIL("pann_hnsw_search: no model (build or set one first)");
  if (ef == 0 || ocap < k || ocap > ef) PANN_FAIL("pann_hnsw_search: need k <= ocap <= ef, ef > 0");
  if (nq == 0) return 0;
  cudaStream_t s = idx->stream;
  if (pann_upload_queries(idx, host_queries, nq, q_stride_bytes)) return 1;
  const uint8_t* dq = idx->queries.as<uint8_t>();
  const uint32_t F = ef;
  DevBuf& B = h->work[11];
  if (B.reserve(4ull * nq * F * 2 + 4ull * nq * 6 + 64)) return 1;
  uint8_t* p = B.as<uint8_t>();
  uint32_t* fr = (uint32_t*)p; p += 4ull * nq * F;
  float* frd = (float*)p; p += 4ull * nq * F;
  uint32_t* nfr = (uint32_t*)p; p += 4ull * nq;
  uint32_t* cmps = (uint32_t*)p; p += 4ull * nq;
  uint32_t* acc = (uint32_t*)p; p += 4ull * nq;
  int* status = (int*)p; p += 4ull * nq;
  uint32_t* ep = (uint32_t*)p; p += 4ull * nq;
  float* epd = (float*)p; p += 4ull * nq;
  PANN_CUDA(cudaMemsetAsync(acc, 0, 4ull * nq, s));
  idx->last_launches = 0;
  cudaEvent_t e0 = idx->ev0, e1 = idx->ev1;
  PANN_CUDA(cudaEventRecord(e0, s));
  const uint32_t top = h->level[h->entrance[0]];
  const uint32_t nent = (uint32_t)h->entrance.size();
  bool have_ep = false;
  for (uint32_t l = top; l >= 1; l--) {   // ef = 1; limit_eval is not passed down (:1553-1556)
    if (hnsw_layer_search(h, l, nq, dq, nullptr, 1, (int64_t)idx->n, have_ep ? nullptr : h->entrance.data(), have_ep ? 1 : nent,
                          have_ep ? ep : nullptr, 1, nullptr, ep, epd, nfr, 1, cmps, status)) return 1;
    pann_hnsw_add_kernel<<<nblk(nq), 256, 0, s>>>(acc, cmps, nq);
    have_ep = true;
  }
  if (hnsw_layer_search(h, 0, nq, dq, nullptr, ef, limit > 0 ? limit : (int64_t)idx->n, have_ep ? nullptr : h->entrance.data(),
                        have_ep ? 1 : nent, have_ep ? ep : nullptr, 1, nullptr, fr, frd, nfr, F, cmps, status)) return 1;
  pann_hnsw_add_kernel<<<nblk(nq), 256, 0, s>>>(acc, cmps, nq);
  PANN_CUDA(cudaEventRecord(e1, s));
  std::vector<uint32_t> hfr((size_t)nq * F), hn(nq);
  std::vector<float> hd((size_t)nq * F);
  PANN_CUDA(cudaMemcpyAsync(hfr.data(), fr, 4ull * nq * F, cudaMemcpyDeviceToHost, s));
  PANN_CUDA(cudaMemcpyAsync(hd.data(), frd, 4ull * nq * F, cudaMemcpyDeviceToHost, s));
  PANN_CUDA(cudaMemcpyAsync(hn.data(), nfr, 4ull * nq, cudaMemcpyDeviceToHost, s));
  if (out_cmps) PANN_CUDA(cudaMemcpyAsync(out_cmps, acc, 4ull * nq, cudaMemcpyDeviceToHost, s));
  PANN_CUDA(cudaStreamSynchronize(s));
  PANN_CUDA(cudaEventElapsedTime(&idx->last_ms, e0, e1));
  for (uint64_t i = 0; i < nq; i++) {
    uint32_t c = std::min<uint32_t>(hn[i], F);
    if (c > k) {  // everything tied with the k-th distance stays (:1597-1599)
      uint32_t kk = k;
      if (k > 0) while (kk < c && !(hd[i * F + k - 1] < hd[i * F + kk])) kk++;
      c = kk;
    }
    if (out_cnt) out_cnt[i] = c;
    for (uint32_t j = 0; j < ocap; j++) {
      out_ids[i * ocap + j] = j < c ? hfr[i * F + j] : 0xffffffffu;
      if (out_dists) out_dists[i * ocap + j] = j < c ? hd[i * F + j] : 0.f;
    }
  }
  return 0;
}

extern "C" int pann_hnsw_info(pann_hnsw* h, uint32_t* top_level, uint64_t* upper_rows, uint32_t* n_entrances, double seconds[4]) {
  if (!h) PANN_FAIL("pann_hnsw_info: null argument");
  if (top_level) *top_level = h->top;
  if (upper_rows) *upper_rows = h->rows;
  if (n_entrances) *n_entrances = (uint32_t)h->entrance.size();
  if (seconds) for (int i = 0; i < 4; i++) seconds[i] = h->last_seconds[i];
  return 0;
}
