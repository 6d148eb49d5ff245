// vamana_build.cu — knn_index::build_index / batch_insert on one B200 (algorithms/vamana/index.h:150-316).
//
// Per batch (the schedule and the insertion order are the reference's, computed on the host, host_util.cpp):
//   phase 1   K1 in build mode (k=0, beam=L, query = base row, returns the visited list)      beam_search.cuh
//             K3 robustPrune of every visited list -> new_out                                 vamana_prune.cuh
//   phase 2   rows written + one (target << 32 | seq) key per new edge                        pann_write_rows_kernel
//   phase 3   radix sort of the keys (= parlay::group_by_key with values in batch order),
//             group heads, then K3 per group: append if it fits, else robustPrune             pann_find_heads_kernel
// All reads of phase 1 see the pre-batch graph; phase 3 touches one row per group.  One GPU.
#include <cub/device/device_radix_sort.cuh>

#include <chrono>

#include <stdio.h>
#include "pann_internal.h"
#include "vamana_prune.cuh"
#include "vamana_build.h"

__global__ void pann_write_rows_kernel(uint32_t* graph, uint32_t maxdeg, uint32_t R, uint64_t n, const uint32_t* perm_lo,
                                       const uint32_t* new_out, const uint32_t* new_deg, uint32_t bs, u64* keys, const uint32_t* rowmap) {
  uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (uint64_t)bs * R) return;
  uint32_t i = (uint32_t)(t / R), pos = (uint32_t)(t % R);
  uint32_t p = perm_lo[i], deg = new_deg[i];
  uint32_t* row = graph + (uint64_t)(rowmap ? rowmap[p] : p) * (maxdeg + 1);
  if (pos == 0) row[0] = deg;
  if (pos < deg) {
    uint32_t nb = new_out[t];
    row[1 + pos] = nb;
    keys[t] = ((u64)nb << 32) | t;
  } else {
    keys[t] = ((u64)n << 32) | t;  // padding: sorts behind every real target
  }
}

__global__ void pann_find_heads_kernel(const u64* sorted, uint32_t ns, uint64_t n, uint32_t* gstart, uint32_t* ngroups) {
  uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= ns) return;
  uint32_t t = (uint32_t)(sorted[j] >> 32);
  if (t >= n) return;
  if (j == 0 || (uint32_t)(sorted[j - 1] >> 32) != t) {
    // (start, length) pairs; the head walks its group (2 members on average) so that K3 does not have to bisect the key
    // list with dependent global loads; a group longer than 64 is left to K3's bisection (length 0)
    uint32_t len = 1;
    while (len <= 64 && j + len < ns && (uint32_t)(sorted[j + len] >> 32) == t) len++;
    const uint32_t g = atomicAdd(ngroups, 1u);
    gstart[2 * g] = j;
    gstart[2 * g + 1] = len <= 64 ? len : 0u;
  }
}

int pann_sort_keys(void* tmp, size_t& tmp_bytes, const u64* in, u64* out, int64_t nk, int bits, cudaStream_t s) {
  return cub::DeviceRadixSort::SortKeys(tmp, tmp_bytes, in, out, nk, 0, bits, s) == cudaSuccess ? 0 : 1;
}

static const void* prune_kernel(int dtype, int metric, int qr, bool gs) {
  switch (dtype) {
    case PANN_DT_F32: return pann_prune_kernel_dt0(metric, qr, gs);
    case PANN_DT_U8: return pann_prune_kernel_dt1(metric, qr, gs);
    default: return pann_prune_kernel_dt2(metric, qr, gs);
  }
}
// tie rows back into the graph after the host sorted them
static __global__ void pann_scatter_rows_kernel(const uint32_t* tie_rows, const uint32_t* tie_graph, uint32_t* graph, uint32_t T, uint32_t maxdeg) {
  const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (w >= T) return;
  const uint32_t* src = tie_graph + (u64)w * (maxdeg + 1);
  uint32_t* dst = graph + (u64)tie_rows[w] * (maxdeg + 1);
  for (uint32_t j = lane; j < src[0]; j += 32) dst[1 + j] = src[1 + j];
}

// Rows whose distances tie: the reference leaves them where libstdc++'s std::sort puts them (vamana/index.h:180-185 ->
// graph.h:111-113), and the order of equal keys under that sort is a property of its algorithm.  One thread per row runs
// that algorithm — introsort: median-of-three quicksort down to 16 elements under a depth limit of 2 lg n, then one
// insertion sort over the row (bits/stl_algo.h: __introsort_loop, __move_median_to_first, __unguarded_partition,
// __final_insertion_sort) — on the (id, distance) pairs, so the row ends up byte for byte as the host's std::sort leaves
// it, the heap-sort fallback of an exhausted depth limit included (nearly sorted rows do run into it).  *fallback is a guard
// that sends every tie row to the host's std::sort.
#include "tie_sort.h"
static __global__ void __launch_bounds__(128) pann_tie_sort_kernel(const uint32_t* __restrict__ tie_rows, const uint32_t* __restrict__ tie_graph,
                                                                  const float* __restrict__ tie_dist, uint32_t* __restrict__ graph, uint32_t T,
                                                                  uint32_t maxdeg, int* fallback) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  TieRow r;
  const uint32_t* src = tie_graph + (u64)t * (maxdeg + 1);
  const float* sd = tie_dist + (u64)t * maxdeg;
  const int n = (int)src[0];
  for (int j = 0; j < n; j++) { r.id[j] = src[1 + j]; r.d[j] = sd[j]; }
  if (!pann_tie_sort_row(r, n)) { atomicOr(fallback, 1); return; }
  uint32_t* dst = graph + (u64)tie_rows[t] * (maxdeg + 1);
  for (int j = 0; j < n; j++) dst[1 + j] = r.id[j];
}

static const void* edge_kernel(int dtype, int metric, int qr) {
  switch (dtype) {
    case PANN_DT_F32: return pann_edge_kernel_dt0(metric, qr);
    case PANN_DT_U8: return pann_edge_kernel_dt1(metric, qr);
    default: return pann_edge_kernel_dt2(metric, qr);
  }
}
static inline uint32_t al16(size_t x) { return (uint32_t)((x + 15) / 16 * 16); }

// launches K3 over A's tasks (shared-memory variant), then the global-scratch variant over whatever it deferred.  No
// host synchronisation: the second launch reads its task count on the device (it is usually zero and the CTAs leave at
// once); its slab has a fixed capacity per warp, and a list beyond even that raises gs_overflow, read once after the build.
#define PANN_PRUNE_SLAB_CAP_DEFAULT 8192u
int pann_launch_prune(pann_index* idx, PruneArgs& A, uint32_t ntasks_bound, DevBuf& big, DevBuf& gscratch, int* gs_overflow, uint32_t slab_cap_min) {
  const PannRowGeom& g = idx->geom;
  const int qrT = pann_qr_of(g);
  if (ntasks_bound == 0) return 0;
  if (big.reserve(8 + 4ull * ntasks_bound)) return 1;
  A.big_count = big.as<uint32_t>();
  A.big_list = big.as<uint32_t>() + 2;
  A.gs_overflow = gs_overflow;
  PANN_CUDA(cudaMemsetAsync(A.big_count, 0, 8, idx->stream));
  const char* rc_env = getenv("PANN_PRUNE_ROW_CACHE");  // 0 / 1 force it off / on
  const bool rc_on = g.stride <= 256 && (rc_env ? rc_env[0] == '1' : (idx->n * (uint64_t)g.stride > (256ull << 20)));
  // reverse-edge groups (mode 1) are the vertex's <= R edges plus a handful of new ones: a smaller list and row cache per
  // warp (80 rows, 128 keys) nearly triples the resident warps of the row-cache path (the rare longer list takes the
  // uncached scan or the slab launch)
  const bool small_lists = rc_on && A.mode == 1 && A.R <= 64;
  // visited lists (mode 0): 152 cached rows, 256 keys, 4 warps per CTA = 8 resident warps per SM.  Measured at 10M (search +
  // prune phase; rows,keys,warps): (160,512,3) 3.01 s, (160,256,3) 3.01 s, (152,256,4) 2.89 s, (144,256,4) 2.90 s, (136,256,4) 2.94 s,
  // (128,256,4) 3.2-3.3 s (too many lists fall out of the cache), (176,256,3) 3.03 s
  A.cap = small_lists ? 128 : 256;
  uint32_t p1_rows = 152, p1_wpc = 4;
  if (!small_lists) if (const char* e = getenv("PANN_PRUNE_P1")) {   // A/B: "rows,keys,warps" of the visited-list launch
    unsigned a = 0, b = 0, c = 0;
    if (sscanf(e, "%u,%u,%u", &a, &b, &c) == 3 && a >= 32 && b >= 64 && c >= 1 && c <= PANN_PRUNE_WARPS) { p1_rows = a; A.cap = b; p1_wpc = c; }
  }
  if (const char* e = getenv("PANN_PRUNE_CAP")) A.cap = std::max<uint32_t>(8u, std::min<uint32_t>(A.cap, (uint32_t)atoi(e)));   // tests: force the slab launch
  A.off_b = al16((size_t)A.cap * 8);
  A.off_q = 2 * A.off_b;
  A.warp_bytes = A.off_q + g.stride;
  A.rc_rows = 0; A.rc_pitch = 0; A.off_rc = 0; A.off_pos = 0;
  int wpc = PANN_PRUNE_WARPS;
  // row cache (vamana_prune.cuh): rows up to 256 bytes, 160 candidates per warp; the pitch is an odd multiple of 16 bytes
  // so that 32 lanes reading 32 different rows spread over all banks
  // It pays once the point set no longer lives in the 126 MB L2 (measured: 10M x 128 B build 36 s -> 15.5 s); while it
  // does (1M x 128 B), the plain path with more resident warps is ~20 % faster.
  if (rc_on) {
    A.rc_rows = small_lists ? 80 : p1_rows;   // measured at 10M (rows, keys): (96,256) 3.14 s, (80,128) 2.54 s, (72,128) 2.83 s, (112,256) 4.26 s
    if (small_lists) if (const char* e = getenv("PANN_PRUNE_RC_ROWS")) A.rc_rows = (uint32_t)atoi(e);
    A.rc_pitch = g.stride + 16;
    if (((A.rc_pitch / 16) & 1) == 0) A.rc_pitch += 16;
    A.off_pos = A.warp_bytes;
    A.off_rc = A.off_pos + al16((size_t)A.cap * 2);
    A.warp_bytes = A.off_rc + A.rc_rows * A.rc_pitch;
    wpc = small_lists ? 4 : (int)p1_wpc;
  }
  // Gram tile for the cached candidate rows (vamana_prune.cuh): 8-bit rows of 128 bytes, candidate-to-candidate distances as
  // warp-level integer MMAs.  MEASURED AND LEFT OFF (profiles/r2_k3_gram_10M_summary.txt, 10M x 128 u8 build): the
  // reverse-prune launch executes 31 % fewer warp instructions (30.4 G vs 44.4 G; the MMAs themselves are < 5 % of them), but
  // the 16-row distance strip adds 7 KB of shared memory per warp — 8 instead of 16 resident warps per SM — and what remains is
  // the O(C^2) rank sort and the per-pick mask bookkeeping: reverse-prune phase 2.46 s -> 3.76 s (3.27 s with 2 warps per CTA),
  // visited-list prune 3.11 s -> 3.77 s.  PANN_PRUNE_GRAM = 1 (reverse-edge lists) / 2 (visited lists as well) turns it on.
  A.gram_tp = 0; A.off_gram = 0; A.off_nrm = 0;
  {
    int gm = 0;
    if (const char* e = getenv("PANN_PRUNE_GRAM")) gm = atoi(e);
    if (rc_on && A.rule != 2 && idx->dtype != PANN_DT_F32 && qrT == PANN_QR && A.rc_rows <= 160 && (gm >= 2 || (gm == 1 && small_lists))) {
      A.gram_tp = ((A.rc_rows + 7) / 8 * 8 + 31) / 32 * 32 + 8;   // = 8 mod 32: the 64-bit stores of a half warp cover all banks once
      A.off_nrm = A.warp_bytes;
      A.off_gram = A.off_nrm + al16((size_t)A.rc_rows * 4);
      A.warp_bytes = A.off_gram + 16 * A.gram_tp * 4;
      if (const char* e = getenv("PANN_PRUNE_GRAM_WARPS")) wpc = atoi(e);
    }
  }
  A.off_stage = A.warp_bytes;
  if (pann_staged_rows_rt(idx->dtype, qrT)) A.warp_bytes += PANN_STAGE_BYTES;
  A.task_map = nullptr; A.ntask_map = 0; A.g_scratch = nullptr;
  {
    const void* fn = prune_kernel(idx->dtype, idx->metric, qrT, false);
    size_t cta_bytes = (size_t)A.warp_bytes * wpc;
    PANN_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cta_bytes));
    int per_sm = 0;
    PANN_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, wpc * 32, cta_bytes));
    if (per_sm < 1) PANN_FAIL("robustPrune: kernel does not fit on an SM");
    long long want = ((long long)ntasks_bound + wpc - 1) / wpc;
    int grid = (int)std::max<long long>(1, std::min<long long>(want, (long long)per_sm * idx->sms));
    void* args[] = {(void*)&A};
    PANN_CUDA(cudaLaunchKernel(fn, dim3(grid), dim3(wpc * 32), args, cta_bytes, idx->stream));
    idx->last_launches++;
  }
  // candidate lists beyond the shared-memory capacity (a hub that collected ~900 reverse edges in one batch) — same
  // kernel, lists in a global scratch slab, bitonic sort
  uint32_t slab_cap = std::max<uint32_t>(PANN_PRUNE_SLAB_CAP_DEFAULT, slab_cap_min);   // the caller's bound on a list (HNSW: a whole batch can point at one hub)
  if (const char* e = getenv("PANN_PRUNE_SLAB_CAP")) slab_cap = std::max<uint32_t>(64u, (uint32_t)atoi(e));
  { uint32_t c = 64; while (c < slab_cap) c <<= 1; slab_cap = c; }   // power of two: the long-list path sorts bitonically in place
  A.cap = slab_cap;
  A.task_map = A.big_list; A.ntask_map = 0;
  A.off_b = 0; A.off_q = 0;
  A.rc_rows = 0;
  A.warp_bytes = g.stride;
  A.off_stage = A.warp_bytes;
  if (pann_staged_rows_rt(idx->dtype, qrT)) A.warp_bytes += PANN_STAGE_BYTES;
  const void* fn = prune_kernel(idx->dtype, idx->metric, qrT, true);
  size_t cta_bytes = (size_t)A.warp_bytes * PANN_PRUNE_WARPS;
  PANN_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cta_bytes));
  int grid = (int)std::max<uint32_t>(1u, std::min<uint32_t>((ntasks_bound + PANN_PRUNE_WARPS - 1) / PANN_PRUNE_WARPS, (uint32_t)idx->sms));
  grid = (int)std::max<size_t>(1, std::min<size_t>((size_t)grid, (1ull << 30) / ((size_t)PANN_PRUNE_WARPS * 2 * A.cap * 8)));   // at most 1 GB of slabs
  if (gscratch.reserve((size_t)grid * PANN_PRUNE_WARPS * 2 * A.cap * 8)) return 1;
  A.g_scratch = gscratch.as<u64>();
  void* args[] = {(void*)&A};
  PANN_CUDA(cudaLaunchKernel(fn, dim3(grid), dim3(PANN_PRUNE_WARPS * 32), args, cta_bytes, idx->stream));
  idx->last_launches++;
  return 0;
}

void pann_host_sort_rows(uint32_t* graph, const float* dists, uint64_t n, uint32_t maxdeg);

// build_index (fresh graph, all points, num_passes, final sort) and batch_insert (a subset into the existing graph) are the same
// loop (vamana/index.h:150-186 calls :188-316 once per pass)
static int vamana_run(pann_index* idx, uint32_t R, uint32_t L, double alpha, int num_passes, int sort_neighbors, const uint32_t* inserts,
                      uint64_t m_in, int random_order, double max_fraction, bool fresh, uint32_t* out_visited, uint32_t* out_cmps,
                      double phase_seconds[3]) {
  if (!idx) PANN_FAIL("pann_vamana_build: null index");
  if (R == 0 || R != idx->maxdeg) PANN_FAIL("pann_vamana_build: R must equal the graph's max degree (BuildParams::max_degree, types.h:210-214)");
  if (L < 1 || L > 65536) PANN_FAIL("pann_vamana_build: L out of range");
  if (num_passes < 1) PANN_FAIL("pann_vamana_build: num_passes must be >= 1");
  PANN_CUDA(cudaSetDevice(idx->device));
  const uint64_t n = idx->n;
  const uint64_t m = inserts ? m_in : n;
  const uint32_t md = idx->maxdeg;
  cudaStream_t s = idx->stream;
  idx->last_launches = 0;
  if (m == 0) { if (phase_seconds) phase_seconds[0] = phase_seconds[1] = phase_seconds[2] = 0; return 0; }
  if (inserts)
    for (uint64_t i = 0; i < m; i++)
      if (inserts[i] >= n) PANN_FAIL("invalid point " + std::to_string(inserts[i]) + " given to batch_insert");   // index.h:193-199

  // insertion order and batch schedule: the reference's, bit for bit (index.h:206-234)
  std::vector<int32_t> perm(m);
  if (random_order) pann_host_random_permutation((int64_t)m, perm.data());
  else for (uint64_t i = 0; i < m; i++) perm[i] = (int32_t)i;
  if (inserts) for (uint64_t i = 0; i < m; i++) perm[i] = (int32_t)inserts[perm[i]];   // shuffled_inserts (:217-218)
  const int64_t nb = pann_host_batch_schedule_f(n, m, max_fraction, nullptr, nullptr, 0);
  std::vector<int64_t> blo(nb), bhi(nb);
  pann_host_batch_schedule_f(n, m, max_fraction, blo.data(), bhi.data(), nb);
  uint64_t max_bs = 0;
  for (int64_t b = 0; b < nb; b++) max_bs = std::max<uint64_t>(max_bs, (uint64_t)(bhi[b] - blo[b]));

  DevBuf d_perm, d_vis_ids, d_vis_dists, d_cnt, d_new_out, d_keys, d_keys2, d_gstart, d_sort_tmp, d_big, d_gscr, d_stats, d_acct;
  auto release_all = [&]() {
    d_acct.release(); d_perm.release(); d_vis_ids.release(); d_vis_dists.release(); d_cnt.release(); d_new_out.release(); d_keys.release();
    d_keys2.release(); d_gstart.release(); d_sort_tmp.release(); d_big.release(); d_gscr.release(); d_stats.release();
  };
  struct Guard { decltype(release_all)& f; ~Guard() { f(); } } guard{release_all};

  if (d_perm.reserve(4 * m)) return 1;
  PANN_CUDA(cudaMemcpyAsync(d_perm.p, perm.data(), 4 * m, cudaMemcpyHostToDevice, s));
  // a fresh build starts from the empty graph (Graph(maxDeg, n), graph.h:134-145)
  if (fresh) PANN_CUDA(cudaMemsetAsync(idx->graph, 0, (size_t)n * (md + 1) * 4, s));
  uint32_t* st_vis = nullptr; uint32_t* st_cmp = nullptr;
  if (out_visited || out_cmps) {
    if (d_stats.reserve(8 * n)) return 1;
    PANN_CUDA(cudaMemsetAsync(d_stats.p, 0, 8 * n, s));
    st_vis = d_stats.as<uint32_t>(); st_cmp = st_vis + n;
  }
  uint32_t vout = std::max<uint32_t>(2 * L, L + 64);
  // per batch: nvis | cmps | status | new_deg | ngroups(1)
  if (d_cnt.reserve(4 * (4 * max_bs + 4)) || d_new_out.reserve(4 * max_bs * R) || d_keys.reserve(8 * max_bs * R) ||
      d_keys2.reserve(8 * max_bs * R) || d_gstart.reserve(8 * max_bs * R))
    return 1;
  int key_bits = 33; while (key_bits < 64 && (n >> (key_bits - 32)) != 0) key_bits++;  // targets 0..n inclusive
  size_t sort_tmp_bytes = 0;
  cub::DeviceRadixSort::SortKeys(nullptr, sort_tmp_bytes, d_keys.as<u64>(), d_keys2.as<u64>(), (int64_t)(max_bs * R), 0, key_bits, s);
  if (d_sort_tmp.reserve(sort_tmp_bytes + 16)) return 1;

  // accounting: [0..2] robustPrune bytes / calls / pair distances (vamana_prune.cuh), [3] = slab-overflow flag (int)
  if (d_acct.reserve(32)) return 1;
  PANN_CUDA(cudaMemsetAsync(d_acct.p, 0, 32, s));
  int* gs_overflow = reinterpret_cast<int*>(d_acct.as<u64>() + 3);
  uint64_t search_bytes = 0;
  // phase timers: four events per batch, read once after the last batch (no host synchronisation inside the loop
  // beyond the one the search needs for its retry flag)
  std::vector<cudaEvent_t> evs((size_t)nb * num_passes * 4);
  for (auto& e : evs) PANN_CUDA(cudaEventCreate(&e));
  size_t evi = 0;
  double ph[3] = {0, 0, 0};
  const uint32_t start = 0;  // knn_index::set_start (index.h:148)
  pann_query_params qp; qp.k = 0; qp.beam = L; qp.cut = 0.0; qp.limit = (int64_t)n; qp.degree_limit = md;  // index.h:250
  int rc = 0;
  for (int pass = 0; pass < num_passes && !rc; pass++) {
    const double a = (pass == num_passes - 1) ? alpha : 1.0;  // index.h:173-178
    for (int64_t b = 0; b < nb && !rc; b++) {
      const uint32_t bs = (uint32_t)(bhi[b] - blo[b]);
      const uint32_t* perm_lo = d_perm.as<uint32_t>() + blo[b];
      uint32_t* c_nvis = d_cnt.as<uint32_t>();
      uint32_t* c_cmps = c_nvis + max_bs;
      int* c_status = (int*)(c_cmps + max_bs);
      uint32_t* c_newdeg = (uint32_t*)(c_status + max_bs);
      uint32_t* c_ngroups = c_newdeg + max_bs;
      cudaEvent_t* ev = &evs[evi]; evi += 4;
      cudaEventRecord(ev[0], s);
      // ---- phase 1: build-mode beam search from the start point, then robustPrune of the visited list
      for (;;) {
        if (d_vis_ids.reserve(4ull * max_bs * vout) || d_vis_dists.reserve(4ull * max_bs * vout)) { rc = 1; break; }
        SearchOutputs o;
        o.vis_ids = d_vis_ids.as<uint32_t>(); o.vis_dists = d_vis_dists.as<float>(); o.n_visited = c_nvis; o.vout = vout;
        o.cmps = c_cmps; o.status = c_status;
        int flag = 0;
        if (pann_run_search(idx, nullptr, bs, idx->geom.stride, perm_lo, &start, 1, &qp, o, &flag)) { rc = 1; break; }
        if (flag & (1 << 1)) {  // a visited list outgrew `vout`: redo the batch's searches with more room (nothing was written)
          if (vout >= (1u << 20)) { pann_set_error("pann_vamana_build: visited list beyond 2^20 entries"); rc = 1; break; }
          vout *= 2;
          continue;
        }
        if (flag) { pann_set_error("pann_vamana_build: beam search failed (status " + std::to_string(flag) + ")"); rc = 1; }
        break;
      }
      if (rc) break;
      search_bytes += idx->last_bytes;
      PruneArgs A;
      memset(&A, 0, sizeof(A));
      A.points = idx->points; A.stride = idx->geom.stride; A.nch = idx->geom.nch;
      A.graph = idx->graph; A.maxdeg = md; A.R = R; A.n = n; A.alpha = a;
      A.perm_lo = perm_lo; A.st_visited = st_vis; A.st_cmps = st_cmp;
      A.mode = 0; A.ntasks = bs;
      A.vis_ids = d_vis_ids.as<uint32_t>(); A.vis_dists = d_vis_dists.as<float>(); A.nvis = c_nvis; A.search_cmps = c_cmps; A.vout = vout;
      A.new_out = d_new_out.as<uint32_t>(); A.new_deg = c_newdeg;
      A.acct = d_acct.as<u64>(); A.rowbytes = idx->dims * idx->geom.elt;
      if (pann_launch_prune(idx, A, bs, d_big, d_gscr, gs_overflow, 0)) { rc = 1; break; }
      // ---- phase 2: rows of the inserted points + reverse-edge keys
      const uint64_t nk = (uint64_t)bs * R;
      pann_write_rows_kernel<<<(unsigned)((nk + 255) / 256), 256, 0, s>>>(idx->graph, md, R, n, perm_lo, d_new_out.as<uint32_t>(),
                                                                       c_newdeg, bs, d_keys.as<u64>(), nullptr);
      idx->last_launches++;
      cudaEventRecord(ev[1], s);
      // ---- phase 3a: group the reverse edges by target, values in batch order (index.h:278-282)
      size_t tb = sort_tmp_bytes;
      if (cub::DeviceRadixSort::SortKeys(d_sort_tmp.p, tb, d_keys.as<u64>(), d_keys2.as<u64>(), (int64_t)nk, 0, key_bits, s) != cudaSuccess) {
        pann_set_error("pann_vamana_build: radix sort failed"); rc = 1; break;
      }
      idx->last_launches += (key_bits + 7) / 8 + 1;
      PANN_CUDA(cudaMemsetAsync(c_ngroups, 0, 4, s));
      pann_find_heads_kernel<<<(unsigned)((nk + 255) / 256), 256, 0, s>>>(d_keys2.as<u64>(), (uint32_t)nk, n, d_gstart.as<uint32_t>(), c_ngroups);
      idx->last_launches++;
      cudaEventRecord(ev[2], s);
      // ---- phase 3b: per target, append or prune (index.h:289-300)
      A.mode = 1; A.ntasks = (uint32_t)nk;
      A.sorted = d_keys2.as<u64>(); A.nsorted = (uint32_t)nk; A.gstart = d_gstart.as<uint32_t>(); A.ngroups = c_ngroups;
      if (pann_launch_prune(idx, A, (uint32_t)std::min<uint64_t>(nk, n), d_big, d_gscr, gs_overflow, 0)) { rc = 1; break; }
      cudaEventRecord(ev[3], s);
    }
  }
  if (!rc) {
    uint64_t acct[4] = {0, 0, 0, 0};
    if (cudaMemcpyAsync(acct, d_acct.p, 32, cudaMemcpyDeviceToHost, s) != cudaSuccess || cudaStreamSynchronize(s) != cudaSuccess) {
      pann_set_error(std::string("pann_vamana_build: ") + cudaGetErrorString(cudaGetLastError())); rc = 1;
    } else if ((int)acct[3] != 0) {
      pann_set_error("pann_vamana_build: a robustPrune candidate list of " + std::to_string((int)acct[3]) +
                     " entries exceeds the long-list slab (set PANN_PRUNE_SLAB_CAP to a larger power of two)");
      rc = 1;
    } else {
      // SURVEY.md §8d B_ins: build search + robustPrune candidate rows + the row written and the reverse-edge row touched
      idx->build_search_bytes = search_bytes;
      idx->build_prune_bytes = acct[0]; idx->build_prune_calls = acct[1]; idx->build_pair_dists = acct[2];
      idx->build_bytes = search_bytes + acct[0] + (uint64_t)num_passes * m * 2ull * 4ull * (md + 1);
    }
    for (size_t b = 0; b + 3 < evi; b += 4) {
      float m[3] = {0, 0, 0};
      for (int i = 0; i < 3; i++) { cudaEventElapsedTime(&m[i], evs[b + i], evs[b + i + 1]); ph[i] += m[i] * 1e-3; }
      if (getenv("PANN_BUILD_VERBOSE")) fprintf(stderr, "  batch %zu: search+prune %.2f ms, group %.2f ms, reverse %.2f ms\n", b / 4, m[0], m[1], m[2]);
    }
  }
  for (auto& e : evs) cudaEventDestroy(e);
  if (rc) return 1;
  PANN_CUDA(cudaGetLastError());

  const auto t_sort0 = std::chrono::steady_clock::now();
  auto since = [&](std::chrono::steady_clock::time_point t0) { return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); };
  if (sort_neighbors) {  // index.h:180-185: per row, by distance to the owner only
    // Rows with distinct distances are sorted on the device (K6, sort mode).  Rows with ties go through std::sort on the
    // host so that they land exactly where the reference's libstdc++ introsort leaves them.
    const char* hs = getenv("PANN_HOST_SORT");
    bool host_all = hs && hs[0] == '1';
    EdgeDistArgs E;
    memset(&E, 0, sizeof(E));
    E.points = idx->points; E.stride = idx->geom.stride; E.nch = idx->geom.nch;
    E.graph = idx->graph; E.maxdeg = md; E.n = n;
    const int qrT = pann_qr_of(idx->geom);
    E.off_q = al16((size_t)md * 4);
    E.off_d = E.off_q + idx->geom.stride;
    E.off_s = E.off_d + al16((size_t)md * 4);
    E.warp_bytes = E.off_s + al16((size_t)md * 4);
    const void* fn = edge_kernel(idx->dtype, idx->metric, qrT);
    size_t cta_bytes = (size_t)E.warp_bytes * PANN_PRUNE_WARPS;
    PANN_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cta_bytes));
    int grid = (int)std::min<uint64_t>((n + PANN_PRUNE_WARPS - 1) / PANN_PRUNE_WARPS, (uint64_t)idx->sms * 16);
    void* args[] = {(void*)&E};
    if (!host_all) {
      const uint32_t cap = (uint32_t)std::min<uint64_t>(std::max<uint64_t>(1024, n / 4), n);
      DevBuf t_cnt, t_rows, t_graph, t_dist;
      if (t_cnt.reserve(4) || t_rows.reserve(4ull * cap) || t_graph.reserve(4ull * cap * (md + 1)) || t_dist.reserve(4ull * cap * md)) return 1;
      PANN_CUDA(cudaMemsetAsync(t_cnt.p, 0, 4, s));
      E.sort = 1; E.tie_count = t_cnt.as<uint32_t>(); E.tie_cap = cap; E.tie_rows = t_rows.as<uint32_t>();
      E.tie_graph = t_graph.as<uint32_t>(); E.tie_dist = t_dist.as<float>();
      PANN_CUDA(cudaLaunchKernel(fn, dim3(grid), dim3(PANN_PRUNE_WARPS * 32), args, cta_bytes, s));
      idx->last_launches++;
      uint32_t T = 0;
      PANN_CUDA(cudaMemcpyAsync(&T, t_cnt.p, 4, cudaMemcpyDeviceToHost, s));
      PANN_CUDA(cudaStreamSynchronize(s));
      if (getenv("PANN_BUILD_VERBOSE")) fprintf(stderr, "  final sort: device pass done at %.3f s (allocations + K6)\n", since(t_sort0));
      if (T > cap) {
        host_all = true;   // rows without ties are already in their final order; sorting them again on the host changes nothing
      } else if (T > 0) {
        bool tie_on_host = md > PANN_TIE_MAXDEG || (hs && hs[0] == '2');   // PANN_HOST_SORT=2: tie rows on the host (A/B, tests)
        if (!tie_on_host) {
          DevBuf fb;
          if (fb.reserve(4)) return 1;
          PANN_CUDA(cudaMemsetAsync(fb.p, 0, 4, s));
          pann_tie_sort_kernel<<<(T + 127) / 128, 128, 0, s>>>(t_rows.as<uint32_t>(), t_graph.as<uint32_t>(), t_dist.as<float>(), idx->graph, T, md, fb.as<int>());
          idx->last_launches++;
          int fell = 0;
          PANN_CUDA(cudaMemcpyAsync(&fell, fb.p, 4, cudaMemcpyDeviceToHost, s));
          PANN_CUDA(cudaStreamSynchronize(s));
          if (getenv("PANN_BUILD_VERBOSE")) fprintf(stderr, "  final sort: tie kernel done at %.3f s, fallback %d\n", since(t_sort0), fell);
          tie_on_host = fell != 0;   // a row ran into std::sort's heap-sort fallback: every tie row again, from its saved original order
        }
        if (tie_on_host) {
        std::vector<uint32_t> hg((size_t)T * (md + 1));
        std::vector<float> hd((size_t)T * md);
        PANN_CUDA(cudaMemcpyAsync(hg.data(), t_graph.p, hg.size() * 4, cudaMemcpyDeviceToHost, s));
        PANN_CUDA(cudaMemcpyAsync(hd.data(), t_dist.p, hd.size() * 4, cudaMemcpyDeviceToHost, s));
        PANN_CUDA(cudaStreamSynchronize(s));
        pann_host_sort_rows(hg.data(), hd.data(), T, md);
        PANN_CUDA(cudaMemcpyAsync(t_graph.p, hg.data(), hg.size() * 4, cudaMemcpyHostToDevice, s));
        pann_scatter_rows_kernel<<<(T * 32 + 255) / 256, 256, 0, s>>>(t_rows.as<uint32_t>(), t_graph.as<uint32_t>(), idx->graph, T, md);
        idx->last_launches++;
        PANN_CUDA(cudaStreamSynchronize(s));
        }
      }
      if (getenv("PANN_BUILD_VERBOSE")) fprintf(stderr, "  final sort: %u of %llu rows have tied distances (sorted as std::sort sorts them), done at %.3f s\n", T, (unsigned long long)n, since(t_sort0));
    }
    if (host_all) {
      DevBuf d_ed;
      if (d_ed.reserve(4ull * n * md)) return 1;
      E.sort = 0; E.out = d_ed.as<float>();
      PANN_CUDA(cudaLaunchKernel(fn, dim3(grid), dim3(PANN_PRUNE_WARPS * 32), args, cta_bytes, s));
      idx->last_launches++;
      std::vector<uint32_t> hg((size_t)n * (md + 1));
      std::vector<float> hd((size_t)n * md);
      PANN_CUDA(cudaMemcpyAsync(hg.data(), idx->graph, hg.size() * 4, cudaMemcpyDeviceToHost, s));
      PANN_CUDA(cudaMemcpyAsync(hd.data(), d_ed.p, hd.size() * 4, cudaMemcpyDeviceToHost, s));
      PANN_CUDA(cudaStreamSynchronize(s));
      d_ed.release();
      pann_host_sort_rows(hg.data(), hd.data(), n, md);
      PANN_CUDA(cudaMemcpyAsync(idx->graph, hg.data(), hg.size() * 4, cudaMemcpyHostToDevice, s));
      PANN_CUDA(cudaStreamSynchronize(s));
    }
  }
  if (out_visited) PANN_CUDA(cudaMemcpy(out_visited, st_vis, 4 * n, cudaMemcpyDeviceToHost));
  if (out_cmps) PANN_CUDA(cudaMemcpy(out_cmps, st_cmp, 4 * n, cudaMemcpyDeviceToHost));
  PANN_CUDA(cudaStreamSynchronize(s));
  if (phase_seconds) { phase_seconds[0] = ph[0]; phase_seconds[1] = ph[1]; phase_seconds[2] = ph[2]; }
  return 0;
}

extern "C" int pann_vamana_build(pann_index* idx, uint32_t R, uint32_t L, double alpha, int num_passes, int sort_neighbors,
                                 uint32_t* out_visited, uint32_t* out_cmps, double phase_seconds[3]) {
  return vamana_run(idx, R, L, alpha, num_passes, sort_neighbors, nullptr, 0, 1, 0.02, true, out_visited, out_cmps, phase_seconds);
}

extern "C" int pann_vamana_batch_insert(pann_index* idx, const uint32_t* inserts, uint64_t m, uint32_t R, uint32_t L, double alpha,
                                        int random_order, double max_fraction, uint32_t* out_visited, uint32_t* out_cmps,
                                        double phase_seconds[3]) {
  if (!inserts && m) PANN_FAIL("pann_vamana_batch_insert: null insert list");
  if (!(max_fraction > 0)) PANN_FAIL("pann_vamana_batch_insert: max_fraction must be positive");
  static const uint32_t none = 0;
  return vamana_run(idx, R, L, alpha, 1, 0, inserts ? inserts : &none, m, random_order, max_fraction, false, out_visited, out_cmps, phase_seconds);
}
