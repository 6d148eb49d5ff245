// pann_internal.h — host-side state behind the C ABI (include/pann.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <vector>

#include "../../include/pann.h"
#include "beam_search.cuh"
#include "beam_search_small.cuh"
#include "pann_common.cuh"

// grow-only device buffer
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes) {
    if (bytes <= cap) return 0;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = bytes + bytes / 4 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) { pann_set_error(std::string("cudaMalloc failed: ") + cudaGetErrorString(e)); return 1; }
    cap = want;
    return 0;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct pann_index {
  int device = 0, dtype = 0, metric = 0;
  int sms = 148;
  size_t smem_optin = 0;
  uint64_t n = 0;
  uint32_t dims = 0, maxdeg = 0;
  PannRowGeom geom{};
  uint8_t* points = nullptr;   // n x geom.stride
  uint32_t* graph = nullptr;   // n x (maxdeg+1)
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  // chunk pipeline of pann_search_batch (created on first use): second compute stream, copy-in / copy-out streams
  cudaStream_t stream2 = nullptr, s_in = nullptr, s_out = nullptr;
  std::vector<cudaEvent_t> pipe_ev;
  float last_ms = 0.f;
  DevBuf bytes_ctr; uint64_t last_bytes = 0;  // algorithmic bytes of the last search (device counter + host copy)
  int last_launches = 0;
  // accounting of the last pann_vamana_build (SURVEY.md §8d B_ins), see pann_index_build_stats
  uint64_t build_bytes = 0, build_search_bytes = 0, build_prune_bytes = 0, build_prune_calls = 0, build_pair_dists = 0;
  // resident query set and results
  DevBuf queries;  uint64_t nq = 0;
  DevBuf results;  // packed: ids[nq*k] dists[nq*k] visited[nq] cmps[nq] status[nq]
  uint64_t res_nq = 0; uint32_t res_k = 0;
  DevBuf scratch_hash, scratch_vis, starts, flag, qmap, misc;
  DevBuf scratch_hash2, scratch_vis2;   // second global-scratch slab: chunk kernels on the second compute stream (pann_search_batch pipeline) run concurrently with the first's
  void* pinned = nullptr; size_t pinned_cap = 0;
  // ground truth results (resident)
  DevBuf gt_ids, gt_dists; uint64_t gt_nq = 0; uint32_t gt_k = 0;
  DevBuf sweep;                // pann_search_sweep work area
  DevBuf gt_scratch[8];        // ground-truth work buffers, kept between calls (cudaMalloc/cudaFree of ~200 MB cost 5 ms per call)
  int gt_mma_launches = 0;     // pann_gt_mma_kernel launches of the last ground-truth call
  int gt_fallback_chunks = 0;  // MMA chunks retaken on the CUDA cores (candidate-list overflow), cumulative
};

struct SearchOutputs {  // device pointers (any may be null)
  uint32_t* ids = nullptr; float* dists = nullptr; uint32_t* visited = nullptr; uint32_t* cmps = nullptr;
  uint32_t* fr_ids = nullptr; float* fr_dists = nullptr; uint32_t* n_frontier = nullptr; uint32_t fcap = 0;
  uint32_t* vis_ids = nullptr; float* vis_dists = nullptr; uint32_t* n_visited = nullptr; uint32_t vout = 0;
  int* status = nullptr;  // nq ints, required
  // optional device->host copies enqueued right after the kernel, ahead of the one host synchronisation of the call
  struct Copy { void* dst; const void* src; size_t bytes; } after[4] = {};
  cudaEvent_t after_kernel = nullptr;  // recorded between the kernel and those copies
  // ---- per-launch inputs beyond the plain call (all optional)
  // per-query start lists on the device: query q starts from d_starts[q * starts_stride .. + (d_nstarts_q ? d_nstarts_q[q] :
  // nstarts)); h_starts of pann_run_search is then ignored and its nstarts is the longest list
  const uint32_t* d_starts = nullptr; uint32_t starts_stride = 0; const uint32_t* d_nstarts_q = nullptr;
  // another graph than the index's own (HNSW's upper layers): rows of (g_maxdeg + 1) words, row of vertex v = g_rowmap[v]
  const uint32_t* g_graph = nullptr; uint32_t g_maxdeg = 0; const uint32_t* g_rowmap = nullptr;
};

// Runs K1 over nq queries (device rows, or base rows self_ids) and retries visited-list overflows with the
// global-scratch variant.  Does not synchronise the host unless a retry is needed; returns the OR of statuses.
int pann_run_search(pann_index* idx, const uint8_t* d_queries, uint64_t nq, uint32_t qstride, const uint32_t* d_self_ids,
                    const uint32_t* h_starts, uint32_t nstarts, const pann_query_params* qp, const SearchOutputs& out,
                    int* status_or);
int pann_pinned(pann_index* idx, size_t bytes, void** p);
int pann_upload_queries(pann_index* idx, const void* host_queries, uint64_t nq, uint64_t q_stride_bytes);   // -> idx->queries (async on idx->stream)
int pann_gen_rows_device(int kind, uint64_t seed, uint64_t stream, uint64_t row0, uint64_t nrows, uint32_t d, uint32_t r,
                         uint32_t nc, void* dev_out, uint64_t stride, cudaStream_t s);
void pann_host_random_permutation(int64_t n, int32_t* out);
void pann_host_random_starts(uint64_t nq, uint64_t n, uint32_t* out);
int64_t pann_host_batch_schedule(uint64_t n, uint64_t m, int64_t* lo, int64_t* hi, int64_t cap);
int64_t pann_host_batch_schedule_f(uint64_t n, uint64_t m, double max_fraction, int64_t* lo, int64_t* hi, int64_t cap);
