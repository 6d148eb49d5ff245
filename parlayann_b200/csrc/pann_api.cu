// pann_api.cu — the C ABI (include/pann.h): index lifetime, query sets, batched beam search.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "pann_internal.h"

static thread_local std::string g_err;
void pann_set_error(const std::string& msg) { g_err = msg; }

extern "C" const char* pann_last_error(void) { return g_err.c_str(); }
extern "C" const char* pann_version(void) { return "parlayann_b200 0.1 (sm_100a)"; }
extern "C" int pann_device_count(int* count) {
  int c = 0;
  cudaError_t e = cudaGetDeviceCount(&c);
  if (e != cudaSuccess) { *count = 0; pann_set_error(std::string("no CUDA device: ") + cudaGetErrorString(e)); return 1; }
  *count = c;
  return 0;
}

extern "C" int pann_host_alloc(void** p, size_t bytes) { PANN_CUDA(cudaHostAlloc(p, bytes, cudaHostAllocDefault)); return 0; }
extern "C" int pann_host_free(void* p) { PANN_CUDA(cudaFreeHost(p)); return 0; }

int pann_pinned(pann_index* idx, size_t bytes, void** p) {
  if (bytes > idx->pinned_cap) {
    if (idx->pinned) cudaFreeHost(idx->pinned);
    idx->pinned = nullptr; idx->pinned_cap = 0;
    size_t want = bytes + bytes / 4 + 4096;
    PANN_CUDA(cudaHostAlloc(&idx->pinned, want, cudaHostAllocDefault));
    idx->pinned_cap = want;
  }
  *p = idx->pinned;
  return 0;
}

static int index_alloc(pann_index** out, int device, int dtype, int metric, uint64_t n, uint32_t dims, uint32_t maxdeg) {
  if (dtype < 0 || dtype > 2) PANN_FAIL("pann_index_create: dtype must be PANN_F32, PANN_U8 or PANN_I8");
  if (metric < 0 || metric > 1) PANN_FAIL("pann_index_create: metric must be PANN_L2 or PANN_MIPS");
  if (n == 0 || dims == 0) PANN_FAIL("pann_index_create: empty point set");
  if (n >= (1ull << 31)) PANN_FAIL("pann_index_create: n must be below 2^31");
  int ndev = 0;
  if (pann_device_count(&ndev) || ndev == 0) PANN_FAIL("pann_index_create: no CUDA device (this library has no CPU fallback)");
  if (device < 0 || device >= ndev) PANN_FAIL("pann_index_create: bad device ordinal");
  PANN_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  PANN_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) PANN_FAIL(std::string("pann_index_create: built for sm_100a (B200); device is ") + prop.name);
  pann_index* idx = new pann_index();
  idx->device = device; idx->dtype = dtype; idx->metric = metric; idx->n = n; idx->dims = dims; idx->maxdeg = maxdeg;
  idx->sms = prop.multiProcessorCount;
  idx->smem_optin = prop.sharedMemPerBlockOptin;
  idx->geom = pann_row_geom(dims, dtype);
  PANN_CUDA(cudaStreamCreateWithFlags(&idx->stream, cudaStreamNonBlocking));
  PANN_CUDA(cudaEventCreate(&idx->ev0));
  PANN_CUDA(cudaEventCreate(&idx->ev1));
  PANN_CUDA(cudaMalloc(&idx->points, (size_t)n * idx->geom.stride));
  PANN_CUDA(cudaMalloc(&idx->graph, (size_t)n * (maxdeg + 1) * sizeof(uint32_t)));
  PANN_CUDA(cudaMemsetAsync(idx->graph, 0, (size_t)n * (maxdeg + 1) * sizeof(uint32_t), idx->stream));
  *out = idx;
  return 0;
}

extern "C" int pann_index_create(pann_index** out, int device, int dtype, int metric, uint64_t n, uint32_t dims,
                                 uint64_t row_stride_bytes, const void* host_points, uint32_t max_degree,
                                 const uint32_t* host_graph) {
  if (!out || !host_points) PANN_FAIL("pann_index_create: null argument");
  pann_index* idx = nullptr;
  if (index_alloc(&idx, device, dtype, metric, n, dims, max_degree)) return 1;
  const size_t rowb = (size_t)dims * idx->geom.elt;
  if (row_stride_bytes < rowb) { pann_index_destroy(idx); PANN_FAIL("pann_index_create: row stride smaller than a row"); }
  if (idx->geom.stride != rowb) PANN_CUDA(cudaMemsetAsync(idx->points, 0, (size_t)n * idx->geom.stride, idx->stream));
  PANN_CUDA(cudaMemcpy2DAsync(idx->points, idx->geom.stride, host_points, row_stride_bytes, rowb, n, cudaMemcpyHostToDevice, idx->stream));
  if (host_graph)
    PANN_CUDA(cudaMemcpyAsync(idx->graph, host_graph, (size_t)n * (max_degree + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, idx->stream));
  PANN_CUDA(cudaStreamSynchronize(idx->stream));
  *out = idx;
  return 0;
}

extern "C" int pann_index_create_synthetic(pann_index** out, int device, int dtype, int metric, uint64_t n, uint32_t dims,
                                           uint32_t max_degree, int gen_kind, uint64_t seed, uint64_t stream,
                                           uint32_t latent_r, uint32_t n_clusters) {
  return pann_index_create_synthetic_shard(out, device, dtype, metric, 0, n, dims, max_degree, gen_kind, seed, stream, latent_r, n_clusters);
}
extern "C" int pann_index_create_synthetic_shard(pann_index** out, int device, int dtype, int metric, uint64_t first_row, uint64_t n,
                                                 uint32_t dims, uint32_t max_degree, int gen_kind, uint64_t seed, uint64_t stream,
                                                 uint32_t latent_r, uint32_t n_clusters) {
  if (!out) PANN_FAIL("pann_index_create_synthetic: null argument");
  int want = gen_kind == PANN_GEN_KIND_U8 ? PANN_U8 : (gen_kind == PANN_GEN_KIND_I8 ? PANN_I8 : PANN_F32);
  if (want != dtype) PANN_FAIL("pann_index_create_synthetic: generator kind does not match dtype");
  pann_index* idx = nullptr;
  if (index_alloc(&idx, device, dtype, metric, n, dims, max_degree)) return 1;
  if (pann_gen_rows_device(gen_kind, seed, stream, first_row, n, dims, latent_r, n_clusters, idx->points, idx->geom.stride, idx->stream)) {
    pann_index_destroy(idx);
    return 1;
  }
  *out = idx;
  return 0;
}

extern "C" void pann_index_destroy(pann_index* idx) {
  if (!idx) return;
  cudaSetDevice(idx->device);
  if (idx->stream) cudaStreamSynchronize(idx->stream);
  if (idx->points) cudaFree(idx->points);
  if (idx->graph) cudaFree(idx->graph);
  idx->queries.release(); idx->results.release(); idx->scratch_hash.release(); idx->scratch_vis.release();
  idx->scratch_hash2.release(); idx->scratch_vis2.release();
  idx->bytes_ctr.release(); idx->starts.release(); idx->flag.release(); idx->qmap.release(); idx->misc.release();
  idx->gt_ids.release(); idx->gt_dists.release(); idx->sweep.release();
  for (auto& b : idx->gt_scratch) b.release();
  if (idx->pinned) cudaFreeHost(idx->pinned);
  for (cudaEvent_t e : idx->pipe_ev) cudaEventDestroy(e);
  if (idx->stream2) cudaStreamDestroy(idx->stream2);
  if (idx->s_in) cudaStreamDestroy(idx->s_in);
  if (idx->s_out) cudaStreamDestroy(idx->s_out);
  if (idx->ev0) cudaEventDestroy(idx->ev0);
  if (idx->ev1) cudaEventDestroy(idx->ev1);
  if (idx->stream) cudaStreamDestroy(idx->stream);
  delete idx;
}

extern "C" int pann_index_set_graph(pann_index* idx, const uint32_t* host_graph) {
  if (!idx || !host_graph) PANN_FAIL("pann_index_set_graph: null argument");
  PANN_CUDA(cudaSetDevice(idx->device));
  PANN_CUDA(cudaMemcpyAsync(idx->graph, host_graph, (size_t)idx->n * (idx->maxdeg + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, idx->stream));
  PANN_CUDA(cudaStreamSynchronize(idx->stream));
  return 0;
}
extern "C" int pann_index_get_graph(pann_index* idx, uint32_t* host_graph_out) {
  if (!idx || !host_graph_out) PANN_FAIL("pann_index_get_graph: null argument");
  PANN_CUDA(cudaSetDevice(idx->device));
  PANN_CUDA(cudaMemcpyAsync(host_graph_out, idx->graph, (size_t)idx->n * (idx->maxdeg + 1) * sizeof(uint32_t), cudaMemcpyDeviceToHost, idx->stream));
  PANN_CUDA(cudaStreamSynchronize(idx->stream));
  return 0;
}
extern "C" int pann_index_get_points(pann_index* idx, uint64_t row0, uint64_t nrows, void* host_out, uint64_t out_stride_bytes) {
  if (!idx || !host_out) PANN_FAIL("pann_index_get_points: null argument");
  if (row0 + nrows > idx->n) PANN_FAIL("pann_index_get_points: row range out of bounds");
  PANN_CUDA(cudaSetDevice(idx->device));
  const size_t rowb = (size_t)idx->dims * idx->geom.elt;
  PANN_CUDA(cudaMemcpy2DAsync(host_out, out_stride_bytes, idx->points + row0 * idx->geom.stride, idx->geom.stride, rowb, nrows, cudaMemcpyDeviceToHost, idx->stream));
  PANN_CUDA(cudaStreamSynchronize(idx->stream));
  return 0;
}
extern "C" int pann_index_info(pann_index* idx, uint64_t* n, uint32_t* dims, uint32_t* max_degree, int* dtype, int* metric) {
  if (!idx) PANN_FAIL("pann_index_info: null index");
  if (n) *n = idx->n;
  if (dims) *dims = idx->dims;
  if (max_degree) *max_degree = idx->maxdeg;
  if (dtype) *dtype = idx->dtype;
  if (metric) *metric = idx->metric;
  return 0;
}
extern "C" int pann_index_last_bytes(pann_index* idx, uint64_t* algorithmic_bytes) {
  if (!idx || !algorithmic_bytes) PANN_FAIL("pann_index_last_bytes: null argument");
  *algorithmic_bytes = idx->last_bytes;
  return 0;
}
extern "C" int pann_index_build_stats(pann_index* idx, uint64_t* algorithmic_bytes, uint64_t* search_bytes, uint64_t* prune_bytes,
                                      uint64_t* prune_calls, uint64_t* pair_distances) {
  if (!idx) PANN_FAIL("pann_index_build_stats: null index");
  if (algorithmic_bytes) *algorithmic_bytes = idx->build_bytes;
  if (search_bytes) *search_bytes = idx->build_search_bytes;
  if (prune_bytes) *prune_bytes = idx->build_prune_bytes;
  if (prune_calls) *prune_calls = idx->build_prune_calls;
  if (pair_distances) *pair_distances = idx->build_pair_dists;
  return 0;
}
extern "C" int pann_index_last_timing(pann_index* idx, float* kernel_ms, int* launches) {
  if (!idx) PANN_FAIL("pann_index_last_timing: null index");
  if (kernel_ms) *kernel_ms = idx->last_ms;
  if (launches) *launches = idx->last_launches;
  return 0;
}

// ---------------------------------------------------------------------------------------------
// K1 planning and launch
static inline uint32_t al16(size_t x) { return (uint32_t)((x + 15) / 16 * 16); }

static const void* beam_kernel(int dtype, int metric, int qr, bool gs) {
  switch (dtype) {
    case PANN_DT_F32: return pann_beam_kernel_dt0(metric, qr, gs);
    case PANN_DT_U8: return pann_beam_kernel_dt1(metric, qr, gs);
    default: return pann_beam_kernel_dt2(metric, qr, gs);
  }
}

static const void* beam_small_kernel(int dtype, int metric, int qr, bool ext) {
  switch (dtype) {
    case PANN_DT_F32: return pann_beam_small_kernel_dt0(metric, qr, ext);
    case PANN_DT_U8: return pann_beam_small_kernel_dt1(metric, qr, ext);
    default: return pann_beam_small_kernel_dt2(metric, qr, ext);
  }
}

static int launch_beam(pann_index* idx, BeamArgs& A, bool gs, uint32_t vcap, cudaStream_t stream = nullptr, int slab = 0) {
  if (!stream) stream = idx->stream;
  const PannRowGeom& g = idx->geom;
  const int qrT = pann_qr_of(g);
  const uint32_t B = (uint32_t)A.beam;
  A.ccap = std::max<uint32_t>(A.maxdeg + B / 8 + 1, A.nstarts);
  A.vcap = vcap;
  // narrow beams without a visited-list output: K1s, frontier in registers (beam_search_small.cuh)
  const char* force_general = getenv("PANN_K1_GENERAL");
  const bool small = !gs && B <= 32 && A.nstarts <= 32 && !A.vis_ids && !(force_general && force_general[0] == '1');
  if (small) {
    // visited ids: 64 in registers, the rest 4 bytes each in shared memory — room for 256 costs 1 KB per warp and keeps
    // the retry with the general kernel for truly long walks
    vcap = (uint32_t)std::min<long long>(std::max<long long>(A.limit, 1), std::max<long long>(256, 4ll * B));
    A.vcap = vcap;
    A.off_cand = 0;
    A.off_scand = al16((size_t)A.ccap * 8);   // 32 keys: the rank-sorted candidate chunk of the batched merge
    A.off_q = A.off_scand + 256;
    A.off_vis = A.off_q + g.stride;
    A.off_hash = A.off_vis + al16((size_t)vcap * 4);
    A.off_stage = A.off_hash + (4u << A.bits);
    A.warp_bytes = A.off_stage + (pann_staged_rows_rt(idx->dtype, qrT) ? PANN_STAGE_BYTES : 0);   // float rows are staged (pann_staged_distance)
    size_t cta_bytes = (size_t)A.warp_bytes * PANN_WARPS_PER_CTA;
    if (const char* e = getenv("PANN_K1_PAD_SMEM")) cta_bytes += (size_t)atoi(e);   // experiments: fewer CTAs per SM, more L1
    const bool ext = A.starts_stride != 0 || A.nstarts_q != nullptr || A.rowmap != nullptr || A.nstarts > B;
    const void* fn = beam_small_kernel(idx->dtype, idx->metric, qrT, ext);
    if (!ext && qrT == PANN_QR && idx->dtype != PANN_DT_F32 && A.maxdeg <= 64) {
      // small batches — at most what the WIDE variant keeps resident — get it: both rounds' gathers of a hop in flight together
      const void* fw = idx->dtype == PANN_DT_U8 ? pann_beam_small_wide_kernel_dt1(idx->metric) : pann_beam_small_wide_kernel_dt2(idx->metric);
      const char* w = getenv("PANN_K1S_WIDE");   // 0 / 1 force it off / on (A/B, tests)
      if (fw && !(w && w[0] == '0')) {
        int per_sm_w = 0;
        PANN_CUDA(cudaFuncSetAttribute(fw, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cta_bytes));
        PANN_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_w, fw, PANN_WARPS_PER_CTA * 32, cta_bytes));
        if (per_sm_w >= 1 && ((w && w[0] == '1') || (long long)A.nq <= (long long)per_sm_w * idx->sms * PANN_WARPS_PER_CTA)) fn = fw;
      }
    }
    PANN_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cta_bytes));
    int per_sm = 0;
    PANN_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, PANN_WARPS_PER_CTA * 32, cta_bytes));
    if (per_sm < 1) PANN_FAIL("beam search: kernel does not fit on an SM");
    long long want = ((long long)A.nq + PANN_WARPS_PER_CTA - 1) / PANN_WARPS_PER_CTA;
    int grid = (int)std::min<long long>(want, (long long)per_sm * idx->sms);
    if (grid < 1) grid = 1;
    void* args[] = {(void*)&A};
    PANN_CUDA(cudaLaunchKernel(fn, dim3(grid), dim3(PANN_WARPS_PER_CTA * 32), args, cta_bytes, stream));
    idx->last_launches++;
    return 0;
  }
  const uint32_t Bcap = std::max<uint32_t>(B, A.nstarts);   // a start list longer than the beam sits in the frontier until the first merge
  A.off_frontB = al16((size_t)Bcap * 8);
  A.off_cand = A.off_frontB + al16((size_t)Bcap * 8);
  A.off_scand = A.off_cand + al16((size_t)A.ccap * 8);
  A.off_nbr = A.off_scand + al16((size_t)A.ccap * 8);
  A.off_q = A.off_nbr + al16((size_t)std::max<uint32_t>(A.maxdeg, 1) * 4);
  A.off_vis = A.off_q + g.stride;
  A.off_hash = A.off_vis + (gs ? 0 : al16((size_t)vcap * 8));
  A.off_stage = A.off_hash + (gs ? 0 : (4u << A.bits));
  A.tune &= ~16u;
  {  // global-scratch variant: both rounds of a hop look the filter up together (beam_search.cuh); needs a 2^bits-bit map per warp
    const char* e = getenv("PANN_K1_TWO_ROUND");
    if (gs && A.bits <= 14 && A.maxdeg <= 64 && !(e && e[0] == '0')) {
      A.tune |= 16u;
      A.off_stage = A.off_hash + al16((size_t)(1u << A.bits) / 8);
    }
  }
  A.warp_bytes = A.off_stage + (pann_staged_rows_rt(idx->dtype, qrT) ? PANN_STAGE_BYTES : 0);   // float rows are staged (pann_staged_distance)
  const size_t cta_bytes = (size_t)A.warp_bytes * PANN_WARPS_PER_CTA;
  if (cta_bytes > idx->smem_optin) PANN_FAIL("beam search: per-CTA shared memory exceeds the device limit (beam too wide for this max degree)");
  const void* fn = beam_kernel(idx->dtype, idx->metric, qrT, gs);
  PANN_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cta_bytes));
  int per_sm = 0;
  PANN_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, PANN_WARPS_PER_CTA * 32, cta_bytes));
  if (per_sm < 1) PANN_FAIL("beam search: kernel does not fit on an SM");
  long long want = ((long long)A.nq + PANN_WARPS_PER_CTA - 1) / PANN_WARPS_PER_CTA;
  int grid = (int)std::min<long long>(want, (long long)per_sm * idx->sms);
  if (grid < 1) grid = 1;
  if (gs) {
    size_t warps = (size_t)grid * PANN_WARPS_PER_CTA;
    // one slab per compute stream: two chunk kernels of the pipelined batch call may be resident at the same time, and a
    // slab is indexed by the launch's own warp numbers (the buffers only ever grow while no kernel of their stream runs:
    // every chunk of a call has the same grid or a smaller one than the first)
    DevBuf& sh = slab ? idx->scratch_hash2 : idx->scratch_hash;
    DevBuf& sv = slab ? idx->scratch_vis2 : idx->scratch_vis;
    if (sh.cap < warps * (4ull << A.bits) || sv.cap < warps * (size_t)vcap * 8) PANN_CUDA(cudaStreamSynchronize(stream));
    if (sh.reserve(warps * (4ull << A.bits))) return 1;
    if (sv.reserve(warps * (size_t)vcap * 8)) return 1;
    A.g_hash = sh.as<uint32_t>();
    A.g_vis = sv.as<u64>();
  }
  void* args[] = {(void*)&A};
  PANN_CUDA(cudaLaunchKernel(fn, dim3(grid), dim3(PANN_WARPS_PER_CTA * 32), args, cta_bytes, stream));
  idx->last_launches++;
  return 0;
}

// validates, uploads the start points, zeroes the counters and the status flag, fills the kernel arguments
struct BeamPlan { BeamArgs A; bool gs; uint32_t vcap; long long vbig; };
#define PANN_MAX_CHUNKS 64
static int beam_plan(pann_index* idx, const uint8_t* d_queries, uint64_t nq, uint32_t qstride, const uint32_t* d_self_ids,
                     const uint32_t* h_starts, uint32_t nstarts, const pann_query_params* qp, const SearchOutputs& out,
                     BeamPlan& P) {
  if (nstarts == 0 || (!h_starts && !out.d_starts)) PANN_FAIL("beam search expects at least one start point");  // beamSearch.h:38-41
  if (qp->beam < 1 || qp->beam > 65536) PANN_FAIL("beam search: beam width out of range");
  if (qp->k < 0) PANN_FAIL("beam search: k must be non-negative");
  if (!out.d_starts)
    for (uint32_t i = 0; i < nstarts; i++)
      if (h_starts[i] >= idx->n) PANN_FAIL("beam search: start point out of range");
  if (nq >= 0x7fffffffull) PANN_FAIL("beam search: more than 2^31 - 1 queries in one call");
  if (idx->flag.reserve(sizeof(int)) || idx->bytes_ctr.reserve(16 + 4 * PANN_MAX_CHUNKS)) return 1;
  PANN_CUDA(cudaMemsetAsync(idx->bytes_ctr.p, 0, 16 + 4 * PANN_MAX_CHUNKS, idx->stream));
  if (!out.d_starts) {
    if (idx->starts.reserve(sizeof(uint32_t) * nstarts)) return 1;
    PANN_CUDA(cudaMemcpyAsync(idx->starts.p, h_starts, sizeof(uint32_t) * nstarts, cudaMemcpyHostToDevice, idx->stream));
  }
  PANN_CUDA(cudaMemsetAsync(idx->flag.p, 0, sizeof(int), idx->stream));

  BeamArgs& A = P.A;
  memset(&A, 0, sizeof(A));
  A.points = idx->points; A.n = idx->n; A.stride = idx->geom.stride; A.nch = idx->geom.nch;
  A.graph = idx->graph; A.maxdeg = idx->maxdeg;
  if (out.g_graph) { A.graph = out.g_graph; A.maxdeg = out.g_maxdeg; A.rowmap = out.g_rowmap; }
  A.queries = d_queries; A.nq = nq; A.qstride = qstride; A.self_ids = d_self_ids;
  A.starts = idx->starts.as<uint32_t>(); A.nstarts = nstarts;
  if (out.d_starts) { A.starts = out.d_starts; A.starts_stride = out.starts_stride; A.nstarts_q = out.d_nstarts_q; }
  const long long beam = qp->beam;
  A.k = (int)qp->k; A.beam = (int)beam;
  A.bits = std::max<int>(10, (int)std::ceil(std::log2((double)(int)(beam * beam))) - 2);  // beamSearch.h:52
  A.limit = qp->limit < 0 ? (long long)idx->n : qp->limit;
  A.degree_limit = qp->degree_limit < 0 ? (int)A.maxdeg : (int)std::min<int64_t>(qp->degree_limit, A.maxdeg);
  A.cut = qp->cut;
  A.cut_rule = (qp->k > 0 && idx->metric == PANN_M_L2) ? 1 : 0;
  A.out_ids = out.ids; A.out_dists = out.dists; A.out_visited = out.visited; A.out_cmps = out.cmps;
  A.fr_ids = out.fr_ids; A.fr_dists = out.fr_dists; A.n_frontier = out.n_frontier; A.fcap = out.fcap;
  A.vis_ids = out.vis_ids; A.vis_dists = out.vis_dists; A.n_visited = out.n_visited; A.vout = out.vout;
  A.status = out.status; A.any_status = idx->flag.as<int>();
  A.bytes_total = idx->bytes_ctr.as<u64>(); A.rowbytes = idx->dims * idx->geom.elt;
  A.qcounter = reinterpret_cast<uint32_t*>(idx->bytes_ctr.as<u64>() + 1);
  { const char* sq = getenv("PANN_STATIC_QUERIES"); if (sq && sq[0] == '1') A.qcounter = nullptr; }
  {  // K1s row prefetch: always while the point set is about L2-sized (most prefetches are L2 hits), else only while draining
    const char* tn = getenv("PANN_K1S_TUNE");
    if (tn) {
      A.tune = (uint32_t)atoi(tn);
    } else {
      int dev = 0, l2 = 0;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, dev);
      A.tune = ((unsigned long long)idx->n * idx->geom.stride <= 2ull * (unsigned long long)l2) ? 1u : 0u;
    }
    { const char* ns = getenv("PANN_NO_STAGE"); if (ns && ns[0] == '1') A.tune |= 4u; }   // A/B: float rows one lane per row
  }

  // shared-memory variant while the hash filter is <= 8 KB per warp (beam <= 90): at 16 KB only 8 warps fit on an SM and
  // the search is latency-bound — with the filter and the visited list in a global slab (L2-resident: 16 KB x ~3500
  // warps) 24 warps are resident (10M build, search phase: 5.2 s -> 4.1 s).  The visited-list capacity is a guess that
  // the retry below makes safe.
  const char* force = getenv("PANN_FORCE_GLOBAL_SCRATCH");
  bool gs = A.bits >= 12 || (force && force[0] == '1');
  if (force && force[0] == '0') gs = A.bits > 12;
  const long long vbig = std::min<long long>(A.limit, std::max<long long>(8 * beam + 4096, 16384));
  uint32_t vcap = gs ? (uint32_t)vbig : (uint32_t)std::min<long long>(A.limit, std::max<long long>(2 * beam, beam + 64));
  if (vcap < 1) vcap = 1;
  P.gs = gs; P.vcap = vcap; P.vbig = vbig;
  return 0;
}

int pann_run_search(pann_index* idx, const uint8_t* d_queries, uint64_t nq, uint32_t qstride, const uint32_t* d_self_ids,
                    const uint32_t* h_starts, uint32_t nstarts, const pann_query_params* qp, const SearchOutputs& out,
                    int* status_or) {
  *status_or = 0;
  if (nq == 0) {
    if (nstarts == 0 || (!h_starts && !out.d_starts)) PANN_FAIL("beam search expects at least one start point");
    return 0;
  }
  BeamPlan P;
  if (beam_plan(idx, d_queries, nq, qstride, d_self_ids, h_starts, nstarts, qp, out, P)) return 1;
  BeamArgs& A = P.A;
  const bool gs = P.gs;
  const uint32_t vcap = P.vcap;
  const long long vbig = P.vbig;
  if (launch_beam(idx, A, gs, vcap)) return 1;
  if (out.after_kernel) PANN_CUDA(cudaEventRecord(out.after_kernel, idx->stream));
  for (const auto& c : out.after)
    if (c.dst && c.bytes) PANN_CUDA(cudaMemcpyAsync(c.dst, c.src, c.bytes, cudaMemcpyDeviceToHost, idx->stream));
  int flag = 0;
  PANN_CUDA(cudaMemcpyAsync(&flag, idx->flag.p, sizeof(int), cudaMemcpyDeviceToHost, idx->stream));
  PANN_CUDA(cudaMemcpyAsync(&idx->last_bytes, idx->bytes_ctr.p, 8, cudaMemcpyDeviceToHost, idx->stream));
  PANN_CUDA(cudaStreamSynchronize(idx->stream));
  if ((flag & (1 << 1)) && !gs) {
    // some visited lists outgrew shared memory: redo exactly those queries with the global-scratch variant
    std::vector<int> st(nq);
    PANN_CUDA(cudaMemcpy(st.data(), out.status, sizeof(int) * nq, cudaMemcpyDeviceToHost));
    std::vector<uint32_t> redo;
    for (uint64_t i = 0; i < nq; i++) if (st[i] == 1) redo.push_back((uint32_t)i);
    if (idx->qmap.reserve(sizeof(uint32_t) * redo.size())) return 1;
    PANN_CUDA(cudaMemcpyAsync(idx->qmap.p, redo.data(), sizeof(uint32_t) * redo.size(), cudaMemcpyHostToDevice, idx->stream));
    PANN_CUDA(cudaMemsetAsync(idx->flag.p, 0, sizeof(int), idx->stream));
    A.qmap = idx->qmap.as<uint32_t>();
    A.qcounter = nullptr;  // the counter was used by the first launch
    A.nq = redo.size();
    if (launch_beam(idx, A, true, (uint32_t)vbig)) return 1;
    int flag2 = 0;
    if (out.after_kernel) PANN_CUDA(cudaEventRecord(out.after_kernel, idx->stream));
    for (const auto& c : out.after)  // the retried queries changed the result block: copy it again
      if (c.dst && c.bytes) PANN_CUDA(cudaMemcpyAsync(c.dst, c.src, c.bytes, cudaMemcpyDeviceToHost, idx->stream));
    PANN_CUDA(cudaMemcpyAsync(&flag2, idx->flag.p, sizeof(int), cudaMemcpyDeviceToHost, idx->stream));
    PANN_CUDA(cudaStreamSynchronize(idx->stream));
    flag = (flag & ~(1 << 1)) | flag2;
  }
  *status_or = flag;
  return 0;
}

static int check_status(int flag) {
  if (flag & (1 << 3)) PANN_FAIL("beam search: more start points than the beam width (the reference overflows here, beamSearch.h:74-76)");
  if (flag & (1 << 1)) PANN_FAIL("beam search: visited list exceeded the scratch capacity");
  if (flag & (1 << 2)) PANN_FAIL("Error: beam search does not return enough candidates");  // beamSearch.h:416-419
  return 0;
}

static int validate_qp(const pann_query_params* qp) {
  if (!qp) PANN_FAIL("null query params");
  if (qp->k > qp->beam) PANN_FAIL("Error: beam search parameter Q = " + std::to_string(qp->beam) + " same size or smaller than k = " + std::to_string(qp->k));  // beamSearch.h:549-553
  return 0;
}

// packed result block: ids | dists | visited | cmps | status
struct ResLayout { size_t ids, dists, visited, cmps, status, total; };
static ResLayout res_layout(uint64_t nq, uint32_t k) {
  ResLayout L;
  L.ids = 0; L.dists = L.ids + 4ull * nq * k; L.visited = L.dists + 4ull * nq * k; L.cmps = L.visited + 4ull * nq;
  L.status = L.cmps + 4ull * nq; L.total = L.status + 4ull * nq;
  return L;
}

static bool is_pinned_host(const void* p) {
  if (!p) return true;
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeHost;
}

// host_out (may be null): {ids, dists, visited, cmps} host buffers of the batch call.  When all of them are pinned the
// results go straight there behind the kernel and the call has ONE host synchronisation; otherwise the caller fetches.
static int search_common(pann_index* idx, const uint8_t* d_q, uint64_t nq, uint32_t qstride, const uint32_t* starts,
                         uint32_t nstarts, const pann_query_params* qp, void* const* host_out = nullptr, bool* delivered = nullptr,
                         const SearchOutputs* per_query = nullptr) {
  ResLayout L = res_layout(nq, (uint32_t)qp->k);
  if (idx->results.reserve(L.total + 16)) return 1;
  idx->res_nq = nq; idx->res_k = (uint32_t)qp->k;
  uint8_t* r = idx->results.as<uint8_t>();
  SearchOutputs o;
  if (per_query) { o.d_starts = per_query->d_starts; o.starts_stride = per_query->starts_stride; o.d_nstarts_q = per_query->d_nstarts_q; }
  o.ids = (uint32_t*)(r + L.ids); o.dists = (float*)(r + L.dists); o.visited = (uint32_t*)(r + L.visited);
  o.cmps = (uint32_t*)(r + L.cmps); o.status = (int*)(r + L.status);
  int flag = 0;
  idx->last_launches = 0;
  if (delivered) *delivered = false;
  if (host_out && is_pinned_host(host_out[0]) && is_pinned_host(host_out[1]) && is_pinned_host(host_out[2]) && is_pinned_host(host_out[3])) {
    const size_t kb = 4ull * nq * (uint32_t)qp->k;
    o.after[0] = {host_out[0], o.ids, kb}; o.after[1] = {host_out[1], o.dists, kb};
    o.after[2] = {host_out[2], o.visited, 4ull * nq}; o.after[3] = {host_out[3], o.cmps, 4ull * nq};
    if (delivered) *delivered = true;
  }
  o.after_kernel = idx->ev1;
  PANN_CUDA(cudaEventRecord(idx->ev0, idx->stream));
  if (pann_run_search(idx, d_q, nq, qstride, nullptr, starts, nstarts, qp, o, &flag)) return 1;   // synchronises once
  PANN_CUDA(cudaEventElapsedTime(&idx->last_ms, idx->ev0, idx->ev1));
  return check_status(flag);
}

// Large host batches (pinned, contiguous rows): the batch is cut into chunks whose H2D copy, search and D2H copies
// run on separate streams, so the copies hide under the neighbouring chunks' kernels.  Chunk kernels alternate between
// two compute streams: the next chunk's CTAs fill the SMs as the previous chunk drains, so chunking costs no tail.
// Returns 2 if some visited list outgrew shared memory (the caller redoes the batch on the plain path).
#define PANN_PIPE_CHUNK (128u * 1024u)
static int search_pipelined(pann_index* idx, const uint8_t* host_queries, uint64_t nq, const uint32_t* starts, uint32_t nstarts,
                            const pann_query_params* qp, void* const* host_out) {
  const uint32_t k = (uint32_t)qp->k;
  const uint64_t nchunks = std::min<uint64_t>((nq + PANN_PIPE_CHUNK - 1) / PANN_PIPE_CHUNK, PANN_MAX_CHUNKS);
  const uint64_t per = ((nq + nchunks - 1) / nchunks + 3) / 4 * 4;
  if (!idx->stream2) {
    PANN_CUDA(cudaStreamCreateWithFlags(&idx->stream2, cudaStreamNonBlocking));
    PANN_CUDA(cudaStreamCreateWithFlags(&idx->s_in, cudaStreamNonBlocking));
    PANN_CUDA(cudaStreamCreateWithFlags(&idx->s_out, cudaStreamNonBlocking));
  }
  while (idx->pipe_ev.size() < 2 * nchunks + 1) {
    cudaEvent_t e;
    PANN_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    idx->pipe_ev.push_back(e);
  }
  const size_t stride = idx->geom.stride;
  if (idx->queries.reserve((size_t)nq * stride + 16)) return 1;
  idx->nq = nq;
  ResLayout L = res_layout(nq, k);
  if (idx->results.reserve(L.total + 16)) return 1;
  idx->res_nq = nq; idx->res_k = k;
  uint8_t* r = idx->results.as<uint8_t>();
  SearchOutputs o;
  o.ids = (uint32_t*)(r + L.ids); o.dists = (float*)(r + L.dists); o.visited = (uint32_t*)(r + L.visited);
  o.cmps = (uint32_t*)(r + L.cmps); o.status = (int*)(r + L.status);
  idx->last_launches = 0;
  PANN_CUDA(cudaEventRecord(idx->ev0, idx->stream));
  BeamPlan P;
  if (beam_plan(idx, idx->queries.as<uint8_t>(), nq, (uint32_t)stride, nullptr, starts, nstarts, qp, o, P)) return 1;
  cudaEvent_t ready = idx->pipe_ev[2 * nchunks];   // counters zeroed, start points uploaded (ordered on idx->stream)
  PANN_CUDA(cudaEventRecord(ready, idx->stream));
  PANN_CUDA(cudaStreamWaitEvent(idx->stream2, ready, 0));
  uint8_t* dq = idx->queries.as<uint8_t>();
  for (uint64_t c = 0; c < nchunks; c++) {
    const uint64_t c0 = c * per, cn = std::min<uint64_t>(per, nq - std::min(nq, c0));
    if (!cn) continue;
    PANN_CUDA(cudaMemcpyAsync(dq + c0 * stride, host_queries + c0 * stride, cn * stride, cudaMemcpyHostToDevice, idx->s_in));
    PANN_CUDA(cudaEventRecord(idx->pipe_ev[2 * c], idx->s_in));
  }
  cudaStream_t cs_last = idx->stream;
  for (uint64_t c = 0; c < nchunks; c++) {
    const uint64_t c0 = c * per, cn = std::min<uint64_t>(per, nq - std::min(nq, c0));
    if (!cn) continue;
    cudaStream_t cs = (c & 1) ? idx->stream2 : idx->stream;
    cs_last = cs;
    PANN_CUDA(cudaStreamWaitEvent(cs, idx->pipe_ev[2 * c], 0));
    BeamArgs A = P.A;
    A.queries = dq + c0 * stride; A.nq = cn;
    A.out_ids = o.ids + c0 * k; A.out_dists = o.dists + c0 * k; A.out_visited = o.visited + c0; A.out_cmps = o.cmps + c0;
    A.status = o.status + c0;
    if (A.qcounter) A.qcounter = reinterpret_cast<uint32_t*>(idx->bytes_ctr.as<uint8_t>() + 16) + c;   // one query counter per chunk
    if (launch_beam(idx, A, P.gs, P.vcap, cs, (int)(c & 1))) return 1;
    PANN_CUDA(cudaEventRecord(idx->pipe_ev[2 * c + 1], cs));
    PANN_CUDA(cudaStreamWaitEvent(idx->s_out, idx->pipe_ev[2 * c + 1], 0));
    PANN_CUDA(cudaMemcpyAsync((uint32_t*)host_out[0] + c0 * k, A.out_ids, 4ull * cn * k, cudaMemcpyDeviceToHost, idx->s_out));
    if (host_out[1]) PANN_CUDA(cudaMemcpyAsync((float*)host_out[1] + c0 * k, A.out_dists, 4ull * cn * k, cudaMemcpyDeviceToHost, idx->s_out));
    if (host_out[2]) PANN_CUDA(cudaMemcpyAsync((uint32_t*)host_out[2] + c0, A.out_visited, 4ull * cn, cudaMemcpyDeviceToHost, idx->s_out));
    if (host_out[3]) PANN_CUDA(cudaMemcpyAsync((uint32_t*)host_out[3] + c0, A.out_cmps, 4ull * cn, cudaMemcpyDeviceToHost, idx->s_out));
  }
  // join: idx->stream waits for the other compute stream, then the flag and the byte counter come back on it
  if (cs_last != idx->stream || nchunks > 1) {
    PANN_CUDA(cudaEventRecord(ready, idx->stream2));
    PANN_CUDA(cudaStreamWaitEvent(idx->stream, ready, 0));
  }
  PANN_CUDA(cudaEventRecord(idx->ev1, idx->stream));
  int flag = 0;
  PANN_CUDA(cudaMemcpyAsync(&flag, idx->flag.p, sizeof(int), cudaMemcpyDeviceToHost, idx->stream));
  PANN_CUDA(cudaMemcpyAsync(&idx->last_bytes, idx->bytes_ctr.p, 8, cudaMemcpyDeviceToHost, idx->stream));
  PANN_CUDA(cudaStreamSynchronize(idx->stream));
  PANN_CUDA(cudaStreamSynchronize(idx->s_out));
  PANN_CUDA(cudaEventElapsedTime(&idx->last_ms, idx->ev0, idx->ev1));
  if (flag & (1 << 1)) return 2;
  return check_status(flag);
}

int pann_upload_queries(pann_index* idx, const void* host_queries, uint64_t nq, uint64_t q_stride_bytes) {
  const size_t rowb = (size_t)idx->dims * idx->geom.elt;
  if (q_stride_bytes < rowb) PANN_FAIL("query stride smaller than a row");
  if (idx->queries.reserve((size_t)nq * idx->geom.stride + 16)) return 1;
  if (idx->geom.stride != rowb) PANN_CUDA(cudaMemsetAsync(idx->queries.p, 0, (size_t)nq * idx->geom.stride, idx->stream));
  if (q_stride_bytes == rowb && idx->geom.stride == rowb)
    PANN_CUDA(cudaMemcpyAsync(idx->queries.p, host_queries, (size_t)nq * rowb, cudaMemcpyHostToDevice, idx->stream));
  else
    PANN_CUDA(cudaMemcpy2DAsync(idx->queries.p, idx->geom.stride, host_queries, q_stride_bytes, rowb, nq, cudaMemcpyHostToDevice, idx->stream));
  idx->nq = nq;
  return 0;
}

extern "C" int pann_queries_upload(pann_index* idx, const void* host_queries, uint64_t nq, uint64_t q_stride_bytes) {
  if (!idx || !host_queries) PANN_FAIL("pann_queries_upload: null argument");
  PANN_CUDA(cudaSetDevice(idx->device));
  if (pann_upload_queries(idx, host_queries, nq, q_stride_bytes)) return 1;
  PANN_CUDA(cudaStreamSynchronize(idx->stream));
  return 0;
}
extern "C" int pann_queries_synthetic(pann_index* idx, uint64_t nq, int gen_kind, uint64_t seed, uint64_t stream,
                                      uint32_t latent_r, uint32_t n_clusters) {
  if (!idx) PANN_FAIL("pann_queries_synthetic: null index");
  PANN_CUDA(cudaSetDevice(idx->device));
  if (idx->queries.reserve((size_t)nq * idx->geom.stride + 16)) return 1;
  if (pann_gen_rows_device(gen_kind, seed, stream, 0, nq, idx->dims, latent_r, n_clusters, idx->queries.p, idx->geom.stride, idx->stream)) return 1;
  idx->nq = nq;
  return 0;
}
extern "C" int pann_queries_download(pann_index* idx, void* host_out, uint64_t out_stride_bytes) {
  if (!idx || !host_out) PANN_FAIL("pann_queries_download: null argument");
  PANN_CUDA(cudaSetDevice(idx->device));
  const size_t rowb = (size_t)idx->dims * idx->geom.elt;
  PANN_CUDA(cudaMemcpy2DAsync(host_out, out_stride_bytes, idx->queries.p, idx->geom.stride, rowb, idx->nq, cudaMemcpyDeviceToHost, idx->stream));
  PANN_CUDA(cudaStreamSynchronize(idx->stream));
  return 0;
}

extern "C" int pann_search_resident(pann_index* idx, const uint32_t* starts, uint32_t nstarts, const pann_query_params* qp) {
  if (!idx) PANN_FAIL("pann_search_resident: null index");
  if (validate_qp(qp)) return 1;
  if (idx->nq == 0) PANN_FAIL("pann_search_resident: no resident query set");
  PANN_CUDA(cudaSetDevice(idx->device));
  return search_common(idx, idx->queries.as<uint8_t>(), idx->nq, idx->geom.stride, starts, nstarts, qp);
}

static int fetch_results(pann_index* idx, uint32_t* out_ids, float* out_dists, uint32_t* out_visited, uint32_t* out_cmps) {
  const uint64_t nq = idx->res_nq; const uint32_t k = idx->res_k;
  ResLayout L = res_layout(nq, k);
  void* pin = nullptr;
  if (pann_pinned(idx, L.status, &pin)) return 1;
  PANN_CUDA(cudaMemcpyAsync(pin, idx->results.p, L.status, cudaMemcpyDeviceToHost, idx->stream));
  PANN_CUDA(cudaStreamSynchronize(idx->stream));
  const uint8_t* h = (const uint8_t*)pin;
  if (out_ids) memcpy(out_ids, h + L.ids, 4ull * nq * k);
  if (out_dists) memcpy(out_dists, h + L.dists, 4ull * nq * k);
  if (out_visited) memcpy(out_visited, h + L.visited, 4ull * nq);
  if (out_cmps) memcpy(out_cmps, h + L.cmps, 4ull * nq);
  return 0;
}

extern "C" int pann_search_fetch(pann_index* idx, uint32_t* out_ids, float* out_dists, uint32_t* out_visited, uint32_t* out_cmps) {
  if (!idx) PANN_FAIL("pann_search_fetch: null index");
  PANN_CUDA(cudaSetDevice(idx->device));
  return fetch_results(idx, out_ids, out_dists, out_visited, out_cmps);
}

extern "C" int pann_search_device_result(pann_index* idx, const uint32_t** d_ids, const float** d_dists, const uint32_t** d_visited,
                                         const uint32_t** d_cmps, uint64_t* nq, uint32_t* k) {
  if (!idx) PANN_FAIL("pann_search_device_result: null index");
  if (!idx->results.p || idx->res_nq == 0) PANN_FAIL("pann_search_device_result: no search results on the device");
  ResLayout L = res_layout(idx->res_nq, idx->res_k);
  const uint8_t* r = idx->results.as<uint8_t>();
  if (d_ids) *d_ids = (const uint32_t*)(r + L.ids);
  if (d_dists) *d_dists = (const float*)(r + L.dists);
  if (d_visited) *d_visited = (const uint32_t*)(r + L.visited);
  if (d_cmps) *d_cmps = (const uint32_t*)(r + L.cmps);
  if (nq) *nq = idx->res_nq;
  if (k) *k = idx->res_k;
  return 0;
}

extern "C" int pann_search_batch(pann_index* idx, const void* host_queries, uint64_t nq, uint64_t q_stride_bytes,
                                 const uint32_t* starts, uint32_t nstarts, const pann_query_params* qp,
                                 uint32_t* out_ids, float* out_dists, uint32_t* out_visited, uint32_t* out_cmps) {
  if (!idx || (!host_queries && nq)) PANN_FAIL("pann_search_batch: null argument");
  if (validate_qp(qp)) return 1;
  PANN_CUDA(cudaSetDevice(idx->device));
  if (nq == 0) return 0;
  void* const host_out[4] = {out_ids, out_dists, out_visited, out_cmps};
  {
    const size_t rowb = (size_t)idx->dims * idx->geom.elt;
    const char* np = getenv("PANN_NO_PIPELINE");
    if (nq >= 2ull * PANN_PIPE_CHUNK && out_ids && q_stride_bytes == rowb && idx->geom.stride == rowb && !(np && np[0] == '1') &&
        is_pinned_host(host_queries) && is_pinned_host(out_ids) && is_pinned_host(out_dists) && is_pinned_host(out_visited) && is_pinned_host(out_cmps)) {
      const int rc = search_pipelined(idx, (const uint8_t*)host_queries, nq, starts, nstarts, qp, host_out);
      if (rc != 2) return rc;   // 2: a visited list outgrew shared memory somewhere: take the plain path, which retries those queries
    }
  }
  if (pann_upload_queries(idx, host_queries, nq, q_stride_bytes)) return 1;
  bool delivered = false;
  if (search_common(idx, idx->queries.as<uint8_t>(), nq, idx->geom.stride, starts, nstarts, qp, host_out, &delivered)) return 1;
  return delivered ? 0 : fetch_results(idx, out_ids, out_dists, out_visited, out_cmps);
}

// per-query start lists from the host -> idx->starts (nq x stride) [+ counts behind them]; fills the SearchOutputs fields
static int upload_starts(pann_index* idx, uint64_t nq, const uint32_t* starts, uint32_t stride, const uint32_t* nstarts_q, SearchOutputs& o) {
  if (!starts || stride == 0) PANN_FAIL("beam search expects at least one start point");  // beamSearch.h:38-41
  for (uint64_t i = 0; i < nq; i++) {
    const uint32_t c = nstarts_q ? nstarts_q[i] : stride;
    if (c == 0) PANN_FAIL("beam search expects at least one start point");
    if (c > stride) PANN_FAIL("beam search: a per-query start count exceeds the stride of the start array");
    for (uint32_t j = 0; j < c; j++)
      if (starts[i * stride + j] >= idx->n) PANN_FAIL("beam search: start point out of range");
  }
  const size_t sb = 4ull * nq * stride;
  if (idx->starts.reserve(sb + (nstarts_q ? 4ull * nq : 0) + 16)) return 1;
  PANN_CUDA(cudaMemcpyAsync(idx->starts.p, starts, sb, cudaMemcpyHostToDevice, idx->stream));
  o.d_starts = idx->starts.as<uint32_t>(); o.starts_stride = stride;
  if (nstarts_q) {
    uint32_t* dn = reinterpret_cast<uint32_t*>(idx->starts.as<uint8_t>() + sb);
    PANN_CUDA(cudaMemcpyAsync(dn, nstarts_q, 4ull * nq, cudaMemcpyHostToDevice, idx->stream));
    o.d_nstarts_q = dn;
  }
  return 0;
}

extern "C" int pann_search_batch_starts(pann_index* idx, const void* host_queries, uint64_t nq, uint64_t q_stride_bytes,
                                        const uint32_t* starts, uint32_t starts_stride, const uint32_t* nstarts_q,
                                        const pann_query_params* qp, uint32_t* out_ids, float* out_dists,
                                        uint32_t* out_visited, uint32_t* out_cmps) {
  if (!idx || (!host_queries && nq)) PANN_FAIL("pann_search_batch_starts: null argument");
  if (validate_qp(qp)) return 1;
  PANN_CUDA(cudaSetDevice(idx->device));
  if (nq == 0) return 0;
  SearchOutputs pq;
  if (upload_starts(idx, nq, starts, starts_stride, nstarts_q, pq)) return 1;
  if (pann_upload_queries(idx, host_queries, nq, q_stride_bytes)) return 1;
  void* const host_out[4] = {out_ids, out_dists, out_visited, out_cmps};
  bool delivered = false;
  if (search_common(idx, idx->queries.as<uint8_t>(), nq, idx->geom.stride, nullptr, starts_stride, qp, host_out, &delivered, &pq)) return 1;
  return delivered ? 0 : fetch_results(idx, out_ids, out_dists, out_visited, out_cmps);
}

static int search_full_impl(pann_index* idx, const void* host_queries, uint64_t nq, uint64_t q_stride_bytes,
                            const uint32_t* self_ids, const uint32_t* starts, uint32_t nstarts, uint32_t starts_stride,
                            const uint32_t* nstarts_q, const pann_query_params* qp, uint32_t fcap, uint32_t* fr_ids,
                            float* fr_dists, uint32_t* n_frontier, uint32_t vcap, uint32_t* vis_ids, float* vis_dists,
                            uint32_t* n_visited, uint32_t* out_cmps) {
  if (!idx || !qp) PANN_FAIL("pann_search_full: null argument");
  if (!self_ids && !host_queries) PANN_FAIL("pann_search_full: need queries or self ids");
  PANN_CUDA(cudaSetDevice(idx->device));
  if (nq == 0) return 0;
  const uint8_t* d_q = nullptr;
  uint32_t* d_self = nullptr;
  size_t need = 4ull * nq * fcap * 2 + 4ull * nq * vcap * 2 + 4ull * nq * 4 + (self_ids ? 4ull * nq : 0) + 256;
  if (idx->misc.reserve(need)) return 1;
  uint8_t* m = idx->misc.as<uint8_t>();
  SearchOutputs o;
  o.fcap = fcap; o.vout = vcap;
  o.fr_ids = (uint32_t*)m; m += 4ull * nq * fcap;
  o.fr_dists = (float*)m; m += 4ull * nq * fcap;
  o.vis_ids = (uint32_t*)m; m += 4ull * nq * vcap;
  o.vis_dists = (float*)m; m += 4ull * nq * vcap;
  o.n_frontier = (uint32_t*)m; m += 4ull * nq;
  o.n_visited = (uint32_t*)m; m += 4ull * nq;
  o.cmps = (uint32_t*)m; m += 4ull * nq;
  o.status = (int*)m; m += 4ull * nq;
  if (self_ids) {
    for (uint64_t i = 0; i < nq; i++) if (self_ids[i] >= idx->n) PANN_FAIL("pann_search_full: self id out of range");
    d_self = (uint32_t*)m;
    PANN_CUDA(cudaMemcpyAsync(d_self, self_ids, 4ull * nq, cudaMemcpyHostToDevice, idx->stream));
  } else {
    if (pann_upload_queries(idx, host_queries, nq, q_stride_bytes)) return 1;
    d_q = idx->queries.as<uint8_t>();
  }
  int flag = 0;
  idx->last_launches = 0;
  if (starts_stride) {
    if (upload_starts(idx, nq, starts, starts_stride, nstarts_q, o)) return 1;
    nstarts = starts_stride;
  }
  if (pann_run_search(idx, d_q, nq, idx->geom.stride, d_self, starts, nstarts, qp, o, &flag)) return 1;
  if (check_status(flag & ~(1 << 2))) return 1;
  if (fr_ids) PANN_CUDA(cudaMemcpy(fr_ids, o.fr_ids, 4ull * nq * fcap, cudaMemcpyDeviceToHost));
  if (fr_dists) PANN_CUDA(cudaMemcpy(fr_dists, o.fr_dists, 4ull * nq * fcap, cudaMemcpyDeviceToHost));
  if (vis_ids) PANN_CUDA(cudaMemcpy(vis_ids, o.vis_ids, 4ull * nq * vcap, cudaMemcpyDeviceToHost));
  if (vis_dists) PANN_CUDA(cudaMemcpy(vis_dists, o.vis_dists, 4ull * nq * vcap, cudaMemcpyDeviceToHost));
  if (n_frontier) PANN_CUDA(cudaMemcpy(n_frontier, o.n_frontier, 4ull * nq, cudaMemcpyDeviceToHost));
  if (n_visited) PANN_CUDA(cudaMemcpy(n_visited, o.n_visited, 4ull * nq, cudaMemcpyDeviceToHost));
  if (out_cmps) PANN_CUDA(cudaMemcpy(out_cmps, o.cmps, 4ull * nq, cudaMemcpyDeviceToHost));
  return 0;
}

extern "C" int pann_search_full(pann_index* idx, const void* host_queries, uint64_t nq, uint64_t q_stride_bytes,
                                const uint32_t* self_ids, const uint32_t* starts, uint32_t nstarts,
                                const pann_query_params* qp, uint32_t fcap, uint32_t* fr_ids, float* fr_dists,
                                uint32_t* n_frontier, uint32_t vcap, uint32_t* vis_ids, float* vis_dists,
                                uint32_t* n_visited, uint32_t* out_cmps) {
  return search_full_impl(idx, host_queries, nq, q_stride_bytes, self_ids, starts, nstarts, 0, nullptr, qp, fcap, fr_ids, fr_dists,
                          n_frontier, vcap, vis_ids, vis_dists, n_visited, out_cmps);
}
extern "C" int pann_search_full_starts(pann_index* idx, const void* host_queries, uint64_t nq, uint64_t q_stride_bytes,
                                       const uint32_t* self_ids, const uint32_t* starts, uint32_t starts_stride,
                                       const uint32_t* nstarts_q, const pann_query_params* qp, uint32_t fcap, uint32_t* fr_ids,
                                       float* fr_dists, uint32_t* n_frontier, uint32_t vcap, uint32_t* vis_ids,
                                       float* vis_dists, uint32_t* n_visited, uint32_t* out_cmps) {
  if (!starts_stride) PANN_FAIL("beam search expects at least one start point");
  return search_full_impl(idx, host_queries, nq, q_stride_bytes, self_ids, starts, 0, starts_stride, nstarts_q, qp, fcap, fr_ids,
                          fr_dists, n_frontier, vcap, vis_ids, vis_dists, n_visited, out_cmps);
}

static const void* rerank_kernel(int dtype, int metric) {
  switch (dtype) {
    case PANN_DT_F32: return pann_rerank_kernel_dt0(metric);
    case PANN_DT_U8: return pann_rerank_kernel_dt1(metric);
    default: return pann_rerank_kernel_dt2(metric);
  }
}

extern "C" int pann_search_rerank_batch(pann_index* q_idx, pann_index* base_idx, const void* host_qqueries, uint64_t qq_stride_bytes,
                                        const void* host_queries, uint64_t q_stride_bytes, uint64_t nq, const uint32_t* starts,
                                        uint32_t nstarts, const pann_query_params* qp, int64_t rerank_factor, uint32_t* out_ids,
                                        float* out_dists, uint32_t* out_visited, uint32_t* out_cmps) {
  if (!q_idx || !base_idx || (nq && (!host_qqueries || !host_queries)) || !out_ids) PANN_FAIL("pann_search_rerank_batch: null argument");
  if (validate_qp(qp)) return 1;
  if (q_idx->device != base_idx->device) PANN_FAIL("pann_search_rerank_batch: both indexes must live on the same device");
  if (q_idx->n != base_idx->n || q_idx->dims != base_idx->dims) PANN_FAIL("pann_search_rerank_batch: the quantised and the exact point sets disagree on n or dims");
  if (q_idx->metric != base_idx->metric) PANN_FAIL("pann_search_rerank_batch: the two indexes use different metrics");
  if (rerank_factor < 1) PANN_FAIL("pann_search_rerank_batch: rerank_factor must be positive");
  if (qp->beam > 2048) PANN_FAIL("pann_search_rerank_batch: beam widths above 2048 are not supported on the re-rank path");
  PANN_CUDA(cudaSetDevice(q_idx->device));
  if (nq == 0) return 0;
  // 1. the search over the quantised copy, whole frontier kept on the device
  if (pann_upload_queries(q_idx, host_qqueries, nq, qq_stride_bytes)) return 1;
  if (pann_upload_queries(base_idx, host_queries, nq, q_stride_bytes)) return 1;
  PANN_CUDA(cudaStreamSynchronize(base_idx->stream));
  const uint32_t k = (uint32_t)qp->k, fcap = (uint32_t)qp->beam;
  size_t need = 4ull * nq * fcap * 2 + 4ull * nq * 4 + 8ull * nq * k + 256;
  if (q_idx->misc.reserve(need)) return 1;
  uint8_t* m = q_idx->misc.as<uint8_t>();
  SearchOutputs o;
  o.fcap = fcap;
  o.fr_ids = (uint32_t*)m; m += 4ull * nq * fcap;
  o.fr_dists = (float*)m; m += 4ull * nq * fcap;
  o.n_frontier = (uint32_t*)m; m += 4ull * nq;
  o.visited = (uint32_t*)m; m += 4ull * nq;
  o.cmps = (uint32_t*)m; m += 4ull * nq;
  o.status = (int*)m; m += 4ull * nq;
  uint32_t* d_ids = (uint32_t*)m; m += 4ull * nq * k;
  float* d_dists = (float*)m; m += 4ull * nq * k;
  int flag = 0;
  q_idx->last_launches = 0;
  PANN_CUDA(cudaEventRecord(q_idx->ev0, q_idx->stream));
  if (pann_run_search(q_idx, q_idx->queries.as<uint8_t>(), nq, q_idx->geom.stride, nullptr, starts, nstarts, qp, o, &flag)) return 1;
  if (check_status(flag)) return 1;   // includes "fewer than k results" (beamSearch.h:416-419)
  // 2. exact re-rank (beamSearch.h:426-444)
  RerankArgs R;
  memset(&R, 0, sizeof(R));
  R.points = base_idx->points; R.stride = base_idx->geom.stride; R.nch = base_idx->geom.nch;
  R.queries = base_idx->queries.as<uint8_t>(); R.qstride = base_idx->geom.stride; R.nq = nq;
  R.fr_ids = o.fr_ids; R.n_frontier = o.n_frontier; R.fcap = fcap;
  R.k = (int)k; R.rerank_factor = rerank_factor;
  R.out_ids = d_ids; R.out_dists = d_dists;
  R.off_q = al16((size_t)fcap * 8);
  R.off_stage = R.off_q + base_idx->geom.stride;
  R.warp_bytes = R.off_stage + (pann_staged_rows_rt(base_idx->dtype, 0) ? PANN_STAGE_BYTES : 0);
  const size_t cta_bytes = (size_t)R.warp_bytes * PANN_WARPS_PER_CTA;
  if (cta_bytes > q_idx->smem_optin) PANN_FAIL("pann_search_rerank_batch: rows too long for the re-rank kernel's shared memory");
  const void* fn = rerank_kernel(base_idx->dtype, base_idx->metric);
  PANN_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cta_bytes));
  const int grid = (int)std::min<uint64_t>((nq + PANN_WARPS_PER_CTA - 1) / PANN_WARPS_PER_CTA, (uint64_t)q_idx->sms * 8);
  void* args[] = {(void*)&R};
  PANN_CUDA(cudaLaunchKernel(fn, dim3(grid), dim3(PANN_WARPS_PER_CTA * 32), args, cta_bytes, q_idx->stream));
  q_idx->last_launches++;
  PANN_CUDA(cudaEventRecord(q_idx->ev1, q_idx->stream));
  PANN_CUDA(cudaMemcpyAsync(out_ids, d_ids, 4ull * nq * k, cudaMemcpyDeviceToHost, q_idx->stream));
  if (out_dists) PANN_CUDA(cudaMemcpyAsync(out_dists, d_dists, 4ull * nq * k, cudaMemcpyDeviceToHost, q_idx->stream));
  if (out_visited) PANN_CUDA(cudaMemcpyAsync(out_visited, o.visited, 4ull * nq, cudaMemcpyDeviceToHost, q_idx->stream));
  if (out_cmps) PANN_CUDA(cudaMemcpyAsync(out_cmps, o.cmps, 4ull * nq, cudaMemcpyDeviceToHost, q_idx->stream));
  PANN_CUDA(cudaStreamSynchronize(q_idx->stream));
  PANN_CUDA(cudaEventElapsedTime(&q_idx->last_ms, q_idx->ev0, q_idx->ev1));
  return 0;
}

extern "C" int pann_random_permutation(int64_t n, int32_t* out) {
  if (n < 0 || !out) PANN_FAIL("pann_random_permutation: bad argument");
  pann_host_random_permutation(n, out);
  return 0;
}
extern "C" int pann_random_starts(uint64_t nq, uint64_t n, uint32_t* out) {
  if (!out || n == 0 || n >= (1ull << 31)) PANN_FAIL("pann_random_starts: bad argument");
  pann_host_random_starts(nq, n, out);
  return 0;
}
extern "C" int64_t pann_batch_schedule(uint64_t n, uint64_t m, int64_t* lo, int64_t* hi, int64_t cap) {
  return pann_host_batch_schedule(n, m, lo, hi, cap);
}
