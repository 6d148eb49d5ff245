// harness.cu — the evaluation loop of the reference's search_and_parse (algorithms/utils/check_nn_recall.h:181-268) on the
// device: for every QueryParams of a sweep, the batched search over the resident query set, the tie-aware recall against a
// resident ground truth (checkRecall, :83-109) and the per-query counter statistics (stats<indexType>::statistics,
// stats.h:84-92: wrapped 32-bit sum / n and the 0.99-quantile element).  Results of all settings come back in ONE copy;
// no ids, distances or counters travel to the host in between.
#include <cub/device/device_radix_sort.cuh>

#include "pann_internal.h"

struct SweepAcc { int correct; uint32_t sum_visited, sum_cmps, pad; };

// one warp per query: ground-truth set = first k ids plus every later id whose distance equals the k-th's; count the
// reported ids that are in it
__global__ void pann_recall_kernel(const uint32_t* ids, const uint32_t* visited, const uint32_t* cmps, uint32_t k, const int32_t* gt_ids,
                                   const float* gt_d, uint32_t gtk, uint64_t nq, SweepAcc* acc) {
  const uint64_t q = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (q >= nq) return;
  const float last = gt_d[q * gtk + k - 1];
  int c = 0;
  for (uint32_t l = lane; l < gtk; l += 32) {
    if (l < k || gt_d[q * gtk + l] == last) {
      const uint32_t g = (uint32_t)gt_ids[q * gtk + l];
      bool hit = false;
      for (uint32_t j = 0; j < k; j++) hit |= ids[q * k + j] == g;
      c += hit;
    }
  }
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if (lane == 0) {
    atomicAdd(&acc->correct, c);
    atomicAdd(&acc->sum_visited, visited[q]);   // wraps like the reference's uint32 reduce
    atomicAdd(&acc->sum_cmps, cmps[q]);
  }
}

__global__ void pann_sweep_finish_kernel(const SweepAcc* acc, const uint32_t* sorted_visited, const uint32_t* sorted_cmps, uint64_t nq, uint32_t k,
                                         float device_ms, int status, pann_sweep_result* out) {
  if (threadIdx.x || blockIdx.x) return;
  const uint32_t tail = (uint32_t)(.99 * (double)(float)nq);                       // stats.h:89
  out->recall = (double)((float)acc->correct / (float)(k * nq));                   // check_nn_recall.h:108
  out->avg_cmps = (uint32_t)(long long)((double)(int)acc->sum_cmps / (double)nq);  // stats.h:88
  out->avg_visited = (uint32_t)(long long)((double)(int)acc->sum_visited / (double)nq);
  out->tail_cmps = sorted_cmps[tail < nq ? tail : nq - 1];
  out->tail_visited = sorted_visited[tail < nq ? tail : nq - 1];
  out->device_ms = device_ms;
  out->status = status;
}

extern "C" int pann_groundtruth_upload(pann_index* idx, const int32_t* gt_ids, const float* gt_dists, uint64_t nq, uint32_t gtk) {
  if (!idx || !gt_ids || !gt_dists || !nq || !gtk) PANN_FAIL("pann_groundtruth_upload: bad argument");
  PANN_CUDA(cudaSetDevice(idx->device));
  if (idx->gt_ids.reserve(4ull * nq * gtk) || idx->gt_dists.reserve(4ull * nq * gtk)) return 1;
  PANN_CUDA(cudaMemcpyAsync(idx->gt_ids.p, gt_ids, 4ull * nq * gtk, cudaMemcpyHostToDevice, idx->stream));
  PANN_CUDA(cudaMemcpyAsync(idx->gt_dists.p, gt_dists, 4ull * nq * gtk, cudaMemcpyHostToDevice, idx->stream));
  PANN_CUDA(cudaStreamSynchronize(idx->stream));
  idx->gt_nq = nq; idx->gt_k = gtk;
  return 0;
}

extern "C" int pann_search_sweep(pann_index* idx, const uint32_t* starts, uint32_t nstarts, const pann_query_params* qps, uint32_t nset,
                                 pann_sweep_result* out) {
  if (!idx || !qps || !out || !nset) PANN_FAIL("pann_search_sweep: bad argument");
  PANN_CUDA(cudaSetDevice(idx->device));
  const uint64_t nq = idx->nq;
  if (nq == 0) PANN_FAIL("pann_search_sweep: no resident query set");
  if (idx->gt_nq != nq || !idx->gt_ids.p) PANN_FAIL("pann_search_sweep: no resident ground truth for this query set (pann_groundtruth_resident or pann_groundtruth_upload)");
  cudaStream_t s = idx->stream;
  size_t tmp_bytes = 0;
  cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int64_t)nq, 0, 32, s);
  DevBuf& W = idx->sweep;
  const size_t need = sizeof(pann_sweep_result) * nset + sizeof(SweepAcc) * nset + 8ull * nq + tmp_bytes + 256;
  if (W.reserve(need)) return 1;
  uint8_t* p = W.as<uint8_t>();
  pann_sweep_result* d_out = (pann_sweep_result*)p; p += (sizeof(pann_sweep_result) * nset + 15) / 16 * 16;
  SweepAcc* d_acc = (SweepAcc*)p; p += sizeof(SweepAcc) * nset;
  uint32_t* s_vis = (uint32_t*)p; p += 4ull * nq;
  uint32_t* s_cmp = (uint32_t*)p; p += 4ull * nq;
  void* d_tmp = p;
  PANN_CUDA(cudaMemsetAsync(d_acc, 0, sizeof(SweepAcc) * nset, s));
  PANN_CUDA(cudaMemsetAsync(d_out, 0, sizeof(pann_sweep_result) * nset, s));
  int launches = 0;
  for (uint32_t i = 0; i < nset; i++) {
    const pann_query_params& qp = qps[i];
    if (qp.k < 1 || (uint64_t)qp.k > idx->gt_k) PANN_FAIL(std::to_string(qp.k) + "@" + std::to_string(qp.k) + " too large for ground truth data of size " + std::to_string(idx->gt_k));  // check_nn_recall.h:33-37
    int status = 0;
    if (pann_search_resident(idx, starts, nstarts, &qp)) {   // a setting the reference would abort on (k > beam, fewer than k results): reported per setting
      status = 1;
      pann_sweep_finish_kernel<<<1, 32, 0, s>>>(d_acc + i, s_vis, s_cmp, 1, (uint32_t)qp.k, 0.f, status, d_out + i);
      continue;
    }
    launches += idx->last_launches;
    const uint32_t *d_ids, *d_vis, *d_cmps;
    uint64_t rn; uint32_t rk;
    if (pann_search_device_result(idx, &d_ids, nullptr, &d_vis, &d_cmps, &rn, &rk)) return 1;
    pann_recall_kernel<<<(unsigned)((nq * 32 + 255) / 256), 256, 0, s>>>(d_ids, d_vis, d_cmps, rk, idx->gt_ids.as<int32_t>(), idx->gt_dists.as<float>(),
                                                                       idx->gt_k, nq, d_acc + i);
    size_t tb = tmp_bytes;
    cub::DeviceRadixSort::SortKeys(d_tmp, tb, d_vis, s_vis, (int64_t)nq, 0, 32, s);
    tb = tmp_bytes;
    cub::DeviceRadixSort::SortKeys(d_tmp, tb, d_cmps, s_cmp, (int64_t)nq, 0, 32, s);
    pann_sweep_finish_kernel<<<1, 32, 0, s>>>(d_acc + i, s_vis, s_cmp, nq, rk, idx->last_ms, 0, d_out + i);
    launches += 2;
  }
  PANN_CUDA(cudaMemcpyAsync(out, d_out, sizeof(pann_sweep_result) * nset, cudaMemcpyDeviceToHost, s));
  PANN_CUDA(cudaStreamSynchronize(s));
  PANN_CUDA(cudaGetLastError());
  idx->last_launches = launches;
  return 0;
}
