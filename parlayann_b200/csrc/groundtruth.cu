// groundtruth.cu — K4: brute-force exact k-NN (data_tools/compute_groundtruth.cpp:22-60 + write_ibin's per-row
// sort :68-73), and K5: the k-way merge of per-shard results.
//
// v1 layout: the base set is walked in chunks; per chunk
//   pann_gt_dist_kernel    64 queries x 64 base rows per CTA, 4x4 register tile per thread, operands staged
//                          k-major in shared memory; 8-bit types accumulate exactly in int32 (dp4a), floats with
//                          one rounded multiply and one rounded add per dimension in ascending order — the same
//                          operation sequence the reference's scalar loop performs (euclidian_point.h:83-90,
//                          mips_point.h:59-65), so float distances are bit-identical to the CPU's;
//   pann_gt_select_kernel  one CTA per query: threshold-filtered scan of its distance row into a shared-memory
//                          candidate buffer, bitonic shrink to the best k when the buffer fills and at chunk end.
// Results are rows sorted by (distance, id).
#include "pann_internal.h"
#include "gt_mma.cuh"
#include <algorithm>

#define GT_TQ 64
#define GT_TB 64
#define GT_KW 32        // 32-bit words of each row staged per step (128 bytes)
#define GT_SEL_THREADS 256
#define GT_SEL_CAP 2048  // candidate buffer (keys) per query CTA

template <int DT, int METRIC>
__global__ void __launch_bounds__(256) pann_gt_dist_kernel(const uint8_t* __restrict__ base, uint64_t b0, uint32_t nbc,
                                                           const uint8_t* __restrict__ queries, uint32_t nq, uint32_t stride,
                                                           uint32_t nwords, float* __restrict__ D, uint32_t ldD) {
  __shared__ __align__(16) uint32_t Qs[GT_KW][GT_TQ];
  __shared__ __align__(16) uint32_t Bs[GT_KW][GT_TB];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const uint32_t q0 = blockIdx.y * GT_TQ, r0 = blockIdx.x * GT_TB;
  float facc[4][4];
  int iacc[4][4];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) { facc[i][j] = 0.f; iacc[i][j] = 0; }
  // loader mapping: thread t stages 16 bytes: row (t & 63), 16-byte chunk (t >> 6) and (t >> 6) + 4 of the step
  const int lrow = threadIdx.x & 63, lch = threadIdx.x >> 6;
  const uint32_t qrow = min(q0 + lrow, nq - 1), brow = min(r0 + lrow, nbc - 1);
  const uint8_t* qp = queries + (uint64_t)qrow * stride;
  const uint8_t* bp = base + (b0 + brow) * (uint64_t)stride;
  for (uint32_t w0 = 0; w0 < nwords; w0 += GT_KW) {
#pragma unroll
    for (int h = 0; h < 2; h++) {
      int c = lch + 4 * h;  // chunk within the step: words 4c .. 4c+3
      uint32_t w = w0 + 4 * c;
      uint4 qv = make_uint4(0, 0, 0, 0), bv = make_uint4(0, 0, 0, 0);
      if (w < nwords) { qv = __ldg(reinterpret_cast<const uint4*>(qp + 4ull * w)); bv = __ldg(reinterpret_cast<const uint4*>(bp + 4ull * w)); }
      Qs[4 * c + 0][lrow] = qv.x; Qs[4 * c + 1][lrow] = qv.y; Qs[4 * c + 2][lrow] = qv.z; Qs[4 * c + 3][lrow] = qv.w;
      Bs[4 * c + 0][lrow] = bv.x; Bs[4 * c + 1][lrow] = bv.y; Bs[4 * c + 2][lrow] = bv.z; Bs[4 * c + 3][lrow] = bv.w;
    }
    __syncthreads();
    const int kw_end = min((uint32_t)GT_KW, nwords - w0);
#pragma unroll 4
    for (int kw = 0; kw < kw_end; kw++) {
      const uint4 q4 = *reinterpret_cast<const uint4*>(&Qs[kw][ty * 4]);
      const uint4 b4 = *reinterpret_cast<const uint4*>(&Bs[kw][tx * 4]);
      const uint32_t qa[4] = {q4.x, q4.y, q4.z, q4.w}, ba[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) {
          if constexpr (DT == PANN_DT_F32) {
            const float qf = __uint_as_float(qa[i]), bf = __uint_as_float(ba[j]);
            if constexpr (METRIC == PANN_M_L2) {
              const float t = __fsub_rn(qf, bf);
              facc[i][j] = __fadd_rn(facc[i][j], __fmul_rn(t, t));
            } else {
              facc[i][j] = __fadd_rn(facc[i][j], __fmul_rn(qf, bf));
            }
          } else if constexpr (DT == PANN_DT_U8) {
            if constexpr (METRIC == PANN_M_L2) { unsigned x = __vabsdiffu4(qa[i], ba[j]); iacc[i][j] = (int)__dp4a(x, x, (unsigned)iacc[i][j]); }
            else iacc[i][j] = (int)__dp4a(qa[i], ba[j], (unsigned)iacc[i][j]);
          } else {
            if constexpr (METRIC == PANN_M_L2) { unsigned x = __vabsdiffs4(qa[i], ba[j]); iacc[i][j] = (int)__dp4a(x, x, (unsigned)iacc[i][j]); }
            else iacc[i][j] = __dp4a((int)qa[i], (int)ba[j], iacc[i][j]);
          }
        }
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const uint32_t q = q0 + ty * 4 + i;
    if (q >= nq) continue;
    float o[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      float v = (DT == PANN_DT_F32) ? facc[i][j] : (float)iacc[i][j];
      o[j] = (METRIC == PANN_M_MIPS) ? -v : v;
    }
    float* dst = D + (uint64_t)q * ldD + r0 + tx * 4;
    if (r0 + tx * 4 + 3 < nbc && (ldD & 3) == 0) *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
    else
      for (int j = 0; j < 4; j++) if (r0 + tx * 4 + j < nbc) dst[j] = o[j];
  }
}

// in-place ascending bitonic sort of `m` (power of two) keys in shared memory by the whole CTA
__device__ __forceinline__ void gt_bitonic(u64* a, int m) {
  for (int k = 2; k <= m; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = threadIdx.x; t < m; t += blockDim.x) {
        int p = t ^ j;
        if (p > t) {
          u64 x = a[t], y = a[p];
          bool up = (t & k) == 0;
          if ((x > y) == up) { a[t] = y; a[p] = x; }
        }
      }
      __syncthreads();
    }
}

// state per query: K keys (sorted, unused slots = ~0) in `state`; D row of this chunk: nbc distances of base rows b0..
__global__ void __launch_bounds__(GT_SEL_THREADS) pann_gt_select_kernel(const float* __restrict__ D, uint32_t ldD, uint32_t nbc,
                                                                        uint64_t b0, uint32_t k, u64* __restrict__ state) {
  __shared__ u64 buf[GT_SEL_CAP];
  __shared__ int s_count;
  __shared__ u64 s_tau;
  const uint32_t q = blockIdx.x;
  u64* st = state + (uint64_t)q * k;
  for (int i = threadIdx.x; i < GT_SEL_CAP; i += blockDim.x) buf[i] = i < (int)k ? st[i] : ~0ull;
  if (threadIdx.x == 0) { s_count = (int)k; s_tau = st[k - 1]; }
  __syncthreads();
  const float* row = D + (uint64_t)q * ldD;
  const int SEG = GT_SEL_THREADS * 4;  // at most SEG appends between capacity checks
  for (uint32_t s0 = 0; s0 < nbc; s0 += SEG) {
    const u64 tau = s_tau;
#pragma unroll
    for (int u = 0; u < 4; u++) {
      uint32_t j = s0 + u * GT_SEL_THREADS + threadIdx.x;
      if (j < nbc) {
        u64 key = pann_key(row[j], (uint32_t)(b0 + j));
        if (key < tau) buf[atomicAdd(&s_count, 1)] = key;
      }
    }
    __syncthreads();
    const int cnt = s_count;  // read once between two barriers: the next segment's appends must not be seen by some threads only
    __syncthreads();
    if (cnt > GT_SEL_CAP - SEG) {  // shrink: keep the best k, tighten the threshold
      for (int i = cnt + threadIdx.x; i < GT_SEL_CAP; i += blockDim.x) buf[i] = ~0ull;
      __syncthreads();
      gt_bitonic(buf, GT_SEL_CAP);
      if (threadIdx.x == 0) { s_count = (int)k; s_tau = buf[k - 1]; }
      __syncthreads();
    }
  }
  {
    int cnt = s_count;
    for (int i = cnt + threadIdx.x; i < GT_SEL_CAP; i += blockDim.x) buf[i] = ~0ull;
    __syncthreads();
    int m = 2; while (m < cnt) m <<= 1;
    gt_bitonic(buf, m);
    for (int i = threadIdx.x; i < (int)k; i += blockDim.x) st[i] = buf[i];
  }
}

__global__ void pann_gt_init_kernel(u64* state, uint64_t total) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < total) state[i] = ~0ull;
}
__global__ void pann_gt_finish_kernel(const u64* state, uint64_t total, int64_t id_offset, int32_t* ids, float* dists) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  u64 key = state[i];
  if (key == ~0ull) { ids[i] = -1; dists[i] = __int_as_float(0x7f800000); return; }   // empty slot (bounded runs only): sorts last in K5
  ids[i] = (int32_t)((int64_t)pann_key_id(key) + id_offset);
  dists[i] = pann_key_dist(key);
}

// K5: one CTA per query, the union of S sorted shard rows -> best k under (dist, id).  ids are global here.
__global__ void __launch_bounds__(GT_SEL_THREADS) pann_gt_merge_kernel(const int32_t* ids, const float* dists, uint32_t S, uint64_t nq,
                                                                       uint32_t k, int32_t* out_ids, float* out_dists) {
  __shared__ u64 buf[GT_SEL_CAP];
  const uint64_t q = blockIdx.x;
  const int tot = (int)(S * k);
  int m = 2; while (m < tot) m <<= 1;
  for (int i = threadIdx.x; i < m; i += blockDim.x) {
    u64 key = ~0ull;
    if (i < tot) { uint64_t src = ((uint64_t)(i / k) * nq + q) * k + (i % k); if (ids[src] >= 0) key = pann_key(dists[src], (uint32_t)ids[src]); }
    buf[i] = key;
  }
  __syncthreads();
  gt_bitonic(buf, m);
  for (int i = threadIdx.x; i < (int)k; i += blockDim.x) {
    const bool empty = buf[i] == ~0ull;
    out_ids[q * k + i] = empty ? -1 : (int32_t)pann_key_id(buf[i]);
    out_dists[q * k + i] = empty ? __int_as_float(0x7f800000) : pann_key_dist(buf[i]);
  }
}

// ---------------------------------------------------------------------------------------------
template <int DT>
static void launch_dist(int metric, dim3 grid, cudaStream_t s, const uint8_t* base, uint64_t b0, uint32_t nbc, const uint8_t* q,
                        uint32_t nq, uint32_t stride, uint32_t nwords, float* D, uint32_t ldD) {
  if (metric == PANN_M_L2) pann_gt_dist_kernel<DT, PANN_M_L2><<<grid, 256, 0, s>>>(base, b0, nbc, q, nq, stride, nwords, D, ldD);
  else pann_gt_dist_kernel<DT, PANN_M_MIPS><<<grid, 256, 0, s>>>(base, b0, nbc, q, nq, stride, nwords, D, ldD);
}

// ---------------------------------------------------------------------------------------------
// K4 v2 host side (gt_mma.cuh): TMA descriptors, chunk schedule, overflow fallback
#include <cudaTypedefs.h>
static PFN_cuTensorMapEncodeTiled_v12000 gt_encode_fn() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  }
  return fn;
}
// rows of `stride` bytes seen as a 2-D byte tensor; boxes of 128 bytes x box_rows, 128-byte swizzle, zero fill outside
static int gt_make_tmap(CUtensorMap* tm, const void* ptr, uint64_t rows, uint32_t stride, uint32_t box_rows) {
  auto fn = gt_encode_fn();
  if (!fn) PANN_FAIL("pann_groundtruth: cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t dims[2] = {stride, rows};
  cuuint64_t strides[1] = {stride};
  cuuint32_t box[2] = {GTM_KB_BYTES, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) PANN_FAIL("pann_groundtruth: cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
  return 0;
}

static int gt_shrink_threads() { const char* e = getenv("PANN_GT_SHRINK_THREADS"); return e ? atoi(e) : 128; }
template <int DT, int METRIC>
static void gt_launch_shrink(pann_index* idx, cudaStream_t s, const uint8_t* d_q, uint32_t nq, uint32_t k, u64* state, const uint32_t* cand_ids,
                             uint32_t* cand_cnt, uint32_t nsub, uint32_t subcap, uint32_t keys_cap, const void* qnorm, const float* bnmax, void* thr, int* overflow, const float* cap) {
  const int threads = keys_cap == GTS_KEYS_SMALL ? gt_shrink_threads() : 256;
  const size_t stage_b = DT == PANN_DT_F32 ? (size_t)(threads / 32) * PANN_STAGE_BYTES : 0;   // per-warp row staging (float rows)
  const size_t sm = 8ull * keys_cap + idx->geom.stride + stage_b;
  cudaFuncSetAttribute(pann_gt_shrink_kernel<DT, METRIC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(8ull * GTS_KEYS + idx->geom.stride + 8 * PANN_STAGE_BYTES));
  cudaFuncSetAttribute(pann_gt_shrink_kernel<DT, METRIC>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  pann_gt_shrink_kernel<DT, METRIC><<<nq, threads, sm, s>>>(idx->points, idx->geom.stride, idx->geom.nch, d_q, k, state, cand_ids, cand_cnt, nsub, subcap, keys_cap, qnorm, bnmax, thr, overflow, cap);
}
template <int DT, int METRIC>
static void gt_launch_thr(cudaStream_t s, const u64* state, uint32_t nq, uint32_t k, const void* qnorm, const float* bnmax, void* thr, const float* cap) {
  pann_gt_thr_kernel<DT, METRIC><<<(nq + 255) / 256, 256, 0, s>>>(state, nq, k, qnorm, bnmax, thr, cap);
}
#define GT_DISPATCH(idx, CALL)                                                                              \
  do {                                                                                                      \
    const bool l2__ = (idx)->metric == PANN_M_L2;                                                           \
    if ((idx)->dtype == PANN_DT_F32) { if (l2__) CALL(PANN_DT_F32, PANN_M_L2); else CALL(PANN_DT_F32, PANN_M_MIPS); } \
    else if ((idx)->dtype == PANN_DT_U8) { if (l2__) CALL(PANN_DT_U8, PANN_M_L2); else CALL(PANN_DT_U8, PANN_M_MIPS); } \
    else { if (l2__) CALL(PANN_DT_I8, PANN_M_L2); else CALL(PANN_DT_I8, PANN_M_MIPS); }                   \
  } while (0)

// v1 over base rows [lo, hi): distance chunks + select into `state`
static void gt_v1_range(pann_index* idx, cudaStream_t s, const uint8_t* d_q, uint64_t nq, uint32_t k, u64* state, float* D, uint64_t cb,
                        uint64_t lo, uint64_t hi) {
  const uint32_t nwords = idx->geom.stride / 4;
  for (uint64_t b0 = lo; b0 < hi; b0 += cb) {
    const uint32_t nbc = (uint32_t)std::min<uint64_t>(cb, hi - b0);
    dim3 grid((nbc + GT_TB - 1) / GT_TB, (unsigned)((nq + GT_TQ - 1) / GT_TQ));
    if (idx->dtype == PANN_DT_F32) launch_dist<PANN_DT_F32>(idx->metric, grid, s, idx->points, b0, nbc, d_q, (uint32_t)nq, idx->geom.stride, nwords, D, (uint32_t)cb);
    else if (idx->dtype == PANN_DT_U8) launch_dist<PANN_DT_U8>(idx->metric, grid, s, idx->points, b0, nbc, d_q, (uint32_t)nq, idx->geom.stride, nwords, D, (uint32_t)cb);
    else launch_dist<PANN_DT_I8>(idx->metric, grid, s, idx->points, b0, nbc, d_q, (uint32_t)nq, idx->geom.stride, nwords, D, (uint32_t)cb);
    pann_gt_select_kernel<<<(unsigned)nq, GT_SEL_THREADS, 0, s>>>(D, (uint32_t)cb, nbc, b0, k, state);
    idx->last_launches += 2;
  }
}
// ground truth of the device-resident query rows d_q against the index's base rows; results stay in idx->gt_*.
// Sharded jobs run it in two legs: rows [0, s) (the sample), then — after the shards have merged their
// samples — rows [s, n) continuing from the kept state with d_cap = per-query upper bounds of the FINAL k-th distance
// (device, nq floats; every threshold is min(own k-th, cap)) and seen_rows = the number of rows that bound was computed
// over (it sets the chunk schedule: survivors of a chunk of c rows ~ k c / rows seen).  The cap only tightens the
// superset filter; the kept keys are exact, so the merged result does not change.
struct GtLeg { uint64_t row_lo = 0, row_hi = 0; bool keep_state = false; const float* d_cap = nullptr; uint64_t seen_rows = 0; };
static int gt_run(pann_index* idx, const uint8_t* d_q, uint64_t nq, uint32_t k, int64_t id_offset, const GtLeg& leg = GtLeg()) {
  if (k == 0 || k > GT_SEL_CAP / 2) PANN_FAIL("pann_groundtruth: k must be in 1.." + std::to_string(GT_SEL_CAP / 2));
  const bool whole = leg.row_lo == 0 && leg.row_hi == 0 && !leg.keep_state;
  if (idx->n < k && whole) PANN_FAIL("pann_groundtruth: fewer base points than k");
  if (nq == 0) { idx->gt_nq = 0; idx->gt_k = k; return 0; }
  if ((int64_t)idx->n + id_offset > (int64_t)INT32_MAX) PANN_FAIL("pann_groundtruth: ids do not fit the .ibin format's int32");
  const uint64_t nb = leg.row_hi ? std::min<uint64_t>(leg.row_hi, idx->n) : idx->n;   // rows [row0, nb)
  const uint64_t row0 = leg.row_lo;
  if (row0 > nb) PANN_FAIL("pann_groundtruth: row range out of order");
  if (row0 % GTM_N) PANN_FAIL("pann_groundtruth: row_lo must be a multiple of " + std::to_string(GTM_N) + " (the MMA tile)");
  if (leg.keep_state && (idx->gt_nq != nq || idx->gt_k != k || idx->gt_scratch[0].cap < 8ull * nq * k))
    PANN_FAIL("pann_groundtruth: no kept state of this shape to continue from (run the sample leg first)");
  const float* d_cap = leg.d_cap;
  const uint64_t seen_rows = leg.seen_rows;
  cudaStream_t s = idx->stream;
  const uint32_t stride = idx->geom.stride;
  const bool isint = idx->dtype != PANN_DT_F32;

  // ---- plan: tensor-core path (K4 v2) for everything after a seed chunk, CUDA-core path (v1) for the seed chunk,
  //      for shapes the MMA kernel does not take, and for a chunk whose candidate lists overflowed
  const uint32_t kb = (stride + GTM_KB_BYTES - 1) / GTM_KB_BYTES;
  int nstages = (int)((212 * 1024 - (long long)kb * GTM_A_STAGE_BYTES) / GTM_B_STAGE_BYTES);
  nstages = std::min(nstages, GTM_STAGES);
  uint64_t seed = std::max<uint64_t>(512, ((2ull * k + GTM_N - 1) / GTM_N) * GTM_N);  // sample that sets the first thresholds
  if (const char* e = getenv("PANN_GT_SEED")) seed = std::max<uint64_t>(((uint64_t)atoll(e) + GTM_N - 1) / GTM_N * GTM_N, ((2ull * k + GTM_N - 1) / GTM_N) * GTM_N);
  const char* force_v1 = getenv("PANN_GT_V1");
  bool use_mma = !(force_v1 && force_v1[0] == '1') && stride >= GTM_KB_BYTES && nstages >= 2 && nb - row0 >= 2 * seed && nq < (1ull << 31);
  // growth of the MMA chunks: expected survivors of a chunk of c rows after s rows have been seen ~ k c / s
  double growth = std::min(4.0, 1.0 + (double)GTM_CAND_CAP / (3.0 * k));
  if (const char* e = getenv("PANN_GT_GROWTH")) growth = std::max(1.25, std::min(growth, atof(e)));

  // v1 distance chunk (~1 GiB for the whole job, only the seed rows when the MMA path runs)
  uint64_t cb = ((1ull << 28) / nq) / GT_TB * GT_TB;
  cb = std::max<uint64_t>(cb, 4096);
  cb = std::min<uint64_t>(cb, (nb + GT_TB - 1) / GT_TB * GT_TB);
  // the (up to 1 GiB) v1 distance block is per call; everything else lives in the index and is reused by later calls
  DevBuf D;
  DevBuf &state = idx->gt_scratch[0], &state_bak = idx->gt_scratch[1], &bnorm = idx->gt_scratch[2], &qnorm = idx->gt_scratch[3],
         &thr = idx->gt_scratch[4], &cand_ids = idx->gt_scratch[5], &cand_cnt = idx->gt_scratch[6], &misc = idx->gt_scratch[7];
  auto fail = [&]() { D.release(); return 1; };
  if ((!use_mma && D.reserve(4ull * nq * cb)) || state.reserve(8ull * nq * k)) return fail();   // the MMA path needs D only for a fallback
  if (idx->gt_ids.reserve(4ull * nq * k) || idx->gt_dists.reserve(4ull * nq * k)) return fail();
  const uint64_t nb_pad = (nb + GTM_N - 1) / GTM_N * GTM_N;
  CUtensorMap tmQ, tmB;
  if (use_mma) {
    if (state_bak.reserve(8ull * nq * k) || bnorm.reserve(4ull * nb_pad) || qnorm.reserve(4ull * nq) || thr.reserve(4ull * nq) ||
        cand_ids.reserve(4ull * nq * GTM_CAND_CAP) || cand_cnt.reserve(4ull * nq * GTM_MAX_SUB) || misc.reserve(16)) return fail();
    if (gt_make_tmap(&tmQ, d_q, nq, stride, GTM_M) || gt_make_tmap(&tmB, idx->points, nb, stride, GTM_N)) return fail();
  }
  int* h_flag = nullptr;
  if (use_mma && pann_pinned(idx, 64, (void**)&h_flag)) return fail();
  const uint64_t tot = nq * k;
  idx->gt_mma_launches = 0;
  idx->last_launches = 0;
  cudaEventRecord(idx->ev0, s);
  if (!leg.keep_state) {
    pann_gt_init_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, s>>>(state.as<u64>(), tot);
    idx->last_launches++;
  }
  if (!use_mma) {
    if (nb > row0) gt_v1_range(idx, s, d_q, nq, k, state.as<u64>(), D.as<float>(), cb, row0, nb);
  } else {
    float* bnmax = misc.as<float>();
    int* overflow = misc.as<int>() + 1;
    cudaMemsetAsync(misc.p, 0, 16, s);
    cudaMemsetAsync(cand_cnt.p, 0, 4ull * nq * GTM_MAX_SUB, s);
    const int store_zero = idx->metric == PANN_M_MIPS ? 1 : 0;
    if (idx->dtype == PANN_DT_F32) {
      pann_gt_norm_kernel<PANN_DT_F32><<<(unsigned)((nb_pad + 255) / 256), 256, 0, s>>>(idx->points, nb, nb_pad, stride, idx->geom.nch, bnorm.p, store_zero, bnmax, 1);
      pann_gt_norm_kernel<PANN_DT_F32><<<(unsigned)((nq + 255) / 256), 256, 0, s>>>(d_q, nq, nq, stride, idx->geom.nch, qnorm.p, 0, nullptr, 0);
    } else if (idx->dtype == PANN_DT_U8) {
      pann_gt_norm_kernel<PANN_DT_U8><<<(unsigned)((nb_pad + 255) / 256), 256, 0, s>>>(idx->points, nb, nb_pad, stride, idx->geom.nch, bnorm.p, store_zero, nullptr, 1);
      pann_gt_norm_kernel<PANN_DT_U8><<<(unsigned)((nq + 255) / 256), 256, 0, s>>>(d_q, nq, nq, stride, idx->geom.nch, qnorm.p, 0, nullptr, 0);
    } else {
      pann_gt_norm_kernel<PANN_DT_I8><<<(unsigned)((nb_pad + 255) / 256), 256, 0, s>>>(idx->points, nb, nb_pad, stride, idx->geom.nch, bnorm.p, store_zero, nullptr, 1);
      pann_gt_norm_kernel<PANN_DT_I8><<<(unsigned)((nq + 255) / 256), 256, 0, s>>>(d_q, nq, nq, stride, idx->geom.nch, qnorm.p, 0, nullptr, 0);
    }
    idx->last_launches += 2;
    // first thresholds: "everything passes" — the first chunk is just the first `seed` rows, all of which become
    // candidates; from then on the chunks grow geometrically and the thresholds do the filtering
#define GT_THR(DT, M) gt_launch_thr<DT, M>(s, state.as<u64>(), (uint32_t)nq, k, qnorm.p, bnmax, thr.p, d_cap)
    if (d_cap || leg.keep_state) GT_DISPATCH(idx, GT_THR);    // from the kept state and / or the caps
    else if (isint) pann_gt_thr_init_kernel<PANN_DT_U8><<<(unsigned)((nq + 255) / 256), 256, 0, s>>>((uint32_t)nq, thr.p);
    else pann_gt_thr_init_kernel<PANN_DT_F32><<<(unsigned)((nq + 255) / 256), 256, 0, s>>>((uint32_t)nq, thr.p);
    idx->last_launches++;
    // instruction descriptor: D = s32 / f32, A and B = u8 / s8 / tf32, K-major both, N = 256, M = 128
    uint32_t idesc = 0;
    if (isint) { const uint32_t f = idx->dtype == PANN_DT_I8 ? 1u : 0u; idesc = (2u << 4) | (f << 7) | (f << 10); }
    else idesc = (1u << 4) | (2u << 7) | (2u << 10);
    idesc |= (uint32_t)(GTM_N >> 3) << 17;
    idesc |= (uint32_t)(GTM_M >> 4) << 24;
    const size_t smem = 1024 + (size_t)kb * GTM_A_STAGE_BYTES + (size_t)nstages * GTM_B_STAGE_BYTES + 256 + GTM_BN * GTM_N * 4;
    const void* kfn = isint ? (const void*)pann_gt_mma_kernel<true> : (const void*)pann_gt_mma_kernel<false>;
    if (cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
      fail(); PANN_FAIL("pann_groundtruth: cannot reserve shared memory for the MMA kernel");
    }
    const uint32_t nqb = (uint32_t)((nq + GTM_M - 1) / GTM_M);
    const bool cold = !d_cap && !leg.keep_state;   // thresholds still "everything passes": the first chunk is the seed
    uint64_t lo = row0;
    while (lo < nb) {
      const uint64_t seen = lo - row0 + (cold ? 0 : std::max<uint64_t>(seen_rows, leg.keep_state ? row0 : 0));
      uint64_t c = seen == 0 ? seed : (uint64_t)((growth - 1.0) * (double)seen);
      c = std::max<uint64_t>(c / GTM_N * GTM_N, GTM_N);
      const uint64_t hi = std::min<uint64_t>(nb, lo + c);
      const uint64_t tiles = (hi - lo + GTM_N - 1) / GTM_N;
      uint64_t S = std::max<uint64_t>(1, std::min<uint64_t>(std::min<uint64_t>(tiles, GTM_MAX_SUB / (GTM_EPI_WARPS / 4)), (4ull * idx->sms + nqb - 1) / nqb));
      const uint32_t tps = (uint32_t)((tiles + S - 1) / S);
      S = (tiles + tps - 1) / tps;
      cudaMemcpyAsync(state_bak.p, state.p, 8ull * nq * k, cudaMemcpyDeviceToDevice, s);
      GtMmaArgs A;
      A.nq = (uint32_t)nq; A.kb = kb; A.nstages = (uint32_t)nstages; A.row_lo = lo; A.row_hi = hi; A.tiles_per_seg = tps;
      A.bnorm = bnorm.p; A.thr = thr.p; A.cand_ids = cand_ids.as<uint32_t>(); A.cand_cnt = cand_cnt.as<uint32_t>();
      { const char* e = getenv("PANN_GT_DBG"); A.dbg = e ? (uint32_t)atoi(e) : 0u; }
      A.subcap = (uint32_t)(GTM_CAND_CAP / ((GTM_EPI_WARPS / 4) * S));
      if (isint) pann_gt_mma_kernel<true><<<dim3(nqb, (unsigned)S), GTM_THREADS, smem, s>>>(tmQ, tmB, A, idesc);
      else pann_gt_mma_kernel<false><<<dim3(nqb, (unsigned)S), GTM_THREADS, smem, s>>>(tmQ, tmB, A, idesc);
      idx->gt_mma_launches++;
#define GT_SHRINK(DT, M) gt_launch_shrink<DT, M>(idx, s, d_q, (uint32_t)nq, k, state.as<u64>(), cand_ids.as<uint32_t>(), cand_cnt.as<uint32_t>(), (uint32_t)((GTM_EPI_WARPS / 4) * S), A.subcap, keys_cap, qnorm.p, bnmax, thr.p, overflow, d_cap)
      // small tier first (8 KB of key slots, many CTAs per SM); queries with more keys raise bit 1 and get the large tier
      uint32_t keys_cap = (k + (seen == 0 ? seed : 0) <= GTS_KEYS_SMALL / 2) ? GTS_KEYS_SMALL : GTS_KEYS;
      GT_DISPATCH(idx, GT_SHRINK);
      idx->last_launches += 2;
      cudaMemcpyAsync(h_flag, overflow, sizeof(int), cudaMemcpyDeviceToHost, s);
      cudaError_t e = cudaStreamSynchronize(s);
      if (e != cudaSuccess) { fail(); PANN_FAIL(std::string("pann_groundtruth (MMA path): ") + cudaGetErrorString(e)); }
      if (*h_flag == 2 && keys_cap != GTS_KEYS) {
        cudaMemsetAsync(overflow, 0, sizeof(int), s);
        keys_cap = GTS_KEYS;
        GT_DISPATCH(idx, GT_SHRINK);
        idx->last_launches++;
        cudaMemcpyAsync(h_flag, overflow, sizeof(int), cudaMemcpyDeviceToHost, s);
        e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) { fail(); PANN_FAIL(std::string("pann_groundtruth (MMA path): ") + cudaGetErrorString(e)); }
      }
      if (*h_flag) {
        // some query passed more than GTM_CAND_CAP rows of this chunk (rows ordered so that the threshold stays loose):
        // take the chunk again on the CUDA cores from the saved state
        cudaMemcpyAsync(state.p, state_bak.p, 8ull * nq * k, cudaMemcpyDeviceToDevice, s);
        cudaMemsetAsync(overflow, 0, sizeof(int), s);
        cudaMemsetAsync(cand_cnt.p, 0, 4ull * nq * GTM_MAX_SUB, s);
        if (D.reserve(4ull * nq * cb)) return fail();
        gt_v1_range(idx, s, d_q, nq, k, state.as<u64>(), D.as<float>(), cb, lo, hi);
        GT_DISPATCH(idx, GT_THR);
        idx->last_launches++;
        idx->gt_fallback_chunks++;
      }
      lo = hi;
    }
  }
  pann_gt_finish_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, s>>>(state.as<u64>(), tot, id_offset, idx->gt_ids.as<int32_t>(), idx->gt_dists.as<float>());
  idx->last_launches++;
  cudaEventRecord(idx->ev1, s);
  cudaError_t e = cudaEventSynchronize(idx->ev1);
  if (e == cudaSuccess) e = cudaGetLastError();
  fail();
  if (e != cudaSuccess) PANN_FAIL(std::string("pann_groundtruth: ") + cudaGetErrorString(e));
  cudaEventElapsedTime(&idx->last_ms, idx->ev0, idx->ev1);
  idx->gt_nq = nq; idx->gt_k = k;
  return 0;
}

extern "C" int pann_groundtruth_resident(pann_index* idx, uint32_t k, int64_t id_offset) {
  if (!idx) PANN_FAIL("pann_groundtruth_resident: null index");
  if (idx->nq == 0) PANN_FAIL("pann_groundtruth_resident: no resident query set");
  PANN_CUDA(cudaSetDevice(idx->device));
  return gt_run(idx, idx->queries.as<uint8_t>(), idx->nq, k, id_offset);
}
extern "C" int pann_groundtruth_resident_range(pann_index* idx, uint32_t k, int64_t id_offset, uint64_t row_lo, uint64_t row_hi,
                                               int keep_state, const float* d_tau_cap, uint64_t seen_rows) {
  if (!idx) PANN_FAIL("pann_groundtruth_resident_range: null index");
  if (idx->nq == 0) PANN_FAIL("pann_groundtruth_resident_range: no resident query set");
  PANN_CUDA(cudaSetDevice(idx->device));
  GtLeg leg;
  leg.row_lo = row_lo; leg.row_hi = row_hi; leg.keep_state = keep_state != 0; leg.d_cap = d_tau_cap; leg.seen_rows = seen_rows;
  return gt_run(idx, idx->queries.as<uint8_t>(), idx->nq, k, id_offset, leg);
}
extern "C" int pann_groundtruth_fetch(pann_index* idx, int32_t* out_ids, float* out_dists) {
  if (!idx) PANN_FAIL("pann_groundtruth_fetch: null index");
  PANN_CUDA(cudaSetDevice(idx->device));
  const size_t b = 4ull * idx->gt_nq * idx->gt_k;
  if (out_ids) PANN_CUDA(cudaMemcpyAsync(out_ids, idx->gt_ids.p, b, cudaMemcpyDeviceToHost, idx->stream));
  if (out_dists) PANN_CUDA(cudaMemcpyAsync(out_dists, idx->gt_dists.p, b, cudaMemcpyDeviceToHost, idx->stream));
  PANN_CUDA(cudaStreamSynchronize(idx->stream));
  return 0;
}
extern "C" int pann_groundtruth(pann_index* idx, const void* host_queries, uint64_t nq, uint64_t q_stride_bytes, uint32_t k,
                                int64_t id_offset, int32_t* out_ids, float* out_dists) {
  if (!idx || (!host_queries && nq)) PANN_FAIL("pann_groundtruth: null argument");
  PANN_CUDA(cudaSetDevice(idx->device));
  if (nq) {
    if (pann_queries_upload(idx, host_queries, nq, q_stride_bytes)) return 1;
  }
  if (gt_run(idx, idx->queries.as<uint8_t>(), nq, k, id_offset)) return 1;
  if (nq == 0) return 0;
  return pann_groundtruth_fetch(idx, out_ids, out_dists);
}

extern "C" int pann_groundtruth_merge(int device, uint32_t nshards, uint64_t nq, uint32_t k, const int32_t* ids, const float* dists,
                                      int32_t* out_ids, float* out_dists) {
  if (!ids || !dists || !out_ids || !out_dists) PANN_FAIL("pann_groundtruth_merge: null argument");
  if (nshards == 0 || k == 0 || (uint64_t)nshards * k > GT_SEL_CAP) PANN_FAIL("pann_groundtruth_merge: nshards * k must be in 1.." + std::to_string(GT_SEL_CAP));
  if (nq == 0) return 0;
  PANN_CUDA(cudaSetDevice(device));
  const size_t in_b = 4ull * nshards * nq * k, out_b = 4ull * nq * k;
  int32_t *d_ids = nullptr, *d_oi = nullptr; float *d_d = nullptr, *d_od = nullptr;
  PANN_CUDA(cudaMalloc(&d_ids, in_b)); PANN_CUDA(cudaMalloc(&d_d, in_b)); PANN_CUDA(cudaMalloc(&d_oi, out_b)); PANN_CUDA(cudaMalloc(&d_od, out_b));
  PANN_CUDA(cudaMemcpy(d_ids, ids, in_b, cudaMemcpyHostToDevice));
  PANN_CUDA(cudaMemcpy(d_d, dists, in_b, cudaMemcpyHostToDevice));
  pann_gt_merge_kernel<<<(unsigned)nq, GT_SEL_THREADS>>>(d_ids, d_d, nshards, nq, k, d_oi, d_od);
  PANN_CUDA(cudaGetLastError());
  PANN_CUDA(cudaMemcpy(out_ids, d_oi, out_b, cudaMemcpyDeviceToHost));
  PANN_CUDA(cudaMemcpy(out_dists, d_od, out_b, cudaMemcpyDeviceToHost));
  cudaFree(d_ids); cudaFree(d_d); cudaFree(d_oi); cudaFree(d_od);
  return 0;
}

// Device-pointer variants for the multi-GPU path (base rows sharded over ranks): the local top-k stays
// on the device, the ranks exchange it with an NCCL all-gather into one [nshards, nq, k] block, and K5 merges that
// block in place on every rank.  `stream` is a cudaStream_t (0 = the legacy default stream).
extern "C" int pann_groundtruth_device_result(pann_index* idx, const int32_t** d_ids, const float** d_dists, uint64_t* nq, uint32_t* k) {
  if (!idx) PANN_FAIL("pann_groundtruth_device_result: null index");
  if (idx->gt_nq == 0) PANN_FAIL("pann_groundtruth_device_result: no ground truth on the device");
  if (d_ids) *d_ids = idx->gt_ids.as<int32_t>();
  if (d_dists) *d_dists = idx->gt_dists.as<float>();
  if (nq) *nq = idx->gt_nq;
  if (k) *k = idx->gt_k;
  return 0;
}
extern "C" int pann_groundtruth_merge_device(int device, uint32_t nshards, uint64_t nq, uint32_t k, const int32_t* d_ids, const float* d_dists,
                                             int32_t* d_out_ids, float* d_out_dists, void* stream) {
  if (!d_ids || !d_dists || !d_out_ids || !d_out_dists) PANN_FAIL("pann_groundtruth_merge_device: null argument");
  if (nshards == 0 || k == 0 || (uint64_t)nshards * k > GT_SEL_CAP) PANN_FAIL("pann_groundtruth_merge: nshards * k must be in 1.." + std::to_string(GT_SEL_CAP));
  if (nq == 0) return 0;
  PANN_CUDA(cudaSetDevice(device));
  pann_gt_merge_kernel<<<(unsigned)nq, GT_SEL_THREADS, 0, (cudaStream_t)stream>>>(d_ids, d_dists, nshards, nq, k, d_out_ids, d_out_dists);
  PANN_CUDA(cudaGetLastError());
  return 0;
}


extern "C" int pann_index_groundtruth_stats(pann_index* idx, int* mma_launches, int* fallback_chunks) {
  if (!idx) PANN_FAIL("pann_index_groundtruth_stats: null index");
  if (mma_launches) *mma_launches = idx->gt_mma_launches;
  if (fallback_chunks) *fallback_chunks = idx->gt_fallback_chunks;
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// pann_groundtruth_static — exact k-NN of device-resident query rows on the CUDA-core kernels above (K4 v1: distance
// chunk -> select), with EVERY byte of working memory in one file-scope device array: no cudaMalloc / cudaFree, no
// synchronisation, only kernel launches on the caller's stream.  For callers that may not allocate (graph capture,
// serving loops).  No counterpart in the reference.
//   scratch layout: [ top-k state nq*k*8 | query rows in the index's 16-byte-chunk layout nq*stride | distance block nq*cb*4 ]
// All calls on a device share the array: calls that overlap on the GPU (two streams, two threads) race.  Nothing in it
// has to survive a call — the state is re-initialised, the query rows and every distance that is read are rewritten.
#define PANN_STATIC_SCRATCH_BYTES (16u << 20)
#define PANN_STATIC_MAX_DEVICES 64
static __device__ __align__(256) uint8_t pann_static_scratch[PANN_STATIC_SCRATCH_BYTES];
static uint8_t* g_static_scratch_ptr[PANN_STATIC_MAX_DEVICES];   // device addresses of the array, one per device (set-up)

__global__ void pann_static_stage_rows_kernel(const uint8_t* __restrict__ src, uint64_t src_stride, uint32_t rowb, uint8_t* __restrict__ dst,
                                              uint32_t stride, uint64_t total) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;   // one byte of the padded layout per thread
  if (i >= total) return;
  const uint64_t r = i / stride;
  const uint32_t c = (uint32_t)(i - r * stride);
  dst[i] = c < rowb ? src[r * src_stride + c] : (uint8_t)0;
}

// the array's device address on the current device: the symbol itself is only a host-side stand-in
static int pann_static_scratch_address(int device, uint8_t** out) {
  if (device < 0 || device >= PANN_STATIC_MAX_DEVICES) PANN_FAIL("pann_groundtruth_static: device ordinal out of range");
  if (!g_static_scratch_ptr[device]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, pann_static_scratch);
    if (e != cudaSuccess || !p) PANN_FAIL(std::string("pann_groundtruth_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
    g_static_scratch_ptr[device] = static_cast<uint8_t*>(p);
  }
  *out = g_static_scratch_ptr[device];
  return 0;
}

extern "C" int pann_groundtruth_static_limit(uint64_t* scratch_bytes) {
  if (scratch_bytes) *scratch_bytes = PANN_STATIC_SCRATCH_BYTES;
  return 0;
}

extern "C" int pann_groundtruth_static(pann_index* idx, const void* d_queries, uint64_t nq, uint64_t q_stride_bytes, uint32_t k,
                                       int64_t id_offset, int32_t* d_out_ids, float* d_out_dists, void* stream) {
  if (!idx || !d_queries || !d_out_ids || !d_out_dists) PANN_FAIL("pann_groundtruth_static: null argument");
  if (k == 0 || k > GT_SEL_CAP / 2) PANN_FAIL("pann_groundtruth_static: k must be in 1.." + std::to_string(GT_SEL_CAP / 2));
  if (idx->n < k) PANN_FAIL("pann_groundtruth_static: fewer base points than k");
  if (nq == 0) PANN_FAIL("pann_groundtruth_static: no queries");
  if ((int64_t)idx->n + id_offset > (int64_t)INT32_MAX) PANN_FAIL("pann_groundtruth_static: ids do not fit int32");
  const uint32_t stride = idx->geom.stride;
  const uint32_t rowb = idx->dims * idx->geom.elt;
  if (q_stride_bytes < rowb) PANN_FAIL("pann_groundtruth_static: query stride smaller than a row");
  // the fixed part and at least one 64-row distance chunk must fit the array
  const uint64_t state_b = (8ull * nq * k + 255) / 256 * 256, q_b = ((uint64_t)nq * stride + 255) / 256 * 256;
  const uint64_t need = state_b + q_b + 4ull * nq * GT_TB;
  if (nq > (1ull << 20) || need > PANN_STATIC_SCRATCH_BYTES)
    PANN_FAIL("pann_groundtruth_static: " + std::to_string(nq) + " queries x k=" + std::to_string(k) + " x " + std::to_string(stride) +
              "-byte rows need " + std::to_string(need) + " bytes of scratch; the static array holds " + std::to_string(PANN_STATIC_SCRATCH_BYTES) +
              " (nq * (8 k + row bytes + 256) must stay below it): split the batch or use pann_groundtruth");
  PANN_CUDA(cudaSetDevice(idx->device));
  uint8_t* scratch = nullptr;
  if (pann_static_scratch_address(idx->device, &scratch)) return 1;
  cudaStream_t s = (cudaStream_t)stream;
  u64* state = reinterpret_cast<u64*>(scratch);
  uint8_t* qrows = scratch + state_b;
  float* D = reinterpret_cast<float*>(scratch + state_b + q_b);
  uint64_t cb = ((PANN_STATIC_SCRATCH_BYTES - state_b - q_b) / (4ull * nq)) / GT_TB * GT_TB;   // rows per distance chunk
  cb = std::min<uint64_t>(cb, (idx->n + GT_TB - 1) / GT_TB * GT_TB);
  const uint64_t tot = nq * k, qbytes = nq * stride;
  pann_static_stage_rows_kernel<<<(unsigned)((qbytes + 255) / 256), 256, 0, s>>>(static_cast<const uint8_t*>(d_queries), q_stride_bytes, rowb, qrows, stride, qbytes);
  pann_gt_init_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, s>>>(state, tot);
  gt_v1_range(idx, s, qrows, nq, k, state, D, cb, 0, idx->n);
  pann_gt_finish_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, s>>>(state, tot, id_offset, d_out_ids, d_out_dists);
  PANN_CUDA(cudaGetLastError());
  return 0;
}
