// beam_search.cuh — K1: batched greedy beam search over the fixed-degree graph, one query per warp.
//
// Behavioural contract: parlayANN::filtered_beam_search (algorithms/utils/beamSearch.h:22-214) with
// use_filtering == false, bit for bit: same visit order, same lossy hash filter (direct mapped,
// overwrite on miss, :52-59), same "skip the merge until beam/8 candidates" rule (:162-168), same cut
// rule with its float-rounded threshold (:190-195), same counters (visited, full_dist_cmps).
//
// How it is laid out for the GPU (not a translation of the CPU code):
//  * the frontier is ONE sorted array of 64-bit keys (sortable-float distance, id, visited flag); the
//    reference's separate unvisited_frontier is implied by the flags: "next to visit" is the first
//    unflagged entry, and offset+1 < remain <=> another unflagged entry exists;
//  * a hop reads the graph row with one coalesced warp load, runs the hash filter for 32 neighbours at
//    once (__match_any resolves same-slot collisions in list order), compacts the survivors and gathers
//    their vectors ONE ROW PER LANE: 32 rows per pass, every lane streams its own row with back-to-back
//    128-bit loads (all of a row's sectors are requested at once) and accumulates in the reference's own
//    dimension order — no cross-lane reduction, so float distances are bit-identical to the CPU's;
//  * the merge is rank based: candidates are rank-sorted, de-duplicated, and each element's slot in the
//    union is (its rank) + (lower_bound in the other list) — no serial set_union;
//  * frontier, candidates, the visited list and the hash filter live in shared memory per warp; for
//    wide beams (hash filter > 16 KB) the filter and the visited list move to a per-warp global
//    scratch slab that stays L2 resident.
#pragma once
#include "pann_common.cuh"

#define PANN_WARPS_PER_CTA 4
#define PANN_FULL 0xffffffffu

#define PANN_STAGE_SLOT 9  // K1s row staging (pann_staged_distance): uint4 per staged row block = 8 chunks + 1 skew
#define PANN_STAGE_BYTES (32 * PANN_STAGE_SLOT * 16)
struct BeamArgs {
  // index
  const uint8_t* points; uint64_t n; uint32_t stride, nch;
  const uint32_t* graph; uint32_t maxdeg;
  // queries: either rows of `queries` (qstride bytes apart, 16-byte padded) or base rows self_ids[i]
  const uint8_t* queries; uint64_t nq; uint32_t qstride;
  const uint32_t* self_ids;
  // start points: one list for the whole batch (starts_stride == 0), or one list per query — query q starts from
  // starts[q * starts_stride .. + (nstarts_q ? nstarts_q[q] : nstarts)) (beamSearchRandom, beamSearch.h:327-332; HNSW's
  // per-query entry points, HNSW.hpp:1570-1613).  nstarts = the largest list (sizes the buffers).
  const uint32_t* starts; uint32_t nstarts; uint32_t starts_stride; const uint32_t* nstarts_q;
  // optional vertex id -> graph row (compacted upper HNSW layers, where only some vertices have a row); null = identity
  const uint32_t* rowmap;
  // QueryParams
  int k, beam, bits, degree_limit, cut_rule;  // cut_rule = (k > 0 && metric is L2)
  long long limit;
  double cut;
  // capacities and per-warp shared-memory layout (byte offsets)
  uint32_t ccap, vcap;
  uint32_t off_frontB, off_cand, off_scand, off_nbr, off_q, off_vis, off_hash, off_stage, warp_bytes;  // off_stage: K1s row staging (float rows)
  // outputs (any may be null)
  uint32_t* out_ids; float* out_dists;       // nq * k
  uint32_t* out_visited; uint32_t* out_cmps; // nq
  uint32_t* fr_ids; float* fr_dists; uint32_t* n_frontier; uint32_t fcap;   // full frontier
  uint32_t* vis_ids; float* vis_dists; uint32_t* n_visited; uint32_t vout;  // full visited list (capacity vout)
  int* status;                               // nq: 0 ok, 1 visited overflow, 2 fewer than k results, 3 too many starts
  int* any_status;                           // OR of all non-zero statuses (one word)
  const uint32_t* qmap;                      // optional: launch index -> query index (retries of a subset)
  // algorithmic bytes of the batch (B_q = cmps*row + sum_hops 4*(1+nbrs read) + row + 4k), one atomic per query
  u64* bytes_total; uint32_t rowbytes;
  uint32_t* qcounter;                        // optional: dynamic query distribution (zeroed before the launch)
  uint32_t tune;                             // K1s: bit 0 = always prefetch neighbour rows to L2, bit 1 = never (default: only while the launch drains); bit 2 = float rows not staged (A/B)
  // global scratch (GS variant): per resident warp
  uint32_t* g_hash; u64* g_vis;
};

// Counting and bisecting on RAW keys (flag bit included): for a threshold key t with bit 0 clear,
//   x < t        <=>  (x >> 1) <  (t >> 1)          and          x <= (t | 1)  <=>  (x >> 1) <= (t >> 1),
// so no shifts are needed.  Short lists are counted, not bisected (n is warp-uniform at every call site): the loads are
// independent and each element costs one subtract-with-borrow chain (3 integer adds), two accumulators deep.
__device__ __forceinline__ void pann_borrow_count(int& neg, u64 x, u64 t) {  // neg -= (x < t)
  const unsigned xl = (unsigned)x, xh = (unsigned)(x >> 32), tl = (unsigned)t, th = (unsigned)(t >> 32);
  asm("{\n\t.reg .u32 tt;\n\tsub.cc.u32 tt, %1, %3;\n\tsubc.cc.u32 tt, %2, %4;\n\tsubc.u32 %0, %0, 0;\n\t}"
      : "+r"(neg) : "r"(xl), "r"(xh), "r"(tl), "r"(th));
}
__device__ __forceinline__ int pann_count_below_any(const u64* a, int n, u64 t) {  // #{j < n : a[j] < t}, any order
  int n0 = 0, n1 = 0, j = 0;
#pragma unroll 1   // lists are short (about one beam): a deeper unroll only bloats the code and runs its remainder loops
  for (; j + 4 <= n; j += 4) {
    const u64 x0 = a[j], x1 = a[j + 1], x2 = a[j + 2], x3 = a[j + 3];
    pann_borrow_count(n0, x0, t); pann_borrow_count(n1, x1, t); pann_borrow_count(n0, x2, t); pann_borrow_count(n1, x3, t);
  }
#pragma unroll 1
  for (; j < n; j++) pann_borrow_count(n0, a[j], t);
  return -(n0 + n1);
}
__device__ __forceinline__ int pann_count_below(const u64* a, int n, u64 t) {  // the same for a SORTED list
  if (n <= 40) return pann_count_below_any(a, n, t);
  int lo = 0, hi = n;
  while (lo < hi) { int mid = (lo + hi) >> 1; if (a[mid] < t) lo = mid + 1; else hi = mid; }
  return lo;
}
// The same for a sorted list in GLOBAL memory (the visited list of the global-scratch variant): a binary search there is
// ~8 dependent L2 round trips; seven pivots per round, loaded together, make it three.
__device__ __forceinline__ int pann_count_below_far(const u64* a, int n, u64 t) {
  int lo = 0, hi = n;   // the count lies in [lo, hi]: a[0..lo) < t, a[hi..n) >= t
  while (hi - lo > 8) {
    const int step = (hi - lo + 7) >> 3;
    u64 p[7];
#pragma unroll
    for (int j = 0; j < 7; j++) { const int ix = lo + (j + 1) * step - 1; p[j] = ix < hi ? a[ix] : ~0ull; }
    int c = 0;
#pragma unroll
    for (int j = 0; j < 7; j++) c += (p[j] < t) ? 1 : 0;   // the pivots are sorted: the first c are below t
    const int nlo = lo + c * step;
    if (c < 7 && nlo + step - 1 < hi) hi = nlo + step - 1;
    lo = nlo;
  }
  int r = lo;
#pragma unroll
  for (int j = 0; j < 8; j++) r += (lo + j < hi && a[lo + j] < t) ? 1 : 0;
  return r;
}
// t = a (dist,id) key with bit 0 clear
__device__ __forceinline__ int pann_lower_bound(const u64* a, int n, u64 t) { return pann_count_below(a, n, t); }       // #{x>>1 <  t>>1}
__device__ __forceinline__ int pann_upper_bound(const u64* a, int n, u64 t) { return pann_count_below(a, n, t + 2); }   // #{x>>1 <= t>>1}

// Integer rows of exactly 128 bytes (QR == 8, uint8 / int8): FOUR lanes share a row.  `a` is this lane's row id and bit l
// of keepmask says whether lane l has one.  Lane 4r + s loads chunks s and s + 4 of the rows of lanes 8g + r, g = 0..3
// (two 128-bit loads 64 bytes apart: the four lanes of a row touch its 128-byte line with two requests instead of
// eight, which is what the L1 data stage counts); the partial sums are folded across the quad by a transposing
// reduction (3 shuffles for 32 rows) and one more shuffle hands every lane the sum of its own row.  qsm = the query
// row in shared memory.  Integer sums are exact in any order, so the result is bit-identical to the
// sequential loop; float rows keep that loop (their order is part of the result).
__device__ __forceinline__ uint4 pann_lds16_volatile(const uint4* p) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"((unsigned)__cvta_generic_to_shared(p)));
  return v;
}
// two 128-bit loads 64 bytes apart, issued only if `on` (the registers are left undefined otherwise: callers ignore them)
__device__ __forceinline__ void pann_ldg16x2_if(const uint8_t* p, unsigned on, uint4& v0, uint4& v1) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %9, 0;\n\t"
      "@p ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%8];\n\t"
      "@p ld.global.nc.v4.u32 {%4,%5,%6,%7}, [%8+64];\n\t}"
      : "=r"(v0.x), "=r"(v0.y), "=r"(v0.z), "=r"(v0.w), "=r"(v1.x), "=r"(v1.y), "=r"(v1.z), "=r"(v1.w)
      : "l"(p), "r"(on));
}
template <int DT, int QR> __host__ __device__ constexpr bool pann_quad_rows() { return QR == 8 && DT != PANN_DT_F32; }
template <int DT, int METRIC, class RG>
__device__ __forceinline__ float pann_quad_distance(const RG& A, uint32_t a, unsigned keepmask, const uint4* qsm, int lane) {
  const int sub = lane & 3, rowi = lane >> 2;
  // all eight loads go out before anything is consumed (volatile asm keeps them in program order: left to itself the
  // compiler, short of registers, interleaves loads and arithmetic and serialises the two halves of the rows)
  uint4 v[4][2];
#pragma unroll
  for (int g = 0; g < 4; g++) {
    const int src = g * 8 + rowi;
    const uint32_t ag = __shfl_sync(PANN_FULL, a, src);
#if !defined(PANN_QUAD_ZERO)   // unwanted rows read row 0 (an L1 hit) instead of zeroed registers, see beam_search_small.cuh
    const uint8_t* row0 = A.points + (u64)(((keepmask >> src) & 1u) ? ag : 0u) * A.stride + sub * 16;
    v[g][0] = pann_ldg16(row0);
    v[g][1] = pann_ldg16(row0 + 64);
#elif defined(PANN_QUAD_ASM)
    pann_ldg16x2_if(A.points + (u64)ag * A.stride + sub * 16, (keepmask >> src) & 1u, v[g][0], v[g][1]);
#else
    v[g][0] = v[g][1] = make_uint4(0u, 0u, 0u, 0u);
    if ((keepmask >> src) & 1u) {
      const uint8_t* row = A.points + (u64)ag * A.stride + sub * 16;
      v[g][0] = pann_ldg16(row);
      v[g][1] = pann_ldg16(row + 64);
    }
#endif
  }
  const uint4 q0 = pann_lds16_volatile(qsm + sub), q1 = pann_lds16_volatile(qsm + sub + 4);
  DistAcc<DT, METRIC> acc[4];
#pragma unroll
  for (int g = 0; g < 4; g++) { acc[g].add(v[g][0], q0); acc[g].add(v[g][1], q1); }
  unsigned part[4];
#pragma unroll
  for (int g = 0; g < 4; g++) part[g] = (unsigned)acc[g].acc;
  const bool b1 = (lane & 2) != 0, b0 = (lane & 1) != 0;
  const unsigned r0 = (b1 ? part[2] : part[0]) + __shfl_xor_sync(PANN_FULL, b1 ? part[0] : part[2], 2);
  const unsigned r1 = (b1 ? part[3] : part[1]) + __shfl_xor_sync(PANN_FULL, b1 ? part[1] : part[3], 2);
  const unsigned tot = (b0 ? r1 : r0) + __shfl_xor_sync(PANN_FULL, b0 ? r0 : r1, 1);
  // lane 4r + s now holds the sum of the row of lane 8s + r; the row of lane l = 8g + r sits in lane 4r + g
  const unsigned mine = __shfl_sync(PANN_FULL, tot, (lane & 7) * 4 + (lane >> 3));
  DistAcc<DT, METRIC> res;
  res.acc = (decltype(res.acc))mine;
  return res.result();
}

// Float rows (any row whose sum order is part of the result): a row must be summed by ONE lane in element order, but a
// lane that also LOADS its row alone makes every 128-bit load of the warp touch 32 different lines (32 L1 wavefronts
// for 512 useful bytes).  Here the warp loads the rows together and hands them over through shared memory, 128 bytes
// of every row at a time: lane 8*rr + cc loads chunk cc of the current 128-byte block of rows 4*i + rr (i = 0..7), so a
// load instruction covers four whole lines (4 wavefronts), the block is parked in `stage` (32 slots of 144 bytes: the
// 16-byte skew makes both the quarter-warp stores and the one-row-per-lane reads bank-conflict free), and lane l adds
// the eight chunks of its own row in order into its running sum.  The next block's loads are in flight during the sums.
// Same operation sequence per row as pann_lane_distance, so the result is bit-identical.
template <int DT, int METRIC, class RG>
__device__ __forceinline__ float pann_staged_distance(const RG& A, uint32_t a, unsigned keepmask, const uint4* qsm, uint4* stage, int lane) {
  const int cc = lane & 7, rr = lane >> 3;
  const int nch = (int)A.nch;
  const bool on = (keepmask >> lane) & 1u;
  uint32_t rid[8];   // ids, not pointers: 8 registers less, one wide multiply-add per load more
  unsigned onbits = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const int src = 4 * i + rr;
    rid[i] = __shfl_sync(PANN_FULL, a, src);
    onbits |= ((keepmask >> src) & 1u) << i;
  }
  const uint8_t* pbase = A.points + cc * 16;
  uint4 v[8];
#pragma unroll
  for (int i = 0; i < 8; i++) {
    v[i] = make_uint4(0u, 0u, 0u, 0u);
    if (((onbits >> i) & 1u) && cc < nch) v[i] = pann_ldg16(pbase + (u64)rid[i] * A.stride);
  }
  DistAcc<DT, METRIC> acc;
  for (int c0 = 0; c0 < nch; c0 += 8) {
#pragma unroll
    for (int i = 0; i < 8; i++) stage[(4 * i + rr) * PANN_STAGE_SLOT + cc] = v[i];
    __syncwarp();
    if (c0 + 8 < nch) {
#pragma unroll
      for (int i = 0; i < 8; i++)
        if (((onbits >> i) & 1u) && c0 + 8 + cc < nch) v[i] = pann_ldg16(pbase + (u64)rid[i] * A.stride + (c0 + 8) * 16);
    }
    if (on) {
      const uint4* mine = stage + lane * PANN_STAGE_SLOT;
      if (c0 + 8 <= nch) {
#pragma unroll
        for (int j = 0; j < 8; j++) acc.add(mine[j], qsm[c0 + j]);
      } else {
        for (int j = 0; c0 + j < nch; j++) acc.add(mine[j], qsm[c0 + j]);
      }
    }
    __syncwarp();
  }
  return acc.result();
}
#ifdef PANN_STAGE_INT   // A/B: stage integer rows of other lengths than 128 bytes as well (their sums are exact in any order)
template <int DT, int QR> __host__ __device__ constexpr bool pann_staged_rows() { return QR == 0; }
inline bool pann_staged_rows_rt(int dtype, int qr) { return qr == 0; }
#else
template <int DT, int QR> __host__ __device__ constexpr bool pann_staged_rows() { return QR == 0 && DT == PANN_DT_F32; }
inline bool pann_staged_rows_rt(int dtype, int qr) { return qr == 0 && dtype == PANN_DT_F32; }
#endif
// Distances from the query to `cnt` rows, one row per lane, 32 rows per pass; row r is pann_key_id(keyOf(r))
// (a 64-bit tag whose id field names the row, so callers can carry a (dist,id) key through).  Calls
// emit(tag, dist, valid) warp-uniformly once per pass; lane l carries row base + l, so ballot-ranked stores in
// emit keep the rows in order.  The query is read from shared memory (qsm, broadcast reads).  QR > 0: rows have
// exactly QR chunks and the loop is unrolled with all QR loads in flight; QR == 0: any row length, 8 loads in flight.
// RG: anything with points / stride / nch (BeamArgs, PruneArgs, ...).
// QUAD: use pann_quad_distance where the row shape allows it (the beam search does; the prune kernel keeps the
// one-row-per-lane loop, whose 40 registers buy it 48 resident warps per SM).
// stage: per-warp PANN_STAGE_BYTES of shared memory, or null: float rows then go through pann_staged_distance.
template <int DT, int METRIC, int QR, bool QUAD = false, class RG, class KeyOf, class Emit>
__device__ __forceinline__ void pann_row_distances_g(const RG& A, KeyOf keyOf, int cnt, const uint4* qsm, int lane,
                                                     Emit emit, uint4* stage = nullptr) {
  for (int base = 0; base < cnt; base += 32) {
    const int r = base + lane;
    const bool ok = r < cnt;
    u64 tag = 0;
    if constexpr (pann_staged_rows<DT, QR>()) {
      if (stage) {
        if (ok) tag = keyOf(r);
        const float d = pann_staged_distance<DT, METRIC>(A, pann_key_id(tag), __ballot_sync(PANN_FULL, ok), qsm, stage, lane);
        emit(tag, d, ok);
        continue;
      }
    }
    if constexpr (QUAD && pann_quad_rows<DT, QR>()) {
      if (ok) tag = keyOf(r);
      const float d = pann_quad_distance<DT, METRIC>(A, pann_key_id(tag), __ballot_sync(PANN_FULL, ok), qsm, lane);
      emit(tag, d, ok);
      continue;
    }
    DistAcc<DT, METRIC> acc;
    if (ok) {
      tag = keyOf(r);
      const uint8_t* row = A.points + (u64)pann_key_id(tag) * A.stride;
      if constexpr (QR > 0) {
        uint4 v[QR];
#pragma unroll
        for (int c = 0; c < QR; c++) v[c] = pann_ldg16(row + c * 16);
#pragma unroll
        for (int c = 0; c < QR; c++) acc.add(v[c], qsm[c]);
      } else {
        const int nch = (int)A.nch;
        for (int c0 = 0; c0 < nch; c0 += 8) {
          uint4 v[8];
#pragma unroll
          for (int j = 0; j < 8; j++) if (c0 + j < nch) v[j] = pann_ldg16(row + (c0 + j) * 16);
#pragma unroll
          for (int j = 0; j < 8; j++) if (c0 + j < nch) acc.add(v[j], qsm[c0 + j]);
        }
      }
    }
    emit(tag, acc.result(), ok);
  }
}
// the same over a plain id list (shared memory): emit(id, dist, valid)
template <int DT, int METRIC, int QR, class RG, class Emit>
__device__ __forceinline__ void pann_row_distances(const RG& A, const uint32_t* ids, int cnt, const uint4* qsm, int lane,
                                                   Emit emit, uint4* stage = nullptr) {
  pann_row_distances_g<DT, METRIC, QR, true>(
      A, [&](int r) { return (u64)ids[r] << 1; }, cnt, qsm, lane,
      [&](u64 tag, float d, bool ok) { emit(pann_key_id(tag), d, ok); }, stage);
}
// the query row into shared memory
__device__ __forceinline__ void pann_load_query(const uint8_t* qrow, uint32_t nch, uint4* qsm, int lane) {
  __syncwarp();
  for (int ch = lane; ch < (int)nch; ch += 32) qsm[ch] = pann_ldg16(qrow + ch * 16);
  __syncwarp();
}

template <int DT, int METRIC, int QR, bool GS>
#ifndef PANN_K1_MINB
#define PANN_K1_MINB 6  // 80 registers, no spills; 10M build, search phase: 6.16 s (8 CTAs/SM) -> 5.63 (7) -> 5.57 (6)
#endif
__global__ void __launch_bounds__(PANN_WARPS_PER_CTA * 32, QR > 0 ? PANN_K1_MINB : 4) pann_beam_search_kernel(const BeamArgs A) {
  extern __shared__ __align__(16) uint8_t pann_smem[];
  const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
  const unsigned lt_mask = (1u << lane) - 1u;
  uint8_t* ws = pann_smem + (size_t)wic * A.warp_bytes;
  u64* bufA = reinterpret_cast<u64*>(ws);
  u64* bufB = reinterpret_cast<u64*>(ws + A.off_frontB);
  u64* cand = reinterpret_cast<u64*>(ws + A.off_cand);
  u64* scand = reinterpret_cast<u64*>(ws + A.off_scand);
  uint32_t* nbr = reinterpret_cast<uint32_t*>(ws + A.off_nbr);
  uint4* qsm = reinterpret_cast<uint4*>(ws + A.off_q);
  uint4* stage = (pann_staged_rows<DT, QR>() && !(A.tune & 4u)) ? reinterpret_cast<uint4*>(ws + A.off_stage) : nullptr;
  const long long wglobal = (long long)blockIdx.x * PANN_WARPS_PER_CTA + wic;
  u64* vis;
  uint32_t* hashf;
  if constexpr (GS) {
    vis = A.g_vis + (u64)wglobal * A.vcap;
    hashf = A.g_hash + ((u64)wglobal << A.bits);
  } else {
    vis = reinterpret_cast<u64*>(ws + A.off_vis);
    hashf = reinterpret_cast<uint32_t*>(ws + A.off_hash);
  }
  const uint32_t hmask = (1u << A.bits) - 1u;
  const int B = A.beam;
  const long long wstride = (long long)gridDim.x * PANN_WARPS_PER_CTA;

  // the first query of a warp is its own index; further ones come from a shared counter, so a warp that drew short
  // searches takes more of them (query lengths vary by 2-3x)
  auto next_query = [&](long long ql) -> long long {
    if (!A.qcounter) return ql + wstride;
    unsigned nx = 0;
    if (lane == 0) nx = atomicAdd(A.qcounter, 1u);
    return wstride + (long long)__shfl_sync(PANN_FULL, nx, 0);
  };
  if constexpr (GS) {
    if (A.tune & 16u) {   // the two-round filter's slot bitmap: all zero between hops
      uint32_t* fb = reinterpret_cast<uint32_t*>(ws + A.off_hash);
      for (int i = lane; i < (1 << A.bits) / 32; i += 32) fb[i] = 0u;
      __syncwarp();
    }
  }
  for (long long ql = wglobal; ql < (long long)A.nq; ql = next_query(ql)) {
    const long long q = A.qmap ? (long long)A.qmap[ql] : ql;
    // ---- per-query init: hash filter := -1 (beamSearch.h:53), query chunks into registers / shared memory
    {
      uint4 ones = make_uint4(~0u, ~0u, ~0u, ~0u);
      uint4* h4 = reinterpret_cast<uint4*>(hashf);
      for (int i = lane; i < (1 << A.bits) / 4; i += 32) h4[i] = ones;
    }
    const long long self = A.self_ids ? (long long)A.self_ids[q] : -1;
    const uint8_t* qrow = A.self_ids ? A.points + (u64)self * A.stride : A.queries + (u64)q * A.qstride;
    pann_load_query(qrow, A.nch, qsm, lane);

    u64* front = bufA;
    u64* nf = bufB;
    int F = 0, V = 0, C = 0, status = 0;
    const uint32_t* qstarts = A.starts + (u64)q * A.starts_stride;
    const uint32_t ns = A.nstarts_q ? A.nstarts_q[q] : A.nstarts;
    unsigned cmps = ns;        // beamSearch.h:84
    unsigned nread = 0;        // neighbour ids read from graph rows

    // ---- starting points (:64-70): distances, hash-filter insertion in list order, sort.  More starts than the beam
    // width (HNSW's ef = 1 descent from several entrances) leave a frontier longer than the beam until the first merge
    // truncates it (:185) — the reference's unvisited_frontier buffer overflows there (:74-76); this follows what the
    // code does with that buffer made large enough.  With the cut rule on, :195's max(.., old size) would keep it
    // longer than the beam: that corner is refused.
    if (ns == 0 || ns > A.nstarts || ns > A.ccap || ((int)ns > B && A.cut_rule && A.k < B)) {
      status = 3;
    } else {
      for (uint32_t s0 = 0; s0 < ns; s0 += A.maxdeg) {
        int cnt = min((int)A.maxdeg, (int)(ns - s0));
        for (int i = lane; i < cnt; i += 32) nbr[i] = qstarts[s0 + i];
        __syncwarp();
        pann_row_distances<DT, METRIC, QR>(A, nbr, cnt, qsm, lane, [&](uint32_t id, float d, bool ok) {
          unsigned m = __ballot_sync(PANN_FULL, ok);
          if (ok) cand[C + __popc(m & lt_mask)] = pann_key(d, id);
          C += __popc(m);
        }, stage);
        __syncwarp();
      }
      if (lane == 0)
        for (uint32_t i = 0; i < ns; i++) { uint32_t a = qstarts[i]; hashf[(uint32_t)pann_hash64_2(a) & hmask] = a; }
      for (int i = lane; i < C; i += 32) {  // rank sort into the frontier
        u64 ci = cand[i];
        int r = 0;
        for (int j = 0; j < C; j++) { u64 cj = cand[j]; r += (cj < ci) || (cj == ci && j < i); }
        front[r] = ci;
      }
      F = C;
      C = 0;
      __syncwarp();
    }

    // ---- main loop (:107-209)
    while (status == 0) {
      // next to visit = first unflagged frontier entry  (unvisited_frontier[offset], :109)
      int pos = -1, nflag = 0;  // nflag = #frontier entries already visited
      for (int base = 0; base < F; base += 32) {
        const int i = base + lane;
        const unsigned m = __ballot_sync(PANN_FULL, i < F && !(front[i] & 1ull));
        nflag += min(32, F - base) - __popc(m);
        if (m && pos < 0) pos = base + __ffs(m) - 1;
      }
      if (pos < 0 || (long long)V >= A.limit) break;  // remain <= offset, or visit limit reached (:107)
      const u64 cur = front[pos];
      // every visited vertex still sits in the frontier (flagged entries are distinct members of `visited`): then a
      // candidate that is not in the frontier cannot be in `visited`, and the merge skips that lookup
      const bool vis_in_front = (nflag == V);
      // the graph row is requested now (degree + the first 64 slots; slots past the degree are stale but in-row)
      // so that its latency overlaps the visited-list insertion below
      const uint32_t* grow = A.graph + (u64)(A.rowmap ? __ldg(A.rowmap + pann_key_id(cur)) : pann_key_id(cur)) * (A.maxdeg + 1);
      const int deg = (int)__ldg(grow);
      const uint32_t a0 = (lane < (int)A.maxdeg) ? __ldg(grow + 1 + lane) : 0u;
      const uint32_t a1 = (32 + lane < (int)A.maxdeg) ? __ldg(grow + 33 + lane) : 0u;
      __syncwarp();
      if (lane == 0) front[pos] = cur | 1ull;
      // visited.insert(upper_bound(cur)) (:112-113): one pass from the top, entries above cur move up by one
      if (V >= (int)A.vcap) { status = 1; break; }
      {
        const u64 kc = cur | 1ull;  // x > kc  <=>  (x >> 1) > (cur >> 1)   (entries of vis have bit 0 clear)
        int ip = V;
        for (int top = V; top > 0; top -= 32) {
          const int i = top - 1 - lane;
          const u64 x = (i >= 0) ? vis[i] : 0ull;
          const bool gt = (i >= 0) && (x > kc);
          const unsigned m = __ballot_sync(PANN_FULL, gt);
          if (gt) vis[i + 1] = x;  // every lane read its entry before the ballot; targets are this block's own slots + 1
          ip -= __popc(m);
          __syncwarp();
          if (__popc(m) < min(32, top)) break;
        }
        if (lane == 0) vis[ip] = cur & ~1ull;
        V++;
      }
      const bool full = (F == B);  // :115
      const float cutoff = full ? pann_key_dist(front[F - 1]) : 2147483648.0f;  // (float)INT_MAX (:150-152)
      // neighbours that pass the hash filter, in list order (:130-136)
      const int ne = min(deg, A.degree_limit);
      nread += ne;
      int S = 0;
      __syncwarp();
      bool filtered = false;
      if constexpr (GS) {
        if (A.tune & 16u) {
          // The filter lives in global memory here (an L2 round trip per lookup).  Both 32-neighbour rounds look their slots up
          // at once; what round 0 writes in between is accounted for by a per-warp bitmap of the slots it touches: a round-1
          // neighbour whose slot round 0 overwrote finds another vertex's id there (ids of a row are distinct), i.e. "not seen"
          // — exactly what the sequential walk over the row sees (:54-58, :130-136).  max degree <= 64, bitmap of 2^bits bits.
          uint32_t* fb = reinterpret_cast<uint32_t*>(ws + A.off_hash);
          const bool act0 = lane < ne, act1 = 32 + lane < ne;
          const uint32_t loc0 = act0 ? ((uint32_t)pann_hash64_2(a0) & hmask) : (0x80000000u | lane);
          const uint32_t loc1 = act1 ? ((uint32_t)pann_hash64_2(a1) & hmask) : (0x80000000u | lane);
          const unsigned peers0 = __match_any_sync(PANN_FULL, loc0), peers1 = __match_any_sync(PANN_FULL, loc1);
          const unsigned prev0 = peers0 & lt_mask, prev1 = peers1 & lt_mask;
          const uint32_t fp0 = __shfl_sync(PANN_FULL, a0, prev0 ? (31 - __clz(prev0)) : lane);
          const uint32_t fp1 = __shfl_sync(PANN_FULL, a1, prev1 ? (31 - __clz(prev1)) : lane);
          uint32_t s0 = 0xffffffffu, s1 = 0xffffffffu;
          if (act0 && !prev0) s0 = hashf[loc0];
          if (act1 && !prev1) s1 = hashf[loc1];
          if (act0) atomicOr(&fb[loc0 >> 5], 1u << (loc0 & 31u));
          __syncwarp();
          const bool hit1 = act1 && ((fb[loc1 >> 5] >> (loc1 & 31u)) & 1u);
          __syncwarp();
          if (act0) fb[loc0 >> 5] = 0u;
          const bool seen0 = act0 && ((prev0 ? fp0 : s0) == a0);
          const bool seen1 = act1 && !hit1 && ((prev1 ? fp1 : s1) == a1);
          __syncwarp();   // every lookup has been consumed: the stores may go
          if (act0 && (peers0 >> lane) <= 1u) hashf[loc0] = a0;
          __syncwarp();
          if (act1 && (peers1 >> lane) <= 1u) hashf[loc1] = a1;
          const bool keep0 = act0 && !seen0 && ((long long)a0 != self), keep1 = act1 && !seen1 && ((long long)a1 != self);
          const unsigned km0 = __ballot_sync(PANN_FULL, keep0), km1 = __ballot_sync(PANN_FULL, keep1);
          if (keep0) nbr[__popc(km0 & lt_mask)] = a0;
          if (keep1) nbr[__popc(km0) + __popc(km1 & lt_mask)] = a1;
          S = __popc(km0) + __popc(km1);
          __syncwarp();
          filtered = true;
        }
      }
      if (!filtered)
      for (int base = 0; base < ne; base += 32) {
        int i = base + lane;
        bool act = i < ne;
        uint32_t a = base == 0 ? a0 : (base == 32 ? a1 : (act ? __ldg(grow + 1 + i) : 0u));
        uint32_t loc = act ? ((uint32_t)pann_hash64_2(a) & hmask) : (0x80000000u | lane);  // inactive lanes never match
        unsigned peers = __match_any_sync(PANN_FULL, loc);
        unsigned prev = peers & lt_mask;
        uint32_t from_prev = __shfl_sync(PANN_FULL, a, prev ? (31 - __clz(prev)) : lane);
        bool seen = false;
        if (act) {
          uint32_t slot = prev ? from_prev : hashf[loc];
          seen = (slot == a);
        }
        __syncwarp();
        if (act && (peers >> lane) <= 1u) hashf[loc] = a;  // last lane of each slot group leaves its id there
        bool keep = act && !seen && ((long long)a != self);
        unsigned km = __ballot_sync(PANN_FULL, keep);
        if (keep) nbr[S + __popc(km & lt_mask)] = a;
        S += __popc(km);
        __syncwarp();
      }
      cmps += S;  // :155
      if (S + C > (int)A.ccap) { status = 1; break; }  // cannot happen (ccap >= maxdeg + beam/8); guards memory
      pann_row_distances<DT, METRIC, QR>(A, nbr, S, qsm, lane, [&](uint32_t id, float d, bool ok) {
        bool take = ok && (d < cutoff);  // :153-158
        unsigned m = __ballot_sync(PANN_FULL, take);
        if (take) cand[C + __popc(m & lt_mask)] = pann_key(d, id);
        C += __popc(m);
      }, stage);
      __syncwarp();
      // skip the merge while there are too few candidates (:162-168); candidates are kept
      if (C == 0) continue;
      if (A.limit >= 2ll * B && C < B / 8) {
        bool more = false;
        for (int base = 0; base < F && !more; base += 32) {
          int i = base + lane;
          more = __any_sync(PANN_FULL, i < F && !(front[i] & 1ull));
        }
        if (more) continue;
      }
      // ---- merge (:173-208).  (a) rank sort of the candidates by strict "less": equal keys (the same vertex reached
      //      twice, :174's std::unique) land in the same slot, and the slot they leave empty keeps the sentinel
      for (int i = lane; i < C; i += 32) scand[i] = ~0ull;
      __syncwarp();
      for (int i = lane; i < C; i += 32) {
        const u64 ci = cand[i];
        scand[pann_count_below_any(cand, C, ci)] = ci;
      }
      __syncwarp();
      // (b) drop the empty slots and entries already in the frontier (set_union keeps one),
      //     place the rest: slot = rank among kept candidates + #frontier entries below it
      int Cv = 0;
      unsigned occ = 0;  // B <= 32: slots of the new frontier taken by candidates
      for (int base = 0; base < C; base += 32) {
        int i = base + lane;
        bool valid = false;
        int lbf = 0;
        u64 c = 0;
        if (i < C) {
          c = scand[i];
          if (c != ~0ull) {
            lbf = pann_count_below(front, F, c);
            valid = !(lbf < F && (front[lbf] & ~1ull) == c);
          }
        }
        unsigned m = __ballot_sync(PANN_FULL, valid);
        int vr = Cv + __popc(m & lt_mask);
        bool placed = false;
        if (valid) {
          cand[vr] = c;
          int np = vr + lbf;
          if (np < B) {
            u64 fl = 0ull;
            if (!vis_in_front) {  // re-entering an already visited vertex keeps it visited
              int vp = GS ? pann_count_below_far(vis, V, c) : pann_count_below(vis, V, c);
              fl = (vp < V && vis[vp] == c) ? 1ull : 0ull;
            }
            nf[np] = c | fl;
            placed = true;
            lbf = np;
          }
        }
        if (B <= 32) occ |= __reduce_or_sync(PANN_FULL, placed ? (1u << lbf) : 0u);
        Cv += __popc(m);
      }
      __syncwarp();
      // (c) the frontier entries fill the remaining slots in order
      if (B <= 32) {
        const int nfs0 = min(B, F + Cv);
        if (lane < nfs0 && !((occ >> lane) & 1u)) nf[lane] = front[__popc(~occ & lt_mask)];
      } else {
        for (int i = lane; i < F; i += 32) {
          u64 f = front[i];
          int np = i + pann_count_below(cand, Cv, f & ~1ull);
          if (np < B) nf[np] = f;
        }
      }
      int nfs = min(B, F + Cv);  // :185
      C = 0;
      __syncwarp();
      // (d) cut rule (:190-195): threshold (float)(cut * (double)dist[k]), never below the old size
      if (A.cut_rule && nfs > A.k) {
        float thr = (float)(A.cut * (double)pann_key_dist(nf[A.k]));
        const u64 tk = ((u64)pann_fsort(thr) << 32) | 1ull;  // x <= tk  <=>  (x >> 1) <= ((thr, id 0) >> 1)
        int c = 0;
        for (int base = 0; base < nfs; base += 32) {  // warp-uniform key: count by ballot
          const int i = base + lane;
          const int pc = __popc(__ballot_sync(PANN_FULL, i < nfs && nf[i] <= tk));
          c += pc;
          if (pc < 32) break;
        }
        nfs = max(c, F);
      }
      u64* t = front; front = nf; nf = t;
      F = nfs;
    }

    // ---- results
    if (status == 0 && F < A.k) status = 2;  // beamSearch.h:416-419
    if (A.out_ids)
      for (int j = lane; j < A.k; j += 32) {
        bool ok = j < F;
        A.out_ids[q * A.k + j] = ok ? pann_key_id(front[j]) : 0xffffffffu;
        if (A.out_dists) A.out_dists[q * A.k + j] = ok ? pann_key_dist(front[j]) : 0.f;
      }
    if (A.fr_ids)
      for (int j = lane; j < F && j < (int)A.fcap; j += 32) {
        A.fr_ids[q * A.fcap + j] = pann_key_id(front[j]);
        A.fr_dists[q * A.fcap + j] = pann_key_dist(front[j]);
      }
    if (A.vis_ids) {
      if (V > (int)A.vout && status == 0) status = 1;
      for (int j = lane; j < V && j < (int)A.vout; j += 32) {
        A.vis_ids[q * A.vout + j] = pann_key_id(vis[j]);
        if (A.vis_dists) A.vis_dists[q * A.vout + j] = pann_key_dist(vis[j]);
      }
    }
    if (lane == 0) {
      if (A.out_visited) A.out_visited[q] = V;
      if (A.out_cmps) A.out_cmps[q] = cmps;
      if (A.n_frontier) A.n_frontier[q] = F;
      if (A.n_visited) A.n_visited[q] = V;
      if (A.status) A.status[q] = status;
      if (status && A.any_status) atomicOr(A.any_status, 1 << status);
      if (A.bytes_total)
        atomicAdd(A.bytes_total, (u64)cmps * A.rowbytes + 4ull * (nread + (unsigned)V) + A.rowbytes + 4ull * (unsigned)A.k);
    }
    __syncwarp();
  }
}

// kernel lookup, one translation unit per element type (beam_search_{f32,u8,i8}.cu)
const void* pann_beam_kernel_dt0(int metric, int qr, bool gs);
const void* pann_beam_kernel_dt1(int metric, int qr, bool gs);
const void* pann_beam_kernel_dt2(int metric, int qr, bool gs);
