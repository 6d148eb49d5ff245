// beam_search_small.cuh — K1s: the beam search for narrow beams (beam <= 32, the regime of the recall-0.95 operating
// points), one query per warp, with the whole frontier in REGISTERS: lane i holds frontier entry i.
//
// Same behavioural contract as pann_beam_search_kernel (beam_search.cuh; parlayANN::filtered_beam_search,
// algorithms/utils/beamSearch.h:22-214) and bit-identical results; what differs is how a hop is executed:
//  * next-to-visit, cutoff, cut rule and results are ballots / shuffles on the register frontier — no shared-memory
//    round trips;
//  * the neighbours are not compacted: lane l filters and scores neighbour l (then l+32) itself, so each round's row
//    gathers are issued right after its filter step and overlap the next round's hashing;
//  * the merge is a run of ballot-ranked insertions (one per surviving candidate, usually < beam of them): position =
//    popc(ballot(entry < candidate)), duplicates found by ballot, entries above the position move up by one lane with
//    a shuffle.  This equals set_union + truncate (:173-185) because both are "the beam smallest distinct keys";
//  * the visited list is an unsorted id list (membership is all the search needs); callers that want the sorted
//    visited list (the build) use the general kernel.
// Shared memory per warp: hash filter (4 KB at beam <= 64), pending candidates, query row, visited ids.
#pragma once
#include "beam_search.cuh"

// distance from the query (shared memory) to base row `a`, computed by this lane alone, chunks in ascending order
template <int DT, int METRIC, int QR, class RG>
__device__ __forceinline__ float pann_lane_distance(const RG& A, uint32_t a, bool on, const uint4* qsm) {
  DistAcc<DT, METRIC> acc;
  if (on) {
    const uint8_t* row = A.points + (u64)a * A.stride;
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
  return acc.result();
}
__device__ __forceinline__ void pann_prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ u64 pann_shfl64(u64 x, int src) {
  const uint32_t lo = __shfl_sync(PANN_FULL, (uint32_t)x, src), hi = __shfl_sync(PANN_FULL, (uint32_t)(x >> 32), src);
  return ((u64)hi << 32) | lo;
}
__device__ __forceinline__ u64 pann_shfl_up64(u64 x) {
  const uint32_t lo = __shfl_up_sync(PANN_FULL, (uint32_t)x, 1), hi = __shfl_up_sync(PANN_FULL, (uint32_t)(x >> 32), 1);
  return ((u64)hi << 32) | lo;
}

// ---- shared memory by 32-bit address (no generic pointers in the hop loop: they cost a register pair each and an
// address-space conversion wherever the compiler cannot prove the window)
__device__ __forceinline__ u64 pann_lds64(uint32_t a) { u64 v; asm volatile("ld.shared.u64 %0, [%1];" : "=l"(v) : "r"(a)); return v; }
__device__ __forceinline__ void pann_sts64(uint32_t a, u64 v) { asm volatile("st.shared.u64 [%0], %1;" ::"r"(a), "l"(v) : "memory"); }
__device__ __forceinline__ uint32_t pann_lds32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ void pann_sts32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ uint4 pann_lds128(uint32_t a) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ void pann_sts128(uint32_t a, uint4 v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// #{j < n : keys[j] < t} over a shared-memory key list (16-byte aligned), two keys per 128-bit broadcast load
__device__ __forceinline__ int pann_count_below_s(uint32_t keys, int n, u64 t) {
  int n0 = 0, n1 = 0, j = 0;
#pragma unroll 1
  for (; j + 2 <= n; j += 2) {
    const uint4 v = pann_lds128(keys + 8u * j);
    pann_borrow_count(n0, ((u64)v.y << 32) | v.x, t);
    pann_borrow_count(n1, ((u64)v.w << 32) | v.z, t);
  }
  if (j < n) pann_borrow_count(n0, pann_lds64(keys + 8u * j), t);
  return -(n0 + n1);
}
// the same over a list padded with ~0 keys to a multiple of four: four keys per iteration, no tail
__device__ __forceinline__ int pann_count_below_s4(uint32_t keys, int n4, u64 t) {
  int n0 = 0, n1 = 0;
#pragma unroll 1
  for (int j = 0; j < n4; j += 4) {
    const uint4 v = pann_lds128(keys + 8u * j), w = pann_lds128(keys + 8u * j + 16u);
    pann_borrow_count(n0, ((u64)v.y << 32) | v.x, t);
    pann_borrow_count(n1, ((u64)v.w << 32) | v.z, t);
    pann_borrow_count(n0, ((u64)w.y << 32) | w.x, t);
    pann_borrow_count(n1, ((u64)w.w << 32) | w.z, t);
  }
  return -(n0 + n1);
}
__device__ __forceinline__ u64 pann_shfl_xor64(u64 x, int d) {
  const uint32_t lo = __shfl_xor_sync(PANN_FULL, (uint32_t)x, d), hi = __shfl_xor_sync(PANN_FULL, (uint32_t)(x >> 32), d);
  return ((u64)hi << 32) | lo;
}
// pann_quad_distance with the query row given by its 32-bit shared address
template <int DT, int METRIC, class RG>
__device__ __forceinline__ float pann_quad_distance_s(const RG& A, uint32_t a, unsigned keepmask, uint32_t qs, int lane) {
  const int sub = lane & 3, rowi = lane >> 2;
  uint4 v[4][2];
#pragma unroll
  for (int g = 0; g < 4; g++) {
    const int src = g * 8 + rowi;
    const uint32_t ag = __shfl_sync(PANN_FULL, a, src);
#if !defined(PANN_QUAD_ZERO)   // default (1M queries on 10M x 128 u8: 47.2 -> 45.7 ms): unwanted rows read row 0 (an L1 hit) instead of leaving zeroed registers: no predicates, no 32 zeroing moves per round
    const uint8_t* row = A.points + (u64)(((keepmask >> src) & 1u) ? ag : 0u) * A.stride + sub * 16;
    v[g][0] = pann_ldg16(row);
    v[g][1] = pann_ldg16(row + 64);
#elif defined(PANN_QUAD_ASM)
    pann_ldg16x2_if(A.points + (u64)ag * A.stride + sub * 16, (keepmask >> src) & 1u, v[g][0], v[g][1]);
#else
#ifndef PANN_QUAD_UNDEF   // leaving the unwanted rows' registers undefined (or predicating the loads in inline asm) makes ptxas spill ~400 bytes
    v[g][0] = v[g][1] = make_uint4(0u, 0u, 0u, 0u);
#endif
    if ((keepmask >> src) & 1u) {
      const uint8_t* row = A.points + (u64)ag * A.stride + sub * 16;
      v[g][0] = pann_ldg16(row);
      v[g][1] = pann_ldg16(row + 64);
    }
#endif
  }
  const uint4 q0 = pann_lds128(qs + 16u * sub), q1 = pann_lds128(qs + 16u * sub + 64u);
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
  const unsigned mine = __shfl_sync(PANN_FULL, tot, (lane & 7) * 4 + (lane >> 3));
  DistAcc<DT, METRIC> res;
  res.acc = (decltype(res.acc))mine;
  return res.result();
}

// The same in two steps, for the WIDE kernel: all loads of BOTH 32-neighbour rounds go out before anything is summed.
template <class RG>
__device__ __forceinline__ void pann_quad_gather_s(const RG& A, uint32_t a, unsigned keepmask, int lane, uint4 (&v)[4][2]) {
  const int sub = lane & 3, rowi = lane >> 2;
#pragma unroll
  for (int g = 0; g < 4; g++) {
    const int src = g * 8 + rowi;
    const uint32_t ag = __shfl_sync(PANN_FULL, a, src);
    const uint8_t* row = A.points + (u64)(((keepmask >> src) & 1u) ? ag : 0u) * A.stride + sub * 16;
    v[g][0] = pann_ldg16(row);
    v[g][1] = pann_ldg16(row + 64);
  }
}
template <int DT, int METRIC>
__device__ __forceinline__ float pann_quad_reduce_s(const uint4 (&v)[4][2], uint32_t qs, int lane) {
  const int sub = lane & 3;
  const uint4 q0 = pann_lds128(qs + 16u * sub), q1 = pann_lds128(qs + 16u * sub + 64u);
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
  const unsigned mine = __shfl_sync(PANN_FULL, tot, (lane & 7) * 4 + (lane >> 3));
  DistAcc<DT, METRIC> res;
  res.acc = (decltype(res.acc))mine;
  return res.result();
}

// EXT: per-query start lists and the vertex -> row map (beamSearchRandom, HNSW layers); a separate instantiation so that
// the plain kernel — which sits at its register budget — is not touched by them.
//
// The merge (beamSearch.h:173-185: sort candidates, unique, set_union with the frontier, truncate to the beam) is ONE
// batched step per 32 candidates: the candidates are rank-sorted through shared memory into DESCENDING lane order, the
// lane-wise minimum with the (ascending) register frontier is then the 32 smallest keys of the union as a bitonic
// sequence, and five compare-exchange stages sort it.  Keys of the same vertex are equal up to the flag bit, so a vertex
// that is reached twice ends up in adjacent lanes: that (rare) case, and two equal candidates, fall back to the
// one-at-a-time insertion below, which resolves duplicates exactly as set_union does.
// "Already visited" is a 1024-bit signature across the warp (one register per lane) with an exact list behind it.
// WIDE (small batches: the launch cannot fill the GPU, so a walk's latency is the run time): a hop's two 32-neighbour
// rounds are filtered first, then the row gathers of BOTH go out together and are summed afterwards — one memory round trip
// per hop instead of two, at 45 more registers (4 CTAs per SM instead of 6, which a small batch does not need anyway).
template <int DT, int METRIC, int QR, bool EXT, bool WIDE = false>
#ifndef PANN_K1S_WIDE_MINB
#define PANN_K1S_WIDE_MINB 4   // 124 registers, 16 warps per SM (3: 134 registers; the same times up to 1 776 queries, and 2 368 queries 0.282 vs 0.240 ms)
#endif
#ifndef PANN_K1S_MINB
#define PANN_K1S_MINB 6  // 78 registers, 24 warps per SM.  Measured with the batched merge, 1M queries on 10M x 128 u8 (CTAs per SM: ms): 8: 53.3, 7: 47.5, 6: 46.6, 5: 54.0, 4: 58.9 — and 1 000 queries (latency bound): 0.33 ms at 7, 0.27 ms at 6
#endif
__global__ void __launch_bounds__(PANN_WARPS_PER_CTA * 32, WIDE ? PANN_K1S_WIDE_MINB : (QR > 0 ? PANN_K1S_MINB : 4)) pann_beam_search_small_kernel(const BeamArgs A) {
  extern __shared__ __align__(16) uint8_t pann_smem[];
  const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
  const unsigned lt_mask = (1u << lane) - 1u;
  const uint32_t ws = (uint32_t)__cvta_generic_to_shared(pann_smem) + (uint32_t)wic * A.warp_bytes;  // this warp's slab
  const uint32_t s_cand = ws + A.off_cand, s_srt = ws + A.off_scand, s_q = ws + A.off_q, s_vis = ws + A.off_vis, s_hash = ws + A.off_hash;
  const uint32_t hmask = (1u << A.bits) - 1u;
  const int B = A.beam;
  // query indices fit 32 bits (the host checks nq < 2^31): narrow scalars keep the register file for the gathers
  const int wglobal = (int)blockIdx.x * PANN_WARPS_PER_CTA + wic;
  const int wstride = (int)gridDim.x * PANN_WARPS_PER_CTA;
  const int nq = (int)A.nq;
  const int limit = (int)(A.limit < 0x7fffffffll ? A.limit : 0x7fffffffll);
  constexpr bool QUAD = pann_quad_rows<DT, QR>();
  // generic views for the row shapes that do not take the four-lanes-per-row path
  const uint4* qsm = reinterpret_cast<const uint4*>(pann_smem + (size_t)wic * A.warp_bytes + A.off_q);
  uint4* stage = reinterpret_cast<uint4*>(pann_smem + (size_t)wic * A.warp_bytes + A.off_stage);
  auto next_query = [&](int ql) -> int {
    if (!A.qcounter) return ql + wstride;
    unsigned nx = 0;
    if (lane == 0) nx = atomicAdd(A.qcounter, 1u);
    return (int)min((unsigned)wstride + __shfl_sync(PANN_FULL, nx, 0), 0x7fffffffu);
  };
  auto distance = [&](uint32_t a, unsigned mask) -> float {
    if constexpr (QUAD) return pann_quad_distance_s<DT, METRIC>(A, a, mask, s_q, lane);
    else if constexpr (pann_staged_rows<DT, QR>()) return pann_staged_distance<DT, METRIC>(A, a, mask, qsm, stage, lane);
    else return pann_lane_distance<DT, METRIC, QR>(A, a, (mask >> lane) & 1u, qsm);
  };
  for (int ql = wglobal; ql < nq; ql = next_query(ql)) {
    const int q = A.qmap ? (int)A.qmap[ql] : ql;
    {
      const uint4 ones = make_uint4(~0u, ~0u, ~0u, ~0u);
      for (int i = lane; i < (1 << A.bits) / 4; i += 32) pann_sts128(s_hash + 16u * i, ones);
    }
    const uint32_t self = A.self_ids ? A.self_ids[q] : 0xffffffffu;  // ids are < 2^31
    {
      const uint8_t* qrow = A.self_ids ? A.points + (u64)self * A.stride : A.queries + (u64)q * A.qstride;
      __syncwarp();
      for (int ch = lane; ch < (int)A.nch; ch += 32) pann_sts128(s_q + 16u * ch, pann_ldg16(qrow + ch * 16));
      __syncwarp();
    }

    // row prefetch policy (see the main loop): always, or — for the queries handed out last — once half the warps retired
    const bool always_pf = (A.tune & 1u) != 0;
#ifdef PANN_NO_DRAIN
    const bool drain_watch = false;
#else
    const bool drain_watch = !always_pf && A.qcounter && !(A.tune & 2u) && (ql + wstride >= nq);
#endif
    u64 fk = ~0ull;  // frontier entry `lane` ((dist,id) key, bit 0 = visited), ~0 beyond the frontier
    int F = 0, V = 0, C = 0, status = 0;
    uint32_t vsig = 0;   // bit (h & 31) of lane (h >> 5) is set for every visited id, h = 10-bit hash of the id
    const uint32_t* qstarts = EXT ? A.starts + (u64)q * A.starts_stride : A.starts;
    const int ns = (int)((EXT && A.nstarts_q) ? A.nstarts_q[q] : A.nstarts);
    unsigned cmps = ns, nread = 0;
    auto vhash = [](uint32_t id) -> uint32_t { return (id * 2654435761u) >> 22; };
    // is `id` (warp-uniform) in the visited list?  exact
    auto visited_exact = [&](uint32_t id) -> bool {
      bool found = false;
      for (int b0 = 0; b0 < V; b0 += 32) found |= (b0 + lane < V) && (pann_lds32(s_vis + 4u * (b0 + lane)) == id);
      return __any_sync(PANN_FULL, found);
    };

    // ---- starting points (:64-70).  More starts than the beam width: see pann_beam_search_kernel.
    if (ns < 1 || ns > 32 || (ns > B && A.cut_rule && A.k < B)) {
      status = 3;
    } else {
      const bool on = lane < ns;
      const uint32_t a = on ? qstarts[lane] : 0u;
      const float d = distance(a, __ballot_sync(PANN_FULL, on));
      const u64 key = on ? pann_key(d, a) : ~0ull;
      if (lane == 0)
        for (int i = 0; i < ns; i++) { const uint32_t s = qstarts[i]; pann_sts32(s_hash + 4u * ((uint32_t)pann_hash64_2(s) & hmask), s); }
      int r = 0;
      for (int j = 0; j < ns; j++) {
        const u64 kj = pann_shfl64(key, j);
        r += (kj < key) || (kj == key && j < lane);
      }
      if (on) pann_sts64(s_srt + 8u * r, key);
      __syncwarp();
      if (on) fk = pann_lds64(s_srt + 8u * lane);
      F = ns;
      __syncwarp();
    }

    // ---- main loop (:107-209)
    while (status == 0) {
      const unsigned unv = __ballot_sync(PANN_FULL, lane < F && !(fk & 1ull));
      if (!unv || V >= limit) break;
      const int pos = __ffs(unv) - 1;
      const bool more_after = __popc(unv) > 1;           // offset + 1 < remain (:164)
      const u64 cur = pann_shfl64(fk, pos);
      const uint32_t cur_id = pann_key_id(cur);
      const uint32_t* grow = A.graph + (u64)((EXT && A.rowmap) ? __ldg(A.rowmap + cur_id) : cur_id) * (A.maxdeg + 1);
      uint32_t drain_cnt = 0;  // read next to the graph row, so its latency is hidden under that load
      if (drain_watch) asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(drain_cnt) : "l"(A.qcounter));
      const int deg = (int)__ldg(grow);
      const uint32_t a0 = (lane < (int)A.maxdeg) ? __ldg(grow + 1 + lane) : 0u;
      const uint32_t a1 = (32 + lane < (int)A.maxdeg) ? __ldg(grow + 33 + lane) : 0u;
      if (lane == pos) fk |= 1ull;
      if (V >= (int)A.vcap) { status = 1; break; }
      {  // visited: exact id list in shared memory + the signature
        const uint32_t h = vhash(cur_id);
        if (lane == 0) pann_sts32(s_vis + 4u * V, cur_id);
        if (lane == (int)(h >> 5)) vsig |= 1u << (h & 31u);
        V++;
      }
      const bool full = (F == B);
      const float cutoff = full ? pann_key_dist(pann_shfl64(fk, F - 1)) : 2147483648.0f;  // (float)INT_MAX (:150-152)
      // Early drop (beams < 16, where every hop with a candidate merges): with the cut rule on and more than k entries, a
      // candidate that sorts after the whole frontier AND after the cut threshold of the CURRENT k-th entry is inserted
      // and cut again by the very same merge (:190-195; the new threshold can only be lower), so it never matters.
      u64 drop_above = ~0ull;
      if (A.cut_rule && F > A.k && B < 16) {
        const float thr0 = (float)(A.cut * (double)pann_key_dist(pann_shfl64(fk, A.k)));
        const u64 tk0 = ((u64)pann_fsort(thr0) << 32) | 1ull;
        const u64 worst = pann_shfl64(fk, F - 1) | 1ull;
        drop_above = tk0 > worst ? tk0 : worst;
      }
      const int ne = min(deg, A.degree_limit);
      nread += ne;
      // L2 prefetch of every neighbour row, issued before the filter runs, takes one memory latency out of the hop.  It
      // pays when the memory system has slack — rows mostly L2-resident (tune bit 0, set by the host for small point
      // sets), or the launch is draining: every warp that runs out of queries still bumps the query counter, so
      // counter - (nq - warps) = retired warps (only the last `warps` queries of a launch look at it).  Under a saturated gather stream (10M points, all warps busy) the extra
      // requests cost more than the latency they hide (measured: -40 % throughput), so it stays off there.
      const bool pf = always_pf || (drain_watch && ((long long)drain_cnt + wstride - nq) * 2 >= wstride);
      if (pf) {
        if (lane < ne) pann_prefetch_l2(A.points + (u64)a0 * A.stride);
        if (32 + lane < ne) pann_prefetch_l2(A.points + (u64)a1 * A.stride);
      }
      if constexpr (WIDE && QUAD) {
        // both rounds through the hash filter first (in list order, as the reference walks the row), then all row gathers
        // of the hop in flight together; the host launches this variant only for max degree <= 64
        uint32_t aa[2];
        bool kp[2];
        unsigned km[2];
#pragma unroll
        for (int rd = 0; rd < 2; rd++) {
          const int i = rd * 32 + lane;
          const bool act = i < ne;
          const uint32_t a = rd == 0 ? a0 : a1;
          const uint32_t loc = act ? ((uint32_t)pann_hash64_2(a) & hmask) : (0x80000000u | lane);
          const unsigned peers = __match_any_sync(PANN_FULL, loc);
          const unsigned prev = peers & lt_mask;
          const uint32_t from_prev = __shfl_sync(PANN_FULL, a, prev ? (31 - __clz(prev)) : lane);
          bool seen = false;
          if (act) {
            const uint32_t slot = prev ? from_prev : pann_lds32(s_hash + 4u * loc);
            seen = (slot == a);
          }
          __syncwarp();
          if (act && (peers >> lane) <= 1u) pann_sts32(s_hash + 4u * loc, a);
          __syncwarp();
          kp[rd] = act && !seen && (a != self);
          km[rd] = __ballot_sync(PANN_FULL, kp[rd]);
          aa[rd] = a;
        }
        cmps += __popc(km[0]) + __popc(km[1]);  // :155
        uint4 v0[4][2], v1[4][2];
        pann_quad_gather_s(A, aa[0], km[0], lane, v0);
        if (km[1]) pann_quad_gather_s(A, aa[1], km[1], lane, v1);
#pragma unroll
        for (int rd = 0; rd < 2; rd++) {
          if (rd == 1 && !km[1]) break;
          const float d = rd == 0 ? pann_quad_reduce_s<DT, METRIC>(v0, s_q, lane) : pann_quad_reduce_s<DT, METRIC>(v1, s_q, lane);
          const u64 key = pann_key(d, aa[rd]);
          const bool take = kp[rd] && (d < cutoff) && (key <= drop_above);  // :153-158
          const unsigned m = __ballot_sync(PANN_FULL, take);
          if (C + __popc(m) > (int)A.ccap) { status = 1; break; }
          if (take) pann_sts64(s_cand + 8u * (C + __popc(m & lt_mask)), key);
          C += __popc(m);
          __syncwarp();
        }
      } else
      for (int base = 0; base < ne; base += 32) {
        const int i = base + lane;
        const bool act = i < ne;
        const uint32_t a = base == 0 ? a0 : (base == 32 ? a1 : (act ? __ldg(grow + 1 + i) : 0u));
        const uint32_t loc = act ? ((uint32_t)pann_hash64_2(a) & hmask) : (0x80000000u | lane);
        const unsigned peers = __match_any_sync(PANN_FULL, loc);
        const unsigned prev = peers & lt_mask;
        const uint32_t from_prev = __shfl_sync(PANN_FULL, a, prev ? (31 - __clz(prev)) : lane);
        bool seen = false;
        if (act) {
          const uint32_t slot = prev ? from_prev : pann_lds32(s_hash + 4u * loc);
          seen = (slot == a);
        }
        __syncwarp();
        if (act && (peers >> lane) <= 1u) pann_sts32(s_hash + 4u * loc, a);
        const bool keep = act && !seen && (a != self);
        const unsigned keepmask = __ballot_sync(PANN_FULL, keep);
        cmps += __popc(keepmask);  // :155
        const float d = distance(a, keepmask);
        const u64 key = pann_key(d, a);
        const bool take = keep && (d < cutoff) && (key <= drop_above);  // :153-158
        const unsigned m = __ballot_sync(PANN_FULL, take);
        if (C + __popc(m) > (int)A.ccap) { status = 1; break; }  // cannot happen (ccap >= maxdeg + beam/8); guards memory
        if (take) {
          pann_sts64(s_cand + 8u * (C + __popc(m & lt_mask)), key);
        }
        C += __popc(m);
        __syncwarp();
      }
      if (status) break;
      if (C == 0) continue;
      if (A.limit >= 2ll * B && C < B / 8 && more_after) continue;  // :162-168, candidates are kept
      // ---- merge (:173-208), 32 candidates at a time
      const int F_old = F;
      for (int c0 = 0; c0 < C; c0 += 32) {
        const int n = min(32, C - c0);
        const uint32_t s_chunk = s_cand + 8u * c0;
        u64 ck = lane < n ? pann_lds64(s_chunk + 8u * lane) : ~0ull;
#ifndef PANN_K1S_RANK2
        // pad the (last) chunk with ~0 keys to a multiple of four for the rank loop; up to three slots past the list: inside
        // the candidate array, or the head of s_srt, which is dead until it is rewritten below
        if (lane >= n && lane < ((n + 3) & ~3)) pann_sts64(s_chunk + 8u * lane, ~0ull);
        __syncwarp();
#endif
        {  // a vertex that comes back after it was visited stays visited (it is in `visited`, so set_difference drops it, :203-208)
          const uint32_t cid = pann_key_id(ck);
          const uint32_t h = vhash(cid);
          const uint32_t w = __shfl_sync(PANN_FULL, vsig, (int)(h >> 5));
          unsigned hm = __ballot_sync(PANN_FULL, lane < n && ((w >> (h & 31u)) & 1u));
          while (hm) {
            const int j = __ffs(hm) - 1;
            hm &= hm - 1;
            if (visited_exact(__shfl_sync(PANN_FULL, cid, j)) && lane == j) ck |= 1ull;
          }
        }
#ifndef PANN_K1S_RANK2
        const int r = pann_count_below_s4(s_chunk, (n + 3) & ~3, ck & ~1ull);   // rank among the chunk's keys (raw keys: bit 0 clear)
#else
        const int r = pann_count_below_s(s_chunk, n, ck & ~1ull);
#endif
        pann_sts64(s_srt + 8u * lane, ~0ull);
        __syncwarp();
        if (lane < n) pann_sts64(s_srt + 8u * (31 - r), ck);          // descending in lane order
        __syncwarp();
        const u64 x = pann_lds64(s_srt + 8u * lane);
        bool slow = __popc(__ballot_sync(PANN_FULL, x != ~0ull)) != n;   // two equal candidates shared a slot
        u64 v = x < fk ? x : fk;                                     // the 32 smallest of frontier + chunk, bitonic
#pragma unroll
        for (int dd = 16; dd > 0; dd >>= 1) {
          const u64 p = pann_shfl_xor64(v, dd);
          const bool up = (lane & dd) != 0;
          v = ((p < v) != up) ? p : v;                               // lower lane keeps the minimum, upper the maximum
        }
        {
          const u64 pv = pann_shfl_up64(v);
          slow |= __any_sync(PANN_FULL, lane > 0 && v != ~0ull && ((v ^ pv) >> 1) == 0ull);   // the same vertex twice
        }
        if (!slow) {
          fk = (lane < B) ? v : ~0ull;
          F = min(B, F + n);
        } else {
          // one candidate at a time: position by ballot, duplicates skipped, entries above move up one lane
          for (int j = 0; j < n; j++) {
            const u64 c = pann_shfl64(ck, j);
            const int lb = __popc(__ballot_sync(PANN_FULL, (fk | 1ull) < (c & ~1ull)));
            const unsigned dup = __ballot_sync(PANN_FULL, (fk & ~1ull) == (c & ~1ull));
            if (dup || lb >= B) continue;
            const u64 upk = pann_shfl_up64(fk);
            if (lane == lb) fk = c;
            else if (lane > lb) fk = upk;
            if (lane >= B) fk = ~0ull;
            F = min(F + 1, B);
          }
          if (EXT && F > B) { F = B; if (lane >= B) fk = ~0ull; }   // a start list longer than the beam: the merge truncates (:185)
        }
        __syncwarp();
      }
      C = 0;
      // cut rule (:190-195)
      int nfs = F;
      if (A.cut_rule && nfs > A.k) {
        const float thr = (float)(A.cut * (double)pann_key_dist(pann_shfl64(fk, A.k)));
        const u64 tk = ((u64)pann_fsort(thr) << 32) | 1ull;
        const int c = __popc(__ballot_sync(PANN_FULL, lane < nfs && fk <= tk));
        nfs = max(c, F_old);
        if (lane >= nfs) fk = ~0ull;
      }
      F = nfs;
      __syncwarp();
    }

    // ---- results
    if (status == 0 && F < A.k) status = 2;  // beamSearch.h:416-419
    if (A.out_ids && lane < A.k) {
      const bool ok = lane < F;
      A.out_ids[(u64)q * A.k + lane] = ok ? pann_key_id(fk) : 0xffffffffu;
      if (A.out_dists) A.out_dists[(u64)q * A.k + lane] = ok ? pann_key_dist(fk) : 0.f;
    }
    if (A.fr_ids && lane < F && lane < (int)A.fcap) {
      A.fr_ids[(u64)q * A.fcap + lane] = pann_key_id(fk);
      A.fr_dists[(u64)q * A.fcap + lane] = pann_key_dist(fk);
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

// K7: the re-rank step of beam_search_rerank when the searched copy is quantised (beamSearch.h:426-444): exact distances
// from the exact query to the first min(k * rerank_factor, |frontier|) frontier entries, sorted by (dist, id), first k
// out.  One warp per query; one exact row per lane, summed in the reference's own order (floats included; float rows are
// loaded by the whole warp and handed over through the shared-memory stage, see pann_staged_distance).
struct RerankArgs {
  const uint8_t* points; uint32_t stride, nch;   // exact base rows (names as in BeamArgs: pann_lane_distance reads them)
  const uint8_t* queries; uint32_t qstride;      // exact query rows
  uint64_t nq;
  const uint32_t* fr_ids; const uint32_t* n_frontier; uint32_t fcap;
  int k; long long rerank_factor;
  uint32_t* out_ids; float* out_dists;
  uint32_t off_q, off_stage, warp_bytes;   // off_stage: float rows staged (pann_staged_distance)
};
template <int DT, int METRIC>
__global__ void __launch_bounds__(PANN_WARPS_PER_CTA * 32) pann_rerank_kernel(const RerankArgs A) {
  extern __shared__ __align__(16) uint8_t pann_smem[];
  const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
  uint8_t* ws = pann_smem + (size_t)wic * A.warp_bytes;
  u64* keys = reinterpret_cast<u64*>(ws);
  uint4* qsm = reinterpret_cast<uint4*>(ws + A.off_q);
  for (long long q = (long long)blockIdx.x * PANN_WARPS_PER_CTA + wic; q < (long long)A.nq; q += (long long)gridDim.x * PANN_WARPS_PER_CTA) {
    pann_load_query(A.queries + (u64)q * A.qstride, A.nch, qsm, lane);
    const int nfr = (int)A.n_frontier[q];
    const int nchk = (int)min((long long)A.k * A.rerank_factor, (long long)min(nfr, (int)A.fcap));
    for (int base = 0; base < nchk; base += 32) {
      const int i = base + lane;
      const bool on = i < nchk;
      const uint32_t id = on ? A.fr_ids[(u64)q * A.fcap + i] : 0u;
      float d;
      if constexpr (pann_staged_rows<DT, 0>())
        d = pann_staged_distance<DT, METRIC>(A, id, __ballot_sync(PANN_FULL, on), qsm, reinterpret_cast<uint4*>(ws + A.off_stage), lane);
      else
        d = pann_lane_distance<DT, METRIC, 0>(A, id, on, qsm);
      if (on) keys[i] = pann_key(d, id);
    }
    __syncwarp();
    for (int i = lane; i < nchk; i += 32) {   // rank sort: only the first k ranks are written
      const u64 ki = keys[i];
      const int r = pann_count_below_any(keys, nchk, ki);   // keys are distinct (distinct ids)
      if (r < A.k) {
        A.out_ids[(u64)q * A.k + r] = pann_key_id(ki);
        if (A.out_dists) A.out_dists[(u64)q * A.k + r] = pann_key_dist(ki);
      }
    }
    __syncwarp();
  }
}

const void* pann_beam_small_kernel_dt0(int metric, int qr, bool ext);
const void* pann_beam_small_kernel_dt1(int metric, int qr, bool ext);
const void* pann_beam_small_kernel_dt2(int metric, int qr, bool ext);
const void* pann_beam_small_wide_kernel_dt0(int metric);
const void* pann_beam_small_wide_kernel_dt1(int metric);
const void* pann_beam_small_wide_kernel_dt2(int metric);
const void* pann_rerank_kernel_dt0(int metric);
const void* pann_rerank_kernel_dt1(int metric);
const void* pann_rerank_kernel_dt2(int metric);
