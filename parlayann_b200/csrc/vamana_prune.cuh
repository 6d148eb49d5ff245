// vamana_prune.cuh — K3: robustPrune, one warp per pruned vertex.
//
// Behavioural contract: knn_index::robustPrune (algorithms/vamana/index.h:63-120 and the ids-only overload
// :124-137) and the reverse-edge step of batch_insert (:289-300) including add_neighbors_without_repeats
// (:140-146): same candidate order ((dist,id) total order, adjacent duplicate ids dropped), same greedy
// alpha-occlusion scan with the test  alpha * (double)d(p*,p') <= (double)d(p,p')  evaluated in double, same
// distance-comparison counts.
//
// GPU layout (not a translation of the CPU loop): the candidate list is one array of 64-bit (dist,id) keys
// in shared memory; the list is kept COMPACT — after each pick the survivors are rewritten in place by
// ballot-ranked stores, so every pass of the occlusion scan is a dense batch of 128-bit row gathers
// (same per-lane row accumulation as K1, so float distances are bit-identical between the kernels — and to
// the CPU's — and duplicate ids always sort next to each other).
#pragma once
#include "beam_search.cuh"

#define PANN_PRUNE_WARPS 4
#ifndef PANN_PRUNE_MINB
#define PANN_PRUNE_MINB 3   // without a min-blocks bound ptxas pins the kernel at an occupancy step (80 / 48 registers, spilling)
#endif
#if PANN_PRUNE_MINB > 0
#define PANN_PRUNE_BOUNDS __launch_bounds__(PANN_PRUNE_WARPS * 32, PANN_PRUNE_MINB)
#else
#define PANN_PRUNE_BOUNDS __launch_bounds__(PANN_PRUNE_WARPS * 32)   // A/B: the round-1 bound
#endif

struct PruneArgs {
  // row geometry (same member names as BeamArgs: pann_row_distances_g reads them)
  const uint8_t* points; uint32_t stride, nch;
  uint32_t* graph; uint32_t maxdeg, R;
  uint64_t n;
  double alpha;
  // rule 0: Vamana's robustPrune (vamana/index.h:63-120).  rule 1: HNSW's prune_heuristic (HNSW/HNSW.hpp:430-486): a later
  // candidate c dies when d(c, r) < alpha_f * d(c, u) for a kept r (float product, strict <).  rule 2: HNSW's reverse-edge
  // trim (HNSW.hpp:963-987): the R closest of old + new neighbours, no occlusion test; and in mode 1 the list that fits is
  // old neighbours followed by the new ones (:989), not Vamana's new-then-old-without-repeats.
  int rule; float alpha_f;
  const uint32_t* rowmap;   // optional vertex id -> graph row (HNSW's compacted upper layers); null = identity
  int mode;                 // 0: phase 1 (visited lists -> new_out), 1: phase 3 (reverse-edge groups -> graph rows)
  uint32_t ntasks;          // mode 0: batch size; mode 1: upper bound (real count read from *ngroups)
  const uint32_t* perm_lo;  // insertion order, already offset to the batch: task i inserts perm_lo[i]
  // mode 0 inputs: K1 build-mode outputs
  const uint32_t* vis_ids; const float* vis_dists; const uint32_t* nvis; const uint32_t* search_cmps; uint32_t vout;
  uint32_t* new_out; uint32_t* new_deg;  // ntasks x R, ntasks
  // mode 1 inputs: sorted (target << 32 | seq) keys, group heads
  const u64* sorted; uint32_t nsorted; const uint32_t* gstart; const uint32_t* ngroups;   // gstart: (first key, length or 0) per group
  // statistics (stats<indexType>, may be null)
  uint32_t* st_visited; uint32_t* st_cmps;
  // tasks whose candidate list exceeds `cap` are deferred to the global-scratch launch
  uint32_t cap; uint32_t* big_list; uint32_t* big_count;  // big_count[0] = #deferred, [1] = their largest candidate count
  const uint32_t* task_map; uint32_t ntask_map;  // GS launch: the deferred tasks; their number is read on the DEVICE from big_count[0] (ntask_map is unused)
  u64* g_scratch;                                 // GS launch: per warp 2 x cap keys
  int* gs_overflow;                               // GS launch: set when a list exceeds even the slab capacity (checked once, after the build)
  // accounting for the build's roofline (SURVEY.md §8d B_ins): [0] += C * row bytes + 4 (R+1) per robustPrune call (C = candidates
  // after unique), [1] += 1 per call, [2] += pair distances of the call (the reference's own count); may be null
  u64* acct; uint32_t rowbytes;
  uint32_t off_b, off_q, off_stage, warp_bytes;   // off_stage: float rows staged for the one-lane sums (pann_staged_distance)
  // row cache (shared memory): the rows of up to rc_rows candidates, rc_pitch bytes apart, and their slot numbers
  uint32_t rc_rows, rc_pitch, off_rc, off_pos;
  // Gram tile (8-bit rows of 128 bytes in the row cache): gram_tp = pitch in words of the 16-row distance strip (0 = off)
  uint32_t gram_tp, off_gram, off_nrm;
};

// D(16x8, s32) += A(16x32, 8-bit, row) * B(32x8, 8-bit, col) on the tensor cores (warp-level MMA)
template <int DT>
__device__ __forceinline__ void pann_mma_16x8x32(int (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  if constexpr (DT == PANN_DT_U8)
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  else
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// does the picked vertex (at distance d_sc from candidate c) remove c, which is at distance d_c from the pruned vertex?
__device__ __forceinline__ bool pann_occludes(const PruneArgs& A, float d_sc, float d_c) {
  if (A.rule == 0) return A.alpha * (double)d_sc <= (double)d_c;   // vamana/index.h:111, evaluated in double
  return d_sc < d_c * A.alpha_f;                                    // HNSW.hpp:456-469, float
}

template <int DT, int METRIC, int QR, bool GS>
__global__ void PANN_PRUNE_BOUNDS pann_prune_kernel(const PruneArgs A) {
  extern __shared__ __align__(16) uint8_t pann_smem[];
  const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
  const unsigned lt_mask = (1u << lane) - 1u;
  uint8_t* ws = pann_smem + (size_t)wic * A.warp_bytes;
  const int wpc = blockDim.x >> 5;  // warps per CTA: PANN_PRUNE_WARPS, fewer when the row cache needs the shared memory
  const long long wglobal = (long long)blockIdx.x * wpc + wic;
  const long long wstride = (long long)gridDim.x * wpc;
  u64 *bufA, *bufB;
  if constexpr (GS) {
    bufA = A.g_scratch + (u64)wglobal * 2 * A.cap;
    bufB = bufA + A.cap;
  } else {
    bufA = reinterpret_cast<u64*>(ws);
    bufB = reinterpret_cast<u64*>(ws + A.off_b);
  }
  uint4* qsm = reinterpret_cast<uint4*>(ws + A.off_q);
  uint4* stage = pann_staged_rows<DT, QR>() ? reinterpret_cast<uint4*>(ws + A.off_stage) : nullptr;
  uint32_t* cid = reinterpret_cast<uint32_t*>(bufB);  // candidate ids before they get their distances
  const uint32_t ntasks = GS ? *A.big_count : (A.mode == 1 ? *A.ngroups : A.ntasks);

  for (long long tl = wglobal; tl < (long long)ntasks; tl += wstride) {
    const uint32_t task = GS ? A.task_map[tl] : (uint32_t)tl;
    // ---- who is pruned, and its candidates
    uint32_t p, nc, g0 = 0;
    if (A.mode == 0) {
      p = A.perm_lo[task];
      nc = A.nvis[task];
    } else {
      g0 = A.gstart[2 * task];
      p = (uint32_t)(A.sorted[g0] >> 32);
      nc = A.gstart[2 * task + 1];
      if (nc == 0) {   // a long group: its end = first key of a larger target (keys are sorted)
        uint32_t lo = g0 + 1, hi = A.nsorted;
        while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if ((uint32_t)(A.sorted[mid] >> 32) <= p) lo = mid + 1; else hi = mid; }
        nc = lo - g0;
      }
    }
    uint32_t* grow = A.graph + (u64)(A.rowmap ? A.rowmap[p] : p) * (A.maxdeg + 1);
    const uint32_t deg = grow[0];
    const uint32_t C0 = nc + deg;
    if (C0 > A.cap) {
      if constexpr (GS) {  // beyond the slab as well: reported, the host fails the build (PANN_PRUNE_SLAB_CAP raises the capacity)
        if (lane == 0) atomicExch(A.gs_overflow, (int)C0);
      } else {             // defer to the global-scratch launch
        if (lane == 0) { A.big_list[atomicAdd(A.big_count, 1u)] = task; atomicMax(A.big_count + 1, C0); }
      }
      continue;
    }
    // ---- reverse edges that fit: new ones first, then the old ones not among them (index.h:291-294, :140-146)
    if (A.mode == 1 && C0 <= A.R && A.rule == 2) {   // HNSW: nbh_v.insert(end, new ones) (HNSW.hpp:989)
      for (uint32_t i = lane; i < nc; i += 32) grow[1 + deg + i] = A.perm_lo[(uint32_t)A.sorted[g0 + i] / A.R];
      if (lane == 0) grow[0] = C0;
      __syncwarp();
      continue;
    }
    if (A.mode == 1 && C0 <= A.R) {
      for (uint32_t i = lane; i < nc; i += 32) cid[i] = A.perm_lo[(uint32_t)A.sorted[g0 + i] / A.R];
      __syncwarp();
      uint32_t m = nc;
      for (uint32_t base = 0; base < deg; base += 32) {
        uint32_t i = base + lane;
        uint32_t x = i < deg ? grow[1 + i] : 0u;
        bool keep = i < deg;
        if (keep) for (uint32_t j = 0; j < nc; j++) if (cid[j] == x) { keep = false; break; }
        unsigned km = __ballot_sync(PANN_FULL, keep);
        __syncwarp();
        if (keep) cid[m + __popc(km & lt_mask)] = x;
        m += __popc(km);
      }
      __syncwarp();
      // the row is rewritten only after all of it has been read
      for (uint32_t i = lane; i < m; i += 32) grow[1 + i] = cid[i];
      if (lane == 0) grow[0] = m;
      __syncwarp();
      continue;
    }
    // ---- query row = the pruned vertex
    const uint8_t* prow = A.points + (u64)p * A.stride;
    auto load_q = [&](const uint8_t* row) { pann_load_query(row, A.nch, qsm, lane); };
    load_q(prow);
    unsigned comps = 0;
    int C = 0;  // keys in bufA
    // candidates that already carry a distance (mode 0: the visited list, sorted by K1)
    if (A.mode == 0) {
      const uint32_t* vi = A.vis_ids + (u64)task * A.vout;
      const float* vd = A.vis_dists + (u64)task * A.vout;
      for (uint32_t i = lane; i < nc; i += 32) bufA[i] = pann_key(vd[i], vi[i]);
      C = (int)nc;
    }
    // candidates that need one: mode 1 group members (:131-134), then the current out-neighbours (:72-77)
    {
      uint32_t m = 0;
      if (A.mode == 1) { for (uint32_t i = lane; i < nc; i += 32) cid[i] = A.perm_lo[(uint32_t)A.sorted[g0 + i] / A.R]; m = nc; }
      for (uint32_t i = lane; i < deg; i += 32) cid[m + i] = grow[1 + i];
      m += deg;
      __syncwarp();
      comps += m;
      pann_row_distances_g<DT, METRIC, QR, true>(
          A, [&](int r) { return (u64)cid[r] << 1; }, (int)m, qsm, lane, [&](u64 tag, float d, bool ok) {
            unsigned bm = __ballot_sync(PANN_FULL, ok);
            if (ok) bufA[C + __popc(bm & lt_mask)] = pann_key(d, pann_key_id(tag));
            C += __popc(bm);
          }, stage);
      __syncwarp();
    }
    // ---- sort by (dist,id) (:80-83).  Mode 0 without out-neighbours is sorted already.
    u64* S = bufB;
    if (A.mode == 0 && deg == 0) {
      S = bufA;
    } else {
      const u64* src = bufB;
      if constexpr (GS) {
        // long lists (a hub that collected thousands of reverse edges in one batch): bitonic sort in place, O(C log^2 C)
        // instead of the rank sort's O(C^2).  Equal keys are identical (same vertex, same distance): order is moot.
        int msz = 2; while (msz < C) msz <<= 1;   // <= cap: the host rounds the scratch slab up to a power of two
        for (int i = C + lane; i < msz; i += 32) bufA[i] = ~0ull;
        __syncwarp();
        for (int k = 2; k <= msz; k <<= 1)
          for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = lane; i < (msz >> 1); i += 32) {
              const int t = 2 * i - (i & (j - 1)), q = t + j;
              const u64 x = bufA[t], y = bufA[q];
              if ((x > y) == ((t & k) == 0)) { bufA[t] = y; bufA[q] = x; }
            }
            __syncwarp();
          }
        src = bufA;
      } else {
        // rank sort by strict "less" (four keys per iteration, borrow-counted): equal keys — the same vertex reached twice, with
        // the same distance — land in the same slot, and the slot they leave empty keeps the sentinel that unique drops below
        for (int i = lane; i < C; i += 32) bufB[i] = ~0ull;
        __syncwarp();
        for (int i = lane; i < C; i += 32) {
          const u64 ci = bufA[i];
          bufB[pann_count_below_any(bufA, C, ci)] = ci;
        }
        __syncwarp();
      }
      // std::unique on ids (:86-88): the same id has the same distance, so duplicates are adjacent
      int w = 0;
      for (int base = 0; base < C; base += 32) {
        int i = base + lane;
        u64 c = i < C ? src[i] : 0;
        bool keep = i < C && c != ~0ull && (i == 0 || pann_key_id(src[i - 1]) != pann_key_id(c));
        unsigned km = __ballot_sync(PANN_FULL, keep);
        __syncwarp();
        if (keep) bufB[w + __popc(km & lt_mask)] = c;
        w += __popc(km);
        __syncwarp();
      }
      C = w;
    }
    // ---- greedy occlusion scan (:95-116); S[start, n) = live candidates in order
    uint32_t* outp = A.mode == 0 ? A.new_out + (u64)task * A.R : grow + 1;
    int start = 0, n = C;
    uint32_t m = 0;
    if (A.rule == 2) {   // the R closest, in order (HNSW.hpp:976-980)
      m = (uint32_t)min(C, (int)A.R);
      for (uint32_t i = lane; i < m; i += 32) outp[i] = pann_key_id(S[i]);
    } else if (!GS && A.rc_rows && C <= (int)A.rc_rows) {
      // Row cache: every candidate row is fetched ONCE (all gathers in flight together) into shared memory; the scan
      // below — up to R picks, each a pass over the live candidates — then never leaves the SM.  Without it each pick
      // costs two dependent trips to L2/HBM (the picked row, then the live rows), ~130 per pruned vertex.
      uint4* rc = reinterpret_cast<uint4*>(ws + A.off_rc);
      uint16_t* P = reinterpret_cast<uint16_t*>(ws + A.off_pos);  // cache slot of S[i], compacted with S
      const uint32_t pw = A.rc_pitch / 16;
      for (int i = lane; i < C * (int)A.nch; i += 32) {
        const int r = i / (int)A.nch, ch = i - r * (int)A.nch;
        rc[r * pw + ch] = pann_ldg16(A.points + (u64)pann_key_id(S[r]) * A.stride + ch * 16);
      }
      for (int i = lane; i < C; i += 32) P[i] = (uint16_t)i;
      __syncwarp();
      if constexpr (QR == 8 && DT != PANN_DT_F32) {
        if (A.gram_tp) {
          // ---- Gram tile.  With every candidate row on chip the occlusion scan is a dense contraction: the picks walk the
          // sorted list, and each needs its distance to every later live candidate — for lists that lose few members
          // (reverse-edge groups: 64 old edges plus a handful of new ones, at most R survive) that is half of the full
          // C x C matrix.  It is computed on the tensor cores sixteen rows at a time, G = X X^T as 16x8x32 integer MMAs
          // straight from the row cache (A and B fragments are both 4-byte reads of cached rows: pitch 144 bytes puts the
          // eight rows of a fragment on disjoint banks), turned into the exact integer distance |xi|^2 + |xj|^2 - 2 G
          // (the same integer the dp4a loop produces, so the float handed to the alpha test is bit-identical) and parked
          // in a 16-row strip; the scan then reads one strip row per pick.  Candidates are not compacted any more: a
          // bit mask per 32 candidates says who is alive.
          const uint8_t* rcb = reinterpret_cast<const uint8_t*>(rc);
          int* T = reinterpret_cast<int*>(ws + A.off_gram);
          int* nrm = reinterpret_cast<int*>(ws + A.off_nrm);
          const int tp = (int)A.gram_tp, pitch = (int)A.rc_pitch;
          const int g = lane >> 2, t4 = (lane & 3) * 4;
          constexpr int NCH = 5;                       // C <= 160
          unsigned alive[NCH];
          float kd[NCH];
#pragma unroll
          for (int c = 0; c < NCH; c++) {
            const int j = 32 * c + lane;
            alive[c] = 32 * c < C ? (C - 32 * c >= 32 ? 0xffffffffu : ((1u << (C - 32 * c)) - 1u)) : 0u;
            kd[c] = j < C ? pann_key_dist(S[j]) : 0.f;
            if (METRIC == PANN_M_L2 && j < C) {        // |x_j|^2
              const uint4* row = rc + (uint32_t)j * pw;
              DistAcc<DT, PANN_M_MIPS> nn;
#pragma unroll
              for (int ch = 0; ch < 8; ch++) { const uint4 v = row[ch]; nn.add(v, v); }
              nrm[j] = (int)nn.acc;
            }
          }
          __syncwarp();
          // (the masks are registers: picked by a warp-uniform index through an unrolled select, never indexed dynamically)
          auto mask_of = [&](int c) -> unsigned {
            unsigned r = 0;
#pragma unroll
            for (int x = 0; x < NCH; x++) r = (x == c) ? alive[x] : r;
            return r;
          };
          const int NB = (C + 7) >> 3;
          bool done = false;
          for (int b = 0; 16 * b < C && !done; b++) {
            if (((mask_of(b >> 1) >> (16 * (b & 1))) & 0xffffu) == 0u) continue;    // nobody left to pick in these rows
            const int r0 = min(16 * b + g, C - 1), r1 = min(16 * b + g + 8, C - 1);
            uint32_t af[4][4];
#pragma unroll
            for (int ks = 0; ks < 4; ks++) {
              af[ks][0] = *reinterpret_cast<const uint32_t*>(rcb + r0 * pitch + ks * 32 + t4);
              af[ks][1] = *reinterpret_cast<const uint32_t*>(rcb + r1 * pitch + ks * 32 + t4);
              af[ks][2] = *reinterpret_cast<const uint32_t*>(rcb + r0 * pitch + ks * 32 + 16 + t4);
              af[ks][3] = *reinterpret_cast<const uint32_t*>(rcb + r1 * pitch + ks * 32 + 16 + t4);
            }
            const int n0 = METRIC == PANN_M_L2 ? nrm[r0] : 0, n1 = METRIC == PANN_M_L2 ? nrm[r1] : 0;
            for (int nb = 2 * b; nb < NB; nb++) {      // only columns at or behind the strip's own rows are ever read
              const int cr = min(8 * nb + g, C - 1);
              int acc[4] = {0, 0, 0, 0};
#pragma unroll
              for (int ks = 0; ks < 4; ks++) {
                const uint32_t b0 = *reinterpret_cast<const uint32_t*>(rcb + cr * pitch + ks * 32 + t4);
                const uint32_t b1 = *reinterpret_cast<const uint32_t*>(rcb + cr * pitch + ks * 32 + 16 + t4);
                pann_mma_16x8x32<DT>(acc, af[ks], b0, b1);
              }
              const int c0 = 8 * nb + (lane & 3) * 2;  // columns c0, c0 + 1 of rows g and g + 8
              if (METRIC == PANN_M_L2) {
                const int m0 = nrm[min(c0, C - 1)], m1 = nrm[min(c0 + 1, C - 1)];
                acc[0] = n0 + m0 - 2 * acc[0]; acc[1] = n0 + m1 - 2 * acc[1];
                acc[2] = n1 + m0 - 2 * acc[2]; acc[3] = n1 + m1 - 2 * acc[3];
              }
              *reinterpret_cast<int2*>(T + g * tp + c0) = make_int2(acc[0], acc[1]);
              *reinterpret_cast<int2*>(T + (g + 8) * tp + c0) = make_int2(acc[2], acc[3]);
            }
            __syncwarp();
            const int iend = min(16 * b + 16, C);
            for (int i = 16 * b; i < iend; i++) {
              if (!((mask_of(i >> 5) >> (i & 31)) & 1u)) continue;
              const uint32_t ps_id = pann_key_id(S[i]);
              if (ps_id == p) continue;
              if (lane == 0) outp[m] = ps_id;
              m++;
              const int* Ti = T + (i & 15) * tp;
#pragma unroll
              for (int c = 0; c < NCH; c++) {
                if (32 * c + 31 <= i || 32 * c >= C) continue;     // warp-uniform: no later candidate in this chunk
                const int j = 32 * c + lane;
                const unsigned later = alive[c] & (i >= 32 * c ? (0xfffffffeu << (i - 32 * c)) : 0xffffffffu);
                comps += __popc(later);                              // one distance per live later candidate (:108)
                const bool on = (later >> lane) & 1u;
                const int v = on ? Ti[j] : 0;
                const float d = METRIC == PANN_M_L2 ? (float)v : -((float)v);
                const bool kill = on && pann_occludes(A, d, kd[c]);  // :111
                alive[c] &= ~__ballot_sync(PANN_FULL, kill);
              }
              if (m >= A.R) { done = true; break; }
            }
            __syncwarp();
          }
          start = n = 0;   // the scan below has nothing left to do
        }
      }
      while (m < A.R && start < n) {
        const u64 ps = S[start];
        const uint4* qrow = rc + (uint32_t)P[start] * pw;
        start++;
        const uint32_t ps_id = pann_key_id(ps);
        if (ps_id == p) continue;
        if (lane == 0) outp[m] = ps_id;
        m++;
        const int cnt = n - start;
        if (cnt == 0) break;
        comps += cnt;  // one distance per live later candidate (:108)
        int w = start;
        // rows of exactly QR chunks: the picked row lives in registers for the whole pass and the candidate's chunks are
        // all loaded before the arithmetic (this path runs at ~6 warps per SM, shared memory being the limit, so the
        // instruction-level parallelism inside the warp is what hides the shared-memory latency)
        uint4 qv[QR > 0 ? QR : 1];
        if constexpr (QR > 0) {
#pragma unroll
          for (int c = 0; c < QR; c++) qv[c] = qrow[c];
        }
        for (int base = 0; base < cnt; base += 32) {
          const int r = base + lane;
          const bool ok = r < cnt;
          u64 key = 0;
          uint16_t slot = 0;
          DistAcc<DT, METRIC> acc;
          if (ok) {
            key = S[start + r];
            slot = P[start + r];
            const uint4* row = rc + (uint32_t)slot * pw;
            if constexpr (QR > 0) {
              uint4 v[QR];
#pragma unroll
              for (int c = 0; c < QR; c++) v[c] = row[c];
#pragma unroll
              for (int c = 0; c < QR; c++) acc.add(v[c], qv[c]);
            } else {
              for (int c = 0; c < (int)A.nch; c++) acc.add(row[c], qrow[c]);
            }
          }
          const float d = acc.result();
          const bool keep = ok && !pann_occludes(A, d, pann_key_dist(key));  // :111
          const unsigned km = __ballot_sync(PANN_FULL, keep);
          if (keep) { const int o = w + __popc(km & lt_mask); S[o] = key; P[o] = slot; }
          w += __popc(km);
          __syncwarp();
        }
        n = w;
      }
    } else {
    while (m < A.R && start < n) {
      const u64 ps = S[start];
      start++;
      const uint32_t ps_id = pann_key_id(ps);
      if (ps_id == p) continue;
      if (lane == 0) outp[m] = ps_id;
      m++;
      const int cnt = n - start;
      if (cnt == 0) break;
      comps += cnt;  // one distance per live later candidate (:108)
      load_q(A.points + (u64)ps_id * A.stride);
      int w = start;
      pann_row_distances_g<DT, METRIC, QR, true>(
          A, [&](int r) { return S[start + r]; }, cnt, qsm, lane, [&](u64 key, float d, bool ok) {
            bool keep = ok && !pann_occludes(A, d, pann_key_dist(key));  // :111
            unsigned km = __ballot_sync(PANN_FULL, keep);
            if (keep) S[w + __popc(km & lt_mask)] = key;
            w += __popc(km);
          }, stage);
      __syncwarp();
      n = w;
    }
    }
    if (lane == 0) {
      if (A.acct) {
        atomicAdd(A.acct, (u64)C * A.rowbytes + 4ull * (A.R + 1));
        atomicAdd(A.acct + 1, 1ull);
        atomicAdd(A.acct + 2, (u64)comps);
      }
      if (A.mode == 0) {
        A.new_deg[task] = m;
        if (A.st_cmps) A.st_cmps[p] += A.search_cmps[task] + comps;
        if (A.st_visited) A.st_visited[p] += nc;
      } else {
        grow[0] = m;
        if (A.st_cmps) A.st_cmps[p] += comps;
      }
    }
    __syncwarp();
  }
}

// K6: distance of every edge to its owner, one warp per graph row (build_index's final per-row sort,
// algorithms/vamana/index.h:180-185, compares exactly these values).
//   sort == 0: out = n x maxdeg floats, the caller sorts.
//   sort == 1: rows whose distances are all distinct have exactly one sorted order and are rewritten in place (rank
//     sort in shared memory).  The reference's comparator looks at the distance only and std::sort is not stable, so a
//     row WITH ties must end up where libstdc++'s introsort leaves it: such rows are copied (edges + distances) to a
//     compact record list for the host, which runs that very std::sort on them (a few percent of the rows).
struct EdgeDistArgs {
  const uint8_t* points; uint32_t stride, nch;
  uint32_t* graph; uint32_t maxdeg; uint64_t n;
  float* out;
  uint32_t off_q, off_d, off_s, warp_bytes;
  int sort;
  uint32_t* tie_count; uint32_t tie_cap; uint32_t* tie_rows; uint32_t* tie_graph; float* tie_dist;
};
template <int DT, int METRIC, int QR>
__global__ void __launch_bounds__(PANN_PRUNE_WARPS * 32, 4) pann_edge_dist_kernel(const EdgeDistArgs A) {
  extern __shared__ __align__(16) uint8_t pann_smem[];
  const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
  uint8_t* ws = pann_smem + (size_t)wic * A.warp_bytes;
  uint32_t* ids = reinterpret_cast<uint32_t*>(ws);
  uint4* qsm = reinterpret_cast<uint4*>(ws + A.off_q);
  float* dsm = reinterpret_cast<float*>(ws + A.off_d);
  uint32_t* sorted = reinterpret_cast<uint32_t*>(ws + A.off_s);
  const long long wglobal = (long long)blockIdx.x * PANN_PRUNE_WARPS + wic;
  const long long wstride = (long long)gridDim.x * PANN_PRUNE_WARPS;
  for (long long i = wglobal; i < (long long)A.n; i += wstride) {
    uint32_t* grow = A.graph + (u64)i * (A.maxdeg + 1);
    const int deg = (int)grow[0];
    const uint8_t* row = A.points + (u64)i * A.stride;
    pann_load_query(row, A.nch, qsm, lane);
    for (int j = lane; j < deg; j += 32) ids[j] = grow[1 + j];
    __syncwarp();
    float* o = A.sort ? dsm : A.out + (u64)i * A.maxdeg;
    int w = 0;
    pann_row_distances_g<DT, METRIC, QR, true>(
        A, [&](int r) { return (u64)ids[r] << 1; }, deg, qsm, lane, [&](u64, float d, bool ok) {
          unsigned bm = __ballot_sync(PANN_FULL, ok);
          if (ok) o[w + __popc(bm & ((1u << lane) - 1u))] = d;
          w += __popc(bm);
        });
    __syncwarp();
    if (A.sort) {
      bool tie = false;
      for (int j = lane; j < deg; j += 32) {
        const float dj = dsm[j];
        int r = 0;
        for (int k = 0; k < deg; k++) {
          const float dk = dsm[k];
          r += (dk < dj) || (dk == dj && k < j);
          tie |= (dk == dj) && (k != j);
        }
        sorted[r] = ids[j];
      }
      __syncwarp();
      if (!__any_sync(PANN_FULL, tie)) {
        for (int j = lane; j < deg; j += 32) grow[1 + j] = sorted[j];
      } else {
        uint32_t slot = 0;
        if (lane == 0) slot = atomicAdd(A.tie_count, 1u);
        slot = __shfl_sync(PANN_FULL, slot, 0);
        if (slot < A.tie_cap) {
          uint32_t* tg = A.tie_graph + (u64)slot * (A.maxdeg + 1);
          float* td = A.tie_dist + (u64)slot * A.maxdeg;
          if (lane == 0) { A.tie_rows[slot] = (uint32_t)i; tg[0] = (uint32_t)deg; }
          for (int j = lane; j < deg; j += 32) { tg[1 + j] = ids[j]; td[j] = dsm[j]; }
        }
      }
    }
    __syncwarp();
  }
}
const void* pann_prune_kernel_dt0(int metric, int qr, bool gs);
const void* pann_prune_kernel_dt1(int metric, int qr, bool gs);
const void* pann_prune_kernel_dt2(int metric, int qr, bool gs);
const void* pann_edge_kernel_dt0(int metric, int qr);
const void* pann_edge_kernel_dt1(int metric, int qr);
const void* pann_edge_kernel_dt2(int metric, int qr);
