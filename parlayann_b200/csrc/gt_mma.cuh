// gt_mma.cuh — K4 v2: brute-force ground truth as a tcgen05/TMA GEMM with a fused per-row threshold filter.
//
// compute_groundtruth (data_tools/compute_groundtruth.cpp:22-60) scans every base row per query.  Here a CTA owns 128
// queries (the A operand, M = 128, loaded once by TMA) and streams 256-row tiles of the base set (B operand, N = 256)
// through a 4-stage TMA pipeline; one elected thread issues tcgen05.mma (kind::i8 for uint8/int8 rows — exact int32 dot
// products; kind::tf32 for float rows) into a double-buffered 2 x 256-column TMEM accumulator; sixteen epilogue warps read
// the accumulators back with tcgen05.ld and keep only the rows that can still enter the query's top-k:
//      |b|^2 - 2 q.b  <=  T_q          (L2;  T_q = tau_q + margin_q - |q|^2,  tau_q = current k-th distance)
//            - 2 q.b  <=  T_q          (MIPS; T_q = 2 (tau_q + margin_q))
// Survivors are appended (row id only) to a per-query candidate list.  pann_gt_shrink_kernel then recomputes the
// EXACT distance of every candidate with the reference's own operation order (DistAcc, as K1), merges with the
// running top-k and tightens tau_q.  The tensor-core pass therefore only has to be a superset filter: for 8-bit rows it
// is exact (margin 0), for float rows margin_q bounds the tf32 truncation error (2^-10 per operand), so the final
// result is bit-identical to the CUDA-core path and to the reference.  No distance matrix ever goes to HBM.
#pragma once
#include <cuda.h>
#include <type_traits>

#include "pann_internal.h"
#include "beam_search.cuh"   // pann_staged_distance / pann_quad_distance: the coalesced row gathers of K1

#define GTM_M 128          // queries per CTA (UMMA M)
#define GTM_N 256          // base rows per tile (UMMA N)
#define GTM_KB_BYTES 128   // bytes of K per smem stage = one 128B swizzle atom
#define GTM_STAGES 4
#define GTM_BN 8             // ring of per-tile |b|^2 rows (1 KB each)
#define GTM_EPI_WARPS 16    // epilogue warps: 4 TMEM lane quadrants x 4 column quarters
#define GTM_THREADS (64 + 32 * GTM_EPI_WARPS)   // warp 0: TMA producer, warp 1: MMA issuer, then the epilogue warps
#define GTM_CAND_CAP 4096  // candidate ids per query per chunk, split into private sub-lists (one per segment CTA and column quarter)
#define GTM_MAX_SUB 64      // sub-lists per query (2 x segments): no atomics, every epilogue thread owns its list and counter
#define GTM_A_STAGE_BYTES (GTM_M * GTM_KB_BYTES)   // 16 KB
#define GTM_B_STAGE_BYTES (GTM_N * GTM_KB_BYTES)   // 32 KB
#define GTS_KEYS 8192       // shrink kernel, large tier: key slots in shared memory (>= k + GTM_CAND_CAP, k <= 1024)
#define GTS_KEYS_SMALL 1024 // small tier (8 KB: eight CTAs per SM); queries with more keys set flag bit 1 and wait for the large tier

// ---------------------------------------------------------------------------------------------------------------
// PTX wrappers (sm_100a)
__device__ __forceinline__ uint32_t gtm_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void gtm_mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(gtm_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void gtm_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(gtm_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void gtm_mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(gtm_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void gtm_mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(gtm_smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void gtm_tma_load_2d(void* smem_dst, const CUtensorMap* tm, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(gtm_smem_u32(smem_dst)),
      "l"((uint64_t)tm), "r"(c0), "r"(c1), "r"(gtm_smem_u32(bar))
      : "memory");
}
// 1-D bulk copy global -> shared (bytes % 16 == 0, both sides 16-byte aligned), completion counted on an mbarrier
__device__ __forceinline__ void gtm_bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(gtm_smem_u32(smem_dst)),
               "l"((uint64_t)gsrc), "r"(bytes), "r"(gtm_smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void gtm_tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void gtm_tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void gtm_tcgen05_commit(uint64_t* bar) {  // implies fence::before_thread_sync
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(gtm_smem_u32(bar)) : "memory");
}
// K-major operand tile in shared memory, 128-byte swizzle: rows of 128 bytes, 8-row groups 1024 bytes apart
__device__ __forceinline__ uint64_t gtm_smem_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3ffff) >> 4);        // start address, bits [0,14)
  d |= (uint64_t)1 << 16;                             // leading byte offset (unused for swizzled K-major), bits [16,30)
  d |= (uint64_t)(1024 >> 4) << 32;                   // stride byte offset = 1024, bits [32,46)
  d |= (uint64_t)1 << 46;                             // descriptor version (sm_100), bits [46,48)
  d |= (uint64_t)2 << 61;                             // SWIZZLE_128B, bits [61,64)
  return d;
}
template <bool ISINT>
__device__ __forceinline__ void gtm_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  const uint32_t z = 0;
  if constexpr (ISINT) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(z), "r"(z), "r"(z), "r"(z)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(z), "r"(z), "r"(z), "r"(z)
        : "memory");
  }
}
// 32 consecutive accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void gtm_tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
        "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
        "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr));
}
__device__ __forceinline__ uint4 gtm_lds128(uint32_t saddr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(saddr));
  return r;
}
__device__ __forceinline__ void gtm_tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// the same 32 columns written from registers (accumulator pre-initialisation)
__device__ __forceinline__ void gtm_tmem_st32(uint32_t taddr, const uint4 (&b)[8]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(b[0].x), "r"(b[0].y), "r"(b[0].z), "r"(b[0].w), "r"(b[1].x), "r"(b[1].y), "r"(b[1].z), "r"(b[1].w), "r"(b[2].x), "r"(b[2].y), "r"(b[2].z),
      "r"(b[2].w), "r"(b[3].x), "r"(b[3].y), "r"(b[3].z), "r"(b[3].w), "r"(b[4].x), "r"(b[4].y), "r"(b[4].z), "r"(b[4].w), "r"(b[5].x), "r"(b[5].y),
      "r"(b[5].z), "r"(b[5].w), "r"(b[6].x), "r"(b[6].y), "r"(b[6].z), "r"(b[6].w), "r"(b[7].x), "r"(b[7].y), "r"(b[7].z), "r"(b[7].w)
      : "memory");
}
__device__ __forceinline__ void gtm_tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// mbarrier operations on a 32-bit shared address (the epilogue loop keeps the converted addresses in registers)
__device__ __forceinline__ void gtm_mbar_arrive_a(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void gtm_mbar_wait_a(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ int gtm_max3(int a, int b, int c) { return __vimax3_s32(a, b, c); }
__device__ __forceinline__ float gtm_max3(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }

// ---------------------------------------------------------------------------------------------------------------
// the PANN_GT_DBG breakdown switches (profiles/r2_gt_breakdown.txt) are compiled in only with -DPANN_GT_DEBUG: as run-time tests
// inside the column loop they were ~10 % of the epilogue's instructions
#ifdef PANN_GT_DEBUG
#define GTM_DBG(A, bits) ((A).dbg & (bits))
#else
#define GTM_DBG(A, bits) (false)
#endif
struct GtMmaArgs {
  uint32_t nq, kb, nstages;    // queries; 128-byte k-blocks per row; B pipeline depth (<= GTM_STAGES)
  uint64_t row_lo, row_hi;     // base rows of this chunk; row_lo is a multiple of GTM_N
  uint32_t tiles_per_seg;      // blockIdx.y walks tiles [y * tiles_per_seg, (y+1) * tiles_per_seg) of the chunk
  const void* bnorm;           // per base row: |b|^2 (L2) or 0 (MIPS); float or int32; padded with +inf / 2^30 to a tile multiple
  const void* thr;             // per query: T_q (float or int32)
  uint32_t* cand_ids;          // [nq][nsub][subcap]  (nsub = (GTM_EPI_WARPS / 4) * gridDim.y, nsub * subcap <= GTM_CAND_CAP)
  uint32_t* cand_cnt;          // [nq][GTM_MAX_SUB]
  uint32_t subcap;
  uint32_t dbg;                // experiments only (PANN_GT_DBG): 1 = skip the filter arithmetic, 2 = skip the TMEM loads and stores as well, 4 = max tree but no appends, 8 = no accumulator pre-initialisation
};

template <bool ISINT>
__global__ void __launch_bounds__(GTM_THREADS, 1)
pann_gt_mma_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmB, const GtMmaArgs A, const uint32_t idesc) {
  extern __shared__ __align__(1024) uint8_t gtm_smem_raw[];
  // carve (1024-byte aligned tiles): A[kb] | B[STAGES] | barriers | tmem base
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(gtm_smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = base;
  uint8_t* sB = sA + (size_t)A.kb * GTM_A_STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB + (size_t)A.nstages * GTM_B_STAGE_BYTES);
  uint64_t* full = bars;                       // [STAGES]  TMA -> MMA
  uint64_t* empty = bars + GTM_STAGES;         // [STAGES]  MMA -> TMA
  uint64_t* a_full = bars + 2 * GTM_STAGES;    // queries landed
  uint64_t* t_full = a_full + 1;               // [2] accumulator ready (MMA -> epilogue)
  uint64_t* t_empty = t_full + 2;              // [2] accumulator drained (epilogue -> MMA)
  uint64_t* bn_full = t_empty + 2;             // [GTM_BN] |b|^2 of a tile landed (bulk copy -> epilogue)
  uint64_t* bn_empty = bn_full + GTM_BN;       // [GTM_BN] slot read by all epilogue warps (epilogue -> producer)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bn_empty + GTM_BN);
  uint32_t* bn_s = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(bars) + 256);   // [GTM_BN][GTM_N] ring of per-tile |b|^2 rows

  const int widx = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t q0 = blockIdx.x * GTM_M;
  const uint64_t chunk_tiles = (A.row_hi - A.row_lo + GTM_N - 1) / GTM_N;
  const uint64_t t_lo = (uint64_t)blockIdx.y * A.tiles_per_seg;
  const uint64_t t_hi = min(chunk_tiles, t_lo + A.tiles_per_seg);
  const uint32_t ntiles = t_lo < t_hi ? (uint32_t)(t_hi - t_lo) : 0u;
  const uint64_t seg_row0 = A.row_lo + t_lo * GTM_N;

  if (widx == 0 && lane == 0) {
    for (int i = 0; i < GTM_STAGES; i++) { gtm_mbar_init(&full[i], 1); gtm_mbar_init(&empty[i], 1); }
    gtm_mbar_init(a_full, 1);
    for (int i = 0; i < 2; i++) { gtm_mbar_init(&t_full[i], 1); gtm_mbar_init(&t_empty[i], GTM_EPI_WARPS); }
    for (int i = 0; i < GTM_BN; i++) { gtm_mbar_init(&bn_full[i], 1); gtm_mbar_init(&bn_empty[i], GTM_EPI_WARPS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (widx == 1) {  // TMEM: all 512 columns (two 256-column accumulators); this warp also frees them
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(gtm_smem_u32(tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  gtm_tcgen05_fence_before();
  __syncthreads();
  gtm_tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (widx == 0) {
    // ===== TMA producer =====
    if (lane == 0 && ntiles > 0) {
      gtm_mbar_expect_tx(a_full, A.kb * GTM_A_STAGE_BYTES);
      for (uint32_t j = 0; j < A.kb; j++) gtm_tma_load_2d(sA + (size_t)j * GTM_A_STAGE_BYTES, &tmQ, (int)(j * GTM_KB_BYTES), (int)q0, a_full);
      uint32_t stage = 0, phase = 0, bs = 0, bphase = 0;
      for (uint32_t t = 0; t < ntiles; t++) {
        const int row0 = (int)(seg_row0 + (uint64_t)t * GTM_N);
        // the tile's |b|^2 row (1 KB) goes into its own small ring, so the operand pipeline never waits for the epilogue
        gtm_mbar_wait(&bn_empty[bs], bphase ^ 1);
        gtm_mbar_expect_tx(&bn_full[bs], GTM_N * 4);
        gtm_bulk_load(bn_s + bs * GTM_N, reinterpret_cast<const uint32_t*>(A.bnorm) + (uint64_t)row0, GTM_N * 4, &bn_full[bs]);
        if (++bs == GTM_BN) { bs = 0; bphase ^= 1; }
        for (uint32_t j = 0; j < A.kb; j++) {
          gtm_mbar_wait(&empty[stage], phase ^ 1);
          gtm_mbar_expect_tx(&full[stage], GTM_B_STAGE_BYTES);
          gtm_tma_load_2d(sB + (size_t)stage * GTM_B_STAGE_BYTES, &tmB, (int)(j * GTM_KB_BYTES), row0, &full[stage]);
          if (++stage == A.nstages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (widx == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0 && ntiles > 0) {
      gtm_mbar_wait(a_full, 0);
      uint32_t stage = 0, phase = 0, as = 0, aphase = 0;
      for (uint32_t t = 0; t < ntiles; t++) {
        gtm_mbar_wait(&t_empty[as], aphase);   // drained AND pre-initialised with -|b|^2 / 2 by the epilogue warps
        gtm_tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + as * GTM_N;
        for (uint32_t j = 0; j < A.kb; j++) {
          gtm_mbar_wait(&full[stage], phase);
          gtm_tcgen05_fence_after();
          const uint64_t adesc = gtm_smem_desc(gtm_smem_u32(sA + (size_t)j * GTM_A_STAGE_BYTES));
          const uint64_t bdesc = gtm_smem_desc(gtm_smem_u32(sB + (size_t)stage * GTM_B_STAGE_BYTES));
#pragma unroll
          for (uint32_t k4 = 0; k4 < 4; k4++)  // UMMA K = 32 bytes: advance the start address inside the swizzle atom
            gtm_mma<ISINT>(tmem_d, adesc + 2 * k4, bdesc + 2 * k4, idesc, 1u);
          gtm_tcgen05_commit(&empty[stage]);
          if (++stage == A.nstages) { stage = 0; phase ^= 1; }
        }
        gtm_tcgen05_commit(&t_full[as]);
        as ^= 1;
        if (as == 0) aphase ^= 1;
      }
    }
  } else {
    // ===== epilogue: GTM_EPI_WARPS warps; warp w owns TMEM lanes 32*(w%4).. and one column quarter =====
    // The accumulator is PRE-INITIALISED with -|b_j|^2 / 2 per column (tcgen05.st from the tile's row of the ring) and the MMA
    // always accumulates, so what comes back is s_j = q.b_j - |b_j|^2 / 2 and the filter "|b_j|^2 - 2 q.b_j <= T_q" is
    // s_j >= -T_q / 2: per 32 columns one TMEM load, a 3-input max tree and one compare — no per-element arithmetic, no
    // |b|^2 reads in the test (ncu of the previous epilogue: 2.3 CUDA-core instructions per accumulator element, tensor pipe
    // 23 % busy; profiles/r1_k4_mma_full_summary.txt).
    const uint32_t quad = widx & 3, part = (uint32_t)(widx - 2) >> 2;   // TMEM lanes 32*quad.., columns 64*part..
    const uint32_t q = q0 + quad * 32 + lane;
    const bool qok = q < A.nq;
    typedef typename std::conditional<ISINT, int, float>::type acc_t;
    acc_t S;   // pass mark on s_j
    if constexpr (ISINT) S = qok ? ((1 - reinterpret_cast<const int*>(A.thr)[q]) >> 1) : INT_MAX;   // ceil(-T / 2)
    else S = qok ? -0.5f * reinterpret_cast<const float*>(A.thr)[q] : INFINITY;
    const uint32_t sub = blockIdx.y * (GTM_EPI_WARPS / 4) + part;
    uint32_t* my_ids = A.cand_ids + (size_t)(qok ? q : 0) * GTM_CAND_CAP + (size_t)sub * A.subcap;
    uint32_t cnt = 0;   // appended so far (may run past subcap: the shrink kernel reports the overflow)
    const uint32_t bn_s_addr = gtm_smem_u32(bn_s);
    const uint32_t a_tfull = gtm_smem_u32(t_full), a_tempty = gtm_smem_u32(t_empty), a_bnfull = gtm_smem_u32(bn_full), a_bnempty = gtm_smem_u32(bn_empty);
    constexpr uint32_t NC = GTM_N / 32 / (GTM_EPI_WARPS / 4);   // 32-column groups per warp and tile
    const uint32_t tmem_mine = tmem_base + ((quad * 32) << 16) + part * (GTM_N / (GTM_EPI_WARPS / 4));
    // accumulator stage `st` := -|b|^2 / 2 of tile u (its row sits in ring slot u % GTM_BN), then both are handed on
    auto preinit = [&](uint32_t u, uint32_t st) {
      const uint32_t slot = u % GTM_BN;
      gtm_mbar_wait_a(a_bnfull + 8 * slot, (u / GTM_BN) & 1);
      if (!GTM_DBG(A, 2 | 8)) {
#pragma unroll 1
        for (uint32_t c = 0; c < NC; c++) {
          uint4 bn[8];
          const uint32_t bnp = bn_s_addr + (slot * GTM_N + part * (GTM_N / (GTM_EPI_WARPS / 4)) + c * 32) * 4;   // broadcast reads
#pragma unroll
          for (int i = 0; i < 8; i++) bn[i] = gtm_lds128(bnp + 16 * i);
          gtm_tmem_st32(tmem_mine + st * GTM_N + c * 32, bn);
        }
        gtm_tmem_st_wait();
      }
      gtm_tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) { gtm_mbar_arrive_a(a_tempty + 8 * st); gtm_mbar_arrive_a(a_bnempty + 8 * slot); }
    };
    if (ntiles > 0) preinit(0, 0);
    if (ntiles > 1) preinit(1, 1);
    uint32_t as = 0, aphase = 0;
    for (uint32_t t = 0; t < ntiles; t++) {
      const uint64_t row0 = seg_row0 + (uint64_t)t * GTM_N;
      gtm_mbar_wait_a(a_tfull + 8 * as, aphase);
      gtm_tcgen05_fence_after();
      unsigned hits[NC];
#pragma unroll
      for (uint32_t c = 0; c < NC; c++) hits[c] = 0u;
#pragma unroll
      for (uint32_t c = 0; c < NC; c++) {
        uint32_t v[32];
        if (GTM_DBG(A, 2)) continue;
        gtm_tmem_ld32(tmem_mine + as * GTM_N + c * 32, v);
        gtm_tmem_ld_wait();
        if (GTM_DBG(A, 1)) { if (v[0] == 0x7fffffffu && v[31] == 0x7ffffffeu) cnt++; continue; }
        acc_t w[32];
#pragma unroll
        for (int j = 0; j < 32; j++) {
          if constexpr (ISINT) w[j] = (int)v[j];
          else w[j] = __uint_as_float(v[j]);
        }
        // almost never >= S once the threshold has tightened: the maximum of the 32 first (3-input maxima: 16 operations)
        acc_t m9[11];
#pragma unroll
        for (int g = 0; g < 10; g++) m9[g] = gtm_max3(w[3 * g], w[3 * g + 1], w[3 * g + 2]);
        m9[10] = w[30] > w[31] ? w[30] : w[31];
        const acc_t m3a = gtm_max3(m9[0], m9[1], m9[2]), m3b = gtm_max3(m9[3], m9[4], m9[5]), m3c = gtm_max3(m9[6], m9[7], m9[8]);
        const acc_t mx = gtm_max3(gtm_max3(m3a, m3b, m3c), m9[9], m9[10]);
        if (GTM_DBG(A, 4)) { if (mx == (acc_t)12345) cnt++; continue; }   // experiments: the max tree without the append path
        if (mx >= S) {   // which of the 32 columns: a bit mask now, the global stores after the stage has been handed on
          unsigned m = 0;
#pragma unroll
          for (int j = 0; j < 32; j++) m |= (w[j] >= S ? 1u : 0u) << j;
          hits[c] = m;
        }
      }
      // this stage's next tenant is tile t + 2: initialise it now (its |b|^2 row has been in the ring for a while), which
      // also releases the stage to the MMA warp
      if (t + 2 < ntiles) {
        preinit(t + 2, as);
      } else {
        gtm_tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) gtm_mbar_arrive_a(a_tempty + 8 * as);
      }
      // the candidates of this tile go out now, behind the barrier arrivals: a store that is still in flight when a warp
      // arrives makes the (releasing) arrive wait for it, and that wait sat on the accumulator hand-over — measured on
      // 10 000 x 10M x 128 u8: 26.3 ms with the stores inside the column loop
#pragma unroll
      for (uint32_t c = 0; c < NC; c++) {
        unsigned m = hits[c];
        while (m) {
          const int j = __ffs(m) - 1;
          m &= m - 1;
          if (cnt < A.subcap) my_ids[cnt] = (uint32_t)(row0 + part * (GTM_N / (GTM_EPI_WARPS / 4)) + c * 32 + j);
          cnt++;
        }
      }
      as ^= 1;
      if (as == 0) aphase ^= 1;
    }
    if (qok) A.cand_cnt[(size_t)q * GTM_MAX_SUB + sub] = cnt;
  }
  gtm_tcgen05_fence_before();
  __syncthreads();
  if (widx == 1) {
    gtm_tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------------------------
// norms: |x|^2 per row (int32 exact for 8-bit rows, fp32 for floats).  store_zero: MIPS keeps 0 in the table (the test
// has no |b|^2 term) but still reports the maximum norm for the float margin.  Rows n..n_pad get the pad value that
// fails every test (+inf / 2^30).
template <int DT>
__global__ void pann_gt_norm_kernel(const uint8_t* __restrict__ rows, uint64_t n, uint64_t n_pad, uint32_t stride, uint32_t nch,
                                    void* __restrict__ out, int store_zero, float* __restrict__ maxnorm, int as_init) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_pad) return;
  // what is stored is the INITIAL VALUE of the row's accumulator column (see pann_gt_mma_kernel): -|b|^2 / 2, so that the
  // MMA leaves q.b - |b|^2 / 2 behind and the epilogue compares it with -T_q / 2 without touching |b|^2 again.  Integers:
  // -floor(|b|^2 / 2), the test being relaxed by the dropped half (a superset filter; the shrink kernel is exact).
  // as_init == 0 (query rows): the plain |q|^2.
  if (i >= n) {
    if constexpr (DT == PANN_DT_F32) reinterpret_cast<float*>(out)[i] = -INFINITY;
    else reinterpret_cast<int*>(out)[i] = -(1 << 29);
    return;
  }
  const uint4* r = reinterpret_cast<const uint4*>(rows + i * stride);
  if constexpr (DT == PANN_DT_F32) {
    float s = 0.f;
    for (uint32_t c = 0; c < nch; c++) {
      const uint4 v = __ldg(r + c);
      const float a = __uint_as_float(v.x), b = __uint_as_float(v.y), cc = __uint_as_float(v.z), d = __uint_as_float(v.w);
      s += a * a + b * b + cc * cc + d * d;
    }
    reinterpret_cast<float*>(out)[i] = store_zero ? 0.f : (as_init ? -0.5f * s : s);
    if (maxnorm) atomicMax(reinterpret_cast<int*>(maxnorm), __float_as_int(s));  // s >= 0: int order = float order
  } else {
    int s = 0;
    for (uint32_t c = 0; c < nch; c++) {
      const uint4 v = __ldg(r + c);
      if constexpr (DT == PANN_DT_U8) { s = __dp4a(v.x, v.x, (unsigned)s); s = __dp4a(v.y, v.y, (unsigned)s); s = __dp4a(v.z, v.z, (unsigned)s); s = __dp4a(v.w, v.w, (unsigned)s); }
      else { s = __dp4a((int)v.x, (int)v.x, s); s = __dp4a((int)v.y, (int)v.y, s); s = __dp4a((int)v.z, (int)v.z, s); s = __dp4a((int)v.w, (int)v.w, s); }
    }
    reinterpret_cast<int*>(out)[i] = store_zero ? 0 : (as_init ? -(s >> 1) : s);
  }
}

// T_q from the current k-th key of query q (see the header comment); conservative by construction.  `cap` (optional) is
// an upper bound of the query's final k-th distance known from elsewhere (the other shards of a sharded job): the
// threshold uses min(own k-th, cap[q]), so a shard with a loose or still empty top-k filters with the global bound.
template <int DT, int METRIC>
__device__ __forceinline__ void gt_write_thr(u64 kth, uint32_t q, const void* qnorm, const float* bnmax, void* thr, const float* cap = nullptr) {
  bool have = kth != ~0ull;
  float tau_f = have ? pann_key_dist(kth) : 0.f;
  if (cap) { const float c = cap[q]; if (c < 3.0e38f && (!have || c < tau_f)) { tau_f = c; have = true; } }
  if constexpr (DT == PANN_DT_F32) {
    float T = 3.0e38f;
    if (have) {
      const double tau = (double)tau_f;
      const double qn = (double)reinterpret_cast<const float*>(qnorm)[q], bm = (double)*bnmax;
      const double margin = 0.0078125 * sqrt(qn * bm) + 1e-4 * (qn + bm) + 1e-30;
      const double t = (METRIC == PANN_M_L2) ? (tau + margin - qn) : 2.0 * (tau + margin);
      T = __double2float_ru(t);
    }
    reinterpret_cast<float*>(thr)[q] = T;
  } else {
    int T = 1 << 29;
    if (have && fabsf(tau_f) < 5.0e8f) {
      const float tau = tau_f;
      const int ti = (int)tau + (fabsf(tau) >= 16777216.f ? 64 : 0);  // distances are (float)(int): exact below 2^24
      T = (METRIC == PANN_M_L2) ? ti - reinterpret_cast<const int*>(qnorm)[q] : 2 * ti;
    }
    reinterpret_cast<int*>(thr)[q] = T;
  }
}
template <int DT>
__global__ void pann_gt_thr_init_kernel(uint32_t nq, void* thr) {   // "everything passes" (pads still fail: +inf / 2^30)
  const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= nq) return;
  if constexpr (DT == PANN_DT_F32) reinterpret_cast<float*>(thr)[q] = 3.0e38f;
  else reinterpret_cast<int*>(thr)[q] = 1 << 29;
}
template <int DT, int METRIC>
__global__ void pann_gt_thr_kernel(const u64* __restrict__ state, uint32_t nq, uint32_t k, const void* qnorm, const float* bnmax, void* thr, const float* cap) {
  const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q < nq) gt_write_thr<DT, METRIC>(state[(uint64_t)q * k + k - 1], q, qnorm, bnmax, thr, cap);
}

// One CTA per query: exact distances of the candidates (reference operation order), merge with the running top-k
// (bitonic sort of k + candidates keys in shared memory), new threshold.  A list that overflowed sets *overflow.
__device__ __forceinline__ void gtm_bitonic(u64* a, int m) {
  for (int k = 2; k <= m; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < (m >> 1); i += blockDim.x) {   // one compare-exchange per thread and step
        const int t = 2 * i - (i & (j - 1)), p = t + j;
        const u64 x = a[t], y = a[p];
        const bool up = (t & k) == 0;
        if ((x > y) == up) { a[t] = y; a[p] = x; }
      }
      __syncthreads();
    }
}
// The same network with E = m / blockDim.x keys per thread held in registers (thread t owns keys t*E .. t*E+E-1): partners
// closer than E are in the thread, closer than 32 E in the warp (shuffles), and only the few strides beyond that go through
// shared memory — 3 barrier-separated steps instead of 45 for 512 keys on 128 threads.
template <int E>
__device__ __forceinline__ void gtm_bitonic_reg(u64* a, int m) {
  const int t = threadIdx.x;
  u64 v[E];
#pragma unroll
  for (int e = 0; e < E; e++) v[e] = a[t * E + e];
  for (int k = 2; k <= m; k <<= 1) {
    const bool up_t = ((t * E) & k) == 0;                 // direction of this thread's keys once k >= E
    for (int j = k >> 1; j >= 32 * E; j >>= 1) {          // partner in another warp
      __syncthreads();
#pragma unroll
      for (int e = 0; e < E; e++) a[t * E + e] = v[e];
      __syncthreads();
      const int pt = t ^ (j / E);
      const bool take_min = ((t & (j / E)) == 0) == up_t;
#pragma unroll
      for (int e = 0; e < E; e++) {
        const u64 p = a[pt * E + e];
        v[e] = take_min ? (p < v[e] ? p : v[e]) : (p > v[e] ? p : v[e]);
      }
    }
    for (int j = min(k >> 1, 16 * E); j >= E; j >>= 1) {  // partner in this warp
      const int d = j / E;
      const bool take_min = ((t & d) == 0) == up_t;
#pragma unroll
      for (int e = 0; e < E; e++) {
        const uint32_t lo = __shfl_xor_sync(0xffffffffu, (uint32_t)v[e], d), hi = __shfl_xor_sync(0xffffffffu, (uint32_t)(v[e] >> 32), d);
        const u64 p = ((u64)hi << 32) | lo;
        v[e] = take_min ? (p < v[e] ? p : v[e]) : (p > v[e] ? p : v[e]);
      }
    }
#pragma unroll
    for (int jj = E / 2; jj >= 1; jj >>= 1) {             // partner in this thread
      if (jj < k) {
#pragma unroll
        for (int e = 0; e < E; e++) {
          if ((e & jj) == 0) {
            const bool up = k >= E ? up_t : ((e & k) == 0);
            const u64 x = v[e], y = v[e | jj];
            if ((x > y) == up) { v[e] = y; v[e | jj] = x; }
          }
        }
      }
    }
  }
  __syncthreads();
#pragma unroll
  for (int e = 0; e < E; e++) a[t * E + e] = v[e];
  __syncthreads();
}
__device__ __forceinline__ void gtm_sort_keys(u64* a, int m) {
  const int T = (int)blockDim.x;
  if (m == T) gtm_bitonic_reg<1>(a, m);
  else if (m == 2 * T) gtm_bitonic_reg<2>(a, m);
  else if (m == 4 * T) gtm_bitonic_reg<4>(a, m);
  else if (m == 8 * T) gtm_bitonic_reg<8>(a, m);
  else gtm_bitonic(a, m);
}
template <int DT, int METRIC>
__global__ void __launch_bounds__(256, 3) pann_gt_shrink_kernel(const uint8_t* __restrict__ points, uint32_t stride, uint32_t nch,
                                                             const uint8_t* __restrict__ queries, uint32_t k, u64* __restrict__ state,
                                                             const uint32_t* __restrict__ cand_ids, uint32_t* __restrict__ cand_cnt,
                                                             uint32_t nsub, uint32_t subcap, uint32_t keys_cap,
                                                             const void* qnorm, const float* bnmax, void* thr, int* overflow, const float* cap) {
  extern __shared__ __align__(16) uint8_t gts_smem[];
  __shared__ uint32_t s_off[GTM_MAX_SUB + 1];
  __shared__ int s_over;
  u64* buf = reinterpret_cast<u64*>(gts_smem);                       // up to k + GTM_CAND_CAP keys, padded to a power of two
  const uint32_t q = blockIdx.x;
  if (threadIdx.x == 0) s_over = 0;
  __syncthreads();
  if (threadIdx.x < nsub) {
    const uint32_t c = cand_cnt[(size_t)q * GTM_MAX_SUB + threadIdx.x];
    if (c > subcap) s_over = 1;
    s_off[threadIdx.x + 1] = min(c, subcap);
  }
  __syncthreads();
  if (s_over) { if (threadIdx.x == 0) atomicOr(overflow, 1); return; }
  if (threadIdx.x < 32) {   // inclusive scan of the <= 64 sub-list lengths: two per lane (a serial loop here was 4 000 cycles per query)
    const int ln = threadIdx.x;
    uint32_t a = ln < (int)nsub ? s_off[ln + 1] : 0u, b = 32 + ln < (int)nsub ? s_off[33 + ln] : 0u;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t ua = __shfl_up_sync(0xffffffffu, a, d), ub = __shfl_up_sync(0xffffffffu, b, d);
      if (ln >= d) { a += ua; b += ub; }
    }
    b += __shfl_sync(0xffffffffu, a, 31);
    if (ln < (int)nsub) s_off[ln + 1] = a;
    if (32 + ln < (int)nsub) s_off[33 + ln] = b;
    if (ln == 0) s_off[0] = 0;
  }
  __syncthreads();
  const uint32_t cnt = s_off[nsub];
  if (cnt == 0) return;                                              // nothing can have changed (or done by the small tier)
  if (k + cnt > keys_cap) { if (threadIdx.x == 0) atomicOr(overflow, 2); return; }   // left for the launch with more key slots
  if (threadIdx.x < nsub) cand_cnt[(size_t)q * GTM_MAX_SUB + threadIdx.x] = 0;        // done: a later tier skips this query
  int m = 2; while (m < (int)(k + cnt)) m <<= 1;
  uint4* qs = reinterpret_cast<uint4*>(gts_smem + 8ull * keys_cap);
  for (uint32_t c = threadIdx.x; c < nch; c += blockDim.x) qs[c] = __ldg(reinterpret_cast<const uint4*>(queries + (uint64_t)q * stride) + c);
  for (int i = threadIdx.x; i < m; i += blockDim.x) buf[i] = i < (int)k ? state[(uint64_t)q * k + i] : ~0ull;
  __syncthreads();
  // the sub-lists are mostly empty: gather the ids densely first (staged in the key slots), then one dense pass of
  // exact distances
  for (uint32_t ci = threadIdx.x; ci < cnt; ci += blockDim.x) {   // one independent load per candidate: its sub-list by bisection
    uint32_t lo = 0, hi = nsub;                                    // s_off[lo] <= ci < s_off[hi]
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (s_off[mid] <= ci) lo = mid; else hi = mid; }
    buf[k + ci] = cand_ids[(uint64_t)q * GTM_CAND_CAP + (uint64_t)lo * subcap + (ci - s_off[lo])];
  }
  __syncthreads();
  // exact distances, 32 candidates per warp and pass.  One row per THREAD makes every 128-bit load of a warp touch 32 lines
  // (32 L1 wavefronts for 512 bytes): at 3M candidates per launch that alone was half of this kernel's 0.48 ms.  Float rows
  // are therefore loaded by the whole warp, 128 bytes of four rows per instruction, and summed per lane in element order
  // from shared memory (pann_staged_distance); 128-byte 8-bit rows are read four lanes per row (pann_quad_distance).
  {
    struct { const uint8_t* points; uint32_t stride, nch; } RG = {points, stride, nch};
    const int lane = threadIdx.x & 31, wi = threadIdx.x >> 5, nw = blockDim.x >> 5;
    uint4* stage = reinterpret_cast<uint4*>(gts_smem + 8ull * keys_cap + stride) + (size_t)wi * (PANN_STAGE_BYTES / 16);
    for (uint32_t base = wi * 32; base < cnt; base += nw * 32) {
      const uint32_t ci = base + lane;
      const bool ok = ci < cnt;
      const uint32_t id = ok ? (uint32_t)buf[k + ci] : 0u;
      const unsigned mask = __ballot_sync(0xffffffffu, ok);
      float d;
      if constexpr (DT == PANN_DT_F32) {
        d = pann_staged_distance<DT, METRIC>(RG, id, mask, qs, stage, lane);
      } else {
        if (nch == 8) {
          d = pann_quad_distance<DT, METRIC>(RG, id, mask, qs, lane);
        } else {
          DistAcc<DT, METRIC> acc;
          if (ok) {
            const uint4* row = reinterpret_cast<const uint4*>(points + (uint64_t)id * stride);
            for (uint32_t c = 0; c < nch; c++) acc.add(__ldg(row + c), qs[c]);
          }
          d = acc.result();
        }
      }
      if (ok) buf[k + ci] = pann_key(d, id);
    }
  }
  __syncthreads();
  gtm_sort_keys(buf, m);
  for (int i = threadIdx.x; i < (int)k; i += blockDim.x) state[(uint64_t)q * k + i] = buf[i];
  if (threadIdx.x == 0) gt_write_thr<DT, METRIC>(buf[k - 1], q, qnorm, bnmax, thr, cap);
}
