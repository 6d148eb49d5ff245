// pann_common.cuh — shared device helpers: (dist,id) keys, the reference's hash, fused distance accumulators.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>
#include <string>

#define PANN_DT_F32 0
#define PANN_DT_U8 1
#define PANN_DT_I8 2
#define PANN_M_L2 0
#define PANN_M_MIPS 1

typedef unsigned long long u64;

// ---------------------------------------------------------------------------------------------
// error plumbing (host)
void pann_set_error(const std::string& msg);
#define PANN_CUDA(call)                                                                              \
  do {                                                                                               \
    cudaError_t e__ = (call);                                                                        \
    if (e__ != cudaSuccess) {                                                                        \
      pann_set_error(std::string(#call) + " failed: " + cudaGetErrorString(e__) + " (" + __FILE__ + ":" + \
                     std::to_string(__LINE__) + ")");                                                \
      return 1;                                                                                      \
    }                                                                                                \
  } while (0)
#define PANN_FAIL(msg)        \
  do {                        \
    pann_set_error(msg);      \
    return 1;                 \
  } while (0)

// ---------------------------------------------------------------------------------------------
// (distance, id) pairs as one 64-bit key that sorts like the reference's `less`
// (beamSearch.h:46-48: by distance, then id):  [63:32] order-preserving image of the float,
// [31:1] id (ids < 2^31, as the reference's `int p_star` already assumes), [0] a "visited" flag
// that never takes part in comparisons (all comparisons are on key >> 1).
__host__ __device__ __forceinline__ uint32_t pann_fsort(float f) {
#ifdef __CUDA_ARCH__
  uint32_t b = __float_as_uint(f);
#else
  uint32_t b; memcpy(&b, &f, 4);
#endif
  if (b == 0x80000000u) b = 0;  // -0.0 == +0.0 under operator<
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ __forceinline__ float pann_funsort(uint32_t s) {
  uint32_t b = (s & 0x80000000u) ? (s & 0x7fffffffu) : ~s;
#ifdef __CUDA_ARCH__
  return __uint_as_float(b);
#else
  float f; memcpy(&f, &b, 4); return f;
#endif
}
__host__ __device__ __forceinline__ u64 pann_key(float d, uint32_t id) { return ((u64)pann_fsort(d) << 32) | ((u64)id << 1); }
__host__ __device__ __forceinline__ uint32_t pann_key_id(u64 k) { return (uint32_t)(k >> 1) & 0x7fffffffu; }
__host__ __device__ __forceinline__ float pann_key_dist(u64 k) { return pann_funsort((uint32_t)(k >> 32)); }

// parlay::hash64_2 (parlay/utilities.h:145-150) — the visited hash filter must hash exactly like the reference.
__host__ __device__ __forceinline__ u64 pann_hash64_2(u64 x) {
  x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
  x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
  return x ^ (x >> 31);
}

#ifdef __CUDACC__
// ---------------------------------------------------------------------------------------------
// Fused distance accumulators over 16-byte chunks of ONE row, fed in ascending chunk order by one thread
// (euclidian_point.h:45-90, mips_point.h:43-65).  8-bit types accumulate exactly in int32 (dp4a).  Floats do
// exactly what the reference's scalar loop does — per dimension one rounded subtract, one rounded multiply and
// one rounded add, in ascending order, never contracted to an FMA — so float distances are bit-identical to the
// CPU's as the reference is built here (oracle/_ref, -O3 -march=native: the loop stays sequential and uncontracted; pinned
// by the float fixtures of tests/golden, d = 24, 32 and 100).  Against another compiler's build the contract is 1e-5 relative.
template <int DT, int METRIC> struct DistAcc;

template <> struct DistAcc<PANN_DT_U8, PANN_M_L2> {
  unsigned acc = 0;
  __device__ __forceinline__ void add(const uint4& p, const uint4& q) {
    unsigned x;
    x = __vabsdiffu4(p.x, q.x); acc = __dp4a(x, x, acc);
    x = __vabsdiffu4(p.y, q.y); acc = __dp4a(x, x, acc);
    x = __vabsdiffu4(p.z, q.z); acc = __dp4a(x, x, acc);
    x = __vabsdiffu4(p.w, q.w); acc = __dp4a(x, x, acc);
  }
  __device__ __forceinline__ float result() const { return (float)(int)acc; }
};
template <> struct DistAcc<PANN_DT_I8, PANN_M_L2> {
  unsigned acc = 0;
  __device__ __forceinline__ void add(const uint4& p, const uint4& q) {
    unsigned x;
    x = __vabsdiffs4(p.x, q.x); acc = __dp4a(x, x, acc);
    x = __vabsdiffs4(p.y, q.y); acc = __dp4a(x, x, acc);
    x = __vabsdiffs4(p.z, q.z); acc = __dp4a(x, x, acc);
    x = __vabsdiffs4(p.w, q.w); acc = __dp4a(x, x, acc);
  }
  __device__ __forceinline__ float result() const { return (float)(int)acc; }
};
template <> struct DistAcc<PANN_DT_U8, PANN_M_MIPS> {
  unsigned acc = 0;
  __device__ __forceinline__ void add(const uint4& p, const uint4& q) {
    acc = __dp4a(p.x, q.x, acc); acc = __dp4a(p.y, q.y, acc); acc = __dp4a(p.z, q.z, acc); acc = __dp4a(p.w, q.w, acc);
  }
  __device__ __forceinline__ float result() const { return -((float)(int)acc); }
};
template <> struct DistAcc<PANN_DT_I8, PANN_M_MIPS> {
  int acc = 0;
  __device__ __forceinline__ void add(const uint4& p, const uint4& q) {
    acc = __dp4a((int)p.x, (int)q.x, acc); acc = __dp4a((int)p.y, (int)q.y, acc);
    acc = __dp4a((int)p.z, (int)q.z, acc); acc = __dp4a((int)p.w, (int)q.w, acc);
  }
  __device__ __forceinline__ float result() const { return -((float)acc); }
};
template <> struct DistAcc<PANN_DT_F32, PANN_M_L2> {
  float acc = 0.f;
  __device__ __forceinline__ void add1(unsigned p, unsigned q) {
    float t = __fsub_rn(__uint_as_float(q), __uint_as_float(p));
    acc = __fadd_rn(acc, __fmul_rn(t, t));
  }
  __device__ __forceinline__ void add(const uint4& p, const uint4& q) { add1(p.x, q.x); add1(p.y, q.y); add1(p.z, q.z); add1(p.w, q.w); }
  __device__ __forceinline__ float result() const { return acc; }
};
template <> struct DistAcc<PANN_DT_F32, PANN_M_MIPS> {
  float acc = 0.f;
  __device__ __forceinline__ void add1(unsigned p, unsigned q) { acc = __fadd_rn(acc, __fmul_rn(__uint_as_float(q), __uint_as_float(p))); }
  __device__ __forceinline__ void add(const uint4& p, const uint4& q) { add1(p.x, q.x); add1(p.y, q.y); add1(p.z, q.z); add1(p.w, q.w); }
  __device__ __forceinline__ float result() const { return -acc; }
};

__device__ __forceinline__ uint4 pann_ldg16(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
#endif  // __CUDACC__

// row geometry shared by all kernels: rows are re-padded on the device to a multiple of 16 bytes
// (zero padding adds 0 to both L2 and inner-product sums), so every row read is a 128-bit load.
struct PannRowGeom {
  uint32_t dims;
  uint32_t elt;     // bytes per element
  uint32_t stride;  // bytes per device row, multiple of 16
  uint32_t nch;     // 16-byte chunks per row
};
inline PannRowGeom pann_row_geom(uint32_t dims, int dtype) {
  PannRowGeom g;
  g.dims = dims;
  g.elt = dtype == PANN_DT_F32 ? 4 : 1;
  g.stride = (dims * g.elt + 15) / 16 * 16;
  g.nch = g.stride / 16;
  return g;
}
// kernels are specialised for rows of exactly PANN_QR chunks (128 bytes: 128-d uint8/int8), whose query lives in
// registers; every other shape takes the generic variant (QR = 0) with the query in shared memory
#define PANN_QR 8
inline int pann_qr_of(const PannRowGeom& g) { return g.nch == PANN_QR ? PANN_QR : 0; }
