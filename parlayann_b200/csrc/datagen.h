// datagen.h — bit-reproducible synthetic datasets, shared by host and device code.
//
// Clustered low-intrinsic-dimension model (latent r, nc cluster centres):
//     z = C[c] + 0.6 * N(0, I_r);   x = z . A / sqrt(r) + 0.15 * N(0, I_d)
// evaluated entirely in integer arithmetic from a counter-based hash (splitmix64 finaliser)
// and an Irwin-Hall(4) Gaussian approximation, so a row is the same on any CPU or GPU.
// Kinds:  u8 "BIGANN-shaped"  = clip(rint(|x| * 40), 0, 255)
//         i8                  = clip(rint(x * 40), -128, 127)
//         f32 "SIFT-shaped"   = x (exact dyadic rationals)
//         f32 "Deep-shaped"   = x / ||x||  (IEEE double sqrt/div, then rounded to float)
//         f32 "T2I-shaped"    = nc = 0 (no clusters), any d
// This is input synthesis, not part of the reference's hot path.
#pragma once
#include <stdint.h>
#include <math.h>

#if defined(__CUDACC__)
#define PANN_HD __host__ __device__ __forceinline__
#else
#define PANN_HD inline
#endif

enum pann_gen_kind { PANN_GEN_U8 = 0, PANN_GEN_I8 = 1, PANN_GEN_F32 = 2, PANN_GEN_F32_NORM = 3 };

#define PANN_GEN_MAX_R 32
#define PANN_GEN_SIGMA 37837  /* std-dev of gen_gauss(): 65536 * sqrt(4/12) */

PANN_HD uint64_t gen_mix(uint64_t x) {
  x += 0x9e3779b97f4a7c15ull;
  x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
  x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
  return x ^ (x >> 31);
}
PANN_HD uint64_t gen_hash(uint64_t seed, uint64_t stream, uint64_t a, uint64_t b) {
  return gen_mix(gen_mix(gen_mix(seed ^ (stream * 0xd1342543de82ef95ull)) + a) + b * 0x2545f4914f6cdd1dull);
}
// Irwin-Hall(4) on 16-bit uniforms, centred: range [-131070, 131070], sigma = PANN_GEN_SIGMA
PANN_HD int32_t gen_gauss(uint64_t h) {
  int32_t s = (int32_t)(h & 0xffff) + (int32_t)((h >> 16) & 0xffff) + (int32_t)((h >> 32) & 0xffff) + (int32_t)(h >> 48);
  return s - 131070;
}
// model tables (depend on model_seed only; base and query sets share them)
PANN_HD int32_t gen_A(uint64_t model_seed, uint32_t k, uint32_t j) { return gen_gauss(gen_hash(model_seed, 1, k, j)); }
PANN_HD int32_t gen_C(uint64_t model_seed, uint32_t c, uint32_t k) { return gen_gauss(gen_hash(model_seed, 2, c, k)); }

// integer sqrt(r) * SIGMA denominators, r = 1..32 (rounded)
PANN_HD int64_t gen_denom(uint32_t r) {
  // round(37837 * sqrt(r))
  const int32_t t[33] = {0, 37837, 53510, 65536, 75674, 84606, 92681, 100107, 107019, 113511, 119651, 125491,
                         131071, 136423, 141573, 146542, 151348, 156006, 160529, 164928, 169212, 173391, 177471,
                         181460, 185363, 189185, 192932, 196607, 200215, 203758, 207242, 210667, 214038};
  return t[r];
}

// latent vector of one row: z[k], k < r.  `stream` separates base (0) from query (1) rows.
PANN_HD void gen_latent(uint64_t model_seed, uint64_t stream, uint64_t row, uint32_t r, uint32_t nc,
                        const int32_t* Ctab /* nc*r or null */, int32_t* z) {
  uint32_t c = nc ? (uint32_t)(gen_hash(model_seed, 16 + stream, row, 0xc1u) % nc) : 0;
  for (uint32_t k = 0; k < r; k++) {
    int32_t nz = gen_gauss(gen_hash(model_seed, 32 + stream, row, k));
    if (nc) {
      int32_t ck = Ctab ? Ctab[c * r + k] : gen_C(model_seed, c, k);
      z[k] = ck + (int32_t)(((int64_t)nz * 39322) / 65536);  // 0.6 * noise
    } else {
      z[k] = nz;
    }
  }
}
// one coordinate in "sigma units" (value = result / PANN_GEN_SIGMA)
PANN_HD int32_t gen_coord(uint64_t model_seed, uint64_t stream, uint64_t row, uint32_t j, uint32_t r,
                          const int32_t* z, const int32_t* Acol /* A[k*d + j] stride d */, uint32_t d) {
  int64_t acc = 0;
  for (uint32_t k = 0; k < r; k++) acc += (int64_t)z[k] * (int64_t)Acol[(uint64_t)k * d];
  int64_t x = acc / gen_denom(r);
  int32_t e = gen_gauss(gen_hash(model_seed, 64 + stream, row, j));
  x += ((int64_t)e * 9830) / 65536;  // 0.15 * noise
  return (int32_t)x;
}
PANN_HD uint8_t gen_to_u8(int32_t x) {
  int64_t a = x < 0 ? -(int64_t)x : (int64_t)x;
  int64_t v = (a * 40 + PANN_GEN_SIGMA / 2) / PANN_GEN_SIGMA;
  return (uint8_t)(v > 255 ? 255 : v);
}
PANN_HD int8_t gen_to_i8(int32_t x) {
  int64_t a = x < 0 ? -(int64_t)x : (int64_t)x;
  int64_t v = (a * 40 + PANN_GEN_SIGMA / 2) / PANN_GEN_SIGMA;
  if (x < 0) return (int8_t)(v > 128 ? -128 : -v);
  return (int8_t)(v > 127 ? 127 : v);
}
PANN_HD float gen_to_f32(int32_t x) { return (float)x * (1.0f / 32768.0f); }  // exact: |x| < 2^24

// Deep-shaped rows: x / ||x||.  The norm is an exact integer sum; sqrt and the divide are IEEE
// double operations (correctly rounded on CPU and GPU alike), so the float result is reproducible.
PANN_HD void gen_normalize_row(const int32_t* x, uint32_t d, float* out) {
  int64_t ss = 0;
  for (uint32_t j = 0; j < d; j++) ss += (int64_t)x[j] * (int64_t)x[j];
  double nrm = ss > 0 ? sqrt((double)ss) : 1.0;
  for (uint32_t j = 0; j < d; j++) out[j] = (float)((double)x[j] / nrm);
}
