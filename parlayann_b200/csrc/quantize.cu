// quantize.cu — the reference's point conversions as kernels: what its Python builder / index loader do to 4-byte data
// before the hot path runs (python/builder.cpp:45-87, python/graph_index.cpp:82-118).
//   Mips_Point<float>::normalize                          (utils/mips_point.h:115-124)
//   Euclidian_Point<uint8_t>::generate_parameters / translate_point   (utils/euclidian_point.h:104-108, :183-237)
//   Quantized_Mips_Point<8,trim>::generate_parameters / translate_point (utils/mips_point.h:421-462)
// Host arrays in, host arrays out; rows go through the device in chunks.  Every rounding step of the reference is kept
// (float products, the double norm sum in element order, std::round = halves away from zero), so the bytes are the
// reference's (tests/test_quantize_gpu.py against the oracle and the committed goldens).
#include <math.h>
#include <string.h>

#include <algorithm>
#include <cub/cub.cuh>

#include "pann_internal.h"

namespace {

constexpr size_t kChunkBytes = 256ull << 20;   // device staging per call

struct Stage {
  void* in = nullptr; void* out = nullptr; void* aux = nullptr;
  ~Stage() { if (in) cudaFree(in); if (out) cudaFree(out); if (aux) cudaFree(aux); }
};

__global__ void pann_normalize_rows_kernel(float* rows, uint64_t n, uint32_t dims) {
  const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  float* v = rows + r * dims;
  double norm = 0.0;
  for (uint32_t j = 0; j < dims; j++) { const float x = v[j]; norm = __dadd_rn(norm, (double)__fmul_rn(x, x)); }
  norm = sqrt(norm);
  if (norm == 0) norm = 1.0;
  const float inv = __double2float_rn(__ddiv_rn(1.0, norm));
  for (uint32_t j = 0; j < dims; j++) v[j] = __fmul_rn(v[j], inv);
}

// min(0, min x), max(0, max x) and "every value is a non-negative integer" (euclidian_point.h:217-226)
__global__ void pann_minmax_kernel(const float* v, uint64_t len, float* out_min, float* out_max, int* not_int) {
  float mn = 0.f, mx = 0.f;
  int bad = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < len; i += (uint64_t)gridDim.x * blockDim.x) {
    const float x = v[i];
    mn = fminf(mn, x); mx = fmaxf(mx, x);
    bad |= !(x >= 0.f && (x - (float)(long long)x) == 0.f);
  }
  for (int d = 16; d > 0; d >>= 1) {
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, d));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, d));
    bad |= __shfl_xor_sync(0xffffffffu, bad, d);
  }
  if ((threadIdx.x & 31) == 0) {
    // mn <= 0 <= mx: the int order of the bit patterns is the float order for mx, the reverse for mn
    atomicMax(reinterpret_cast<int*>(out_max), __float_as_int(mx));
    atomicMax(reinterpret_cast<unsigned*>(out_min), __float_as_uint(mn));   // negative floats: larger pattern = smaller value
    if (bad) atomicOr(not_int, 1);
  }
}

__global__ void pann_quantize_u8_kernel(const float* v, uint64_t len, float slope, long long offset, int plain, uint8_t* out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= len) return;
  const float x = v[i];
  if (plain) { out[i] = (uint8_t)x; return; }                       // slope == 1 && offset == 0: values[j] = p[j]
  long long r = (long long)roundf(__fmul_rn(x, slope)) - offset;   // std::round: halves away from zero
  r = r < 0 ? 0 : (r > 255 ? 255 : r);
  out[i] = (uint8_t)r;
}

__global__ void pann_quantize_mips8_kernel(const float* v, uint64_t len, float mv, float scale, int8_t* out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= len) return;
  const float x = v[i];
  int q;
  if (x < -mv) q = -127;
  else if (x > mv) q = 127;
  else q = (int)roundf(__fmul_rn(x, scale));
  out[i] = (int8_t)q;
}

// rows [r0, r0 + m) of a strided host array -> dense device rows
int upload_rows(const float* rows, uint64_t r0, uint64_t m, uint32_t dims, uint64_t stride_bytes, float* d) {
  PANN_CUDA(cudaMemcpy2D(d, (size_t)dims * 4, reinterpret_cast<const uint8_t*>(rows) + r0 * stride_bytes, stride_bytes, (size_t)dims * 4, m,
                         cudaMemcpyHostToDevice));
  return 0;
}

int check_args(const char* who, const void* rows, uint64_t n, uint32_t dims, uint64_t stride_bytes) {
  if (!rows && n) PANN_FAIL(std::string(who) + ": null argument");
  if (dims == 0) PANN_FAIL(std::string(who) + ": zero dimensions");
  if (stride_bytes < (uint64_t)dims * 4) PANN_FAIL(std::string(who) + ": row stride smaller than a row");
  return 0;
}

}  // namespace

extern "C" int pann_rows_normalize(int device, float* rows, uint64_t n, uint32_t dims, uint64_t stride_bytes) {
  if (check_args("pann_rows_normalize", rows, n, dims, stride_bytes)) return 1;
  if (n == 0) return 0;
  PANN_CUDA(cudaSetDevice(device));
  const uint64_t per = std::max<uint64_t>(1, kChunkBytes / ((uint64_t)dims * 4));
  Stage st;
  PANN_CUDA(cudaMalloc(&st.in, std::min(per, n) * dims * 4));
  for (uint64_t r0 = 0; r0 < n; r0 += per) {
    const uint64_t m = std::min(per, n - r0);
    if (upload_rows(rows, r0, m, dims, stride_bytes, (float*)st.in)) return 1;
    pann_normalize_rows_kernel<<<(unsigned)((m + 127) / 128), 128>>>((float*)st.in, m, dims);
    PANN_CUDA(cudaGetLastError());
    PANN_CUDA(cudaMemcpy2D(reinterpret_cast<uint8_t*>(rows) + r0 * stride_bytes, stride_bytes, st.in, (size_t)dims * 4, (size_t)dims * 4, m,
                           cudaMemcpyDeviceToHost));
  }
  return 0;
}

extern "C" int pann_rows_quant_params_u8(int device, const float* rows, uint64_t n, uint32_t dims, uint64_t stride_bytes, float* slope,
                                         int64_t* offset) {
  if (check_args("pann_rows_quant_params_u8", rows, n, dims, stride_bytes)) return 1;
  if (!slope || !offset) PANN_FAIL("pann_rows_quant_params_u8: null argument");
  float mn = 0.f, mx = 0.f;
  int not_int = 0;
  if (n) {
    PANN_CUDA(cudaSetDevice(device));
    const uint64_t per = std::max<uint64_t>(1, kChunkBytes / ((uint64_t)dims * 4));
    Stage st;
    PANN_CUDA(cudaMalloc(&st.in, std::min(per, n) * dims * 4));
    PANN_CUDA(cudaMalloc(&st.aux, 16));
    PANN_CUDA(cudaMemset(st.aux, 0, 16));      // min = +0, max = +0 (the reference starts both at 0), flag clear
    float* d_acc = (float*)st.aux;
    for (uint64_t r0 = 0; r0 < n; r0 += per) {
      const uint64_t m = std::min(per, n - r0);
      if (upload_rows(rows, r0, m, dims, stride_bytes, (float*)st.in)) return 1;
      pann_minmax_kernel<<<1184, 256>>>((const float*)st.in, m * dims, d_acc, d_acc + 1, reinterpret_cast<int*>(d_acc + 2));
      PANN_CUDA(cudaGetLastError());
    }
    float h[3];
    PANN_CUDA(cudaMemcpy(h, st.aux, 12, cudaMemcpyDeviceToHost));
    mn = h[0]; mx = h[1]; memcpy(&not_int, &h[2], 4);
  }
  if (!not_int) {                 // euclidian_point.h:227-231
    if (mx < 256) mx = 255;
    mn = 0;
  }
  const float s = 255.0f / (mx - mn);                       // parameters(min, max, dims) (:105-108)
  *slope = s;
  *offset = (int64_t)(int32_t)round((double)(mn * s));      // round(float product), halves away from zero
  return 0;
}

extern "C" int pann_rows_quantize_u8(int device, const float* rows, uint64_t n, uint32_t dims, uint64_t stride_bytes, float slope,
                                     int64_t offset, uint8_t* out, uint64_t out_stride_bytes) {
  if (check_args("pann_rows_quantize_u8", rows, n, dims, stride_bytes)) return 1;
  if ((!out && n) || out_stride_bytes < dims) PANN_FAIL("pann_rows_quantize_u8: bad output buffer");
  if (n == 0) return 0;
  PANN_CUDA(cudaSetDevice(device));
  const uint64_t per = std::max<uint64_t>(1, kChunkBytes / ((uint64_t)dims * 4));
  Stage st;
  PANN_CUDA(cudaMalloc(&st.in, std::min(per, n) * dims * 4));
  PANN_CUDA(cudaMalloc(&st.out, std::min(per, n) * dims));
  const int plain = (slope == 1.0f && offset == 0) ? 1 : 0;
  for (uint64_t r0 = 0; r0 < n; r0 += per) {
    const uint64_t m = std::min(per, n - r0), len = m * dims;
    if (upload_rows(rows, r0, m, dims, stride_bytes, (float*)st.in)) return 1;
    pann_quantize_u8_kernel<<<(unsigned)((len + 255) / 256), 256>>>((const float*)st.in, len, slope, (long long)offset, plain, (uint8_t*)st.out);
    PANN_CUDA(cudaGetLastError());
    PANN_CUDA(cudaMemcpy2D(out + r0 * out_stride_bytes, out_stride_bytes, st.out, dims, dims, m, cudaMemcpyDeviceToHost));
  }
  return 0;
}

extern "C" int pann_rows_quant_params_mips8(int device, const float* rows, uint64_t n, uint32_t dims, uint64_t stride_bytes, int trim,
                                            float* max_val) {
  if (check_args("pann_rows_quant_params_mips8", rows, n, dims, stride_bytes)) return 1;
  if (!max_val) PANN_FAIL("pann_rows_quant_params_mips8: null argument");
  const uint64_t len = n * dims;
  if (len == 0) { *max_val = 1.0f; return 0; }
  if (len >= (1ull << 31)) PANN_FAIL("pann_rows_quant_params_mips8: more than 2^31 values (the device sort takes an int count)");
  PANN_CUDA(cudaSetDevice(device));
  // parlay::sort_inplace of all n * dims values, then two order statistics (mips_point.h:441-458)
  Stage st;
  PANN_CUDA(cudaMalloc(&st.in, len * 4));
  PANN_CUDA(cudaMalloc(&st.out, len * 4));
  if (upload_rows(rows, 0, n, dims, stride_bytes, (float*)st.in)) return 1;
  size_t tb = 0;
  cub::DeviceRadixSort::SortKeys(nullptr, tb, (const float*)st.in, (float*)st.out, (int)len);
  PANN_CUDA(cudaMalloc(&st.aux, tb + 16));
  cub::DeviceRadixSort::SortKeys(st.aux, tb, (const float*)st.in, (float*)st.out, (int)len);
  PANN_CUDA(cudaGetLastError());
  uint64_t lo = 0, hi = len - 1;
  if (trim) {
    const float cutoff = .0001f;
    lo = (uint64_t)(long)(cutoff * (float)len);                      // (long)(cutoff * len): float product
    hi = (uint64_t)(long)((1.0 - (double)cutoff) * (double)(len - 1));  // (long)((1.0 - cutoff) * (len - 1)): double product
  }
  float mn = 0.f, mx = 0.f;
  PANN_CUDA(cudaMemcpy(&mn, (const float*)st.out + lo, 4, cudaMemcpyDeviceToHost));
  PANN_CUDA(cudaMemcpy(&mx, (const float*)st.out + hi, 4, cudaMemcpyDeviceToHost));
  *max_val = std::max(mx, -mn);
  return 0;
}

extern "C" int pann_rows_quantize_mips8(int device, const float* rows, uint64_t n, uint32_t dims, uint64_t stride_bytes, float max_val,
                                        int8_t* out, uint64_t out_stride_bytes) {
  if (check_args("pann_rows_quantize_mips8", rows, n, dims, stride_bytes)) return 1;
  if ((!out && n) || out_stride_bytes < dims) PANN_FAIL("pann_rows_quantize_mips8: bad output buffer");
  if (n == 0) return 0;
  PANN_CUDA(cudaSetDevice(device));
  const uint64_t per = std::max<uint64_t>(1, kChunkBytes / ((uint64_t)dims * 4));
  Stage st;
  PANN_CUDA(cudaMalloc(&st.in, std::min(per, n) * dims * 4));
  PANN_CUDA(cudaMalloc(&st.out, std::min(per, n) * dims));
  const float scale = 127.0f / max_val;                              // (range / 2) / mv, float division
  for (uint64_t r0 = 0; r0 < n; r0 += per) {
    const uint64_t m = std::min(per, n - r0), len = m * dims;
    if (upload_rows(rows, r0, m, dims, stride_bytes, (float*)st.in)) return 1;
    pann_quantize_mips8_kernel<<<(unsigned)((len + 255) / 256), 256>>>((const float*)st.in, len, max_val, scale, (int8_t*)st.out);
    PANN_CUDA(cudaGetLastError());
    PANN_CUDA(cudaMemcpy2D(out + r0 * out_stride_bytes, out_stride_bytes, st.out, dims, dims, m, cudaMemcpyDeviceToHost));
  }
  return 0;
}
