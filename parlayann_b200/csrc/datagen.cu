// datagen.cu — device evaluation of the synthetic-data generator (datagen.h).
// Input synthesis for benchmarks and tests, bit-identical to the CPU evaluation of the same header.
#include "datagen.h"
#include "pann_common.cuh"

__global__ void pann_gen_tables_kernel(uint64_t seed, uint32_t d, uint32_t r, uint32_t nc, int32_t* A, int32_t* Ctab) {
  uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < (uint64_t)r * d) A[t] = gen_A(seed, (uint32_t)(t / d), (uint32_t)(t % d));
  if (t < (uint64_t)nc * r) Ctab[t] = gen_C(seed, (uint32_t)(t / r), (uint32_t)(t % r));
}

// one CTA per row; thread j computes coordinate j (strided); out rows are `stride` bytes apart and
// pre-zeroed by the caller (padding stays 0).
__global__ void pann_gen_rows_kernel(int kind, uint64_t seed, uint64_t stream, uint64_t row0, uint64_t nrows, uint32_t d,
                                     uint32_t r, uint32_t nc, const int32_t* __restrict__ A, const int32_t* __restrict__ Ctab,
                                     uint8_t* out, uint64_t stride, int32_t* xbuf) {
  __shared__ int32_t z[PANN_GEN_MAX_R];
  __shared__ double s_nrm;
  for (uint64_t i = blockIdx.x; i < nrows; i += gridDim.x) {
    uint64_t row = row0 + i;
    if (threadIdx.x == 0) gen_latent(seed, stream, row, r, nc, nc ? Ctab : nullptr, z);
    __syncthreads();
    uint8_t* o = out + i * stride;
    if (kind != PANN_GEN_F32_NORM) {
      for (uint32_t j = threadIdx.x; j < d; j += blockDim.x) {
        int32_t x = gen_coord(seed, stream, row, j, r, z, A + j, d);
        if (kind == PANN_GEN_U8) o[j] = gen_to_u8(x);
        else if (kind == PANN_GEN_I8) ((int8_t*)o)[j] = gen_to_i8(x);
        else ((float*)o)[j] = gen_to_f32(x);
      }
    } else {
      int32_t* xb = xbuf + (uint64_t)blockIdx.x * d;
      for (uint32_t j = threadIdx.x; j < d; j += blockDim.x) xb[j] = gen_coord(seed, stream, row, j, r, z, A + j, d);
      __syncthreads();
      if (threadIdx.x == 0) {  // exact integer norm, then IEEE double sqrt — same as gen_normalize_row
        int64_t ss = 0;
        for (uint32_t j = 0; j < d; j++) ss += (int64_t)xb[j] * (int64_t)xb[j];
        s_nrm = ss > 0 ? sqrt((double)ss) : 1.0;
      }
      __syncthreads();
      for (uint32_t j = threadIdx.x; j < d; j += blockDim.x) ((float*)o)[j] = (float)((double)xb[j] / s_nrm);
    }
    __syncthreads();
  }
}

int pann_gen_rows_device(int kind, uint64_t seed, uint64_t stream, uint64_t row0, uint64_t nrows, uint32_t d, uint32_t r,
                         uint32_t nc, void* dev_out, uint64_t stride, cudaStream_t s) {
  if (r == 0 || r > PANN_GEN_MAX_R) PANN_FAIL("pann_gen: latent_r must be in 1..32");
  int32_t *A = nullptr, *Ct = nullptr, *xb = nullptr;
  PANN_CUDA(cudaMalloc(&A, sizeof(int32_t) * (size_t)r * d));
  PANN_CUDA(cudaMalloc(&Ct, sizeof(int32_t) * ((size_t)nc * r + 1)));
  size_t tt = (size_t)r * (d > nc ? d : nc);
  pann_gen_tables_kernel<<<(unsigned)((tt + 255) / 256), 256, 0, s>>>(seed, d, r, nc, A, Ct);
  int grid = (int)(nrows < 148 * 16 ? (nrows ? nrows : 1) : 148 * 16);
  if (kind == PANN_GEN_F32_NORM) PANN_CUDA(cudaMalloc(&xb, sizeof(int32_t) * (size_t)grid * d));
  PANN_CUDA(cudaMemsetAsync(dev_out, 0, (size_t)nrows * stride, s));
  pann_gen_rows_kernel<<<grid, 128, 0, s>>>(kind, seed, stream, row0, nrows, d, r, nc, A, Ct, (uint8_t*)dev_out, stride, xb);
  PANN_CUDA(cudaGetLastError());
  PANN_CUDA(cudaStreamSynchronize(s));
  cudaFree(A); cudaFree(Ct); if (xb) cudaFree(xb);
  return 0;
}
