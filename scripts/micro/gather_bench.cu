// gather_bench.cu — measurement tool (not part of the library): how fast can one B200 gather random 128-byte rows
// from a table of a given size when nothing else limits it?  Same access pattern as K1s (four lanes per row, two
// 128-bit loads each) and the same residency (128-thread CTAs, 8 per SM), with 32 or 64 rows in flight per warp.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/gather_bench scripts/micro/gather_bench.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 mix(u64 x) { x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull; x = (x ^ (x >> 27)) * 0x94d049bb133111ebull; return x ^ (x >> 31); }

template <int DEPTH>  // DEPTH chunks of 32 rows in flight per warp
__global__ void __launch_bounds__(128, 8) gather(const uint8_t* tab, u64 nrows, u64 nchunks, unsigned* sink, int dependent) {
  const int lane = threadIdx.x & 31, sub = lane & 3, rowi = lane >> 2;
  const u64 warp = (u64)blockIdx.x * 4 + (threadIdx.x >> 5), nwarps = (u64)gridDim.x * 4;
  unsigned acc = 0;
  u64 salt = 0;
  for (u64 c = warp; c < nchunks; c += nwarps * DEPTH) {
    uint4 v[DEPTH][4][2];
#pragma unroll
    for (int d = 0; d < DEPTH; d++)
#pragma unroll
      for (int g = 0; g < 4; g++) {
        const u64 id = mix(((c + d * nwarps) * 32 + g * 8 + rowi) ^ salt) % nrows;
        const uint8_t* row = tab + id * 128 + sub * 16;
        v[d][g][0] = __ldg(reinterpret_cast<const uint4*>(row));
        v[d][g][1] = __ldg(reinterpret_cast<const uint4*>(row + 64));
      }
#pragma unroll
    for (int d = 0; d < DEPTH; d++)
#pragma unroll
      for (int g = 0; g < 4; g++) acc += v[d][g][0].x ^ v[d][g][0].w ^ v[d][g][1].y ^ v[d][g][1].z;
    if (dependent) salt = __shfl_sync(0xffffffffu, acc, 0) & 1;  // next addresses depend on this data: a dependent chain per warp
  }
  if (acc == 0x12345678u) *sink = acc;
}

int main(int argc, char** argv) {
  unsigned* sink; cudaMalloc(&sink, 4);
  const double sizes_gb[] = {0.128, 0.388, 1.28, 3.9, 16.0};
  for (double gb : sizes_gb) {
    const u64 nrows = (u64)(gb * 1e9 / 128);
    uint8_t* tab; if (cudaMalloc(&tab, nrows * 128) != cudaSuccess) { printf("alloc failed\n"); return 1; }
    cudaMemset(tab, 1, nrows * 128);
    const u64 nchunks = 4ull << 20;  // 128M rows = 17 GB of gathers
    for (int dep = 0; dep < 2; dep++)
      for (int depth = 1; depth <= 2; depth++) {
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        float best = 1e30f;
        for (int rep = 0; rep < 3; rep++) {
          cudaEventRecord(e0);
          if (depth == 1) gather<1><<<148 * 8, 128>>>(tab, nrows, nchunks, sink, dep); else gather<2><<<148 * 8, 128>>>(tab, nrows, nchunks, sink, dep);
          cudaEventRecord(e1); cudaEventSynchronize(e1);
          float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
        }
        printf("table %6.3f GB  rows in flight per warp %2d  %s: %8.1f GB/s  (%.2f ms)\n", gb, 32 * depth, dep ? "dependent chunks  " : "independent chunks", nchunks * 32 * 128 / (best * 1e-3) / 1e9, best);
      }
    cudaFree(tab);
  }
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
  return 0;
}
