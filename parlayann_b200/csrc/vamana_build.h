// vamana_build.h — pieces of the batch-insert machinery shared by the Vamana build (vamana_build.cu) and the HNSW build (hnsw.cu)
#pragma once
#include "pann_internal.h"
#include "vamana_prune.cuh"

// rows of the inserted points + one (target << 32 | task * R + position) key per new edge (padding keys sort last)
__global__ void pann_write_rows_kernel(uint32_t* graph, uint32_t maxdeg, uint32_t R, uint64_t n, const uint32_t* perm_lo,
                                       const uint32_t* new_out, const uint32_t* new_deg, uint32_t bs, u64* keys, const uint32_t* rowmap);
// first key of every target in the sorted key list
__global__ void pann_find_heads_kernel(const u64* sorted, uint32_t ns, uint64_t n, uint32_t* gstart, uint32_t* ngroups);
// K3 over A's tasks: shared-memory launch + the long-list slab launch (no host synchronisation)
// slab_cap_min: the caller's bound on a candidate list (0 = the default 8192-key slabs)
int pann_launch_prune(pann_index* idx, PruneArgs& A, uint32_t ntasks_bound, DevBuf& big, DevBuf& gscratch, int* gs_overflow, uint32_t slab_cap_min);
// radix sort of the reverse-edge keys (cub), bits = significant key bits
int pann_sort_keys(void* tmp, size_t& tmp_bytes, const u64* in, u64* out, int64_t nk, int bits, cudaStream_t s);
