// The step before the path: the index work that turns a cell partition into the part-contiguous
// numbering the solver reads — everything of mini::mesh::Shuffler that is the reference's own deterministic code
// (the METIS calls that produce `cell_parts` are third-party and stay with the host):
//   metis::GetNodeParts          include/mini/mesh/metis.hpp:314-335   node -> smallest part among its cells
//   GetNewOrder                  include/mini/mesh/shuffler.hpp:35-49  stable sort of the indices by part
//   ShuffleData                  shuffler.hpp:57-64                    new[i] = old[new_to_old[i]]
//   ShuffleConnectivity          shuffler.hpp:85-105                   node ids renamed (1-based), cells reordered
// All integer work, bit-exact.
#ifndef MINICFD_B200_SHUFFLE_CUH_
#define MINICFD_B200_SHUFFLE_CUH_

#include <cstdint>

namespace mcfd {
#ifdef __CUDACC__

__global__ void FillKernel(long long *a, long long n, long long value) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) a[i] = value;
}

// one thread per (cell, node) incidence; `cell_of_entry` comes from a binary search in the CSR ranges
__global__ void NodePartsKernel(const long long *range, const long long *nodes, const long long *cell_parts,
    long long n_cells, long long n_entries, long long *node_parts) {
  const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_entries) return;
  long long lo = 0, hi = n_cells;          // largest c with range[c] <= j
  while (hi - lo > 1) {
    const long long mid = (lo + hi) >> 1;
    if (range[mid] <= j) lo = mid; else hi = mid;
  }
  atomicMin(&node_parts[nodes[j]], cell_parts[lo]);
}

// Stable counting sort by part.  Chunk = 1024 consecutive indices per CTA.
//   pass 1: hist[part][chunk] = how many of the chunk's indices carry that part
//   (host: exclusive scan of hist in part-major order -> base[part][chunk])
//   pass 2: new position = base[part][chunk] + rank of the index among the chunk's earlier indices of the same part
constexpr int kSortChunk = 1024;

__global__ void __launch_bounds__(256) PartHistogramKernel(const long long *parts, long long n, int n_parts,
    long long n_chunks, unsigned int *hist) {
  const long long base = (long long)blockIdx.x * kSortChunk;
  for (int k = threadIdx.x; k < kSortChunk; k += 256) {
    if (base + k < n) atomicAdd(&hist[(size_t)parts[base + k] * n_chunks + blockIdx.x], 1u);
  }
}

__global__ void __launch_bounds__(32) PartScatterKernel(const long long *parts, long long n, long long n_chunks,
    const unsigned long long *base_of, unsigned int *running, long long *new_to_old, long long *old_to_new) {
  // one warp walks its chunk 32 indices at a time; `running[part * n_chunks + chunk]` counts what it has placed
  const long long chunk = blockIdx.x, first = chunk * kSortChunk;
  const int lane = threadIdx.x;
  for (int k0 = 0; k0 < kSortChunk; k0 += 32) {
    const long long i = first + k0 + lane;
    const bool live = i < n;
    const long long p = live ? parts[i] : -1 - lane;                      // distinct keys for idle lanes
    const unsigned int same = __match_any_sync(0xffffffffu, p);
    const int rank = __popc(same & ((1u << lane) - 1));
    const int leader = __ffs(same) - 1;
    unsigned int before = 0;
    if (live && lane == leader) {
      unsigned int *slot = &running[(size_t)p * n_chunks + chunk];
      before = *slot;
      *slot = before + __popc(same);
    }
    before = __shfl_sync(0xffffffffu, before, leader);
    if (live) {
      const long long at = (long long)base_of[(size_t)p * n_chunks + chunk] + before + rank;
      new_to_old[at] = i;
      old_to_new[i] = at;
    }
    __syncwarp();
  }
}

__global__ void GatherRowsKernel(const double *src, double *dst, const long long *new_to_old, long long n, long long width) {
  const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n * width) return;
  dst[j] = src[new_to_old[j / width] * width + j % width];
}

__global__ void ShuffleConnectivityKernel(const long long *old_cells, long long *new_cells, const long long *old_to_new_nodes,
    const long long *new_to_old_cells, long long n_cells, int npe) {
  const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_cells * npe) return;
  const long long old_nid = old_cells[new_to_old_cells[j / npe] * npe + j % npe];
  new_cells[j] = old_to_new_nodes[old_nid - 1] + 1;
}

#endif  // __CUDACC__
}  // namespace mcfd
#endif  // MINICFD_B200_SHUFFLE_CUH_
