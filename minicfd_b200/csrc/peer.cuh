// peer.cuh — ghost-cell exchange and the time-step minimum through peer-mapped memory.
//
// The reference packs the coefficients of the cells a neighbour needs, sends them with MPI_Isend / MPI_Irecv and unpacks
// them into the ghost cells (Part::ShareGhostCellCoeffs / UpdateGhostCellCoeffs, mesh/part.hpp:1242-1314), once per
// residual; the driver takes MPI_Allreduce(MIN) of the time step once per step (demo/euler/sourceless.cpp:194).
// With one part per GPU on an NVLink / NVSwitch node every GPU can store into every other GPU's memory, so here
//   * PeerPushKernel copies the rows of the send cells straight into the rows of their ghost images in the
//     neighbours' state buffers (plain st.global through a peer pointer — no pack, no message, no unpack) and, when
//     its last block is done, raises one epoch word per neighbour with a system-scope release store;
//   * PeerWaitKernel holds the stream until every neighbour's epoch word has reached the expected value;
//   * PeerDtPublishKernel / PeerDtReduceKernel do the same for one double per rank.
// Why no buffer is overwritten while a neighbour still reads it: a part's ghost rows are read by its inter-cell stage
// kernels only; stage s of part A starts after every neighbour's push of stage s - 1 has arrived, which those
// neighbours issue after their own stage s - 1; A's push of stage s (into the stage's OUTPUT buffer) follows A's stage
// s.  So when A writes the ghost rows of a neighbour's buffer X, that neighbour is past every stage that read X
// as an input before, and its stage s reads a different buffer (input != output).
#ifndef MINICFD_B200_PEER_CUH_
#define MINICFD_B200_PEER_CUH_

namespace mcfd {

constexpr int kPeerMaxNeighbours = 32, kPeerMaxRanks = 64;
constexpr int kPeerWords = kPeerMaxNeighbours + kPeerMaxRanks;   // halo epochs per neighbour slot, dt epochs per rank
constexpr long long kPeerSpinLimit = 40000000;                   // about ten seconds of polling before giving up

#ifdef __CUDACC__
__device__ __forceinline__ void StoreReleaseSys(unsigned long long *p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long LoadAcquireSys(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ bool SpinUntil(const unsigned long long *word, unsigned long long epoch) {
  long long spins = 0;
  while (LoadAcquireSys(word) < epoch) {
    if (++spins > kPeerSpinLimit) return false;
    __nanosleep(200);
  }
  return true;
}

// Rows of the send cells of `src` (tiles, `fields` rows per tile) -> the same rows of their ghost images in buffer
// `buf` of the neighbours.  grid.x runs over the send items (all neighbours back to back), grid.y over the rows.
__global__ void PeerPushKernel(const double *src, int fields, const int *send_cells, const int *dest_cells,
    const int *item_neighbour, int n_items, double *const *peer_state, int buf,
    unsigned long long *const *peer_flag, int n_neighbours, unsigned long long epoch, unsigned int *done) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x < n_items) {
    const int s = send_cells[x], d = dest_cells[x];
    const double *from = src + (size_t)(s >> 5) * fields * kTile + (s & 31);
    double *to = peer_state[item_neighbour[x] * 3 + buf] + (size_t)(d >> 5) * fields * kTile + (d & 31);
    for (int r = blockIdx.y; r < fields; r += gridDim.y) to[(size_t)r * kTile] = from[(size_t)r * kTile];
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int total = gridDim.x * gridDim.y;
    if (atomicAdd(done, 1u) == total - 1) {   // every other block has fenced its stores
      *done = 0;
      __threadfence_system();
      for (int i = 0; i < n_neighbours; ++i) StoreReleaseSys(peer_flag[i], epoch);
    }
  }
}

__global__ void PeerWaitKernel(const unsigned long long *flags, int n, unsigned long long epoch, int *error) {
  const int i = threadIdx.x;
  if (i < n && !SpinUntil(flags + i, epoch)) *error = 1;
}

// my time step -> slot [parity][rank] of every rank (my own included), then the epoch word kept for me there
__global__ void PeerDtPublishKernel(const double *dt_local, int rank, int world, int parity, double *const *dt_slots,
    unsigned long long *const *dt_flags, unsigned long long epoch) {
  const int r = threadIdx.x;
  if (r >= world) return;
  dt_slots[r][parity * world + rank] = *dt_local;
  __threadfence_system();
  StoreReleaseSys(dt_flags[r], epoch);
}

// MPI_Allreduce(MIN): wait for every rank's value of this step, take the minimum (the same bits on every rank)
__global__ void PeerDtReduceKernel(const unsigned long long *dt_flags, const double *slots, int world, int parity,
    unsigned long long epoch, double *dt_out, int *error) {
  __shared__ double v[kPeerMaxRanks];
  const int r = threadIdx.x;
  if (r < world) {
    if (!SpinUntil(dt_flags + r, epoch)) *error = 1;
    v[r] = slots[parity * world + r];
  }
  __syncthreads();
  if (r == 0) {
    double m = v[0];
    for (int i = 1; i < world; ++i) m = fmin(m, v[i]);
    *dt_out = m;
  }
}
#endif  // __CUDACC__

}  // namespace mcfd
#endif  // MINICFD_B200_PEER_CUH_
