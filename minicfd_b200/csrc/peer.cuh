// peer.cuh — ghost-cell exchange and the time-step minimum between parts placed one per GPU of an NVLink node.
//
// The reference packs the coefficients of the cells a neighbour needs, sends them with MPI_Isend / MPI_Irecv and unpacks
// them into the ghost cells (Part::ShareGhostCellCoeffs / UpdateGhostCellCoeffs, mesh/part.hpp:1242-1314), once per
// residual; the driver takes MPI_Allreduce(MIN) of the time step once per step (demo/euler/sourceless.cpp:194).
// Here every part keeps, per state buffer, a SEND stage [row][send item] and a RECEIVE stage [row][ghost item] (rows =
// the kFields rows of a tile, items = the cells of the halo lists, all neighbours back to back).  A delivery is
//   1. the send cells' rows of a state buffer in the send stage — written by the stage kernel itself beside its
//      ordinary stores (node-line kernel), or by PeerPackKernel;
//   2. one 2-D device-to-device copy per neighbour (cudaMemcpy2DAsync through the peer mapping) into that neighbour's
//      receive stage: the copy engines move the bytes over NVLink, no SM is involved.  (Stores issued by SMs through a
//      peer pointer were measured first: about 2 GB/s per SM, and every sending block keeps its registers and shared
//      memory while its stores drain — a tenth of a stage at two GPUs, more with more neighbours.)
//   3. PeerSignalKernel: one system-scope release store of the delivery count into an epoch word at each neighbour.
// The receiver polls its own epoch words for the expected count — inside the node-line stage kernel, whose inter-cell
// blocks run last and read the ghost states straight from the receive stage, or in PeerWaitKernel followed by
// PeerUnpackKernel (receive stage -> ghost rows) for the other stage kernels.  All counts are kept or derived on the
// device (explicit pushes, stages of the finished steps, the stage's place in its step), so a captured step can be
// replayed as a CUDA graph.  PeerDtPublishKernel / PeerDtReduceKernel exchange one double per rank the same way.
//
// Why nothing is overwritten while it is still read.  The stages are triple buffered like the state (index = state
// buffer).  A part's receive stage X is read by its stage that has X as INPUT; the neighbour writes X again when it
// delivers the output of a later stage with the same buffer index, which is at least two stages later; that
// delivery follows a stage of the neighbour which itself had to wait for this part's delivery of the stage in
// between — issued after this part finished reading X.  The send stage X is written by the stage with OUTPUT X and
// read by the copies that follow it; the next writer is two stages later and waits for those copies (an event).
// Outside a step (mcfd_peer_push / mcfd_peer_wait) the same buffer may be delivered twice in a row — the limiter's
// unlimited then limited coefficients, a solution written over the previous one: the second delivery goes to a spare
// fourth slot, so that two deliveries in a row never share a slot; the one before the previous is free by the
// argument above (this rank's wait in between needed the neighbour's next push, enqueued behind its unpack).
#ifndef MINICFD_B200_PEER_CUH_
#define MINICFD_B200_PEER_CUH_

namespace mcfd {

constexpr int kPeerMaxNeighbours = 32, kPeerMaxRanks = 64;
constexpr int kPeerSlots = 4, kPeerSpareSlot = 3;                // stages per state buffer, and one spare (mcfd_peer_push)
constexpr int kPeerWords = kPeerMaxNeighbours + kPeerMaxRanks;   // delivery counts per neighbour slot, dt epochs per rank
constexpr long long kPeerSpinLimit = 40000000;                   // about ten seconds of polling before giving up

// what the device keeps count of (mcfd_ctx::Peer::counts)
enum : int {
  kPeerPlain = 0,      // deliveries by mcfd_peer_push
  kPeerStageBase = 1,  // deliveries that followed the stages of the steps before the current one
  kPeerDtEpoch = 2,    // time steps published
  kPeerCounts = 4
};

#ifdef __CUDACC__
__device__ __forceinline__ void StoreReleaseSys(unsigned long long *p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long LoadAcquireSys(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ bool SpinUntil(const unsigned long long *word, unsigned long long count) {
  long long spins = 0;
  while (LoadAcquireSys(word) < count) {
    if (++spins > kPeerSpinLimit) return false;
    __nanosleep(200);
  }
  return true;
}

// rows of the send cells of `src` (tiles) -> send stage [row][item]; consecutive threads take consecutive items of a row
__global__ void PeerPackKernel(const double *src, int fields, const int *send_cells, int n_items, double *stage) {
  const long long total = (long long)n_items * fields;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int x = (int)(idx % n_items), r = (int)(idx / n_items);
    const int s = send_cells[x];
    stage[idx] = src[((size_t)(s >> 5) * fields + r) * kTile + (s & 31)];
  }
}
// receive stage [row][item] -> the ghost rows of `dst` (tiles)
__global__ void PeerUnpackKernel(double *dst, int fields, const int *recv_cells, int n_items, const double *stage) {
  const long long total = (long long)n_items * fields;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int x = (int)(idx % n_items), r = (int)(idx / n_items);
    const int g = recv_cells[x];
    dst[((size_t)(g >> 5) * fields + r) * kTile + (g & 31)] = stage[idx];
  }
}

// The copies before this kernel on the stream are complete: tell the neighbours.  count = explicit pushes (one more
// with bump_plain) + deliveries of the finished steps + `add` (the host's part: deliveries of the current step).
__global__ void PeerSignalKernel(unsigned long long *const *peer_flag, int n, unsigned long long *counts, int bump_plain, int add) {
  __shared__ unsigned long long count;
  if (threadIdx.x == 0) {
    if (bump_plain) ++counts[kPeerPlain];
    count = counts[kPeerPlain] + counts[kPeerStageBase] + (unsigned long long)add;
  }
  __syncthreads();
  if ((int)threadIdx.x < n) StoreReleaseSys(peer_flag[threadIdx.x], count);
}

__global__ void PeerWaitKernel(const unsigned long long *flags, int n, const unsigned long long *counts, int add, int *error) {
  const int i = threadIdx.x;
  if (i < n && !SpinUntil(flags + i, counts[kPeerPlain] + counts[kPeerStageBase] + (unsigned long long)add)) *error = 1;
}

// my time step -> slot [parity][rank] of every rank (my own included), then the epoch word kept for me there.
// Once per step, so it also keeps kPeerStageBase: `deliveries_before` deliveries followed the stages of the previous step.
__global__ void PeerDtPublishKernel(const double *dt_local, int rank, int world, double *const *dt_slots,
    unsigned long long *const *dt_flags, unsigned long long *counts, int deliveries_before) {
  __shared__ unsigned long long e;
  const int r = threadIdx.x;
  if (r == 0) { e = ++counts[kPeerDtEpoch]; counts[kPeerStageBase] += (unsigned long long)deliveries_before; }
  __syncthreads();
  if (r >= world) return;
  dt_slots[r][(int)(e & 1) * world + rank] = *dt_local;
  __threadfence_system();
  StoreReleaseSys(dt_flags[r], e);
}

// MPI_Allreduce(MIN): wait for every rank's value of this step, take the minimum (the same bits on every rank)
__global__ void PeerDtReduceKernel(const unsigned long long *dt_flags, const double *slots, int world,
    const unsigned long long *counts, double *dt_out, int *error) {
  __shared__ double v[kPeerMaxRanks];
  const int r = threadIdx.x;
  const unsigned long long e = counts[kPeerDtEpoch];
  if (r < world) {
    if (!SpinUntil(dt_flags + r, e)) *error = 1;
    v[r] = slots[(int)(e & 1) * world + r];
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
