// libmcfd.cu — the C ABI of include/minicfd_b200.h on top of kernels.cuh.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/minicfd_b200.h"
#include "kernels.cuh"
#include "format.cuh"
#include "shuffle.cuh"
#include "tables.hpp"

using namespace mcfd;

namespace {

thread_local std::string g_create_error;

// "this function's attributes have been set" — per device: cudaFuncSetAttribute applies to the current device only
struct DeviceFlags {
  bool done[64] = {false};
  static int Current() { int d = 0; cudaGetDevice(&d); return (d >= 0 && d < 64) ? d : 0; }
  bool Here() const { return done[Current()]; }
  void Set() { done[Current()] = true; }
};

template <class T>
struct DevBuf {
  T *p = nullptr;
  size_t n = 0;
  cudaError_t Alloc(size_t count) {
    Free();
    n = count;
    if (count == 0) return cudaSuccess;
    return cudaMalloc(&p, count * sizeof(T));
  }
  void Free() {
    if (p) cudaFree(p);
    p = nullptr;
    n = 0;
  }
  ~DevBuf() { Free(); }
};

}  // namespace

struct mcfd_ctx {
  mcfd_config cfg{};
  int q = 3, n_nodes = 27, fields = 135;
  int cpb = 32, n_warps = 9;
  int device = 0;                               // the device the context lives on; made current by every call
  bool time_dependent = false;                  // some boundary value depends on time (mcfd_set_time_dependent)
  cudaStream_t stream = nullptr;
  std::string error;
  Tables tables{};
  LineTables line_tables{};                     // the node-line kernel's share of `tables`, a kernel parameter
  DevBuf<unsigned char> line_bytes;             // its byte tables once more, in device memory (staged by cp.async)
  bool use_lines = false;                       // inviscid lumped Lobatto schemes: LineStageKernel (stage_lines.cuh)
  bool use_nodes = false;                       // ... or its persistent one-warp-per-node form (stage_nodes.cuh)
  bool stage_kernel_forced = false;             // MCFD_STAGE_KERNEL chose; else mcfd_set_mesh does (all_affine)
  int sm_count = 148;
  HostMesh mesh;
  bool has_mesh = false;
  int64_t n_cells = 0, n_tiles = 0;
  std::vector<int32_t> btag_of_face;           // per boundary face
  std::vector<unsigned char> bkind_host;
  std::map<int32_t, int32_t> tag_kind;
  DevBuf<double> u[3], res, geom, bvals, length, column;
  DevBuf<double> grad, hn, h2s;                 // viscous traces (TraceKernel) and HolderToSharer vectors
  DevBuf<double> cell_nu;                       // MCFD_CELL_VISCOSITY_DDG: [tile][comp][lane]
  DevBuf<double> vtk_values;                    // mcfd_sample_vtk: [cell][27][5]
  // text of the frame's arrays (format.cuh): tokens, their lengths, per-CTA sums and offsets, packed text
  DevBuf<uint4> fmt_slots;
  DevBuf<unsigned char> fmt_lens;
  DevBuf<unsigned int> fmt_sums, fmt_unsure;    // fmt_unsure: [0] = count, then the indices
  DevBuf<unsigned long long> fmt_offsets;
  DevBuf<char> fmt_text;
  DevBuf<double> fmt_pow10, fmt_values;
  DevBuf<double> damping;                       // EnergyBasedViscosity: [tile][r*N + c][lane] damping matrices
  DevBuf<unsigned long long> visc_dt_bits;      // min_dt_ of the last mcfd_update_cell_viscosity
  bool energy_viscosity = false;                // WithViscosity mode: the time step comes from min_dt_
  DevBuf<double> nu_send, nu_recv;              // packed viscosities of the send / ghost cells (5 doubles per cell)
  DevBuf<double> bas, mod, tay, mod_new, tay_new;   // WENO limiter: per-cell bases and modal coefficients
  DevBuf<double> frames, weno_scratch;          // weno::Eigen: face frames, per-CTA candidate scratch
  std::vector<int64_t> global_ids;              // optional, for the next mcfd_set_mesh
  cudaEvent_t trace_done = nullptr;             // local traces written (inner and inter stages may sit on different streams)
  int64_t first_inter = 0;                      // local cells [first_inter, n_local) have a ghost neighbour
  Viscosity vis{2.0, 1.0 / 12, 0.0, 0.0};       // Direct::SetBetaValues / NavierStokes::SetProperty
  double prandtl = 0.708;
  DevBuf<int> nbr, send_cells, recv_cells;
  std::vector<int64_t> send_counts;             // per neighbour, in list order (mcfd_set_halo)
  DevBuf<unsigned char> code, bkind;
  DevBuf<double> send_buf, recv_buf;
  DevBuf<unsigned long long> dt_bits;
  int cur = 0;                                  // index of the current solution in u[]
  // open RK step
  int rk_order = 0, rk_u0 = 0, rk_in = 0;
  double rk_t = 0, rk_dt = 0;
  const double *rk_dt_dev = nullptr;            // device-resident dt of the open step, or null
  DevBuf<double> dt_dev;                        // part-local dt computed by mcfd_time_step_device
  mcfd_time_callback time_cb = nullptr;
  void *time_user = nullptr;
  int64_t launches = 0;
  // kernel timing
  bool timing = false;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> events;
  size_t events_used = 0;
  int64_t inner_blocks = 0, total_blocks = 0;  // CTA-sized blocks without / with ghost neighbours
  // ---- halo through peer-mapped memory (mcfd_peer_*): the send cells' rows go straight into the neighbours' ghost rows
  struct Peer {
    bool connected = false, pushed_once = false;
    int n_neighbours = 0, rank = 0, world = 1;
    DevBuf<unsigned long long> flags;            // [kPeerWords]: delivery counts per neighbour slot, then dt epochs per rank
    DevBuf<unsigned long long> counts;           // [kPeerCounts] peer.cuh: what the device keeps count of
    DevBuf<double> dt_slots;                     // [2][world]: every rank's time step, by step parity
    DevBuf<double> send_stage, recv_stage;       // [kPeerSlots][fields][n_send | n_recv]: packed halo per state buffer + a spare
    int last_slot = -1;                          // where the latest delivery went (and the next reader looks)
    bool last_plain = false;                     // ... and whether it was a mcfd_peer_push
    DevBuf<int> recv_index;                      // [n_ghost] ghost cell -> item of the receive stage
    DevBuf<int> tgt_offset, tgt_item;            // per local cell from first_inter on: its items of the send stage
    std::vector<double *> nbr_recv;              // [n_neighbours] the neighbour's receive stage, at this part's segment
    std::vector<int64_t> nbr_recv_total;         // [n_neighbours] row length of that stage
    std::vector<int64_t> send_offset;            // [n_neighbours + 1] this part's segments of the send stage
    DevBuf<unsigned long long *> flag_dev;       // [n_neighbours] the word to raise at each neighbour
    DevBuf<double *> dt_slot_dev;                // [world] every rank's dt_slots
    DevBuf<unsigned long long *> dt_flag_dev;    // [world] every rank's dt epoch word for this rank
    DevBuf<int> error;                           // set when a wait gave up
    std::vector<void *> opened;                  // cudaIpcOpenMemHandle results
    int deliveries_this_step = 0;                // stage deliveries of the step in progress / just finished (host side)
    cudaStream_t side = nullptr;
    cudaEvent_t ev_main = nullptr, ev_side = nullptr, ev_copy[3] = {nullptr, nullptr, nullptr};
    unsigned long long ev_copy_capture[3] = {~0ull, ~0ull, ~0ull};   // capture the event was recorded in (0: none; ~0: never)
  } peer;
};

namespace {

int Fail(mcfd_ctx *ctx, int status, std::string msg) {
  if (ctx) ctx->error = std::move(msg);
  else g_create_error = std::move(msg);
  return status;
}

// The kernels other than the node-line stage kernel (whose tables are a kernel parameter) read the per-degree tables
// from ONE __constant__ symbol per device, while every context owns its own copy (degree, scheme, launch shape and
// Vincent parameter differ between contexts).  Before a context launches such a kernel the symbol is brought in line
// with it.  A switch waits for everything in flight on the device (kernels of the other shape may still be reading the
// old constants, on any stream) and is complete before the call returns (whatever stream the context uses next finds
// the new ones): contexts of different shape may be used in turn on any streams; they take turns on the device.
// Not during stream capture — run a step of the context first.
Tables g_device_tables[64];
bool g_device_tables_valid[64] = {false};

cudaError_t EnsureTables(mcfd_ctx *ctx) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  if (g_device_tables_valid[dev] && std::memcmp(&g_device_tables[dev], &ctx->tables, sizeof(Tables)) == 0)
    return cudaSuccess;
  g_device_tables_valid[dev] = false;
  e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return e;
  e = cudaMemcpyToSymbol(c_tab, &ctx->tables, sizeof(Tables), 0, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) return e;
  e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return e;
  g_device_tables[dev] = ctx->tables;
  g_device_tables_valid[dev] = true;
  return cudaSuccess;
}

#define MCFD_CUDA(ctx, call)                                                          \
  do {                                                                                \
    cudaError_t e_ = (call);                                                          \
    if (e_ != cudaSuccess)                                                            \
      return Fail(ctx, MCFD_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
  } while (0)

template <int Q, int R, int CPB, int NW, bool kGeneral, int kViscous = kDiffNone>
cudaError_t LaunchStageG(mcfd_ctx *ctx, StageArgs const &a, int n_blocks) {
  size_t smem = StageGeometry<Q, CPB, NW>::kSharedBytes;
  auto kernel = StageKernel<Q, R, CPB, NW, kGeneral, kViscous>;
  static DeviceFlags configured;
  if (!configured.Here()) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    configured.Set();
  }
  kernel<<<n_blocks, NW * 32, smem, ctx->stream>>>(a);
  return cudaGetLastError();
}

// The node-line kernel (stage_lines.cuh): inviscid fr::Lobatto / dg::Lobatto.  Its tables are a kernel parameter.
template <int Q, int R, int CPB, bool kAffine>
cudaError_t LaunchLineA(mcfd_ctx *ctx, StageArgs const &a, int n_blocks) {
  using S = LineShape<Q, CPB>;
  auto kernel = LineStageKernel<Q, R, CPB, kAffine>;
  static DeviceFlags configured;
  if (!configured.Here()) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::kSharedBytes);
    if (e != cudaSuccess) return e;
    // all of the SM's shared memory: the default carve-out holds one CTA of this size, the kernel is built for kMinBlocks
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return e;
    configured.Set();
  }
  kernel<<<n_blocks, S::kThreads, S::kSharedBytes, ctx->stream>>>(a, ctx->line_tables);
  return cudaGetLastError();
}
// the persistent one-warp-per-node kernel: one CTA per SM walks over the tiles [a.block_begin, + n_tiles)
template <int R, bool kAffine>
cudaError_t LaunchNodeA(mcfd_ctx *ctx, StageArgs const &a, int n_tiles) {
  using S = NodeShape;
  auto kernel = NodeStageKernel<R, kAffine>;
  static DeviceFlags configured;
  if (!configured.Here()) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::kSharedBytes);
    if (e != cudaSuccess) return e;
    configured.Set();
  }
  kernel<<<std::min(n_tiles, ctx->sm_count), S::kThreads, S::kSharedBytes, ctx->stream>>>(a, ctx->line_tables, n_tiles);
  return cudaGetLastError();
}
template <int R>
cudaError_t LaunchNodeR(mcfd_ctx *ctx, StageArgs const &a, int n_tiles) {
  return ctx->mesh.all_affine ? LaunchNodeA<R, true>(ctx, a, n_tiles) : LaunchNodeA<R, false>(ctx, a, n_tiles);
}
cudaError_t LaunchNode(mcfd_ctx *ctx, StageArgs const &a, int n_tiles) {
  switch (ctx->cfg.riemann) {
  case MCFD_EXACT: return LaunchNodeR<kExact>(ctx, a, n_tiles);
  case MCFD_HLLC: return LaunchNodeR<kHllc>(ctx, a, n_tiles);
  default: return LaunchNodeR<kAusm>(ctx, a, n_tiles);
  }
}

// a mesh of parallelepipeds only (uniform and sheared boxes) takes the instantiation with the metric per cell
template <int Q, int R, int CPB>
cudaError_t LaunchLineR(mcfd_ctx *ctx, StageArgs const &a, int n_blocks) {
  return ctx->mesh.all_affine ? LaunchLineA<Q, R, CPB, true>(ctx, a, n_blocks) : LaunchLineA<Q, R, CPB, false>(ctx, a, n_blocks);
}
template <int Q, int CPB>
cudaError_t LaunchLine(mcfd_ctx *ctx, StageArgs const &a, int n_blocks) {
  switch (ctx->cfg.riemann) {
  case MCFD_EXACT: return LaunchLineR<Q, kExact, CPB>(ctx, a, n_blocks);
  case MCFD_HLLC: return LaunchLineR<Q, kHllc, CPB>(ctx, a, n_blocks);
  default: return LaunchLineR<Q, kAusm, CPB>(ctx, a, n_blocks);
  }
}

// the general Vincent correction (fr::General) is a separate instantiation: the lumped schemes pay nothing
template <int Q, int R, int CPB, int NW>
cudaError_t LaunchStageT(mcfd_ctx *ctx, StageArgs const &a, int n_blocks) {
  return ctx->cfg.scheme == MCFD_FR_GENERAL ? LaunchStageG<Q, R, CPB, NW, true>(ctx, a, n_blocks)
                                            : LaunchStageG<Q, R, CPB, NW, false>(ctx, a, n_blocks);
}

template <int Q, int CPB, int NW>
cudaError_t LaunchStageQ(mcfd_ctx *ctx, StageArgs const &a, int n_blocks) {
  switch (ctx->cfg.riemann) {
  case MCFD_EXACT: return LaunchStageT<Q, kExact, CPB, NW>(ctx, a, n_blocks);
  case MCFD_HLLC: return LaunchStageT<Q, kHllc, CPB, NW>(ctx, a, n_blocks);
  default: return LaunchStageT<Q, kAusm, CPB, NW>(ctx, a, n_blocks);
  }
}

template <int Q, int CPB, int NW, int kDiff, bool kGeneral = false>
cudaError_t LaunchStageDiffusive(mcfd_ctx *ctx, StageArgs const &a, int n_blocks) {
  switch (ctx->cfg.riemann) {
  case MCFD_EXACT: return LaunchStageG<Q, kExact, CPB, NW, kGeneral, kDiff>(ctx, a, n_blocks);
  case MCFD_HLLC: return LaunchStageG<Q, kHllc, CPB, NW, kGeneral, kDiff>(ctx, a, n_blocks);
  default: return LaunchStageG<Q, kAusm, CPB, NW, kGeneral, kDiff>(ctx, a, n_blocks);
  }
}
template <int Q, int CPB, int NW>
cudaError_t LaunchStageViscous(mcfd_ctx *ctx, StageArgs const &a, int n_blocks) {
  if (a.mode == kModeProbe) {   // damping-matrix probes never solve a Riemann problem: one instantiation serves all
    StageArgs b = a;
    b.mode = kModeResidual;
    return LaunchStageG<Q, kHllc, CPB, NW, false, kDiffProbe>(ctx, b, n_blocks);
  }
  if (ctx->cfg.scheme == MCFD_FR_GENERAL)   // fr::General on the Navier-Stokes Riemann type
    return LaunchStageDiffusive<Q, CPB, NW, kDiffNavierStokes, true>(ctx, a, n_blocks);
  if (ctx->cfg.diffusion == MCFD_NAVIER_STOKES_CELL_VISCOSITY_DDG) return LaunchStageDiffusive<Q, CPB, NW, kDiffNsIsotropic>(ctx, a, n_blocks);
  return ctx->cfg.diffusion == MCFD_NAVIER_STOKES_DDG ? LaunchStageDiffusive<Q, CPB, NW, kDiffNavierStokes>(ctx, a, n_blocks)
                                                      : LaunchStageDiffusive<Q, CPB, NW, kDiffIsotropic>(ctx, a, n_blocks);
}

// Viscous traces (gradient at the nodes, Hessian . normal at the face points) of cells
// [cell_begin, cell_end) of the state `u`.
int LaunchTrace(mcfd_ctx *ctx, const double *u, int64_t cell_begin, int64_t cell_end) {
  if (cell_end <= cell_begin) return MCFD_OK;
  const int tile_begin = (int)(cell_begin / kTile), tile_end = (int)((cell_end + kTile - 1) / kTile);
  cudaError_t e = EnsureTables(ctx);
  if (e != cudaSuccess) return Fail(ctx, MCFD_ERR_CUDA, cudaGetErrorString(e));
  if (ctx->q == 3) {
    if (ctx->cfg.k_local)
      TraceKernel<3, true><<<tile_end - tile_begin, TraceShape<3>::NW * 32, TraceShape<3>::kSharedBytes, ctx->stream>>>(u,
          ctx->geom.p, ctx->grad.p, ctx->hn.p, tile_begin, (int)cell_begin, (int)cell_end);
    else
    TraceKernel<3><<<tile_end - tile_begin, TraceShape<3>::NW * 32, TraceShape<3>::kSharedBytes, ctx->stream>>>(u, ctx->geom.p,
        ctx->grad.p, ctx->hn.p, tile_begin, (int)cell_begin, (int)cell_end);
    e = cudaGetLastError();
  } else if (ctx->cfg.k_local) {
    auto kernel = TraceKernel<4, true>;
    static DeviceFlags configured;
    if (!configured.Here()) {
      e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TraceShape<4>::kSharedBytes);
      if (e != cudaSuccess) return Fail(ctx, MCFD_ERR_CUDA, cudaGetErrorString(e));
      configured.Set();
    }
    kernel<<<tile_end - tile_begin, TraceShape<4>::NW * 32, TraceShape<4>::kSharedBytes, ctx->stream>>>(u, ctx->geom.p, ctx->grad.p,
        ctx->hn.p, tile_begin, (int)cell_begin, (int)cell_end);
    e = cudaGetLastError();
  } else {
    auto kernel = TraceKernel<4>;
    static DeviceFlags configured;
    if (!configured.Here()) {
      e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TraceShape<4>::kSharedBytes);
      if (e != cudaSuccess) return Fail(ctx, MCFD_ERR_CUDA, cudaGetErrorString(e));
      configured.Set();
    }
    kernel<<<tile_end - tile_begin, TraceShape<4>::NW * 32, TraceShape<4>::kSharedBytes, ctx->stream>>>(u, ctx->geom.p, ctx->grad.p,
        ctx->hn.p, tile_begin, (int)cell_begin, (int)cell_end);
    e = cudaGetLastError();
  }
  if (e != cudaSuccess) return Fail(ctx, MCFD_ERR_CUDA, std::string("trace kernel: ") + cudaGetErrorString(e));
  ctx->launches++;
  return MCFD_OK;
}

// dg::General on the modal polynomial: Taylor form of the cells [cell_begin, cell_end) of `u`
int LaunchDgmTaylor(mcfd_ctx *ctx, const double *u, int64_t cell_begin, int64_t cell_end) {
  if (cell_end <= cell_begin) return MCFD_OK;
  const int t0 = (int)(cell_begin / kTile), t1 = (int)((cell_end + kTile - 1) / kTile);
  MCFD_CUDA(ctx, EnsureTables(ctx));
  if (ctx->q == 3) DgmTaylorKernel<3><<<t1 - t0, kTile, 0, ctx->stream>>>(u, ctx->bas.p, ctx->tay.p, t0, (int)cell_begin, (int)cell_end);
  else DgmTaylorKernel<4><<<t1 - t0, kTile, 0, ctx->stream>>>(u, ctx->bas.p, ctx->tay.p, t0, (int)cell_begin, (int)cell_end);
  MCFD_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  return MCFD_OK;
}

template <int Q>
cudaError_t LaunchDgmQ(mcfd_ctx *ctx, DgmArgs const &a, int t0, int n_tiles) {
  switch (ctx->cfg.riemann) {
  case MCFD_EXACT: DgmStageKernel<Q, kExact><<<n_tiles, kTile, 0, ctx->stream>>>(a, t0); break;
  case MCFD_HLLC: DgmStageKernel<Q, kHllc><<<n_tiles, kTile, 0, ctx->stream>>>(a, t0); break;
  default: DgmStageKernel<Q, kAusm><<<n_tiles, kTile, 0, ctx->stream>>>(a, t0); break;
  }
  return cudaGetLastError();
}

// contexts with one viscosity per cell and component: alone, or on top of the Navier-Stokes terms
static bool HasCellViscosity(const mcfd_ctx *ctx) {
  return ctx->cfg.diffusion == MCFD_CELL_VISCOSITY_DDG || ctx->cfg.diffusion == MCFD_NAVIER_STOKES_CELL_VISCOSITY_DDG;
}

// Launch the stage kernel over CTA blocks [block_begin, block_end).
int LaunchStage(mcfd_ctx *ctx, const double *u_in, const double *u0, double *u_out, int mode,
    double dt, int64_t block_begin, int64_t block_end, const StageArgs::PeerStage *fused = nullptr) {
  if (block_end <= block_begin) return MCFD_OK;
  StageArgs a{};
  if (fused) a.peer = *fused;
  a.u_in = u_in; a.u0 = u0; a.u_out = u_out;
  a.geom = ctx->geom.p; a.nbr = ctx->nbr.p; a.code = ctx->code.p;
  a.bkind = ctx->bkind.p; a.bvals = ctx->bvals.p;
  a.n_local = (int)ctx->mesh.n_local;
  a.block_begin = (int)block_begin;
  a.mode = mode; a.dt = dt; a.dt_dev = ctx->rk_order ? ctx->rk_dt_dev : nullptr;
  a.grad = ctx->grad.p; a.hn = ctx->hn.p; a.h2s = ctx->h2s.p; a.vis = ctx->vis; a.cell_nu = ctx->cell_nu.p;
  a.line_bytes = ctx->line_bytes.p;
  int n_blocks = (int)(block_end - block_begin);
  if (!ctx->use_lines) MCFD_CUDA(ctx, EnsureTables(ctx));
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (ctx->timing) {
    if (ctx->events_used == ctx->events.size()) {
      cudaEvent_t a0, a1;
      MCFD_CUDA(ctx, cudaEventCreate(&a0));
      MCFD_CUDA(ctx, cudaEventCreate(&a1));
      ctx->events.emplace_back(a0, a1);
    }
    e0 = ctx->events[ctx->events_used].first;
    e1 = ctx->events[ctx->events_used].second;
    ctx->events_used++;
    MCFD_CUDA(ctx, cudaEventRecord(e0, ctx->stream));
  }
  cudaError_t e;
  if (ctx->cfg.scheme == MCFD_DG_GENERAL) {   // blocks are tiles here
    DgmArgs d;
    d.u_in = u_in; d.u0 = u0; d.u_out = u_out; d.tay = ctx->tay.p; d.geom = ctx->geom.p; d.bas = ctx->bas.p;
    d.h2s = ctx->h2s.p; d.nbr = ctx->nbr.p; d.code = ctx->code.p; d.bkind = ctx->bkind.p; d.bvals = ctx->bvals.p;
    d.cell_begin = (int)(block_begin * kTile);
    d.cell_end = (int)std::min<int64_t>(block_end * kTile, ctx->mesh.n_local);
    d.mode = mode; d.dt = dt; d.dt_dev = ctx->rk_order ? ctx->rk_dt_dev : nullptr;
    MCFD_CUDA(ctx, EnsureTables(ctx));
    e = ctx->q == 3 ? LaunchDgmQ<3>(ctx, d, (int)block_begin, (int)(block_end - block_begin))
                                : LaunchDgmQ<4>(ctx, d, (int)block_begin, (int)(block_end - block_begin));
  } else if (ctx->cfg.diffusion != MCFD_INVISCID) {
    e = (ctx->q == 3) ? LaunchStageViscous<3, kP2ViscousCells, kP2ViscousWarps>(ctx, a, n_blocks)
                      : LaunchStageViscous<4, kP3ViscousCells, kP3ViscousWarps>(ctx, a, n_blocks);
  } else if (ctx->use_nodes) {
    e = LaunchNode(ctx, a, n_blocks);
  } else if (ctx->use_lines) {
    e = LaunchLine<3, kLineP2Cells>(ctx, a, n_blocks);   // degree 3 stays on the slot kernel (use_lines)
  } else {
    e = (ctx->q == 3) ? LaunchStageQ<3, kP2Cells, kP2Warps>(ctx, a, n_blocks)
                      : LaunchStageQ<4, kP3Cells, kP3Warps>(ctx, a, n_blocks);
  }
  if (e != cudaSuccess) return Fail(ctx, MCFD_ERR_CUDA, std::string("stage kernel: ") + cudaGetErrorString(e));
  if (ctx->timing) MCFD_CUDA(ctx, cudaEventRecord(e1, ctx->stream));
  ctx->launches++;
  return MCFD_OK;
}

template <bool kToTiles>
int LaunchConvert(mcfd_ctx *ctx, double *tiles, double *column, const int *cells, int64_t n_items) {
  if (n_items == 0) return MCFD_OK;
  int blocks = (int)((n_items + kTile - 1) / kTile);
  cudaError_t e;
  if (ctx->fields == 50 || ctx->fields == 100) {   // dg::General: 5 x M modal coefficients per cell
    if (ctx->fields == 50) ConvertKernel<50, kToTiles><<<blocks, 256, (size_t)kTile * 51 * sizeof(double), ctx->stream>>>(tiles, column, cells, (int)n_items);
    else ConvertKernel<100, kToTiles><<<blocks, 256, (size_t)kTile * 101 * sizeof(double), ctx->stream>>>(tiles, column, cells, (int)n_items);
    e = cudaGetLastError();
  } else if (ctx->q == 3) {
    constexpr int F = 135;
    size_t smem = (size_t)kTile * (F + 1) * sizeof(double);
    ConvertKernel<F, kToTiles><<<blocks, 256, smem, ctx->stream>>>(tiles, column, cells, (int)n_items);
    e = cudaGetLastError();
  } else {
    constexpr int F = 320;
    size_t smem = (size_t)kTile * (F + 1) * sizeof(double);
    auto kernel = ConvertKernel<F, kToTiles>;
    static DeviceFlags configured;
    if (!configured.Here()) {
      e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return Fail(ctx, MCFD_ERR_CUDA, cudaGetErrorString(e));
      configured.Set();
    }
    kernel<<<blocks, 256, smem, ctx->stream>>>(tiles, column, cells, (int)n_items);
    e = cudaGetLastError();
  }
  if (e != cudaSuccess) return Fail(ctx, MCFD_ERR_CUDA, std::string("convert kernel: ") + cudaGetErrorString(e));
  ctx->launches++;
  return MCFD_OK;
}

// kLocal == true contexts: columns hold U = det(J) u, the device keeps u
int ScaleColumnByDet(mcfd_ctx *ctx, double *column_dev) {
  if (!ctx->cfg.k_local) return MCFD_OK;
  MCFD_CUDA(ctx, EnsureTables(ctx));
  const int nl = (int)ctx->mesh.n_local, n = nl * ctx->n_nodes;
  if (ctx->q == 3) ScaleColumnKernel<3><<<(n + 255) / 256, 256, 0, ctx->stream>>>(column_dev, ctx->geom.p, nl);
  else ScaleColumnKernel<4><<<(n + 255) / 256, 256, 0, ctx->stream>>>(column_dev, ctx->geom.p, nl);
  MCFD_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  return MCFD_OK;
}
int UnscaleTilesByDet(mcfd_ctx *ctx, double *tiles) {
  if (!ctx->cfg.k_local) return MCFD_OK;
  MCFD_CUDA(ctx, EnsureTables(ctx));
  const int nl = (int)ctx->mesh.n_local;
  if (ctx->q == 3) ScaleTilesKernel<3><<<(nl + 127) / 128, 128, 0, ctx->stream>>>(tiles, ctx->geom.p, 0, nl, 1);
  else ScaleTilesKernel<4><<<(nl + 127) / 128, 128, 0, ctx->stream>>>(tiles, ctx->geom.p, 0, nl, 1);
  MCFD_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  return MCFD_OK;
}

int RequireMesh(mcfd_ctx *ctx) {
  if (!ctx) return MCFD_ERR_ARG;
  if (!ctx->has_mesh) return Fail(ctx, MCFD_ERR_STATE, "mcfd_set_mesh has not been called");
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev != ctx->device) MCFD_CUDA(ctx, cudaSetDevice(ctx->device));
  return MCFD_OK;
}

int EnsureColumn(mcfd_ctx *ctx) {
  size_t need = (size_t)ctx->mesh.n_local * ctx->fields;
  if (ctx->column.n < need) MCFD_CUDA(ctx, ctx->column.Alloc(need));
  return MCFD_OK;
}

// row-per-cell host array -> tile layout
template <class T>
std::vector<T> ToTiles(const std::vector<T> &src, int rows, int64_t n_cells, int64_t n_tiles) {
  std::vector<T> out((size_t)n_tiles * rows * kTile, T{});
  for (int64_t c = 0; c < n_cells; ++c)
    for (int r = 0; r < rows; ++r)
      out[((c >> 5) * rows + r) * kTile + (c & 31)] = src[c * rows + r];
  return out;
}

void StageBuffers(mcfd_ctx *ctx, int stage, int *in, int *out, int *mode) {
  // three buffers rotate: u0 stays put for the whole step
  const int u0 = ctx->rk_u0, a = (u0 + 1) % 3, b = (u0 + 2) % 3;
  if (ctx->rk_order == 1) { *in = u0; *out = a; *mode = kModeEuler; return; }
  if (ctx->rk_order == 2) {
    if (stage == 0) { *in = u0; *out = a; *mode = kModeEuler; }
    else { *in = a; *out = b; *mode = kModeRk2b; }
    return;
  }
  if (stage == 0) { *in = u0; *out = a; *mode = kModeEuler; }
  else if (stage == 1) { *in = a; *out = b; *mode = kModeRk3b; }
  else { *in = b; *out = a; *mode = kModeRk3c; }
}

}  // namespace

extern "C" {

int mcfd_create(const mcfd_config *config, mcfd_ctx **ctx_out) {
  if (!config || !ctx_out) return Fail(nullptr, MCFD_ERR_ARG, "null argument");
  if (config->degree != 2 && config->degree != 3)
    return Fail(nullptr, MCFD_ERR_ARG, "degree must be 2 or 3");
  if (config->scheme < MCFD_FR_LOBATTO || config->scheme > MCFD_DG_GENERAL)
    return Fail(nullptr, MCFD_ERR_ARG, "unknown scheme");
  if (config->scheme == MCFD_DG_GENERAL && config->diffusion != MCFD_INVISCID)
    return Fail(nullptr, MCFD_ERR_ARG, "the reference's DG schemes have no diffusive terms");
  if (config->riemann < MCFD_EXACT || config->riemann > MCFD_AUSM)
    return Fail(nullptr, MCFD_ERR_ARG, "unknown Riemann solver");
  if (config->diffusion < MCFD_INVISCID || config->diffusion > MCFD_NAVIER_STOKES_CELL_VISCOSITY_DDG)
    return Fail(nullptr, MCFD_ERR_ARG, "unknown diffusion model");
  // the reference's DG schemes carry no diffusive terms at all (spatial/dg/*.hpp); fr::General has them for any Riemann
  // type, built here for Navier-Stokes
  if (config->diffusion != MCFD_INVISCID && config->scheme == MCFD_DG_LOBATTO)
    return Fail(nullptr, MCFD_ERR_ARG, "the reference's DG schemes have no diffusive terms; use MCFD_FR_LOBATTO or MCFD_FR_GENERAL");
  if (config->k_local != 0 && config->k_local != 1) return Fail(nullptr, MCFD_ERR_ARG, "k_local must be 0 or 1");
  if (config->k_local && (config->scheme == MCFD_DG_LOBATTO || config->scheme == MCFD_DG_GENERAL ||
                          config->diffusion == MCFD_CELL_VISCOSITY_DDG || config->diffusion == MCFD_NAVIER_STOKES_CELL_VISCOSITY_DDG))
    return Fail(nullptr, MCFD_ERR_ARG, "k_local is built for the FR schemes on Euler and Navier-Stokes");
  if ((config->diffusion == MCFD_CELL_VISCOSITY_DDG || config->diffusion == MCFD_NAVIER_STOKES_CELL_VISCOSITY_DDG) &&
      config->scheme != MCFD_FR_LOBATTO)
    return Fail(nullptr, MCFD_ERR_ARG, "the per-cell viscosity is built for MCFD_FR_LOBATTO only");
  int n_dev = 0;
  cudaError_t e = cudaGetDeviceCount(&n_dev);
  if (e != cudaSuccess || n_dev == 0)
    return Fail(nullptr, MCFD_ERR_CUDA, std::string("no CUDA device: ") +
        (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
        " (libminicfd_b200 has no CPU fallback)");
  if (config->device >= 0) {
    e = cudaSetDevice(config->device);
    if (e != cudaSuccess) return Fail(nullptr, MCFD_ERR_CUDA, cudaGetErrorString(e));
  }
  auto ctx = std::make_unique<mcfd_ctx>();
  ctx->cfg = *config;
  e = cudaGetDevice(&ctx->device);
  if (e != cudaSuccess) return Fail(nullptr, MCFD_ERR_CUDA, cudaGetErrorString(e));
  ctx->q = config->degree + 1;
  ctx->n_nodes = ctx->q * ctx->q * ctx->q;
  ctx->fields = 5 * ctx->n_nodes;
  if (config->scheme == MCFD_DG_GENERAL) ctx->fields = 5 * (ctx->q == 3 ? 10 : 20);   // K x M modal coefficients
  const bool viscous = config->diffusion != MCFD_INVISCID;
  ctx->cpb = (config->scheme == MCFD_DG_GENERAL) ? kTile : ((ctx->q == 3) ? (viscous ? kP2ViscousCells : kP2Cells) : (viscous ? kP3ViscousCells : kP3Cells));
  ctx->n_warps = (ctx->q == 3) ? (viscous ? kP2ViscousWarps : kP2Warps) : (viscous ? kP3ViscousWarps : kP3Warps);
  ctx->tables = BuildTables(config->degree, config->scheme, ctx->n_warps, kTile / ctx->cpb, config->fr_c_next);
  ctx->use_lines = !viscous && ctx->q == 3 && (config->scheme == MCFD_FR_LOBATTO || config->scheme == MCFD_DG_LOBATTO);
  if (ctx->use_lines) {
    // "lines" | "nodes" for A/B runs; without it mcfd_set_mesh decides: the node kernel where every cell is a
    // parallelepiped (141 against 155 us per stage on the benchmark box), the node-line kernel elsewhere (164 : 166)
    const char *which = std::getenv("MCFD_STAGE_KERNEL");
    ctx->stage_kernel_forced = which != nullptr;
    ctx->use_nodes = which ? std::string(which) == "nodes" : false;
    cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, ctx->device);
    ctx->cpb = ctx->use_nodes ? kTile : ((ctx->q == 3) ? kLineP2Cells : kLineP3Cells);
    ctx->n_warps = (ctx->q == 3) ? LineShape<3, kLineP2Cells>::NW : LineShape<4, kLineP3Cells>::NW;
    ctx->line_tables = MakeLineTables(ctx->tables, ctx->q, kTile / ctx->cpb);
    const std::vector<unsigned char> bytes = LineTableBytes(ctx->line_tables, ctx->q);
    e = ctx->line_bytes.Alloc(bytes.size());
    if (e == cudaSuccess) e = cudaMemcpy(ctx->line_bytes.p, bytes.data(), bytes.size(), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) return Fail(nullptr, MCFD_ERR_CUDA, cudaGetErrorString(e));
  }
  e = ctx->dt_bits.Alloc(1);
  if (e != cudaSuccess) return Fail(nullptr, MCFD_ERR_CUDA, cudaGetErrorString(e));
  *ctx_out = ctx.release();
  return MCFD_OK;
}

int mcfd_destroy(mcfd_ctx *ctx) {
  if (!ctx) return MCFD_OK;
  for (auto &p : ctx->events) { cudaEventDestroy(p.first); cudaEventDestroy(p.second); }
  if (ctx->trace_done) cudaEventDestroy(ctx->trace_done);
  for (void *p : ctx->peer.opened) cudaIpcCloseMemHandle(p);
  if (ctx->peer.ev_main) cudaEventDestroy(ctx->peer.ev_main);
  if (ctx->peer.ev_side) cudaEventDestroy(ctx->peer.ev_side);
  for (auto e : ctx->peer.ev_copy) if (e) cudaEventDestroy(e);
  if (ctx->peer.side) cudaStreamDestroy(ctx->peer.side);
  delete ctx;
  return MCFD_OK;
}

const char *mcfd_last_error(const mcfd_ctx *ctx) {
  return ctx ? ctx->error.c_str() : g_create_error.c_str();
}

int mcfd_set_stream(mcfd_ctx *ctx, void *cuda_stream) {
  if (!ctx) return MCFD_ERR_ARG;
  ctx->stream = static_cast<cudaStream_t>(cuda_stream);
  return MCFD_OK;
}

int mcfd_set_mesh(mcfd_ctx *ctx, int64_t n_nodes, const double *xyz, int64_t n_local,
    int64_t n_ghost, const int64_t *cell_nodes, int64_t n_bfaces, const int64_t *bface_nodes,
    const int32_t *bface_tags, int64_t n_pairs, const int64_t *periodic_pairs) {
  if (!ctx || !xyz || !cell_nodes || n_local <= 0 || n_ghost < 0)
    return Fail(ctx, MCFD_ERR_ARG, "bad mesh arguments");
  if (n_local + n_ghost > (int64_t)1 << 30) return Fail(ctx, MCFD_ERR_ARG, "too many cells for int32 ids");
  MCFD_CUDA(ctx, cudaSetDevice(ctx->device));
  ctx->mesh = HostMesh();
  if (!ctx->global_ids.empty()) {
    if ((int64_t)ctx->global_ids.size() != n_local + n_ghost)
      return Fail(ctx, MCFD_ERR_ARG, "mcfd_set_cell_global_ids was given a different number of cells");
    ctx->mesh.global_ids = ctx->global_ids;
  }
  if (!ctx->mesh.Build(ctx->cfg.degree, n_nodes, xyz, n_local, n_ghost, cell_nodes, n_bfaces,
          bface_nodes, bface_tags, n_pairs, periodic_pairs))
    return Fail(ctx, MCFD_ERR_MESH, ctx->mesh.error);
  ctx->n_cells = n_local + n_ghost;
  ctx->n_tiles = (ctx->n_cells + kTile - 1) / kTile;
  auto geom = ToTiles(ctx->mesh.coef, kGeomRows, ctx->n_cells, ctx->n_tiles);
  auto nbr = ToTiles(ctx->mesh.nbr, 6, ctx->n_cells, ctx->n_tiles);
  auto code = ToTiles(ctx->mesh.code, 6, ctx->n_cells, ctx->n_tiles);
  MCFD_CUDA(ctx, ctx->geom.Alloc(geom.size()));
  MCFD_CUDA(ctx, ctx->nbr.Alloc(nbr.size()));
  MCFD_CUDA(ctx, ctx->code.Alloc(code.size()));
  MCFD_CUDA(ctx, cudaMemcpy(ctx->geom.p, geom.data(), geom.size() * sizeof(double), cudaMemcpyHostToDevice));
  MCFD_CUDA(ctx, cudaMemcpy(ctx->nbr.p, nbr.data(), nbr.size() * sizeof(int), cudaMemcpyHostToDevice));
  MCFD_CUDA(ctx, cudaMemcpy(ctx->code.p, code.data(), code.size(), cudaMemcpyHostToDevice));
  MCFD_CUDA(ctx, ctx->length.Alloc(n_local));
  MCFD_CUDA(ctx, cudaMemcpy(ctx->length.p, ctx->mesh.length.data(), n_local * sizeof(double), cudaMemcpyHostToDevice));
  size_t state = (size_t)ctx->n_tiles * ctx->fields * kTile;
  for (auto &b : ctx->u) {
    MCFD_CUDA(ctx, b.Alloc(state));
    MCFD_CUDA(ctx, cudaMemset(b.p, 0, state * sizeof(double)));
  }
  ctx->res.Free();
  {
    auto h2s = ToTiles(ctx->mesh.h2s, 18, ctx->n_cells, ctx->n_tiles);
    MCFD_CUDA(ctx, ctx->h2s.Alloc(h2s.size()));
    MCFD_CUDA(ctx, cudaMemcpy(ctx->h2s.p, h2s.data(), h2s.size() * sizeof(double), cudaMemcpyHostToDevice));
  }
  ctx->bas.Free(); ctx->tay.Free(); ctx->frames.Free(); ctx->weno_scratch.Free(); ctx->damping.Free(); ctx->vtk_values.Free(); ctx->energy_viscosity = false;
  if (ctx->cfg.scheme == MCFD_DG_GENERAL) {   // Projection's orthonormal bases on the Gauss-Legendre rule
    const int rows = ctx->q == 3 ? ModalShape<2>::kBasRows : ModalShape<3>::kBasRows;
    auto bas = ToTiles(ctx->mesh.BuildModalBases(ctx->cfg.degree, true), rows, ctx->n_cells, ctx->n_tiles);
    MCFD_CUDA(ctx, ctx->bas.Alloc(bas.size()));
    MCFD_CUDA(ctx, cudaMemcpy(ctx->bas.p, bas.data(), bas.size() * sizeof(double), cudaMemcpyHostToDevice));
    MCFD_CUDA(ctx, ctx->tay.Alloc(state));
    MCFD_CUDA(ctx, cudaMemset(ctx->tay.p, 0, state * sizeof(double)));
  }
  if (ctx->cfg.diffusion != MCFD_INVISCID) {
    MCFD_CUDA(ctx, ctx->grad.Alloc((size_t)ctx->n_tiles * ctx->n_nodes * 15 * kTile));
    MCFD_CUDA(ctx, ctx->hn.Alloc((size_t)ctx->n_tiles * 6 * ctx->q * ctx->q * 15 * kTile));
    MCFD_CUDA(ctx, cudaMemset(ctx->grad.p, 0, ctx->grad.n * sizeof(double)));
    MCFD_CUDA(ctx, cudaMemset(ctx->hn.p, 0, ctx->hn.n * sizeof(double)));
    if (HasCellViscosity(ctx)) {   // starts uniform at the nu of mcfd_set_diffusion; at zero on top of Navier-Stokes
      std::vector<double> nu((size_t)ctx->n_tiles * 5 * kTile, ctx->cfg.diffusion == MCFD_CELL_VISCOSITY_DDG ? ctx->vis.nu : 0.0);
      MCFD_CUDA(ctx, ctx->cell_nu.Alloc(nu.size()));
      MCFD_CUDA(ctx, cudaMemcpy(ctx->cell_nu.p, nu.data(), nu.size() * sizeof(double), cudaMemcpyHostToDevice));
    }
  }
  size_t nb = ctx->mesh.bound_faces.size();
  ctx->btag_of_face.resize(nb);
  for (size_t b = 0; b < nb; ++b) ctx->btag_of_face[b] = ctx->mesh.bound_faces[b].tag;
  ctx->bkind_host.assign(nb, MCFD_BC_INVISCID_WALL);
  for (size_t b = 0; b < nb; ++b) {
    auto it = ctx->tag_kind.find(ctx->btag_of_face[b]);
    if (it != ctx->tag_kind.end()) ctx->bkind_host[b] = (unsigned char)it->second;
  }
  MCFD_CUDA(ctx, ctx->bkind.Alloc(std::max<size_t>(nb, 1)));
  MCFD_CUDA(ctx, ctx->bvals.Alloc(std::max<size_t>(nb, 1) * ctx->q * ctx->q * 5));
  MCFD_CUDA(ctx, cudaMemset(ctx->bvals.p, 0, ctx->bvals.n * sizeof(double)));
  if (nb) MCFD_CUDA(ctx, cudaMemcpy(ctx->bkind.p, ctx->bkind_host.data(), nb, cudaMemcpyHostToDevice));
  if (ctx->use_lines && !ctx->stage_kernel_forced) {
    ctx->use_nodes = ctx->mesh.all_affine;
    ctx->cpb = ctx->use_nodes ? kTile : kLineP2Cells;
  }
  // CTA blocks that contain only inner cells: cells are ordered inner first (part.hpp:746-767);
  // a local cell is "inter" when one of its neighbours is a ghost.
  int64_t first_inter = n_local;
  for (int64_t c = 0; c < n_local && first_inter == n_local; ++c)
    for (int f = 0; f < 6; ++f)
      if (ctx->mesh.nbr[c * 6 + f] >= n_local) { first_inter = c; break; }
  ctx->first_inter = first_inter;
  ctx->total_blocks = (n_local + ctx->cpb - 1) / ctx->cpb;
  // without ghosts the whole part goes in one launch; with ghosts only full blocks of inner cells
  ctx->inner_blocks = (n_ghost == 0) ? ctx->total_blocks : first_inter / ctx->cpb;
  ctx->send_cells.Free(); ctx->recv_cells.Free(); ctx->send_buf.Free(); ctx->recv_buf.Free();
  ctx->cur = 0;
  ctx->has_mesh = true;
  return MCFD_OK;
}

int mcfd_set_cell_global_ids(mcfd_ctx *ctx, int64_t n_cells, const int64_t *ids) {
  if (!ctx || n_cells < 0 || (n_cells > 0 && !ids)) return Fail(ctx, MCFD_ERR_ARG, "bad arguments");
  ctx->global_ids.assign(ids, ids + n_cells);
  return MCFD_OK;
}

int64_t mcfd_get_faces(const mcfd_ctx *ctx, int kind, int64_t *holder, int64_t *sharer,
    int32_t *holder_face, int32_t *sharer_face, int32_t *orientation) {
  if (!ctx || !ctx->has_mesh) return MCFD_ERR_STATE;
  auto const &faces = kind == 0 ? ctx->mesh.local_faces : kind == 1 ? ctx->mesh.ghost_faces
                                                                    : ctx->mesh.bound_faces;
  for (size_t i = 0; i < faces.size(); ++i) {
    if (holder) holder[i] = faces[i].holder;
    if (sharer) sharer[i] = faces[i].sharer;
    if (holder_face) holder_face[i] = faces[i].holder_face;
    if (sharer_face) sharer_face[i] = faces[i].sharer_face;
    if (orientation) orientation[i] = faces[i].orientation;
  }
  return (int64_t)faces.size();
}

int mcfd_get_cell_geometry(const mcfd_ctx *ctx, double *length, double *volume) {
  if (!ctx || !ctx->has_mesh) return MCFD_ERR_STATE;
  if (length) std::memcpy(length, ctx->mesh.length.data(), ctx->mesh.length.size() * sizeof(double));
  if (volume) std::memcpy(volume, ctx->mesh.volume.data(), ctx->mesh.volume.size() * sizeof(double));
  return MCFD_OK;
}

int mcfd_set_halo(mcfd_ctx *ctx, int32_t n_send_parts, const int64_t *send_counts,
    const int64_t *send_cells, int32_t n_recv_parts, const int64_t *recv_counts,
    const int64_t *recv_cells) {
  if (int s = RequireMesh(ctx)) return s;
  if (n_send_parts < 0 || n_recv_parts < 0 || (n_send_parts && !send_counts) || (n_recv_parts && !recv_counts))
    return Fail(ctx, MCFD_ERR_ARG, "bad halo lists");
  int64_t ns = 0, nr = 0;
  for (int i = 0; i < n_send_parts; ++i) ns += send_counts[i];
  for (int i = 0; i < n_recv_parts; ++i) if (recv_counts[i] < 0) return Fail(ctx, MCFD_ERR_ARG, "negative halo count");
  for (int i = 0; i < n_send_parts; ++i) if (send_counts[i] < 0) return Fail(ctx, MCFD_ERR_ARG, "negative halo count");
  for (int i = 0; i < n_recv_parts; ++i) nr += recv_counts[i];
  if ((ns && !send_cells) || (nr && !recv_cells)) return Fail(ctx, MCFD_ERR_ARG, "null halo cell list");
  std::vector<int> s32(ns), r32(nr);
  for (int64_t i = 0; i < ns; ++i) {
    if (send_cells[i] < 0 || send_cells[i] >= ctx->mesh.n_local)
      return Fail(ctx, MCFD_ERR_ARG, "send cell is not a local cell");
    s32[i] = (int)send_cells[i];
  }
  for (int64_t i = 0; i < nr; ++i) {
    if (recv_cells[i] < ctx->mesh.n_local || recv_cells[i] >= ctx->n_cells)
      return Fail(ctx, MCFD_ERR_ARG, "recv cell is not a ghost cell");
    r32[i] = (int)recv_cells[i];
  }
  ctx->send_counts.assign(send_counts, send_counts + n_send_parts);
  ctx->peer.connected = false;
  MCFD_CUDA(ctx, ctx->send_cells.Alloc(ns));
  MCFD_CUDA(ctx, ctx->recv_cells.Alloc(nr));
  MCFD_CUDA(ctx, ctx->send_buf.Alloc(ns * ctx->fields));
  MCFD_CUDA(ctx, ctx->recv_buf.Alloc(nr * ctx->fields));
  if (ns) MCFD_CUDA(ctx, cudaMemcpy(ctx->send_cells.p, s32.data(), ns * sizeof(int), cudaMemcpyHostToDevice));
  if (nr) MCFD_CUDA(ctx, cudaMemcpy(ctx->recv_cells.p, r32.data(), nr * sizeof(int), cudaMemcpyHostToDevice));
  return MCFD_OK;
}

int mcfd_halo_buffers(mcfd_ctx *ctx, double **send_dev, int64_t *send_doubles, double **recv_dev,
    int64_t *recv_doubles) {
  if (int s = RequireMesh(ctx)) return s;
  if (send_dev) *send_dev = ctx->send_buf.p;
  if (send_doubles) *send_doubles = (int64_t)ctx->send_buf.n;
  if (recv_dev) *recv_dev = ctx->recv_buf.p;
  if (recv_doubles) *recv_doubles = (int64_t)ctx->recv_buf.n;
  return MCFD_OK;
}

// The buffer that the next residual / stage reads: the current solution outside a step, the stage
// input inside one.
static double *HaloState(mcfd_ctx *ctx) { return ctx->u[ctx->rk_order ? ctx->rk_in : ctx->cur].p; }

int mcfd_halo_pack(mcfd_ctx *ctx) {
  if (int s = RequireMesh(ctx)) return s;
  return LaunchConvert<false>(ctx, HaloState(ctx), ctx->send_buf.p, ctx->send_cells.p,
      (int64_t)ctx->send_cells.n);
}

int mcfd_halo_unpack(mcfd_ctx *ctx) {
  if (int s = RequireMesh(ctx)) return s;
  return LaunchConvert<true>(ctx, HaloState(ctx), ctx->recv_buf.p, ctx->recv_cells.p,
      (int64_t)ctx->recv_cells.n);
}

int mcfd_set_boundary(mcfd_ctx *ctx, int32_t tag, int32_t bc_kind) {
  if (!ctx) return MCFD_ERR_ARG;
  if (bc_kind < MCFD_BC_INVISCID_WALL || bc_kind > MCFD_BC_NO_SLIP_WALL)
    return Fail(ctx, MCFD_ERR_ARG, "unknown boundary kind");
  ctx->tag_kind[tag] = bc_kind;
  if (ctx->has_mesh) {
    size_t nb = ctx->bkind_host.size();
    for (size_t b = 0; b < nb; ++b)
      if (ctx->btag_of_face[b] == tag) ctx->bkind_host[b] = (unsigned char)bc_kind;
    if (nb) MCFD_CUDA(ctx, cudaMemcpyAsync(ctx->bkind.p, ctx->bkind_host.data(), nb, cudaMemcpyHostToDevice, ctx->stream));
    MCFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  return MCFD_OK;
}

int64_t mcfd_count_boundary_faces(const mcfd_ctx *ctx, int32_t tag) {
  if (!ctx || !ctx->has_mesh) return MCFD_ERR_STATE;
  int64_t n = 0;
  for (int32_t t : ctx->btag_of_face) n += (t == tag);
  return n;
}

int mcfd_get_boundary_points(const mcfd_ctx *ctx, int32_t tag, double *xyz) {
  if (!ctx || !ctx->has_mesh || !xyz) return MCFD_ERR_STATE;
  if (ctx->cfg.scheme == MCFD_DG_GENERAL) ctx->mesh.BoundaryGaussPoints(ctx->tables, ctx->cfg.degree, tag, xyz);
  else ctx->mesh.BoundaryPoints(ctx->tables, ctx->cfg.degree, tag, xyz);
  return MCFD_OK;
}

int mcfd_set_boundary_values(mcfd_ctx *ctx, int32_t tag, const double *values) {
  if (int s = RequireMesh(ctx)) return s;
  if (!values) return Fail(ctx, MCFD_ERR_ARG, "null values");
  const size_t per_face = (size_t)ctx->q * ctx->q * 5;
  // faces of one tag are contiguous in the boundary table (grouped by tag)
  size_t nb = ctx->btag_of_face.size(), first = nb, count = 0;
  for (size_t b = 0; b < nb; ++b)
    if (ctx->btag_of_face[b] == tag) { if (first == nb) first = b; ++count; }
  if (count == 0) return MCFD_OK;
  MCFD_CUDA(ctx, cudaMemcpyAsync(ctx->bvals.p + first * per_face, values,
      count * per_face * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  MCFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // `values` may be reused by the caller
  return MCFD_OK;
}

int mcfd_set_diffusion(mcfd_ctx *ctx, double beta_0, double beta_1, double nu, double prandtl) {
  if (!ctx) return MCFD_ERR_ARG;
  if (ctx->cfg.diffusion == MCFD_INVISCID) return Fail(ctx, MCFD_ERR_STATE, "the context was created inviscid");
  if (!(prandtl > 0) || nu < 0) return Fail(ctx, MCFD_ERR_ARG, "need nu >= 0 and prandtl > 0");
  ctx->vis.beta0 = beta_0;
  ctx->vis.beta1 = beta_1;
  ctx->vis.nu = nu;
  ctx->vis.cond = nu * (kGasR * kGammaOverGm1) / prandtl;   // nu * Cp / Pr (navier_stokes.hpp:76-79)
  ctx->prandtl = prandtl;
  if (ctx->cfg.diffusion == MCFD_CELL_VISCOSITY_DDG && ctx->cell_nu.p) {   // Isotropic::SetProperty(nu): uniform
    std::vector<double> v((size_t)ctx->n_tiles * 5 * kTile, nu);
    MCFD_CUDA(ctx, cudaMemcpyAsync(ctx->cell_nu.p, v.data(), v.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    MCFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  return MCFD_OK;
}

int mcfd_build_damping_matrices(mcfd_ctx *ctx) {
  if (int s = RequireMesh(ctx)) return s;
  if (!HasCellViscosity(ctx))
    return Fail(ctx, MCFD_ERR_STATE, "the context was not created with a per-cell viscosity (MCFD_CELL_VISCOSITY_DDG, MCFD_NAVIER_STOKES_CELL_VISCOSITY_DDG)");
  if (ctx->rk_order) return Fail(ctx, MCFD_ERR_STATE, "an RK step is open");
  const int n = ctx->n_nodes;
  const size_t state = (size_t)ctx->n_tiles * ctx->fields * kTile;
  MCFD_CUDA(ctx, ctx->damping.Alloc((size_t)ctx->n_tiles * n * n * kTile));
  if (!ctx->visc_dt_bits.p) MCFD_CUDA(ctx, ctx->visc_dt_bits.Alloc(1));
  if (ctx->res.n < state) MCFD_CUDA(ctx, ctx->res.Alloc(state));
  // the probes use two of the three state buffers that do not hold the current solution
  double *probe = ctx->u[(ctx->cur + 1) % 3].p;
  const int threads = 256;
  for (int col = 0; col < n; ++col) {   // viscosity.hpp:338-349: one unit nodal vector after the other
    ProbeStateKernel<<<(int)((state + threads - 1) / threads), threads, 0, ctx->stream>>>(probe, ctx->fields,
        (int)ctx->n_tiles, col);
    if (int s = LaunchTrace(ctx, probe, 0, ctx->n_cells)) return s;
    if (int s = LaunchStage(ctx, probe, probe, ctx->res.p, kModeProbe, 0.0, 0, ctx->total_blocks)) return s;
    const int nl = (int)ctx->mesh.n_local;
    if (ctx->q == 3) DampingColumnKernel<3><<<(nl + 127) / 128, 128, 0, ctx->stream>>>(ctx->res.p, ctx->geom.p, ctx->damping.p, col, nl);
    else DampingColumnKernel<4><<<(nl + 127) / 128, 128, 0, ctx->stream>>>(ctx->res.p, ctx->geom.p, ctx->damping.p, col, nl);
    ctx->launches += 2;
  }
  MCFD_CUDA(ctx, cudaGetLastError());
  MCFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->energy_viscosity = true;
  return MCFD_OK;
}

int mcfd_get_damping_matrices(mcfd_ctx *ctx, double *matrices) {
  if (int s = RequireMesh(ctx)) return s;
  if (!ctx->damping.p || !matrices) return Fail(ctx, MCFD_ERR_STATE, "mcfd_build_damping_matrices has not been called");
  const int n = ctx->n_nodes;
  std::vector<double> tiles(ctx->damping.n);
  MCFD_CUDA(ctx, cudaMemcpy(tiles.data(), ctx->damping.p, tiles.size() * sizeof(double), cudaMemcpyDeviceToHost));
  for (int64_t c = 0; c < ctx->mesh.n_local; ++c)
    for (int i = 0; i < n * n; ++i) matrices[c * n * n + i] = tiles[((size_t)(c >> 5) * n * n + i) * kTile + (c & 31)];
  return MCFD_OK;
}

int mcfd_update_cell_viscosity(mcfd_ctx *ctx, double time_scale) {
  if (int s = RequireMesh(ctx)) return s;
  if (!ctx->damping.p) return Fail(ctx, MCFD_ERR_STATE, "mcfd_build_damping_matrices has not been called");
  MCFD_CUDA(ctx, EnsureTables(ctx));
  double *u = HaloState(ctx);
  TimeStepResetKernel<<<1, 1, 0, ctx->stream>>>(ctx->visc_dt_bits.p);
  const int nl = (int)ctx->mesh.n_local, tiles = (nl + kTile - 1) / kTile;
  static DeviceFlags configured;
  if (!configured.Here()) {
    MCFD_CUDA(ctx, cudaFuncSetAttribute(ViscosityKernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ViscosityShape<3>::kSharedBytes));
    MCFD_CUDA(ctx, cudaFuncSetAttribute(ViscosityKernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ViscosityShape<4>::kSharedBytes));
    MCFD_CUDA(ctx, cudaFuncSetAttribute(ViscosityKernel<3>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    MCFD_CUDA(ctx, cudaFuncSetAttribute(ViscosityKernel<4>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    configured.Set();
  }
  if (ctx->q == 3) ViscosityKernel<3><<<tiles, 160, ViscosityShape<3>::kSharedBytes, ctx->stream>>>(u, ctx->geom.p, ctx->nbr.p,
      ctx->code.p, ctx->damping.p, ctx->length.p, nl, time_scale, ctx->cell_nu.p, ctx->visc_dt_bits.p);
  else ViscosityKernel<4><<<tiles, 160, ViscosityShape<4>::kSharedBytes, ctx->stream>>>(u, ctx->geom.p, ctx->nbr.p,
      ctx->code.p, ctx->damping.p, ctx->length.p, nl, time_scale, ctx->cell_nu.p, ctx->visc_dt_bits.p);
  MCFD_CUDA(ctx, cudaGetLastError());
  ctx->launches += 2;
  return MCFD_OK;
}

int mcfd_viscosity_halo_buffers(mcfd_ctx *ctx, double **send_dev, int64_t *send_doubles, double **recv_dev,
    int64_t *recv_doubles) {
  if (int s = RequireMesh(ctx)) return s;
  if (!ctx->cell_nu.p) return Fail(ctx, MCFD_ERR_STATE, "no per-cell viscosity in this context");
  if (ctx->nu_send.n != ctx->send_cells.n * 5) MCFD_CUDA(ctx, ctx->nu_send.Alloc(ctx->send_cells.n * 5));
  if (ctx->nu_recv.n != ctx->recv_cells.n * 5) MCFD_CUDA(ctx, ctx->nu_recv.Alloc(ctx->recv_cells.n * 5));
  if (send_dev) *send_dev = ctx->nu_send.p;
  if (send_doubles) *send_doubles = (int64_t)ctx->nu_send.n;
  if (recv_dev) *recv_dev = ctx->nu_recv.p;
  if (recv_doubles) *recv_doubles = (int64_t)ctx->nu_recv.n;
  return MCFD_OK;
}

static int ViscosityHalo(mcfd_ctx *ctx, int unpack) {
  if (int s = mcfd_viscosity_halo_buffers(ctx, nullptr, nullptr, nullptr, nullptr)) return s;
  const int n = (int)(unpack ? ctx->recv_cells.n : ctx->send_cells.n);
  if (n == 0) return MCFD_OK;
  ViscosityHaloKernel<<<(n * 5 + 127) / 128, 128, 0, ctx->stream>>>(ctx->cell_nu.p, unpack ? ctx->nu_recv.p : ctx->nu_send.p,
      unpack ? ctx->recv_cells.p : ctx->send_cells.p, n, unpack);
  MCFD_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  return MCFD_OK;
}

int mcfd_viscosity_halo_pack(mcfd_ctx *ctx) { return ViscosityHalo(ctx, 0); }
int mcfd_viscosity_halo_unpack(mcfd_ctx *ctx) { return ViscosityHalo(ctx, 1); }

int mcfd_get_cell_viscosity(mcfd_ctx *ctx, double *nu) {
  if (int s = RequireMesh(ctx)) return s;
  if (!ctx->cell_nu.p || !nu) return Fail(ctx, MCFD_ERR_STATE, "no per-cell viscosity in this context");
  MCFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  std::vector<double> tiles(ctx->cell_nu.n);
  MCFD_CUDA(ctx, cudaMemcpy(tiles.data(), ctx->cell_nu.p, tiles.size() * sizeof(double), cudaMemcpyDeviceToHost));
  for (int64_t c = 0; c < ctx->n_cells; ++c)
    for (int k = 0; k < 5; ++k) nu[c * 5 + k] = tiles[((size_t)(c >> 5) * 5 + k) * kTile + (c & 31)];
  return MCFD_OK;
}

int mcfd_set_cell_viscosity(mcfd_ctx *ctx, int64_t n_cells, const double *nu) {
  if (int s = RequireMesh(ctx)) return s;
  if (!HasCellViscosity(ctx))
    return Fail(ctx, MCFD_ERR_STATE, "the context was not created with a per-cell viscosity (MCFD_CELL_VISCOSITY_DDG, MCFD_NAVIER_STOKES_CELL_VISCOSITY_DDG)");
  if (!nu || n_cells != ctx->n_cells) return Fail(ctx, MCFD_ERR_ARG, "need [n_local + n_ghost][5] values");
  std::vector<double> rows((size_t)n_cells * 5);
  std::copy(nu, nu + n_cells * 5, rows.begin());
  auto tiles = ToTiles(rows, 5, ctx->n_cells, ctx->n_tiles);
  MCFD_CUDA(ctx, cudaMemcpyAsync(ctx->cell_nu.p, tiles.data(), tiles.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  MCFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return MCFD_OK;
}

int mcfd_get_jacobian_determinants(const mcfd_ctx *ctx, double *det) {
  if (!ctx || !ctx->has_mesh || !det) return MCFD_ERR_STATE;
  const int q = ctx->q;
  for (int64_t c = 0; c < ctx->mesh.n_local; ++c) {
    const double *g = &ctx->mesh.coef[c * kGeomRows];
    for (int i = 0; i < q; ++i) for (int j = 0; j < q; ++j) for (int k = 0; k < q; ++k) {
      const HostMetric m = ComputeMetricHost(g, ctx->tables.xi[i], ctx->tables.xi[j], ctx->tables.xi[k]);
      *det++ = m.det;
    }
  }
  return MCFD_OK;
}

int mcfd_get_quadrature_points(const mcfd_ctx *ctx, double *xyz) {
  if (!ctx || !ctx->has_mesh || !xyz) return MCFD_ERR_STATE;
  if (ctx->cfg.scheme != MCFD_DG_GENERAL) return MCFD_ERR_STATE;
  ctx->mesh.QuadraturePoints(ctx->tables, ctx->cfg.degree, xyz);
  return MCFD_OK;
}

int mcfd_project(mcfd_ctx *ctx, const double *values) {
  if (int s = RequireMesh(ctx)) return s;
  if (ctx->cfg.scheme != MCFD_DG_GENERAL) return Fail(ctx, MCFD_ERR_STATE, "mcfd_project is for MCFD_DG_GENERAL contexts");
  if (!values) return Fail(ctx, MCFD_ERR_ARG, "null values");
  if (ctx->rk_order) return Fail(ctx, MCFD_ERR_STATE, "an RK step is open");
  const size_t n = (size_t)ctx->mesh.n_local * ctx->n_nodes * 5;
  DevBuf<double> dv;
  MCFD_CUDA(ctx, dv.Alloc(n));
  MCFD_CUDA(ctx, cudaMemcpyAsync(dv.p, values, n * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  MCFD_CUDA(ctx, EnsureTables(ctx));
  const int nl = (int)ctx->mesh.n_local;
  if (ctx->q == 3) DgmProjectKernel<3><<<(nl + 63) / 64, 64, 0, ctx->stream>>>(dv.p, ctx->geom.p, ctx->bas.p, ctx->u[ctx->cur].p, nl);
  else DgmProjectKernel<4><<<(nl + 63) / 64, 64, 0, ctx->stream>>>(dv.p, ctx->geom.p, ctx->bas.p, ctx->u[ctx->cur].p, nl);
  MCFD_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  MCFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return MCFD_OK;
}

int mcfd_set_time_callback(mcfd_ctx *ctx, mcfd_time_callback fn, void *user) {
  if (!ctx) return MCFD_ERR_ARG;
  ctx->time_cb = fn;
  ctx->time_user = user;
  return MCFD_OK;
}

int mcfd_set_time_dependent(mcfd_ctx *ctx, int flag) {
  if (!ctx) return MCFD_ERR_ARG;
  ctx->time_dependent = flag != 0;
  return MCFD_OK;
}

int64_t mcfd_count_local_cells(const mcfd_ctx *ctx) {
  return (ctx && ctx->has_mesh) ? ctx->mesh.n_local : (int64_t)MCFD_ERR_STATE;
}

int64_t mcfd_cell_data_size(const mcfd_ctx *ctx) {
  return (ctx && ctx->has_mesh) ? ctx->mesh.n_local * ctx->fields : (int64_t)MCFD_ERR_STATE;
}

int mcfd_set_solution_device(mcfd_ctx *ctx, const double *column_dev) {
  if (int s = RequireMesh(ctx)) return s;
  if (ctx->rk_order) return Fail(ctx, MCFD_ERR_STATE, "an RK step is open");
  if (int s = LaunchConvert<true>(ctx, ctx->u[ctx->cur].p, const_cast<double *>(column_dev), nullptr, ctx->mesh.n_local))
    return s;
  return UnscaleTilesByDet(ctx, ctx->u[ctx->cur].p);
}

int mcfd_get_solution_device(mcfd_ctx *ctx, double *column_dev) {
  if (int s = RequireMesh(ctx)) return s;
  if (int s = LaunchConvert<false>(ctx, ctx->u[ctx->cur].p, column_dev, nullptr, ctx->mesh.n_local)) return s;
  return ScaleColumnByDet(ctx, column_dev);
}

int mcfd_set_solution(mcfd_ctx *ctx, const double *column) {
  if (int s = RequireMesh(ctx)) return s;
  if (!column) return Fail(ctx, MCFD_ERR_ARG, "null column");
  if (int s = EnsureColumn(ctx)) return s;
  size_t bytes = (size_t)ctx->mesh.n_local * ctx->fields * sizeof(double);
  MCFD_CUDA(ctx, cudaMemcpyAsync(ctx->column.p, column, bytes, cudaMemcpyHostToDevice, ctx->stream));
  if (int s = mcfd_set_solution_device(ctx, ctx->column.p)) return s;
  MCFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return MCFD_OK;
}

int mcfd_get_solution(mcfd_ctx *ctx, double *column) {
  if (int s = RequireMesh(ctx)) return s;
  if (!column) return Fail(ctx, MCFD_ERR_ARG, "null column");
  if (int s = EnsureColumn(ctx)) return s;
  if (int s = mcfd_get_solution_device(ctx, ctx->column.p)) return s;
  size_t bytes = (size_t)ctx->mesh.n_local * ctx->fields * sizeof(double);
  MCFD_CUDA(ctx, cudaMemcpyAsync(column, ctx->column.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  MCFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return MCFD_OK;
}

int mcfd_residual(mcfd_ctx *ctx, double t, double *column, double *column_dev) {
  if (int s = RequireMesh(ctx)) return s;
  if (ctx->rk_order) return Fail(ctx, MCFD_ERR_STATE, "an RK step is open");
  size_t state = (size_t)ctx->n_tiles * ctx->fields * kTile;
  if (ctx->res.n < state) MCFD_CUDA(ctx, ctx->res.Alloc(state));
  if (ctx->time_cb) ctx->time_cb(ctx->time_user, t);
  const double *u = ctx->u[ctx->cur].p;
  if (ctx->cfg.diffusion != MCFD_INVISCID) {
    if (int s = LaunchTrace(ctx, u, 0, ctx->n_cells)) return s;
  }
  if (ctx->cfg.scheme == MCFD_DG_GENERAL) {
    if (int s = LaunchDgmTaylor(ctx, u, 0, ctx->n_cells)) return s;
  }
  if (int s = LaunchStage(ctx, u, u, ctx->res.p, kModeResidual, 0.0, 0, ctx->total_blocks)) return s;
  if (column_dev) {   // kLocal: the column is dU/dt = det du/dt (fr/general.hpp:251-253)
    if (int s = LaunchConvert<false>(ctx, ctx->res.p, column_dev, nullptr, ctx->mesh.n_local)) return s;
    if (int s = ScaleColumnByDet(ctx, column_dev)) return s;
  }
  if (column) {
    if (int s = EnsureColumn(ctx)) return s;
    if (int s = LaunchConvert<false>(ctx, ctx->res.p, ctx->column.p, nullptr, ctx->mesh.n_local)) return s;
    if (int s = ScaleColumnByDet(ctx, ctx->column.p)) return s;
    size_t bytes = (size_t)ctx->mesh.n_local * ctx->fields * sizeof(double);
    MCFD_CUDA(ctx, cudaMemcpyAsync(column, ctx->column.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  }
  MCFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return MCFD_OK;
}

int mcfd_rk_begin(mcfd_ctx *ctx, int order, double t, double dt) {
  if (int s = RequireMesh(ctx)) return s;
  if (order < 1 || order > 3) return Fail(ctx, MCFD_ERR_ARG, "RK order must be 1, 2 or 3");
  if (ctx->rk_order) return Fail(ctx, MCFD_ERR_STATE, "an RK step is already open");
  ctx->rk_order = order;
  ctx->rk_u0 = ctx->cur;
  ctx->rk_in = ctx->cur;
  ctx->rk_t = t;
  ctx->rk_dt = dt;
  ctx->rk_dt_dev = nullptr;
  return MCFD_OK;
}

int mcfd_rk_begin_device(mcfd_ctx *ctx, int order, double t, const double *dt_dev) {
  if (!dt_dev) return Fail(ctx, MCFD_ERR_ARG, "null dt_dev");
  if (ctx && ctx->time_cb && ctx->time_dependent && order > 1)
    return Fail(ctx, MCFD_ERR_STATE, "time-dependent boundary values need the stage times t + dt, t + dt / 2 on the host: "
        "use mcfd_rk_begin / mcfd_rk_step with the time step read back (mcfd_time_step)");
  if (int s = mcfd_rk_begin(ctx, order, t, 0.0)) return s;
  ctx->rk_dt_dev = dt_dev;
  return MCFD_OK;
}

double mcfd_rk_stage_time(const mcfd_ctx *ctx, int stage) {
  // rk.hpp:16-49: RK1 (t); RK2 (t, t+dt); RK3 (t, t+dt, t+dt/2)
  if (stage == 0) return ctx->rk_t;
  if (stage == 1) return ctx->rk_t + ctx->rk_dt;
  return ctx->rk_t + ctx->rk_dt / 2;
}

int mcfd_rk_stage_inner(mcfd_ctx *ctx, int stage) {
  if (int s = RequireMesh(ctx)) return s;
  if (!ctx->rk_order || stage < 0 || stage >= ctx->rk_order)
    return Fail(ctx, MCFD_ERR_STATE, "no such stage in the open RK step");
  int in, out, mode;
  StageBuffers(ctx, stage, &in, &out, &mode);
  if (in != ctx->rk_in) return Fail(ctx, MCFD_ERR_STATE, "stages must be run in order");
  if (ctx->time_cb) ctx->time_cb(ctx->time_user, mcfd_rk_stage_time(ctx, stage));
  if (ctx->cfg.scheme == MCFD_DG_GENERAL) {
    if (int s = LaunchDgmTaylor(ctx, ctx->u[in].p, 0, ctx->mesh.n_local)) return s;
    if (!ctx->trace_done) MCFD_CUDA(ctx, cudaEventCreateWithFlags(&ctx->trace_done, cudaEventDisableTiming));
    MCFD_CUDA(ctx, cudaEventRecord(ctx->trace_done, ctx->stream));
  }
  if (ctx->cfg.diffusion != MCFD_INVISCID) {   // the inner cells' neighbours are all local
    if (int s = LaunchTrace(ctx, ctx->u[in].p, 0, ctx->mesh.n_local)) return s;
    if (!ctx->trace_done) MCFD_CUDA(ctx, cudaEventCreateWithFlags(&ctx->trace_done, cudaEventDisableTiming));
    MCFD_CUDA(ctx, cudaEventRecord(ctx->trace_done, ctx->stream));
  }
  return LaunchStage(ctx, ctx->u[in].p, ctx->u[ctx->rk_u0].p, ctx->u[out].p, mode, ctx->rk_dt, 0,
      ctx->inner_blocks);
}

int mcfd_rk_stage_inter(mcfd_ctx *ctx, int stage) {
  if (int s = RequireMesh(ctx)) return s;
  if (!ctx->rk_order || stage < 0 || stage >= ctx->rk_order)
    return Fail(ctx, MCFD_ERR_STATE, "no such stage in the open RK step");
  int in, out, mode;
  StageBuffers(ctx, stage, &in, &out, &mode);
  if (in != ctx->rk_in) return Fail(ctx, MCFD_ERR_STATE, "stages must be run in order");
  if (ctx->cfg.scheme == MCFD_DG_GENERAL) {
    if (ctx->trace_done) MCFD_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->trace_done, 0));
    if (int s = LaunchDgmTaylor(ctx, ctx->u[in].p, ctx->mesh.n_local, ctx->n_cells)) return s;
  }
  if (ctx->cfg.diffusion != MCFD_INVISCID) {   // the ghost cells hold this stage's values by now
    // the inter cells also read the local traces, which mcfd_rk_stage_inner may have launched on another stream
    if (ctx->trace_done) MCFD_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->trace_done, 0));
    if (int s = LaunchTrace(ctx, ctx->u[in].p, ctx->mesh.n_local, ctx->n_cells)) return s;
  }
  if (int s = LaunchStage(ctx, ctx->u[in].p, ctx->u[ctx->rk_u0].p, ctx->u[out].p, mode, ctx->rk_dt,
          ctx->inner_blocks, ctx->total_blocks)) return s;
  ctx->rk_in = out;
  if (stage == ctx->rk_order - 1) {  // step complete
    ctx->cur = out;
    ctx->rk_order = 0;
  }
  return MCFD_OK;
}

int mcfd_rk_step(mcfd_ctx *ctx, int order, double t, double dt) {
  if (int s = mcfd_rk_begin(ctx, order, t, dt)) return s;
  for (int stage = 0; stage < order; ++stage) {
    if (int s = mcfd_rk_stage_inner(ctx, stage)) { ctx->rk_order = 0; return s; }
    if (int s = mcfd_rk_stage_inter(ctx, stage)) { ctx->rk_order = 0; return s; }
  }
  return MCFD_OK;
}

static const double kCfl[5][9] = {   // fem.hpp:26-37
  {1.000},
  {1.000, 0.333},
  {1.000, 0.333, 0.209, 0.130, 0.089, 0.066, 0.051, 0.040, 0.033},
  {1.256, 0.409, 0.209, 0.130, 0.089, 0.066, 0.051, 0.040, 0.033},
  {1.392, 0.464, 0.235, 0.145, 0.100, 0.073, 0.056, 0.045, 0.037}};

int mcfd_time_step_device(mcfd_ctx *ctx, double dt_guess, int rk_order, double *dt_dev) {
  if (int s = RequireMesh(ctx)) return s;
  if (rk_order < 0 || rk_order > 4) return Fail(ctx, MCFD_ERR_ARG, "bad RK order");
  if (!ctx->dt_dev.p) MCFD_CUDA(ctx, ctx->dt_dev.Alloc(1));
  if (!dt_dev) dt_dev = ctx->dt_dev.p;
  if (ctx->energy_viscosity) {   // WithViscosity::GetTimeStep (with_viscosity.hpp:67-71): CFL * min_dt_ of UpdateProperties
    TimeStepFinishKernel<<<1, 1, 0, ctx->stream>>>(ctx->visc_dt_bits.p, kCfl[rk_order][ctx->cfg.degree], dt_guess,
        ctx->cfg.scheme == MCFD_FR_LOBATTO, dt_dev);
    MCFD_CUDA(ctx, cudaGetLastError());
    ctx->launches += 1;
    return MCFD_OK;
  }
  if (ctx->cfg.scheme == MCFD_DG_GENERAL) {
    if (int s = LaunchDgmTaylor(ctx, ctx->u[ctx->cur].p, 0, ctx->mesh.n_local)) return s;
    TimeStepResetKernel<<<1, 1, 0, ctx->stream>>>(ctx->dt_bits.p);
    const int nl = (int)ctx->mesh.n_local, tiles = (nl + kTile - 1) / kTile;
    if (ctx->q == 3) DgmTimeStepKernel<3><<<tiles, kTile, 0, ctx->stream>>>(ctx->tay.p, ctx->geom.p, ctx->length.p, nl, ctx->dt_bits.p);
    else DgmTimeStepKernel<4><<<tiles, kTile, 0, ctx->stream>>>(ctx->tay.p, ctx->geom.p, ctx->length.p, nl, ctx->dt_bits.p);
    TimeStepFinishKernel<<<1, 1, 0, ctx->stream>>>(ctx->dt_bits.p, kCfl[rk_order][ctx->cfg.degree], dt_guess, 0, dt_dev);
    MCFD_CUDA(ctx, cudaGetLastError());
    ctx->launches += 3;
    return MCFD_OK;
  }
  TimeStepResetKernel<<<1, 1, 0, ctx->stream>>>(ctx->dt_bits.p);
  int n = (int)ctx->mesh.n_local, threads = 256, blocks = (n + kTile - 1) / kTile;
  if (ctx->q == 3) TimeStepKernel<3><<<blocks, threads, 0, ctx->stream>>>(ctx->u[ctx->cur].p, ctx->length.p, n, ctx->dt_bits.p);
  else TimeStepKernel<4><<<blocks, threads, 0, ctx->stream>>>(ctx->u[ctx->cur].p, ctx->length.p, n, ctx->dt_bits.p);
  TimeStepFinishKernel<<<1, 1, 0, ctx->stream>>>(ctx->dt_bits.p, kCfl[rk_order][ctx->cfg.degree], dt_guess,
      ctx->cfg.scheme == MCFD_FR_LOBATTO, dt_dev);
  MCFD_CUDA(ctx, cudaGetLastError());
  ctx->launches += 3;
  return MCFD_OK;
}

int mcfd_time_step(mcfd_ctx *ctx, double dt_guess, int rk_order, double *dt_local) {
  if (!dt_local) return Fail(ctx, MCFD_ERR_ARG, "bad arguments");
  if (int s = mcfd_time_step_device(ctx, dt_guess, rk_order, nullptr)) return s;
  MCFD_CUDA(ctx, cudaMemcpyAsync(dt_local, ctx->dt_dev.p, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  MCFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return MCFD_OK;
}

int mcfd_rk_step_device(mcfd_ctx *ctx, int order, double t, const double *dt_dev) {
  if (int s = mcfd_rk_begin_device(ctx, order, t, dt_dev)) return s;
  for (int stage = 0; stage < order; ++stage) {
    if (int s = mcfd_rk_stage_inner(ctx, stage)) { ctx->rk_order = 0; return s; }
    if (int s = mcfd_rk_stage_inter(ctx, stage)) { ctx->rk_order = 0; return s; }
  }
  return MCFD_OK;
}

// ---- halo and time-step minimum between parts on the GPUs of one node (peer.cuh) ----------------------------------

int mcfd_peer_export(mcfd_ctx *ctx, int32_t world, mcfd_ipc_handle *handles, void **pointers) {
  if (int s = RequireMesh(ctx)) return s;
  if (world < 1 || world > kPeerMaxRanks) return Fail(ctx, MCFD_ERR_ARG, "world size must be 1..64");
  auto &pe = ctx->peer;
  pe.world = world;
  const size_t n_recv = ctx->recv_cells.n, n_send = ctx->send_cells.n;
  MCFD_CUDA(ctx, pe.flags.Alloc(kPeerWords));
  MCFD_CUDA(ctx, cudaMemset(pe.flags.p, 0, kPeerWords * sizeof(unsigned long long)));
  MCFD_CUDA(ctx, pe.counts.Alloc(kPeerCounts));
  MCFD_CUDA(ctx, cudaMemset(pe.counts.p, 0, kPeerCounts * sizeof(unsigned long long)));
  MCFD_CUDA(ctx, pe.dt_slots.Alloc(2 * (size_t)world));
  MCFD_CUDA(ctx, cudaMemset(pe.dt_slots.p, 0, 2 * (size_t)world * sizeof(double)));
  MCFD_CUDA(ctx, pe.recv_stage.Alloc(std::max<size_t>(1, kPeerSlots * n_recv * ctx->fields)));
  MCFD_CUDA(ctx, pe.send_stage.Alloc(std::max<size_t>(1, kPeerSlots * n_send * ctx->fields)));
  MCFD_CUDA(ctx, cudaMemset(pe.recv_stage.p, 0, pe.recv_stage.n * sizeof(double)));
  pe.pushed_once = false;
  pe.last_slot = -1; pe.last_plain = false;
  pe.connected = false;
  pe.deliveries_this_step = 0;
  void *ptrs[3] = {pe.flags.p, pe.dt_slots.p, pe.recv_stage.p};
  for (int i = 0; i < 3; ++i) {
    if (pointers) pointers[i] = ptrs[i];
    if (handles) {
      static_assert(sizeof(cudaIpcMemHandle_t) == sizeof(mcfd_ipc_handle), "handle size");
      cudaIpcMemHandle_t h;
      MCFD_CUDA(ctx, cudaIpcGetMemHandle(&h, ptrs[i]));
      std::memcpy(handles[i].bytes, &h, sizeof(h));
    }
  }
  return MCFD_OK;
}

int mcfd_peer_open(mcfd_ctx *ctx, const mcfd_ipc_handle *handle, void **pointer) {
  if (!ctx || !handle || !pointer) return Fail(ctx, MCFD_ERR_ARG, "null argument");
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle->bytes, sizeof(h));
  void *p = nullptr;
  MCFD_CUDA(ctx, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  ctx->peer.opened.push_back(p);
  *pointer = p;
  return MCFD_OK;
}

int mcfd_peer_connect(mcfd_ctx *ctx, int32_t rank, int32_t world, int32_t n_neighbours, void *const *neighbour_flags,
    const int32_t *slot_at_neighbour, void *const *neighbour_recv, const int64_t *neighbour_recv_items,
    const int64_t *offset_at_neighbour, void *const *world_flags, void *const *world_dt_slots) {
  if (int s = RequireMesh(ctx)) return s;
  auto &pe = ctx->peer;
  if (!pe.flags.p || world != pe.world) return Fail(ctx, MCFD_ERR_STATE, "call mcfd_peer_export with this world size first");
  if (rank < 0 || rank >= world || n_neighbours < 0 || n_neighbours > kPeerMaxNeighbours ||
      n_neighbours != (int)ctx->send_counts.size())
    return Fail(ctx, MCFD_ERR_ARG, "rank / neighbour count does not match mcfd_set_halo");
  if (!world_flags || !world_dt_slots || (n_neighbours && (!neighbour_flags || !slot_at_neighbour || !neighbour_recv ||
                                                            !neighbour_recv_items || !offset_at_neighbour)))
    return Fail(ctx, MCFD_ERR_ARG, "null argument");
  const size_t n_send = ctx->send_cells.n, n_recv = ctx->recv_cells.n;
  const int64_t first = ctx->first_inter, n_local = ctx->mesh.n_local;
  // the halo lists seen from the cells: every local cell's items of the send stage, every ghost cell's item of the
  // receive stage
  std::vector<int> send_host(n_send), recv_host(n_recv);
  if (n_send) MCFD_CUDA(ctx, cudaMemcpy(send_host.data(), ctx->send_cells.p, n_send * sizeof(int), cudaMemcpyDeviceToHost));
  if (n_recv) MCFD_CUDA(ctx, cudaMemcpy(recv_host.data(), ctx->recv_cells.p, n_recv * sizeof(int), cudaMemcpyDeviceToHost));
  std::vector<int> off(n_local - first + 2, 0);
  for (size_t k = 0; k < n_send; ++k) {
    if (send_host[k] < first) return Fail(ctx, MCFD_ERR_STATE, "a send cell precedes the first inter cell");
    ++off[send_host[k] - first + 1];
  }
  for (size_t c = 1; c < off.size(); ++c) off[c] += off[c - 1];
  std::vector<int> items(std::max<size_t>(n_send, 1)), fill(off.begin(), off.end() - 1);
  for (size_t k = 0; k < n_send; ++k) items[fill[send_host[k] - first]++] = (int)k;
  std::vector<int> ridx(std::max<int64_t>(ctx->mesh.n_ghost, 1), 0);
  for (size_t k = 0; k < n_recv; ++k) ridx[recv_host[k] - n_local] = (int)k;
  MCFD_CUDA(ctx, pe.tgt_offset.Alloc(off.size()));
  MCFD_CUDA(ctx, pe.tgt_item.Alloc(items.size()));
  MCFD_CUDA(ctx, pe.recv_index.Alloc(ridx.size()));
  MCFD_CUDA(ctx, cudaMemcpy(pe.tgt_offset.p, off.data(), off.size() * sizeof(int), cudaMemcpyHostToDevice));
  MCFD_CUDA(ctx, cudaMemcpy(pe.tgt_item.p, items.data(), items.size() * sizeof(int), cudaMemcpyHostToDevice));
  MCFD_CUDA(ctx, cudaMemcpy(pe.recv_index.p, ridx.data(), ridx.size() * sizeof(int), cudaMemcpyHostToDevice));
  pe.nbr_recv.assign(n_neighbours, nullptr);
  pe.nbr_recv_total.assign(n_neighbours, 0);
  pe.send_offset.assign(n_neighbours + 1, 0);
  std::vector<unsigned long long *> fl(std::max(n_neighbours, 1), nullptr);
  for (int i = 0; i < n_neighbours; ++i) {
    if (slot_at_neighbour[i] < 0 || slot_at_neighbour[i] >= kPeerMaxNeighbours) return Fail(ctx, MCFD_ERR_ARG, "bad neighbour slot");
    if (offset_at_neighbour[i] < 0 || offset_at_neighbour[i] + ctx->send_counts[i] > neighbour_recv_items[i])
      return Fail(ctx, MCFD_ERR_ARG, "this part's segment does not fit the neighbour's receive stage");
    fl[i] = static_cast<unsigned long long *>(neighbour_flags[i]) + slot_at_neighbour[i];
    pe.nbr_recv[i] = static_cast<double *>(neighbour_recv[i]) + offset_at_neighbour[i];
    pe.nbr_recv_total[i] = neighbour_recv_items[i];
    pe.send_offset[i + 1] = pe.send_offset[i] + ctx->send_counts[i];
  }
  std::vector<double *> ds(world);
  std::vector<unsigned long long *> df(world);
  for (int r = 0; r < world; ++r) {
    ds[r] = static_cast<double *>(world_dt_slots[r]);
    df[r] = static_cast<unsigned long long *>(world_flags[r]) + kPeerMaxNeighbours + rank;
  }
  MCFD_CUDA(ctx, pe.flag_dev.Alloc(fl.size()));
  MCFD_CUDA(ctx, pe.dt_slot_dev.Alloc(world));
  MCFD_CUDA(ctx, pe.dt_flag_dev.Alloc(world));
  MCFD_CUDA(ctx, pe.error.Alloc(1));
  MCFD_CUDA(ctx, cudaMemset(pe.error.p, 0, sizeof(int)));
  MCFD_CUDA(ctx, cudaMemcpy(pe.flag_dev.p, fl.data(), fl.size() * sizeof(void *), cudaMemcpyHostToDevice));
  MCFD_CUDA(ctx, cudaMemcpy(pe.dt_slot_dev.p, ds.data(), world * sizeof(double *), cudaMemcpyHostToDevice));
  MCFD_CUDA(ctx, cudaMemcpy(pe.dt_flag_dev.p, df.data(), world * sizeof(void *), cudaMemcpyHostToDevice));
  if (!ctx->dt_dev.p) MCFD_CUDA(ctx, ctx->dt_dev.Alloc(1));   // not inside a step: cudaMalloc may wait for kernels that poll
  if (!pe.side) {
    int lo = 0, hi = 0;
    MCFD_CUDA(ctx, cudaDeviceGetStreamPriorityRange(&lo, &hi));
    MCFD_CUDA(ctx, cudaStreamCreateWithPriority(&pe.side, cudaStreamNonBlocking, hi));
    MCFD_CUDA(ctx, cudaEventCreateWithFlags(&pe.ev_main, cudaEventDisableTiming));
    MCFD_CUDA(ctx, cudaEventCreateWithFlags(&pe.ev_side, cudaEventDisableTiming));
    for (auto &e : pe.ev_copy) MCFD_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  pe.rank = rank;
  pe.n_neighbours = n_neighbours;
  pe.connected = true;
  return MCFD_OK;
}

// id of the stream capture `stream` is part of, 0 when it is not capturing
static unsigned long long CaptureId(cudaStream_t stream) {
  cudaStreamCaptureStatus status = cudaStreamCaptureStatusNone;
  unsigned long long id = 0;
  if (cudaStreamGetCaptureInfo(stream, &status, &id) != cudaSuccess || status != cudaStreamCaptureStatusActive) return 0;
  return id;
}

// the node-line stage kernel packs its send cells and waits for its ghost states itself
static bool PeerIsFused(const mcfd_ctx *ctx) {
  return ctx->use_lines && ctx->peer.n_neighbours > 0 && ctx->first_inter < ctx->mesh.n_local;
}

static int RequirePeer(mcfd_ctx *ctx) {
  if (int s = RequireMesh(ctx)) return s;
  if (!ctx->peer.connected) return Fail(ctx, MCFD_ERR_STATE, "no peer connection (mcfd_peer_connect)");
  return MCFD_OK;
}

static double *PeerStage3(DevBuf<double> &b, int slot) { return b.p + (size_t)slot * (b.n / kPeerSlots); }

// steps 2 and 3 of a delivery (peer.cuh): the copies of send stage `buf` (a slot) into the neighbours' receive stages
// of the same slot, then the signal.  add: the host's part of the count; bump_plain: this is a delivery by mcfd_peer_push.
static int PeerCopyAndSignal(mcfd_ctx *ctx, cudaStream_t stream, int buf, int bump_plain, int add) {
  auto &pe = ctx->peer;
  const size_t n_send = ctx->send_cells.n;
  const double *from = PeerStage3(pe.send_stage, buf);
  for (int i = 0; i < pe.n_neighbours; ++i) {
    const size_t count = (size_t)ctx->send_counts[i];
    if (!count) continue;
    double *to = pe.nbr_recv[i] + (size_t)buf * (size_t)pe.nbr_recv_total[i] * ctx->fields;
    MCFD_CUDA(ctx, cudaMemcpy2DAsync(to, (size_t)pe.nbr_recv_total[i] * sizeof(double), from + pe.send_offset[i],
        n_send * sizeof(double), count * sizeof(double), (size_t)ctx->fields, cudaMemcpyDeviceToDevice, stream));
  }
  PeerSignalKernel<<<1, kPeerMaxNeighbours, 0, stream>>>(pe.flag_dev.p, pe.n_neighbours, pe.counts.p, bump_plain, add);
  MCFD_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  pe.last_slot = buf;
  pe.last_plain = bump_plain != 0;
  return MCFD_OK;
}

static int LaunchPeerPack(mcfd_ctx *ctx, cudaStream_t stream, int buf, int slot) {
  auto &pe = ctx->peer;
  const int n = (int)ctx->send_cells.n;
  if (!n) return MCFD_OK;
  const long long work = (long long)n * ctx->fields;
  PeerPackKernel<<<(int)std::min<long long>(1184, (work + 255) / 256), 256, 0, stream>>>(ctx->u[buf].p, ctx->fields,
      ctx->send_cells.p, n, PeerStage3(pe.send_stage, slot));
  MCFD_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  return MCFD_OK;
}
static int LaunchPeerUnpack(mcfd_ctx *ctx, cudaStream_t stream, int buf, int slot) {
  auto &pe = ctx->peer;
  const int n = (int)ctx->recv_cells.n;
  if (!n) return MCFD_OK;
  const long long work = (long long)n * ctx->fields;
  PeerUnpackKernel<<<(int)std::min<long long>(1184, (work + 255) / 256), 256, 0, stream>>>(ctx->u[buf].p, ctx->fields,
      ctx->recv_cells.p, n, PeerStage3(pe.recv_stage, slot));
  MCFD_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  return MCFD_OK;
}

int mcfd_peer_push(mcfd_ctx *ctx) {
  if (int s = RequirePeer(ctx)) return s;
  auto &pe = ctx->peer;
  pe.pushed_once = true;
  if (pe.n_neighbours == 0) return MCFD_OK;
  const int buf = ctx->rk_order ? ctx->rk_in : ctx->cur;
  // A second push of the same buffer in a row (the limited state after the unlimited one, a new solution written over
  // the old one) must not land on the slot the neighbour may still be unpacking: it takes the spare slot.  The slot
  // before that one is free again, because this rank's wait in between needed the neighbour's later push, which that
  // neighbour enqueued behind its unpack.
  const int slot = (pe.last_plain && pe.last_slot == buf) ? kPeerSpareSlot : buf;
  if (int s = LaunchPeerPack(ctx, ctx->stream, buf, slot)) return s;
  // the count goes on from the stage deliveries of the last step, which the device adds up at the next step only
  return PeerCopyAndSignal(ctx, ctx->stream, slot, 1, pe.deliveries_this_step);
}

int mcfd_peer_wait(mcfd_ctx *ctx) {
  if (int s = RequirePeer(ctx)) return s;
  auto &pe = ctx->peer;
  if (pe.n_neighbours == 0) return MCFD_OK;
  if (!pe.pushed_once) return Fail(ctx, MCFD_ERR_STATE, "nothing has been pushed yet (mcfd_peer_push)");
  const int buf = ctx->rk_order ? ctx->rk_in : ctx->cur;
  PeerWaitKernel<<<1, kPeerMaxNeighbours, 0, ctx->stream>>>(pe.flags.p, pe.n_neighbours, pe.counts.p,
      pe.deliveries_this_step, pe.error.p);
  MCFD_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  return LaunchPeerUnpack(ctx, ctx->stream, buf, pe.last_slot);
}

static int PeerDtPublish(mcfd_ctx *ctx, double *dt_dev) {
  auto &pe = ctx->peer;
  PeerDtPublishKernel<<<1, kPeerMaxRanks, 0, ctx->stream>>>(dt_dev, pe.rank, pe.world, pe.dt_slot_dev.p,
      pe.dt_flag_dev.p, pe.counts.p, pe.deliveries_this_step);
  pe.deliveries_this_step = 0;
  MCFD_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  return MCFD_OK;
}
static int PeerDtReduce(mcfd_ctx *ctx, double *dt_dev) {
  auto &pe = ctx->peer;
  PeerDtReduceKernel<<<1, kPeerMaxRanks, 0, ctx->stream>>>(pe.flags.p + kPeerMaxNeighbours, pe.dt_slots.p, pe.world,
      pe.counts.p, dt_dev, pe.error.p);
  MCFD_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  return MCFD_OK;
}

int mcfd_peer_min_time_step(mcfd_ctx *ctx, double *dt_dev) {
  if (int s = RequirePeer(ctx)) return s;
  if (!dt_dev) dt_dev = ctx->dt_dev.p;
  if (int s = PeerDtPublish(ctx, dt_dev)) return s;
  return PeerDtReduce(ctx, dt_dev);
}

// One stage of an open RK step with the stage kernels that read ghost ROWS: the inner cells on the caller's stream
// (fem.hpp:249-254) while the high-priority side stream waits for the delivery, unpacks it, advances the inter cells
// (fem.hpp:255-256) and at once delivers their new values.
static int PeerStage(mcfd_ctx *ctx, int stage) {
  auto &pe = ctx->peer;
  cudaStream_t main = ctx->stream;
  MCFD_CUDA(ctx, cudaEventRecord(pe.ev_main, main));
  MCFD_CUDA(ctx, cudaStreamWaitEvent(pe.side, pe.ev_main, 0));
  if (int s = mcfd_rk_stage_inner(ctx, stage)) return s;
  ctx->stream = pe.side;
  int status = mcfd_peer_wait(ctx);
  if (status == MCFD_OK) status = mcfd_rk_stage_inter(ctx, stage);
  const int buf = ctx->rk_order ? ctx->rk_in : ctx->cur;        // the stage's output
  if (status == MCFD_OK) status = LaunchPeerPack(ctx, pe.side, buf, buf);
  if (status == MCFD_OK) status = PeerCopyAndSignal(ctx, pe.side, buf, 0, ++pe.deliveries_this_step);
  ctx->stream = main;
  if (status != MCFD_OK) return status;
  MCFD_CUDA(ctx, cudaEventRecord(pe.ev_side, pe.side));
  MCFD_CUDA(ctx, cudaStreamWaitEvent(main, pe.ev_side, 0));
  return MCFD_OK;
}

// The same stage with the node-line kernel: ONE launch over all blocks on the caller's stream (StageArgs::PeerStage:
// its inter-cell blocks, the last ones, wait for the delivery themselves, read the ghost states from the receive
// stage and leave a packed copy of the send cells), and behind it, on the side stream, the copies and the signal.
static int PeerFusedStage(mcfd_ctx *ctx, int stage) {
  auto &pe = ctx->peer;
  if (!ctx->rk_order || stage < 0 || stage >= ctx->rk_order) return Fail(ctx, MCFD_ERR_STATE, "no such stage in the open RK step");
  int in, out, mode;
  StageBuffers(ctx, stage, &in, &out, &mode);
  if (in != ctx->rk_in) return Fail(ctx, MCFD_ERR_STATE, "stages must be run in order");
  if (ctx->time_cb) ctx->time_cb(ctx->time_user, mcfd_rk_stage_time(ctx, stage));
  StageArgs::PeerStage f{};
  f.on = 1;
  f.first_block = (int)(ctx->first_inter / ctx->cpb);
  f.first_cell = (int)ctx->first_inter;
  f.n_neighbours = pe.n_neighbours;
  f.wait_add = pe.deliveries_this_step;
  f.n_send = (int)ctx->send_cells.n; f.n_recv = (int)ctx->recv_cells.n;
  f.flags = pe.flags.p; f.counts = pe.counts.p;
  f.recv = PeerStage3(pe.recv_stage, pe.last_slot); f.recv_index = pe.recv_index.p;   // `in`, or the spare slot after a push
  f.send = PeerStage3(pe.send_stage, out); f.tgt_offset = pe.tgt_offset.p; f.tgt_item = pe.tgt_item.p;
  f.error = pe.error.p;
  cudaStream_t main = ctx->stream;
  // The copies that last read this send stage, two stages ago.  An event recorded outside the capture under way (or in
  // one that is over) cannot be waited for and need not be: graph launches and plain work on a stream do not overlap.
  if (pe.ev_copy_capture[out] == CaptureId(main)) MCFD_CUDA(ctx, cudaStreamWaitEvent(main, pe.ev_copy[out], 0));
  if (int s = LaunchStage(ctx, ctx->u[in].p, ctx->u[ctx->rk_u0].p, ctx->u[out].p, mode, ctx->rk_dt, 0, ctx->total_blocks, &f))
    return s;
  MCFD_CUDA(ctx, cudaEventRecord(pe.ev_main, main));
  MCFD_CUDA(ctx, cudaStreamWaitEvent(pe.side, pe.ev_main, 0));
  if (int s = PeerCopyAndSignal(ctx, pe.side, out, 0, ++pe.deliveries_this_step)) return s;
  MCFD_CUDA(ctx, cudaEventRecord(pe.ev_copy[out], pe.side));
  pe.ev_copy_capture[out] = CaptureId(main);
  ctx->rk_in = out;
  if (stage == ctx->rk_order - 1) { ctx->cur = out; ctx->rk_order = 0; }
  return MCFD_OK;
}

// Several contexts driven by this thread (one per GPU, or several on one GPU) advance in lockstep: every delivery is
// enqueued before any wait that depends on it, whatever streams and hardware queues the contexts share.
int mcfd_peer_step_group(mcfd_ctx *const *ctxs, int32_t n, int order, double t, double dt_guess) {
  if (!ctxs || n < 1) return MCFD_ERR_ARG;
  for (int c = 0; c < n; ++c) {
    if (int s = RequirePeer(ctxs[c])) return s;
    if (ctxs[c]->cfg.diffusion != MCFD_INVISCID || ctxs[c]->cfg.scheme == MCFD_DG_GENERAL)
      return Fail(ctxs[c], MCFD_ERR_ARG, "mcfd_peer_step is built for the inviscid nodal schemes");
  }
  int dev0 = 0;
  cudaGetDevice(&dev0);
  auto on = [&](mcfd_ctx *ctx) { if (ctx->cfg.device >= 0) cudaSetDevice(ctx->cfg.device); };
  int status = MCFD_OK;
  for (int c = 0; c < n && !status; ++c) {   // the ghost images of a solution that has never been sent
    on(ctxs[c]);
    if (!ctxs[c]->peer.pushed_once) status = mcfd_peer_push(ctxs[c]);
  }
  for (int c = 0; c < n && !status; ++c) {
    on(ctxs[c]);
    status = mcfd_time_step_device(ctxs[c], dt_guess, order, nullptr);
    if (!status) status = PeerDtPublish(ctxs[c], ctxs[c]->dt_dev.p);
  }
  for (int c = 0; c < n && !status; ++c) {
    on(ctxs[c]);
    status = PeerDtReduce(ctxs[c], ctxs[c]->dt_dev.p);
    if (!status) status = mcfd_rk_begin_device(ctxs[c], order, t, ctxs[c]->dt_dev.p);
  }
  for (int stage = 0; stage < order && !status; ++stage)
    for (int c = 0; c < n && !status; ++c) {
      on(ctxs[c]);
      status = PeerIsFused(ctxs[c]) ? PeerFusedStage(ctxs[c], stage) : PeerStage(ctxs[c], stage);
    }
  if (status) for (int c = 0; c < n; ++c) ctxs[c]->rk_order = 0;
  cudaSetDevice(dev0);
  return status;
}

int mcfd_peer_step(mcfd_ctx *ctx, int order, double t, double dt_guess) {
  return mcfd_peer_step_group(&ctx, 1, order, t, dt_guess);
}

int mcfd_peer_join(mcfd_ctx *ctx) {
  if (int s = RequirePeer(ctx)) return s;
  auto &pe = ctx->peer;
  MCFD_CUDA(ctx, cudaEventRecord(pe.ev_side, pe.side));
  MCFD_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, pe.ev_side, 0));
  return MCFD_OK;
}

int mcfd_peer_check(mcfd_ctx *ctx) {
  if (int s = RequirePeer(ctx)) return s;
  MCFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  MCFD_CUDA(ctx, cudaStreamSynchronize(ctx->peer.side));
  int err = 0;
  MCFD_CUDA(ctx, cudaMemcpy(&err, ctx->peer.error.p, sizeof(int), cudaMemcpyDeviceToHost));
  if (err) return Fail(ctx, MCFD_ERR_STATE, "a neighbour's delivery did not arrive within the polling limit");
  return MCFD_OK;
}

int mcfd_limiter_reconstruct(mcfd_ctx *ctx, int kind) {
  if (int s = RequireMesh(ctx)) return s;
  if (kind != 0 && kind != 1) return Fail(ctx, MCFD_ERR_ARG, "limiter kind must be 0 (average) or 1 (majority)");
  if (ctx->cfg.k_local) return Fail(ctx, MCFD_ERR_STATE, "the limiters are built for kLocal == false");
  if (ctx->cfg.scheme == MCFD_DG_GENERAL) return Fail(ctx, MCFD_ERR_STATE, "the limiters are built for the nodal expansions");
  if (ctx->rk_order) return Fail(ctx, MCFD_ERR_STATE, "an RK step is open");
  int n = (int)ctx->mesh.n_local, threads = 128, blocks = (n + threads - 1) / threads;
  MCFD_CUDA(ctx, EnsureTables(ctx));
  if (ctx->q == 3) CellConstantKernel<3><<<blocks, threads, 0, ctx->stream>>>(ctx->u[ctx->cur].p, ctx->geom.p, n, kind);
  else CellConstantKernel<4><<<blocks, threads, 0, ctx->stream>>>(ctx->u[ctx->cur].p, ctx->geom.p, n, kind);
  MCFD_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  return MCFD_OK;
}

}  // extern "C"

namespace {

constexpr int kWenoEigenCtas = 296;   // persistent CTAs of the weno::Eigen kernel, one scratch slot each

template <int Q>
int WenoLazyBatch(mcfd_ctx *ctx, double *u, double w0, double eps, int64_t modal_begin, int64_t modal_end,
    int64_t cell_begin, int64_t cell_end, bool publish, bool eigen) {
  constexpr int M = ModalShape<Q - 1>::M;
  const size_t smem = WenoShape<Q>::kSharedBytes;
  static DeviceFlags configured;
  if (!configured.Here()) {
    MCFD_CUDA(ctx, cudaFuncSetAttribute(ModalKernel<Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    MCFD_CUDA(ctx, cudaFuncSetAttribute(WenoLazyKernel<Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    MCFD_CUDA(ctx, cudaFuncSetAttribute(WenoEigenKernel<Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // several CTAs per SM only fit when the L1/shared split is asked for the shared side
    MCFD_CUDA(ctx, cudaFuncSetAttribute(ModalKernel<Q>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    MCFD_CUDA(ctx, cudaFuncSetAttribute(WenoLazyKernel<Q>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    configured.Set();
  }
  MCFD_CUDA(ctx, EnsureTables(ctx));
  if (modal_end > modal_begin) {   // Extrapolation::UpdateModalCoeff of the cells whose nodal values are new
    const int t0 = (int)(modal_begin / kTile), t1 = (int)((modal_end + kTile - 1) / kTile);
    ModalKernel<Q><<<t1 - t0, 160, WenoShape<Q>::kLazySharedBytes, ctx->stream>>>(u, ctx->geom.p, ctx->bas.p, ctx->mod.p, ctx->tay.p, t0,
        (int)modal_begin, (int)modal_end);
    ctx->launches++;
  }
  if (cell_end > cell_begin) {
    WenoArgs a;
    a.u = u; a.geom = ctx->geom.p; a.bas = ctx->bas.p; a.h2s = ctx->h2s.p; a.nbr = ctx->nbr.p; a.code = ctx->code.p;
    a.mod = ctx->mod.p; a.tay = ctx->tay.p; a.mod_new = ctx->mod_new.p; a.tay_new = ctx->tay_new.p;
    a.cell_begin = (int)cell_begin; a.cell_end = (int)cell_end; a.w0 = w0; a.eps = eps;
    a.comp_stride = M; a.mode_stride = 1; a.modal_state = 0; a.legendre = 0;
    const int t0 = (int)(cell_begin / kTile), t1 = (int)((cell_end + kTile - 1) / kTile);
    if (eigen) {
      WenoEigenArgs e;
      e.w = a; e.frames = ctx->frames.p; e.cand = nullptr; e.blend = nullptr;
      const int grid = std::min(t1 - t0, kWenoEigenCtas);
      WenoEigenKernel<Q><<<grid, 160, smem, ctx->stream>>>(e, ctx->weno_scratch.p, t0, t1);
    } else {
      WenoLazyKernel<Q><<<t1 - t0, 160, WenoShape<Q>::kLazySharedBytes, ctx->stream>>>(a, t0);
    }
    ctx->launches++;
    if (publish) {   // the next batch borrows the limited polynomials of this one (limiter/reconstruct.hpp:41-51)
      const int n = (int)(cell_end - cell_begin), threads = 128, blocks = (n + threads - 1) / threads;
      CopyCellRowsKernel<<<blocks, threads, 0, ctx->stream>>>(ctx->tay_new.p, ctx->tay.p, 5 * M, (int)cell_begin, (int)cell_end);
      CopyCellRowsKernel<<<blocks, threads, 0, ctx->stream>>>(ctx->mod_new.p, ctx->mod.p, 5 * M, (int)cell_begin, (int)cell_end);
      ctx->launches += 2;
    }
  }
  MCFD_CUDA(ctx, cudaGetLastError());
  return MCFD_OK;
}

// weno::Lazy on Projection cells (dg::General): the modal state is limited in place; its Taylor form is all the
// neighbours read, so a batch is: Taylor form of the cells whose coefficients are new -> blend
template <int Q>
int DgmWenoBatch(mcfd_ctx *ctx, double *u, double w0, double eps, int64_t taylor_begin, int64_t taylor_end,
    int64_t cell_begin, int64_t cell_end) {
  constexpr int M = ModalShape<Q - 1>::M;
  const size_t smem = WenoShape<Q>::kSharedBytes;
  static DeviceFlags configured;
  if (!configured.Here()) {
    MCFD_CUDA(ctx, cudaFuncSetAttribute(WenoLazyKernel<Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    MCFD_CUDA(ctx, cudaFuncSetAttribute(WenoLazyKernel<Q>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    configured.Set();
  }
  if (int s = LaunchDgmTaylor(ctx, u, taylor_begin, taylor_end)) return s;
  if (cell_end <= cell_begin) return MCFD_OK;
  WenoArgs a;
  a.u = u; a.geom = ctx->geom.p; a.bas = ctx->bas.p; a.h2s = ctx->h2s.p; a.nbr = ctx->nbr.p; a.code = ctx->code.p;
  a.mod = u; a.tay = ctx->tay.p; a.mod_new = nullptr; a.tay_new = nullptr;
  a.cell_begin = (int)cell_begin; a.cell_end = (int)cell_end; a.w0 = w0; a.eps = eps;
  a.comp_stride = 1; a.mode_stride = 5; a.modal_state = 1; a.legendre = 1;
  (void)M;
  const int t0 = (int)(cell_begin / kTile), t1 = (int)((cell_end + kTile - 1) / kTile);
  WenoLazyKernel<Q><<<t1 - t0, 160, WenoShape<Q>::kLazySharedBytes, ctx->stream>>>(a, t0);
  MCFD_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  return MCFD_OK;
}

}  // namespace

extern "C" {

static int LimiterWeno(mcfd_ctx *ctx, double w0, double eps, int phase, bool eigen);

int mcfd_limiter_weno_lazy(mcfd_ctx *ctx, double w0, double eps, int phase) {
  return LimiterWeno(ctx, w0, eps, phase, false);
}

int mcfd_limiter_weno_eigen(mcfd_ctx *ctx, double w0, double eps, int phase) {
  return LimiterWeno(ctx, w0, eps, phase, true);
}

int mcfd_eigen_matrices(mcfd_ctx *ctx, int64_t n, const double *conservative, const double *frames, double *left,
    double *right) {
  if (!ctx || n <= 0 || !conservative || !frames || !left || !right) return MCFD_ERR_ARG;
  DevBuf<double> du, df, dl, dr;
  MCFD_CUDA(ctx, du.Alloc(5 * n));
  MCFD_CUDA(ctx, df.Alloc(9 * n));
  MCFD_CUDA(ctx, dl.Alloc(25 * n));
  MCFD_CUDA(ctx, dr.Alloc(25 * n));
  MCFD_CUDA(ctx, cudaMemcpy(du.p, conservative, 5 * n * sizeof(double), cudaMemcpyHostToDevice));
  MCFD_CUDA(ctx, cudaMemcpy(df.p, frames, 9 * n * sizeof(double), cudaMemcpyHostToDevice));
  EigenMatricesKernel<<<(int)((n + 127) / 128), 128, 0, ctx->stream>>>((int)n, du.p, df.p, dl.p, dr.p);
  MCFD_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  MCFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  MCFD_CUDA(ctx, cudaMemcpy(left, dl.p, 25 * n * sizeof(double), cudaMemcpyDeviceToHost));
  MCFD_CUDA(ctx, cudaMemcpy(right, dr.p, 25 * n * sizeof(double), cudaMemcpyDeviceToHost));
  return MCFD_OK;
}

// Extrapolation<Hexahedron> on a nodal context: per-cell orthonormal bases, built on the host at first use like the
// reference does at set-up, and the modal / Taylor coefficient tables
static int EnsureExtrapolation(mcfd_ctx *ctx) {
  if (ctx->bas.p) return MCFD_OK;
  const int m = ctx->q == 3 ? 10 : 20, rows = ctx->q == 3 ? ModalShape<2>::kBasRows : ModalShape<3>::kBasRows;
  auto bas = ToTiles(ctx->mesh.BuildModalBases(ctx->cfg.degree), rows, ctx->n_cells, ctx->n_tiles);
  MCFD_CUDA(ctx, ctx->bas.Alloc(bas.size()));
  MCFD_CUDA(ctx, cudaMemcpy(ctx->bas.p, bas.data(), bas.size() * sizeof(double), cudaMemcpyHostToDevice));
  const size_t n = (size_t)ctx->n_tiles * 5 * m * kTile;
  for (DevBuf<double> *b : {&ctx->mod, &ctx->tay, &ctx->mod_new, &ctx->tay_new}) {
    MCFD_CUDA(ctx, b->Alloc(n));
    MCFD_CUDA(ctx, cudaMemset(b->p, 0, n * sizeof(double)));
  }
  return MCFD_OK;
}

static int LimiterWeno(mcfd_ctx *ctx, double w0, double eps, int phase, bool eigen) {
  if (int s = RequireMesh(ctx)) return s;
  if (ctx->cfg.k_local) return Fail(ctx, MCFD_ERR_STATE, "the limiters are built for kLocal == false");
  if (ctx->cfg.scheme == MCFD_DG_GENERAL) {
    if (eigen) return Fail(ctx, MCFD_ERR_STATE, "weno::Eigen is built for the nodal expansions");
    if (phase < -1 || phase > 1) return Fail(ctx, MCFD_ERR_ARG, "phase must be -1 (both), 0 (inner cells) or 1 (inter cells)");
    double *u = HaloState(ctx);
    const int64_t nl = ctx->mesh.n_local, nc = ctx->n_cells, fi = ctx->mesh.n_ghost ? ctx->first_inter : nl;
    auto batch = [&](int64_t tb, int64_t te, int64_t cb, int64_t ce) {
      return ctx->q == 3 ? DgmWenoBatch<3>(ctx, u, w0, eps, tb, te, cb, ce) : DgmWenoBatch<4>(ctx, u, w0, eps, tb, te, cb, ce);
    };
    if (phase != 1) {
      if (int s = batch(0, nl, 0, fi)) return s;
    }
    if (phase != 0) {   // the limited inner cells' Taylor forms are refreshed together with the ghosts'
      if (int s = batch(0, fi, 0, 0)) return s;
      if (int s = batch(nl, nc, fi, nl)) return s;
    }
    return MCFD_OK;
  }
  if (phase < -1 || phase > 1) return Fail(ctx, MCFD_ERR_ARG, "phase must be -1 (both), 0 (inner cells) or 1 (inter cells)");
  if (int s = EnsureExtrapolation(ctx)) return s;
  if (eigen && !ctx->frames.p) {
    auto frames = ToTiles(ctx->mesh.frames, 54, ctx->n_cells, ctx->n_tiles);
    MCFD_CUDA(ctx, ctx->frames.Alloc(frames.size()));
    MCFD_CUDA(ctx, cudaMemcpy(ctx->frames.p, frames.data(), frames.size() * sizeof(double), cudaMemcpyHostToDevice));
    const size_t per_cta = ctx->q == 3 ? WenoEigenShape<3>::kScratchDoubles : WenoEigenShape<4>::kScratchDoubles;
    MCFD_CUDA(ctx, ctx->weno_scratch.Alloc(per_cta * kWenoEigenCtas));
  }
  double *u = HaloState(ctx);
  const int64_t nl = ctx->mesh.n_local, nc = ctx->n_cells, fi = ctx->mesh.n_ghost ? ctx->first_inter : nl;
  auto batch = [&](int64_t mb, int64_t me, int64_t cb, int64_t ce, bool publish) {
    return ctx->q == 3 ? WenoLazyBatch<3>(ctx, u, w0, eps, mb, me, cb, ce, publish, eigen)
                       : WenoLazyBatch<4>(ctx, u, w0, eps, mb, me, cb, ce, publish, eigen);
  };
  if (phase != 1) {   // act(inner cells): needs no ghost cell (limiter/reconstruct.hpp:46-48)
    if (int s = batch(0, nl, 0, fi, fi < nl)) return s;
  }
  if (phase != 0) {   // act(inter cells) after UpdateGhostCellCoeffs (limiter/reconstruct.hpp:49-51)
    if (int s = batch(nl, nc, fi, nl, false)) return s;
  }
  return MCFD_OK;
}

int mcfd_sample_vtk(mcfd_ctx *ctx, int32_t modal, double *points, double *values) {
  if (int s = RequireMesh(ctx)) return s;
  const int64_t nl = ctx->mesh.n_local;
  static const int vtk[27][3] = {
    {-1, -1, -1}, {1, -1, -1}, {1, 1, -1}, {-1, 1, -1}, {-1, -1, 1}, {1, -1, 1}, {1, 1, 1}, {-1, 1, 1},
    {0, -1, -1}, {1, 0, -1}, {0, 1, -1}, {-1, 0, -1}, {0, -1, 1}, {1, 0, 1}, {0, 1, 1}, {-1, 0, 1},
    {-1, -1, 0}, {1, -1, 0}, {1, 1, 0}, {-1, 1, 0},
    {-1, 0, 0}, {1, 0, 0}, {0, -1, 0}, {0, 1, 0}, {0, 0, -1}, {0, 0, 1}, {0, 0, 0}};
  if (points) {   // coordinate().LocalToGlobal(local): static geometry, from the host copy of the trilinear maps
    for (int64_t c = 0; c < nl; ++c) {
      const double *g = &ctx->mesh.coef[c * kGeomRows], *o = &ctx->mesh.c0[c * 3];
      for (int p = 0; p < 27; ++p) {
        const double x = vtk[p][0], e = vtk[p][1], z = vtk[p][2];
        for (int k = 0; k < 3; ++k) {
          points[(c * 27 + p) * 3 + k] = o[k] + g[0 + k] * x + g[3 + k] * e + g[6 + k] * z + g[9 + k] * x * e
              + g[12 + k] * e * z + g[15 + k] * z * x + g[18 + k] * x * e * z;
        }
      }
    }
  }
  if (nl == 0) return MCFD_OK;
  const bool dgm = ctx->cfg.scheme == MCFD_DG_GENERAL;
  if (dgm) modal = 1;
  if (modal && ctx->cfg.k_local) return Fail(ctx, MCFD_ERR_STATE, "the modal polynomials are built for kLocal == false");
  double *u = HaloState(ctx);
  VtkSampleArgs a{};
  a.u = u; a.geom = ctx->geom.p; a.n_local = (int)nl; a.modal = modal ? 1 : 0; a.k_local = ctx->cfg.k_local;
  const int q = ctx->q, m = q == 3 ? 10 : 20;
  if (dgm) {   // Projection: the state is the modal coefficients
    if (int s = LaunchDgmTaylor(ctx, u, 0, nl)) return s;
    a.tay = ctx->tay.p; a.comp_stride = 1; a.mode_stride = 5;
  } else if (modal) {   // Extrapolation: UpdateModalCoeff of the current nodal values first
    if (int s = EnsureExtrapolation(ctx)) return s;
    const int st = q == 3 ? WenoLazyBatch<3>(ctx, u, 0, 0, 0, nl, 0, 0, false, false)
                          : WenoLazyBatch<4>(ctx, u, 0, 0, 0, nl, 0, 0, false, false);
    if (st) return st;
    a.tay = ctx->tay.p; a.comp_stride = m; a.mode_stride = 1;
  }
  for (int r = 0; r < 3; ++r) {   // Lagrange basis on the solution nodes at -1, 0, +1 (basis/lagrange.hpp)
    const double x = r - 1;
    for (int i = 0; i < q; ++i) {
      double v = 1;
      for (int j = 0; j < q; ++j) if (j != i) v *= (x - ctx->tables.xi[j]) / (ctx->tables.xi[i] - ctx->tables.xi[j]);
      a.lag[r][i] = v;
    }
  }
  const size_t n = (size_t)nl * 27 * 5;
  if (ctx->vtk_values.n < n) MCFD_CUDA(ctx, ctx->vtk_values.Alloc(n));
  a.values = ctx->vtk_values.p;
  MCFD_CUDA(ctx, EnsureTables(ctx));
  const int tiles = (int)((nl + kTile - 1) / kTile);
  if (q == 3) VtkSampleKernel<3><<<tiles, 256, 0, ctx->stream>>>(a);
  else VtkSampleKernel<4><<<tiles, 256, 0, ctx->stream>>>(a);
  MCFD_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  if (values) {   // null: the samples stay on the device for mcfd_vtk_component_text
    MCFD_CUDA(ctx, cudaMemcpyAsync(values, ctx->vtk_values.p, n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    MCFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  return MCFD_OK;
}

// ---- text of a frame's arrays (format.cuh) -------------------------------------------------------------------------
static const unsigned int kUnsureCapacity = 1u << 16;

// values_dev[i * stride] (kind 0) or first + i * step (kind 1), i < n, as blank-separated tokens into text[capacity]
static int FormatAndPack(mcfd_ctx *ctx, int kind, const double *values_dev, int64_t stride, int64_t n, int64_t first,
    int64_t step, char *text, int64_t capacity, int64_t *length) {
  if (!text || !length || n < 0) return Fail(ctx, MCFD_ERR_ARG, "null text / length");
  *length = 0;
  if (n == 0) return MCFD_OK;
  if (n > 0x7fffffffll) return Fail(ctx, MCFD_ERR_ARG, "too many values for one array");
  if (!ctx->fmt_pow10.p) {   // 10^k as a double pair, from the host's long double arithmetic
    std::vector<double> t(2 * (kPow10Max - kPow10Min + 1));
    BuildPow10Table(t.data());
    MCFD_CUDA(ctx, ctx->fmt_pow10.Alloc(t.size()));
    MCFD_CUDA(ctx, cudaMemcpy(ctx->fmt_pow10.p, t.data(), t.size() * sizeof(double), cudaMemcpyHostToDevice));
    MCFD_CUDA(ctx, ctx->fmt_unsure.Alloc(1 + kUnsureCapacity));
  }
  const int n_blocks = (int)((n + kPackPerBlock - 1) / kPackPerBlock);
  if (ctx->fmt_slots.n < (size_t)n) {
    MCFD_CUDA(ctx, ctx->fmt_slots.Alloc(n));
    MCFD_CUDA(ctx, ctx->fmt_lens.Alloc(n));
    MCFD_CUDA(ctx, ctx->fmt_text.Alloc((size_t)n * kTokenSlot));
    MCFD_CUDA(ctx, ctx->fmt_sums.Alloc(n_blocks));
    MCFD_CUDA(ctx, ctx->fmt_offsets.Alloc(n_blocks + 1));
  }
  FormatArgs a{};
  a.values = values_dev; a.stride = stride; a.n = n; a.first = first; a.step = step; a.kind = kind;
  a.pow10 = ctx->fmt_pow10.p; a.slots = ctx->fmt_slots.p; a.lens = ctx->fmt_lens.p;
  a.unsure_count = ctx->fmt_unsure.p; a.unsure_index = ctx->fmt_unsure.p + 1; a.unsure_capacity = kUnsureCapacity;
  MCFD_CUDA(ctx, cudaMemsetAsync(ctx->fmt_unsure.p, 0, sizeof(unsigned int), ctx->stream));
  FormatTokensKernel<<<(int)((n + 255) / 256), 256, 0, ctx->stream>>>(a);
  MCFD_CUDA(ctx, cudaGetLastError());
  unsigned int n_unsure = 0;
  MCFD_CUDA(ctx, cudaMemcpyAsync(&n_unsure, ctx->fmt_unsure.p, sizeof n_unsure, cudaMemcpyDeviceToHost, ctx->stream));
  MCFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (n_unsure) {   // rounding too close to call on the device, or beyond its table: snprintf decides
    const bool all = n_unsure > kUnsureCapacity;   // pathological input: the host formats the whole array
    std::vector<unsigned int> index(all ? 0 : n_unsure);
    if (!all) MCFD_CUDA(ctx, cudaMemcpy(index.data(), ctx->fmt_unsure.p + 1, n_unsure * sizeof(unsigned int), cudaMemcpyDeviceToHost));
    std::vector<double> column;
    if (all) {
      column.resize((size_t)n);
      MCFD_CUDA(ctx, cudaMemcpy2D(column.data(), sizeof(double), values_dev, stride * sizeof(double), sizeof(double), n,
          cudaMemcpyDeviceToHost));
    }
    const size_t count = all ? (size_t)n : n_unsure;
    for (size_t j = 0; j < count; ++j) {
      const size_t i = all ? j : index[j];
      double v;
      if (all) v = column[i];
      else MCFD_CUDA(ctx, cudaMemcpy(&v, values_dev + i * stride, sizeof v, cudaMemcpyDeviceToHost));
      char token[32] = {0};
      const unsigned char len = (unsigned char)std::snprintf(token, sizeof token, "%g ", v);
      MCFD_CUDA(ctx, cudaMemcpy(ctx->fmt_slots.p + i, token, kTokenSlot, cudaMemcpyHostToDevice));
      MCFD_CUDA(ctx, cudaMemcpy(ctx->fmt_lens.p + i, &len, 1, cudaMemcpyHostToDevice));
    }
  }
  TokenBlockSumsKernel<<<n_blocks, 256, 0, ctx->stream>>>(ctx->fmt_lens.p, n, ctx->fmt_sums.p);
  TokenScanKernel<<<1, 1024, 0, ctx->stream>>>(ctx->fmt_sums.p, n_blocks, ctx->fmt_offsets.p);
  TokenPackKernel<<<n_blocks, 256, 0, ctx->stream>>>(ctx->fmt_slots.p, ctx->fmt_lens.p, n, ctx->fmt_offsets.p, ctx->fmt_text.p);
  MCFD_CUDA(ctx, cudaGetLastError());
  ctx->launches += 4;
  unsigned long long total = 0;
  MCFD_CUDA(ctx, cudaMemcpyAsync(&total, ctx->fmt_offsets.p + n_blocks, sizeof total, cudaMemcpyDeviceToHost, ctx->stream));
  MCFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *length = (int64_t)total;
  if ((int64_t)total > capacity) return Fail(ctx, MCFD_ERR_ARG, "text buffer too small (16 bytes per value always suffice)");
  MCFD_CUDA(ctx, cudaMemcpy(text, ctx->fmt_text.p, total, cudaMemcpyDeviceToHost));
  return MCFD_OK;
}

int mcfd_format_doubles(mcfd_ctx *ctx, int64_t n, const double *values, char *text, int64_t capacity, int64_t *length) {
  if (!ctx) return MCFD_ERR_ARG;
  if (n > 0 && !values) return Fail(ctx, MCFD_ERR_ARG, "null values");
  if (n > 0) {
    if (ctx->fmt_values.n < (size_t)n) MCFD_CUDA(ctx, ctx->fmt_values.Alloc(n));
    MCFD_CUDA(ctx, cudaMemcpyAsync(ctx->fmt_values.p, values, n * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  }
  return FormatAndPack(ctx, 0, ctx->fmt_values.p, 1, n, 0, 0, text, capacity, length);
}

int mcfd_format_integers(mcfd_ctx *ctx, int64_t n, int64_t first, int64_t step, char *text, int64_t capacity,
    int64_t *length) {
  if (!ctx) return MCFD_ERR_ARG;
  // a token slot holds 14 digits, a sign and the blank
  const long double last = (long double)first + (long double)step * (long double)(n > 0 ? n - 1 : 0);
  if (std::fabs((double)first) >= 1e14 || std::fabs((double)last) >= 1e14)
    return Fail(ctx, MCFD_ERR_ARG, "integers of 10^14 or more are not supported");
  return FormatAndPack(ctx, 1, nullptr, 0, n, first, step, text, capacity, length);
}

int mcfd_vtk_component_text(mcfd_ctx *ctx, int32_t component, char *text, int64_t capacity, int64_t *length) {
  if (int s = RequireMesh(ctx)) return s;
  if (component < 0 || component >= 5) return Fail(ctx, MCFD_ERR_ARG, "component must be 0..4");
  const int64_t n = ctx->mesh.n_local * 27;
  if (ctx->vtk_values.n < (size_t)n * 5) return Fail(ctx, MCFD_ERR_STATE, "call mcfd_sample_vtk first");
  return FormatAndPack(ctx, 0, ctx->vtk_values.p + component, 5, n, 0, 0, text, capacity, length);
}

// ---- the partition's index work (shuffle.cuh): context-free, host arrays in and out ---------------------------------
#define MCFD_CUDA0(what, call)                                                                                   \
  do {                                                                                                           \
    cudaError_t e_ = (call);                                                                                     \
    if (e_ != cudaSuccess) return Fail(nullptr, MCFD_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e_)); \
  } while (0)

}  // extern "C"
template <class T>
static cudaError_t Upload(DevBuf<T> *buf, const T *host, size_t n) {
  cudaError_t e = buf->Alloc(n);
  if (e != cudaSuccess || n == 0) return e;
  return cudaMemcpy(buf->p, host, n * sizeof(T), cudaMemcpyHostToDevice);
}
extern "C" {
static int Blocks(int64_t n, int threads) { return (int)((n + threads - 1) / threads); }

int mcfd_get_node_parts(int32_t device, int64_t n_nodes, int64_t n_cells, const int64_t *range, const int64_t *nodes,
    const int64_t *cell_parts, int64_t n_parts, int64_t *node_parts) {
  const char *me = "mcfd_get_node_parts";
  if (n_nodes < 0 || n_cells < 0 || !range || !nodes || !cell_parts || !node_parts)
    return Fail(nullptr, MCFD_ERR_ARG, std::string(me) + ": null pointer or negative size");
  const int64_t n_entries = range[n_cells] - range[0];
  if (range[0] != 0 || n_entries < 0) return Fail(nullptr, MCFD_ERR_ARG, std::string(me) + ": range must start at 0 and grow");
  for (int64_t j = 0; j < n_entries; ++j)
    if (nodes[j] < 0 || nodes[j] >= n_nodes) return Fail(nullptr, MCFD_ERR_ARG, std::string(me) + ": node index out of range");
  MCFD_CUDA0(me, cudaSetDevice(device));
  DevBuf<long long> d_range, d_nodes, d_parts, d_out;
  MCFD_CUDA0(me, Upload(&d_range, (const long long *)range, (size_t)n_cells + 1));
  MCFD_CUDA0(me, Upload(&d_nodes, (const long long *)nodes, (size_t)n_entries));
  MCFD_CUDA0(me, Upload(&d_parts, (const long long *)cell_parts, (size_t)n_cells));
  MCFD_CUDA0(me, d_out.Alloc((size_t)n_nodes));
  if (n_nodes) FillKernel<<<Blocks(n_nodes, 256), 256>>>(d_out.p, n_nodes, n_parts);   // metis.hpp:319
  if (n_entries) NodePartsKernel<<<Blocks(n_entries, 256), 256>>>(d_range.p, d_nodes.p, d_parts.p, n_cells, n_entries, d_out.p);
  MCFD_CUDA0(me, cudaGetLastError());
  MCFD_CUDA0(me, cudaMemcpy(node_parts, d_out.p, (size_t)n_nodes * sizeof(int64_t), cudaMemcpyDeviceToHost));
  return MCFD_OK;
}

int mcfd_get_new_order(int32_t device, int64_t n, const int64_t *parts, int64_t *new_to_old, int64_t *old_to_new) {
  const char *me = "mcfd_get_new_order";
  if (n < 0 || (n > 0 && (!parts || !new_to_old || !old_to_new)))
    return Fail(nullptr, MCFD_ERR_ARG, std::string(me) + ": null pointer or negative size");
  if (n == 0) return MCFD_OK;
  int64_t n_parts = 0;
  for (int64_t i = 0; i < n; ++i) {
    if (parts[i] < 0) return Fail(nullptr, MCFD_ERR_ARG, std::string(me) + ": negative part");
    n_parts = std::max(n_parts, parts[i] + 1);
  }
  const int64_t n_chunks = (n + kSortChunk - 1) / kSortChunk, table = n_parts * n_chunks;
  if (table > (int64_t)1 << 28) return Fail(nullptr, MCFD_ERR_ARG, std::string(me) + ": parts x chunks table too large");
  MCFD_CUDA0(me, cudaSetDevice(device));
  DevBuf<long long> d_parts, d_n2o, d_o2n;
  DevBuf<unsigned int> d_hist, d_running;
  DevBuf<unsigned long long> d_base;
  MCFD_CUDA0(me, Upload(&d_parts, (const long long *)parts, (size_t)n));
  MCFD_CUDA0(me, d_n2o.Alloc(n));
  MCFD_CUDA0(me, d_o2n.Alloc(n));
  MCFD_CUDA0(me, d_hist.Alloc(table));
  MCFD_CUDA0(me, d_running.Alloc(table));
  MCFD_CUDA0(me, d_base.Alloc(table + 1));
  MCFD_CUDA0(me, cudaMemset(d_hist.p, 0, table * sizeof(unsigned int)));
  MCFD_CUDA0(me, cudaMemset(d_running.p, 0, table * sizeof(unsigned int)));
  PartHistogramKernel<<<(int)n_chunks, 256>>>(d_parts.p, n, (int)n_parts, n_chunks, d_hist.p);
  TokenScanKernel<<<1, 1024>>>(d_hist.p, (int)table, d_base.p);     // exclusive scan, part-major = the sorted order
  PartScatterKernel<<<(int)n_chunks, 32>>>(d_parts.p, n, n_chunks, d_base.p, d_running.p, d_n2o.p, d_o2n.p);
  MCFD_CUDA0(me, cudaGetLastError());
  MCFD_CUDA0(me, cudaMemcpy(new_to_old, d_n2o.p, n * sizeof(int64_t), cudaMemcpyDeviceToHost));
  MCFD_CUDA0(me, cudaMemcpy(old_to_new, d_o2n.p, n * sizeof(int64_t), cudaMemcpyDeviceToHost));
  return MCFD_OK;
}

int mcfd_shuffle_data(int32_t device, int64_t n, int64_t width, const int64_t *new_to_old, double *data) {
  const char *me = "mcfd_shuffle_data";
  if (n < 0 || width <= 0 || (n > 0 && (!new_to_old || !data)))
    return Fail(nullptr, MCFD_ERR_ARG, std::string(me) + ": null pointer or bad size");
  if (n == 0) return MCFD_OK;
  for (int64_t i = 0; i < n; ++i)
    if (new_to_old[i] < 0 || new_to_old[i] >= n) return Fail(nullptr, MCFD_ERR_ARG, std::string(me) + ": index out of range");
  MCFD_CUDA0(me, cudaSetDevice(device));
  DevBuf<long long> d_order;
  DevBuf<double> d_src, d_dst;
  MCFD_CUDA0(me, Upload(&d_order, (const long long *)new_to_old, (size_t)n));
  MCFD_CUDA0(me, Upload(&d_src, (const double *)data, (size_t)(n * width)));
  MCFD_CUDA0(me, d_dst.Alloc(n * width));
  GatherRowsKernel<<<Blocks(n * width, 256), 256>>>(d_src.p, d_dst.p, d_order.p, n, width);
  MCFD_CUDA0(me, cudaGetLastError());
  MCFD_CUDA0(me, cudaMemcpy(data, d_dst.p, n * width * sizeof(double), cudaMemcpyDeviceToHost));
  return MCFD_OK;
}

int mcfd_shuffle_connectivity(int32_t device, int64_t n_cells, int32_t npe, int64_t n_nodes,
    const int64_t *old_to_new_for_nodes, const int64_t *new_to_old_for_cells, int64_t *cell_nodes) {
  const char *me = "mcfd_shuffle_connectivity";
  if (n_cells < 0 || npe <= 0 || n_nodes < 0 || (n_cells > 0 && (!old_to_new_for_nodes || !new_to_old_for_cells || !cell_nodes)))
    return Fail(nullptr, MCFD_ERR_ARG, std::string(me) + ": null pointer or bad size");
  if (n_cells == 0) return MCFD_OK;
  for (int64_t i = 0; i < n_cells * npe; ++i)   // the reference's .at() (shuffler.hpp:94) throws on such input
    if (cell_nodes[i] < 1 || cell_nodes[i] > n_nodes) return Fail(nullptr, MCFD_ERR_ARG, std::string(me) + ": node id out of range (ids are 1-based)");
  for (int64_t i = 0; i < n_cells; ++i)
    if (new_to_old_for_cells[i] < 0 || new_to_old_for_cells[i] >= n_cells) return Fail(nullptr, MCFD_ERR_ARG, std::string(me) + ": cell index out of range");
  MCFD_CUDA0(me, cudaSetDevice(device));
  DevBuf<long long> d_nodes, d_cells, d_old, d_new;
  MCFD_CUDA0(me, Upload(&d_nodes, (const long long *)old_to_new_for_nodes, (size_t)n_nodes));
  MCFD_CUDA0(me, Upload(&d_cells, (const long long *)new_to_old_for_cells, (size_t)n_cells));
  MCFD_CUDA0(me, Upload(&d_old, (const long long *)cell_nodes, (size_t)(n_cells * npe)));
  MCFD_CUDA0(me, d_new.Alloc(n_cells * npe));
  ShuffleConnectivityKernel<<<Blocks(n_cells * npe, 256), 256>>>(d_old.p, d_new.p, d_nodes.p, d_cells.p, n_cells, npe);
  MCFD_CUDA0(me, cudaGetLastError());
  MCFD_CUDA0(me, cudaMemcpy(cell_nodes, d_new.p, n_cells * npe * sizeof(int64_t), cudaMemcpyDeviceToHost));
  return MCFD_OK;
}

int64_t mcfd_launch_count(const mcfd_ctx *ctx) { return ctx ? ctx->launches : 0; }
const char *mcfd_stage_kernel(const mcfd_ctx *ctx) {
  if (!ctx || !ctx->has_mesh) return "?";
  if (ctx->cfg.scheme == MCFD_DG_GENERAL) return "DgmStageKernel";
  if (!ctx->use_lines) return "StageKernel";
  if (ctx->use_nodes) return ctx->mesh.all_affine ? "NodeStageKernel<affine>" : "NodeStageKernel<general>";
  return ctx->mesh.all_affine ? "LineStageKernel<affine>" : "LineStageKernel<general>";
}

int mcfd_enable_kernel_timing(mcfd_ctx *ctx, int enable) {
  if (!ctx) return MCFD_ERR_ARG;
  ctx->timing = enable != 0;
  ctx->events_used = 0;
  return MCFD_OK;
}

int mcfd_kernel_time(mcfd_ctx *ctx, double *total_ms, int64_t *n_launches, int reset) {
  if (!ctx) return MCFD_ERR_ARG;
  MCFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  double total = 0;
  for (size_t i = 0; i < ctx->events_used; ++i) {
    float ms = 0;
    MCFD_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->events[i].first, ctx->events[i].second));
    total += ms;
  }
  if (total_ms) *total_ms = total;
  if (n_launches) *n_launches = (int64_t)ctx->events_used;
  if (reset) ctx->events_used = 0;
  return MCFD_OK;
}

int mcfd_riemann_flux(mcfd_ctx *ctx, int32_t bc_kind, int64_t n, const double *normals,
    const double *u_left, const double *u_right, double *flux) {
  if (!ctx || n <= 0) return MCFD_ERR_ARG;
  DevBuf<double> dn, dl, dr, df;
  MCFD_CUDA(ctx, dn.Alloc(3 * n));
  MCFD_CUDA(ctx, dl.Alloc(5 * n));
  MCFD_CUDA(ctx, dr.Alloc(5 * n));
  MCFD_CUDA(ctx, df.Alloc(5 * n));
  MCFD_CUDA(ctx, cudaMemcpy(dn.p, normals, 3 * n * sizeof(double), cudaMemcpyHostToDevice));
  MCFD_CUDA(ctx, cudaMemcpy(dl.p, u_left, 5 * n * sizeof(double), cudaMemcpyHostToDevice));
  MCFD_CUDA(ctx, cudaMemcpy(dr.p, u_right, 5 * n * sizeof(double), cudaMemcpyHostToDevice));
  int threads = 128, blocks = (int)((n + threads - 1) / threads);
  switch (ctx->cfg.riemann) {
  case MCFD_EXACT: RiemannKernel<kExact><<<blocks, threads, 0, ctx->stream>>>(bc_kind, (int)n, dn.p, dl.p, dr.p, df.p); break;
  case MCFD_HLLC: RiemannKernel<kHllc><<<blocks, threads, 0, ctx->stream>>>(bc_kind, (int)n, dn.p, dl.p, dr.p, df.p); break;
  default: RiemannKernel<kAusm><<<blocks, threads, 0, ctx->stream>>>(bc_kind, (int)n, dn.p, dl.p, dr.p, df.p); break;
  }
  MCFD_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  MCFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  MCFD_CUDA(ctx, cudaMemcpy(flux, df.p, 5 * n * sizeof(double), cudaMemcpyDeviceToHost));
  return MCFD_OK;
}


// ---- time step without any allocation -------------------------------------------------------------
//
// Some hosts of this library must not allocate device memory once they are running: a solver that has
// handed the whole GPU to its own pool at start-up (a later cudaMalloc would fail or fragment that
// pool), a rank that runs under a fixed memory budget checked by its job manager, or a time loop whose
// iterations must not hit the implicit device synchronisation that cudaMalloc/cudaFree carry.
// mcfd_time_step() needs a context, and contexts allocate.  This entry point needs neither: the caller
// passes the solution column, the cell lengths and the output, all already on the device, and the
// reduction's scratch — one partial minimum per block — lives in a fixed array that is part of the
// library image.  Every GPU that loads the library holds that 1 MiB until the process exits.
}  // extern "C"

namespace {

constexpr int kStaticCellsPerBlock = 8;                   // one warp per cell
constexpr int64_t kStaticMaxCells = 1 << 20;              // 1 048 576 cells (141 M DOF at degree 2)
constexpr int kStaticMaxBlocks = (int)(kStaticMaxCells / kStaticCellsPerBlock);

static __device__ __align__(256) double g_static_dt_partials[kStaticMaxBlocks];   // 1 MiB

// Variant of TimeStepKernel for the reference's column layout [cell][node][component]: a warp takes one
// cell (lanes over its nodes), a block leaves min(length / max speed) of its 8 cells in partials[block].
template <int Q>
__global__ void TimeStepPartialsKernel(const double *column, const double *length, int n_cells,
    double *partials) {
  constexpr int N = Q * Q * Q;
  __shared__ double block_dt[kStaticCellsPerBlock];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cell = blockIdx.x * kStaticCellsPerBlock + warp;
  double dt = 1e100;
  if (cell < n_cells) {
    double vmax = 1e-5;
    for (int n = lane; n < N; n += 32) {
      const double *u = column + ((size_t)cell * N + n) * 5;
      Prim q = ConsToPrim(u[0], u[1], u[2], u[3], u[4]);
      double c = SoundSpeed(q.rho, q.p);
      double v = sqrt(q.u * q.u + q.v * q.v + q.w * q.w);
      vmax = fmax(vmax, fmax(c + v, c - v));
    }
    for (int o = 16; o > 0; o >>= 1) vmax = fmax(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
    dt = length[cell] / vmax;
  }
  if (lane == 0) block_dt[warp] = dt;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < kStaticCellsPerBlock; ++w) dt = fmin(dt, block_dt[w]);
    partials[blockIdx.x] = dt;
  }
}

// One block folds the partial minima and applies fem.hpp:433-434 and fr/lobatto.hpp:220-222.
__global__ void TimeStepFinalKernel(const double *partials, int n_partials, double cfl, double dt_guess,
    int fr_doubling, double *dt_out) {
  __shared__ double s[1024];
  double m = 1e100;
  for (int i = threadIdx.x; i < n_partials; i += blockDim.x) m = fmin(m, partials[i]);
  s[threadIdx.x] = m;
  __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) s[threadIdx.x] = fmin(s[threadIdx.x], s[threadIdx.x + o]);
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    double dt = fmin(cfl * s[0], dt_guess);
    if (fr_doubling) dt = fmin(dt_guess, 2.0 * dt);
    *dt_out = dt;
  }
}

// In host code the array's name is only a handle for the runtime, not a device address; the address
// that kernels can use is looked up once per device, with that device current.
constexpr int kMaxDevices = 64;
double *g_static_dt_address[kMaxDevices] = {nullptr};

}  // namespace

extern "C" {

int mcfd_time_step_static(int32_t scheme, int32_t degree, int64_t n_cells, const double *column_dev,
    const double *length_dev, double dt_guess, int32_t rk_order, double *dt_dev, void *cuda_stream) {
  if (!column_dev || !length_dev || !dt_dev) return Fail(nullptr, MCFD_ERR_ARG, "mcfd_time_step_static: null pointer");
  if ((degree != 2 && degree != 3) || (scheme < MCFD_FR_LOBATTO || scheme > MCFD_FR_GENERAL) || rk_order < 0 || rk_order > 4)
    return Fail(nullptr, MCFD_ERR_ARG, "mcfd_time_step_static: unsupported scheme, degree or RK order");
  if (n_cells <= 0) return Fail(nullptr, MCFD_ERR_ARG, "mcfd_time_step_static: n_cells must be positive");
  if (n_cells > kStaticMaxCells)
    return Fail(nullptr, MCFD_ERR_ARG, "mcfd_time_step_static: n_cells = " + std::to_string(n_cells) +
        " exceeds the fixed scratch of " + std::to_string(kStaticMaxCells) +
        " cells; use mcfd_time_step (which allocates) for larger parts");
  int device = -1;
  cudaError_t e = cudaGetDevice(&device);
  if (e != cudaSuccess) return Fail(nullptr, MCFD_ERR_CUDA, std::string("mcfd_time_step_static: ") + cudaGetErrorString(e));
  if (device < 0 || device >= kMaxDevices) return Fail(nullptr, MCFD_ERR_ARG, "mcfd_time_step_static: device ordinal out of range");
  if (!g_static_dt_address[device]) {   // first call on this device
    double *address = nullptr;
    e = cudaGetSymbolAddress((void **)&address, g_static_dt_partials);
    if (e != cudaSuccess || !address)
      return Fail(nullptr, MCFD_ERR_CUDA, std::string("mcfd_time_step_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
    g_static_dt_address[device] = address;
  }
  double *partials = g_static_dt_address[device];
  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
  const int n_blocks = (int)((n_cells + kStaticCellsPerBlock - 1) / kStaticCellsPerBlock);
  if (degree == 2)
    TimeStepPartialsKernel<3><<<n_blocks, kStaticCellsPerBlock * 32, 0, stream>>>(column_dev, length_dev, (int)n_cells, partials);
  else
    TimeStepPartialsKernel<4><<<n_blocks, kStaticCellsPerBlock * 32, 0, stream>>>(column_dev, length_dev, (int)n_cells, partials);
  TimeStepFinalKernel<<<1, 1024, 0, stream>>>(partials, n_blocks, kCfl[rk_order][degree], dt_guess,
      scheme == MCFD_FR_LOBATTO, dt_dev);
  e = cudaGetLastError();
  if (e != cudaSuccess) return Fail(nullptr, MCFD_ERR_CUDA, std::string("mcfd_time_step_static: ") + cudaGetErrorString(e));
  return MCFD_OK;
}

}  // extern "C"
