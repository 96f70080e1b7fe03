// TEST INFRASTRUCTURE — never loaded by the minicfd_b200 package or by libminicfd_b200.so.
//
// Compiles the *device source* of the stage kernel (minicfd_b200/csrc/{physics,kernels}.cuh) for the
// host and runs it thread by thread, so that index logic, tables and arithmetic can be checked
// against the reference on a machine without a GPU (`pytest -m "not gpu"`).  It shares the host-side
// mesh/table construction (tables.hpp) with the library.  It is not an oracle and not a fallback.
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../minicfd_b200/csrc/kernels.cuh"
#include "../../minicfd_b200/csrc/tables.hpp"

using namespace mcfd;

namespace {

template <int Q, int R, int CPB, int NW>
void RunStage(StageArgs a, int n_blocks) {
  using G = StageGeometry<Q, CPB, NW>;
  std::vector<double> smem(G::kSharedBytes / sizeof(double) + 1);
  for (int b = 0; b < n_blocks; ++b) {
    if (a.grad) RunStageBlockOnHost<Q, R, CPB, NW, false, true>(a, b, smem.data());
    else if (c_tab.general) RunStageBlockOnHost<Q, R, CPB, NW, true>(a, b, smem.data());
    else RunStageBlockOnHost<Q, R, CPB, NW, false>(a, b, smem.data());
  }
}

template <int Q, int CPB, int NW>
void RunStageQ(int riemann, StageArgs a, int n_blocks) {
  if (riemann == 0) RunStage<Q, kExact, CPB, NW>(a, n_blocks);
  else if (riemann == 1) RunStage<Q, kHllc, CPB, NW>(a, n_blocks);
  else RunStage<Q, kAusm, CPB, NW>(a, n_blocks);
}

template <class T>
std::vector<T> ToTiles(const std::vector<T> &src, int rows, int64_t n_cells, int64_t n_tiles) {
  std::vector<T> out((size_t)n_tiles * rows * kTile, T{});
  for (int64_t c = 0; c < n_cells; ++c)
    for (int r = 0; r < rows; ++r) out[((c >> 5) * rows + r) * kTile + (c & 31)] = src[c * rows + r];
  return out;
}

}  // namespace

extern "C" {

// One stage on the host.  column_in / column_u0: [n_cells][F] (ghost cells included in column_in),
// column_out: [n_local][F].  bkinds per *tag* (tags 0..15); bvals [n_bfaces][Q*Q][5] in the library's
// boundary-face order.  Returns 0, or -1 with a message in err.
int emu_stage(int degree, int scheme, int riemann, int64_t n_nodes, const double *xyz,
    int64_t n_local, int64_t n_ghost, const int64_t *cell_nodes, int64_t n_bfaces,
    const int64_t *bface_nodes, const int32_t *bface_tags, int64_t n_pairs, const int64_t *pairs,
    const int32_t *tag_kinds, const double *bvals, const double *column_in, const double *column_u0,
    int mode, double dt, double *column_out, char *err, int err_len, double c_next,
    const double *viscous /* null, or beta0, beta1, nu, prandtl */) {
  const int q = degree + 1, F = 5 * q * q * q;
  const int cpb = (q == 3) ? kP2Cells : kP3Cells, nw = (q == 3) ? kP2Warps : kP3Warps;
  c_tab = BuildTables(degree, scheme, nw, kTile / cpb, c_next);
  HostMesh mesh;
  if (!mesh.Build(degree, n_nodes, xyz, n_local, n_ghost, cell_nodes, n_bfaces, bface_nodes,
          bface_tags, n_pairs, pairs)) {
    std::strncpy(err, mesh.error.c_str(), err_len - 1);
    return -1;
  }
  const int64_t nc = n_local + n_ghost, nt = (nc + kTile - 1) / kTile;
  auto geom = ToTiles(mesh.coef, kGeomRows, nc, nt);
  auto nbr = ToTiles(mesh.nbr, 6, nc, nt);
  auto code = ToTiles(mesh.code, 6, nc, nt);
  std::vector<unsigned char> bkind(mesh.bound_faces.size() + 1, 0);
  for (size_t b = 0; b < mesh.bound_faces.size(); ++b) bkind[b] = (unsigned char)tag_kinds[mesh.bound_faces[b].tag];
  std::vector<double> zero_vals((mesh.bound_faces.size() + 1) * q * q * 5, 0.0);
  std::vector<double> uin((size_t)nt * F * kTile, 0.0), u0(uin.size(), 0.0), uout(uin.size(), 0.0);
  for (int64_t c = 0; c < nc; ++c)
    for (int f = 0; f < F; ++f) uin[((c >> 5) * F + f) * kTile + (c & 31)] = column_in[c * F + f];
  for (int64_t c = 0; c < n_local; ++c)
    for (int f = 0; f < F; ++f) u0[((c >> 5) * F + f) * kTile + (c & 31)] = column_u0[c * F + f];
  StageArgs a;
  a.u_in = uin.data(); a.u0 = u0.data(); a.u_out = uout.data();
  a.geom = geom.data(); a.nbr = nbr.data(); a.code = code.data();
  a.bkind = bkind.data(); a.bvals = bvals ? bvals : zero_vals.data();
  a.n_local = (int)n_local; a.block_begin = 0; a.mode = mode; a.dt = dt;
  a.dt_dev = nullptr; a.grad = nullptr; a.hn = nullptr; a.h2s = nullptr;
  std::vector<double> grad, hn, h2s;
  if (viscous) {
    const int N = q * q * q;
    grad.assign((size_t)nt * N * 15 * kTile, 0.0);
    hn.assign((size_t)nt * 6 * q * q * 15 * kTile, 0.0);
    h2s = ToTiles(mesh.h2s, 18, nc, nt);
    if (q == 3) RunTraceOnHost<3>(uin.data(), geom.data(), grad.data(), hn.data(), (int)nc);
    else RunTraceOnHost<4>(uin.data(), geom.data(), grad.data(), hn.data(), (int)nc);
    a.grad = grad.data(); a.hn = hn.data(); a.h2s = h2s.data();
    a.vis.beta0 = viscous[0]; a.vis.beta1 = viscous[1]; a.vis.nu = viscous[2];
    a.vis.cond = viscous[2] * (kGasR * kGammaOverGm1) / viscous[3];
  }
  int n_blocks = (int)((n_local + cpb - 1) / cpb);
  if (q == 3) RunStageQ<3, kP2Cells, kP2Warps>(riemann, a, n_blocks);
  else RunStageQ<4, kP3Cells, kP3Warps>(riemann, a, n_blocks);
  for (int64_t c = 0; c < n_local; ++c)
    for (int f = 0; f < F; ++f) column_out[c * F + f] = uout[((c >> 5) * F + f) * kTile + (c & 31)];
  return 0;
}

// Boundary-face order and flux-point coordinates as the library defines them.
int64_t emu_boundary(int degree, int64_t n_nodes, const double *xyz, int64_t n_local, int64_t n_ghost,
    const int64_t *cell_nodes, int64_t n_bfaces, const int64_t *bface_nodes, const int32_t *bface_tags,
    int64_t *holder, int32_t *holder_face, int32_t *tag, double *points) {
  HostMesh mesh;
  if (!mesh.Build(degree, n_nodes, xyz, n_local, n_ghost, cell_nodes, n_bfaces, bface_nodes,
          bface_tags, 0, nullptr)) return -1;
  for (size_t b = 0; b < mesh.bound_faces.size(); ++b) {
    holder[b] = mesh.bound_faces[b].holder;
    holder_face[b] = mesh.bound_faces[b].holder_face;
    tag[b] = mesh.bound_faces[b].tag;
  }
  if (points) {  // [face][Q*Q][3] in boundary-face order (tags ascending)
    Tables t = BuildTables(degree, 0, degree == 2 ? kP2Warps : kP3Warps, kTile / (degree == 2 ? kP2Cells : kP3Cells));
    const int fq = (degree + 1) * (degree + 1);
    int last = -1;
    for (size_t b = 0; b < mesh.bound_faces.size(); ++b) {
      int tg = mesh.bound_faces[b].tag;
      if (tg == last) continue;
      last = tg;
      mesh.BoundaryPoints(t, degree, tg, points + b * fq * 3);
    }
  }
  return (int64_t)mesh.bound_faces.size();
}

void emu_riemann(int riemann, int bc_kind, const double *n, const double *ul, const double *ur,
    double *flux) {
  if (bc_kind < 0) {
    if (riemann == 0) FluxUpwind<kExact>(ul, ur, n[0], n[1], n[2], flux);
    else if (riemann == 1) FluxUpwind<kHllc>(ul, ur, n[0], n[1], n[2], flux);
    else FluxUpwind<kAusm>(ul, ur, n[0], n[1], n[2], flux);
  } else {
    if (riemann == 0) BoundaryFlux<kExact>(bc_kind, ul, ur, n[0], n[1], n[2], flux);
    else if (riemann == 1) BoundaryFlux<kHllc>(bc_kind, ul, ur, n[0], n[1], n[2], flux);
    else BoundaryFlux<kAusm>(bc_kind, ul, ur, n[0], n[1], n[2], flux);
  }
}

}  // extern "C"
