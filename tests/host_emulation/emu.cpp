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
#include "../../minicfd_b200/csrc/format.cuh"

using namespace mcfd;

namespace {

bool g_ns_base = false;   // viscous[4] == 2: the per-cell viscosities come on top of the Navier-Stokes terms

template <int Q, int R, int CPB, int NW>
void RunStage(StageArgs a, int n_blocks) {
  using G = StageGeometry<Q, CPB, NW>;
  std::vector<double> smem(G::kSharedBytes / sizeof(double) + 1);
  for (int b = 0; b < n_blocks; ++b) {
    if (a.grad && a.cell_nu && g_ns_base) RunStageBlockOnHost<Q, R, CPB, NW, false, kDiffNsIsotropic>(a, b, smem.data());   // over a Navier-Stokes base
    else if (a.grad && a.cell_nu) RunStageBlockOnHost<Q, R, CPB, NW, false, kDiffIsotropic>(a, b, smem.data());
    else if (a.grad && c_tab.general) RunStageBlockOnHost<Q, R, CPB, NW, true, kDiffNavierStokes>(a, b, smem.data());
    else if (a.grad) RunStageBlockOnHost<Q, R, CPB, NW, false, kDiffNavierStokes>(a, b, smem.data());
    else if (c_tab.general) RunStageBlockOnHost<Q, R, CPB, NW, true>(a, b, smem.data());
    else RunStageBlockOnHost<Q, R, CPB, NW, false>(a, b, smem.data());
  }
}

template <int Q, int CPB, bool kAffine>
void RunLineStageA(int riemann, StageArgs a, int64_t n_local) {
  using S = LineShape<Q, CPB>;
  std::vector<double> smem(S::kSharedBytes / sizeof(double) + 1);
  const LineTables tab = MakeLineTables(c_tab, Q, kTile / CPB);
  const std::vector<unsigned char> bytes = LineTableBytes(tab, Q);
  a.line_bytes = bytes.data();
  const int n_blocks = (int)((n_local + CPB - 1) / CPB);
  for (int b = 0; b < n_blocks; ++b) {
    if (riemann == 0) RunLineStageBlockOnHost<Q, kExact, CPB, kAffine>(a, tab, b, smem.data());
    else if (riemann == 1) RunLineStageBlockOnHost<Q, kHllc, CPB, kAffine>(a, tab, b, smem.data());
    else RunLineStageBlockOnHost<Q, kAusm, CPB, kAffine>(a, tab, b, smem.data());
  }
}
// the node-line kernel of the lumped inviscid schemes (stage_lines.cuh), block by block; the affine instantiation
// when every local cell is a parallelepiped, as the library chooses
template <int Q, int CPB>
void RunLineStage(int riemann, StageArgs a, int64_t n_local, bool affine) {
  if (affine) RunLineStageA<Q, CPB, true>(riemann, a, n_local);
  else RunLineStageA<Q, CPB, false>(riemann, a, n_local);
}

// the persistent one-warp-per-node kernel (stage_nodes.cuh), tile by tile
template <bool kAffine>
void RunNodeStageA(int riemann, StageArgs a, int64_t n_local) {
  using S = NodeShape;
  std::vector<double> smem(S::kSharedBytes / sizeof(double) + 1);
  const LineTables tab = MakeLineTables(c_tab, 3, 1);
  const std::vector<unsigned char> bytes = LineTableBytes(tab, 3);
  a.line_bytes = bytes.data();
  const int n_tiles = (int)((n_local + kTile - 1) / kTile);
  for (int t = 0; t < n_tiles; ++t) {
    if (riemann == 0) RunNodeStageTileOnHost<kExact, kAffine>(a, tab, t, smem.data());
    else if (riemann == 1) RunNodeStageTileOnHost<kHllc, kAffine>(a, tab, t, smem.data());
    else RunNodeStageTileOnHost<kAusm, kAffine>(a, tab, t, smem.data());
  }
}

void RunNodeStage(int riemann, StageArgs a, int64_t n_local, bool affine) {
  if (affine) RunNodeStageA<true>(riemann, a, n_local);
  else RunNodeStageA<false>(riemann, a, n_local);
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
    const double *viscous /* null, or beta0, beta1, nu, prandtl, kLocal (0 / 1: the traces interpolate det u) */,
    const double *cell_nu /* null, or [n_cells][5]: the component-wise isotropic model */,
    int line_cells /* 0: the slot kernel (StageKernel); 16 | 32: the node-line kernel with that many cells per block;
                      -1: the library's choice */) {
  const int q = degree + 1, F = 5 * q * q * q;
  if (line_cells < 0) line_cells = q == 3 ? kLineP2Cells : 0;   // degree 3 stays on the slot kernel
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
  StageArgs a{};
  a.u_in = uin.data(); a.u0 = u0.data(); a.u_out = uout.data();
  a.geom = geom.data(); a.nbr = nbr.data(); a.code = code.data();
  a.bkind = bkind.data(); a.bvals = bvals ? bvals : zero_vals.data();
  a.n_local = (int)n_local; a.block_begin = 0; a.mode = mode; a.dt = dt;
  a.dt_dev = nullptr; a.grad = nullptr; a.hn = nullptr; a.h2s = nullptr; a.cell_nu = nullptr;
  std::vector<double> grad, hn, h2s, nu_tiles;
  if (viscous && cell_nu) {
    std::vector<double> rows(cell_nu, cell_nu + nc * 5);
    nu_tiles = ToTiles(rows, 5, nc, nt);
    a.cell_nu = nu_tiles.data();
  }
  if (viscous) {
    const int N = q * q * q;
    grad.assign((size_t)nt * N * 15 * kTile, 0.0);
    hn.assign((size_t)nt * 6 * q * q * 15 * kTile, 0.0);
    h2s = ToTiles(mesh.h2s, 18, nc, nt);
    const bool klocal = viscous[4] == 1;
    g_ns_base = viscous[4] == 2;
    if (q == 3 && klocal) RunTraceOnHost<3, true>(uin.data(), geom.data(), grad.data(), hn.data(), (int)nc);
    else if (q == 3) RunTraceOnHost<3>(uin.data(), geom.data(), grad.data(), hn.data(), (int)nc);
    else if (klocal) RunTraceOnHost<4, true>(uin.data(), geom.data(), grad.data(), hn.data(), (int)nc);
    else RunTraceOnHost<4>(uin.data(), geom.data(), grad.data(), hn.data(), (int)nc);
    a.grad = grad.data(); a.hn = hn.data(); a.h2s = h2s.data();
    a.vis.beta0 = viscous[0]; a.vis.beta1 = viscous[1]; a.vis.nu = viscous[2];
    a.vis.cond = viscous[2] * (kGasR * kGammaOverGm1) / viscous[3];
  }
  int n_blocks = (int)((n_local + cpb - 1) / cpb);
  const bool lines = !viscous && scheme != 2 && line_cells != 0;   // what LaunchStage of the library picks
  if (lines && q == 3 && line_cells == 1) RunNodeStage(riemann, a, n_local, mesh.all_affine);
  else if (lines && q == 3 && line_cells == 16) RunLineStage<3, 16>(riemann, a, n_local, mesh.all_affine);
  else if (lines && q == 3) RunLineStage<3, 32>(riemann, a, n_local, mesh.all_affine);
  else if (lines && line_cells == 32) RunLineStage<4, 32>(riemann, a, n_local, mesh.all_affine);
  else if (lines && line_cells == 8) RunLineStage<4, 8>(riemann, a, n_local, mesh.all_affine);
  else if (lines) RunLineStage<4, 16>(riemann, a, n_local, mesh.all_affine);
  else if (q == 3) RunStageQ<3, kP2Cells, kP2Warps>(riemann, a, n_blocks);
  else RunStageQ<4, kP3Cells, kP3Warps>(riemann, a, n_blocks);
  for (int64_t c = 0; c < n_local; ++c)
    for (int f = 0; f < F; ++f) column_out[c * F + f] = uout[((c >> 5) * F + f) * kTile + (c & 31)];
  return 0;
}

}  // extern "C"

// weno::Lazy limiter on the host: the device functions of kernels.cuh, one (cell, component) at a time, in the
// same two batches as mcfd_limiter_weno_lazy.  column: [n_cells][F] in, [n_local][F] out.
template <int Q>
static void RunWeno(HostMesh const &mesh, std::vector<double> &u, std::vector<double> const &geom,
    std::vector<int> const &nbr, std::vector<unsigned char> const &code, std::vector<double> const &h2s,
    std::vector<double> const &bas, int64_t first_inter, double w0, double eps, bool eigen,
    std::vector<double> const &frames) {
  constexpr int P = Q - 1, N = Q * Q * Q, F = 5 * N, M = ModalShape<P>::M, R = ModalShape<P>::kBasRows;
  const int64_t nl = mesh.n_local, nc = mesh.n_local + mesh.n_ghost, nt = (nc + kTile - 1) / kTile;
  std::vector<double> quad((size_t)nt * N * 4 * kTile, 0.0), mod((size_t)nt * 5 * M * kTile, 0.0), tay(mod), modn(mod), tayn(mod);
  for (int64_t c = 0; c < nc; ++c) {
    const int64_t t = c >> 5; const int l = c & 31;
    for (int n = 0; n < N; ++n) {
      double d[3], w;
      QuadraturePoint<Q>(n, geom.data() + t * kGeomRows * kTile + l, kTile, d, &w);
      for (int a = 0; a < 3; ++a) quad[(t * N * 4 + n * 4 + a) * kTile + l] = d[a];
      quad[(t * N * 4 + n * 4 + 3) * kTile + l] = w;
    }
  }
  auto modal = [&](int64_t b, int64_t e) {
    for (int64_t c = b; c < e; ++c) for (int k = 0; k < 5; ++k) {
      const int64_t t = c >> 5; const int l = c & 31;
      ModalProjectCell<Q>(u.data() + t * F * kTile + k * kTile + l, kTile, quad.data() + t * N * 4 * kTile + l, kTile,
          bas.data() + t * R * kTile + l, kTile, mod.data() + (t * 5 * M + k * M) * kTile + l,
          tay.data() + (t * 5 * M + k * M) * kTile + l, kTile);
    }
  };
  WenoArgs a;
  a.u = u.data(); a.geom = geom.data(); a.bas = bas.data(); a.h2s = h2s.data(); a.nbr = nbr.data(); a.code = code.data();
  a.mod = mod.data(); a.tay = tay.data(); a.mod_new = modn.data(); a.tay_new = tayn.data(); a.w0 = w0; a.eps = eps;
  a.comp_stride = M; a.mode_stride = 1; a.modal_state = 0; a.legendre = 0;
  std::vector<double> scratch((size_t)(7 * 5 * M + 5 * M) * kTile, 0.0);
  auto limit_eigen_tile = [&](int64_t t) {   // WenoEigenKernel's phases, thread by thread
    WenoEigenArgs e;
    e.w = a; e.frames = frames.data(); e.cand = scratch.data(); e.blend = scratch.data() + (size_t)7 * 5 * M * kTile;
    static EigenState<P> st[5][kTile];
    auto q = [&](int l) { return quad.data() + t * N * 4 * kTile + l; };
    auto bs = [&](int l) { return bas.data() + t * R * kTile + l; };
    bool active[kTile];
    for (int l = 0; l < kTile; ++l) {
      const int64_t c = t * kTile + l;
      active[l] = c >= a.cell_begin && c < a.cell_end;
      for (int k = 0; k < 5; ++k) {
        st[k][l].total_volume = 0; st[k][l].adj_cnt = 0; st[k][l].troubled = false;
        for (int i = 0; i < M; ++i) st[k][l].total[i] = 0;
        if (active[l]) st[k][l].troubled = WenoEigenDetect<Q>(a, (int)t, l, q(l), kTile, bs(l), kTile, &st[k][l].adj_cnt);
        if (active[l] && st[k][l].troubled) WenoEigenBorrow<Q>(e, (int)t, l, k, q(l), kTile, bs(l), kTile);
      }
    }
    for (int f = 0; f < 6; ++f) {
      for (int l = 0; l < kTile; ++l) for (int k = 0; k < 5; ++k)
        if (active[l] && st[k][l].troubled && nbr[(t * 6 + f) * kTile + l] >= 0)
          WenoEigenFaceBlend<Q>(e, (int)t, l, k, f, st[k][l].adj_cnt, q(l), kTile, bs(l), kTile);
      for (int l = 0; l < kTile; ++l) for (int k = 0; k < 5; ++k)
        if (active[l] && st[k][l].troubled && nbr[(t * 6 + f) * kTile + l] >= 0)
          WenoEigenFaceAccumulate<Q>(e, (int)t, l, k, f, bs(l), kTile, &st[k][l]);
    }
    for (int l = 0; l < kTile; ++l) for (int k = 0; k < 5; ++k)
      if (active[l]) WenoEigenFinish<Q>(e, (int)t, l, k, q(l), kTile, bs(l), kTile, st[k][l]);
  };
  auto limit = [&](int64_t b, int64_t e) {
    a.cell_begin = (int)b; a.cell_end = (int)e;
    if (eigen) {
      if (e > b) for (int64_t t = b / kTile; t <= (e - 1) / kTile; ++t) limit_eigen_tile(t);
    } else
    for (int64_t c = b; c < e; ++c) for (int k = 0; k < 5; ++k) {
      const int64_t t = c >> 5; const int l = c & 31;
      WenoLazyCell<Q>(a, (int)t, l, k, quad.data() + t * N * 4 * kTile + l, kTile, bas.data() + t * R * kTile + l, kTile);
    }
    for (int64_t c = b; c < e; ++c) for (int r = 0; r < 5 * M; ++r) {
      const size_t i = ((c >> 5) * 5 * M + r) * kTile + (c & 31);
      tay[i] = tayn[i]; mod[i] = modn[i];
    }
  };
  const int64_t fi = mesh.n_ghost ? first_inter : nl;
  modal(0, nl); limit(0, fi);
  modal(nl, nc); limit(fi, nl);
}

extern "C" {

int emu_weno_lazy(int degree, int64_t n_nodes, const double *xyz, int64_t n_local, int64_t n_ghost,
    const int64_t *cell_nodes, int64_t n_bfaces, const int64_t *bface_nodes, const int32_t *bface_tags,
    int64_t n_pairs, const int64_t *pairs, double w0, double eps, const double *column_in, double *column_out,
    int eigen, const int64_t *global_ids) {
  const int q = degree + 1, F = 5 * q * q * q;
  c_tab = BuildTables(degree, 0, (q == 3) ? kP2Warps : kP3Warps, kTile / ((q == 3) ? kP2Cells : kP3Cells));
  HostMesh mesh;
  if (global_ids) mesh.global_ids.assign(global_ids, global_ids + n_local + n_ghost);
  if (!mesh.Build(degree, n_nodes, xyz, n_local, n_ghost, cell_nodes, n_bfaces, bface_nodes, bface_tags, n_pairs, pairs))
    return -1;
  const int64_t nc = n_local + n_ghost, nt = (nc + kTile - 1) / kTile;
  auto geom = ToTiles(mesh.coef, kGeomRows, nc, nt);
  auto nbr = ToTiles(mesh.nbr, 6, nc, nt);
  auto code = ToTiles(mesh.code, 6, nc, nt);
  auto h2s = ToTiles(mesh.h2s, 18, nc, nt);
  auto bas = ToTiles(mesh.BuildModalBases(degree), q == 3 ? ModalShape<2>::kBasRows : ModalShape<3>::kBasRows, nc, nt);
  std::vector<double> u((size_t)nt * F * kTile, 0.0);
  for (int64_t c = 0; c < nc; ++c)
    for (int f = 0; f < F; ++f) u[((c >> 5) * F + f) * kTile + (c & 31)] = column_in[c * F + f];
  int64_t first_inter = n_local;
  for (int64_t c = 0; c < n_local && first_inter == n_local; ++c)
    for (int f = 0; f < 6; ++f) if (mesh.nbr[c * 6 + f] >= n_local) { first_inter = c; break; }
  auto frames = ToTiles(mesh.frames, 54, nc, nt);
  if (q == 3) RunWeno<3>(mesh, u, geom, nbr, code, h2s, bas, first_inter, w0, eps, eigen != 0, frames);
  else RunWeno<4>(mesh, u, geom, nbr, code, h2s, bas, first_inter, w0, eps, eigen != 0, frames);
  for (int64_t c = 0; c < n_local; ++c)
    for (int f = 0; f < F; ++f) column_out[c * F + f] = u[((c >> 5) * F + f) * kTile + (c & 31)];
  return 0;
}

}  // extern "C"

template <int Q, int CPB, int NW>
static void RunEnergy(HostMesh const &mesh, const double *betas, const double *column, double time_scale,
    double *damping_out, double *nu_out, double *min_dt_out) {
  constexpr int N = Q * Q * Q, F = 5 * N;
  using G = StageGeometry<Q, CPB, NW>;
  const int64_t nl = mesh.n_local, nc = nl + mesh.n_ghost, nt = (nc + kTile - 1) / kTile;
  auto geom = ToTiles(mesh.coef, kGeomRows, nc, nt);
  auto nbr = ToTiles(mesh.nbr, 6, nc, nt);
  auto code = ToTiles(mesh.code, 6, nc, nt);
  auto h2s = ToTiles(mesh.h2s, 18, nc, nt);
  std::vector<unsigned char> bkind(mesh.bound_faces.size() + 1, 0);
  std::vector<double> bvals((mesh.bound_faces.size() + 1) * Q * Q * 5, 0.0);
  std::vector<double> u((size_t)nt * F * kTile, 0.0), res(u.size(), 0.0), grad((size_t)nt * N * 15 * kTile, 0.0),
      hn((size_t)nt * 6 * Q * Q * 15 * kTile, 0.0), damping((size_t)nt * N * N * kTile, 0.0), smem(G::kSharedBytes / 8 + 1);
  StageArgs a{};
  a.u_in = u.data(); a.u0 = u.data(); a.u_out = res.data(); a.geom = geom.data(); a.nbr = nbr.data(); a.code = code.data();
  a.bkind = bkind.data(); a.bvals = bvals.data(); a.n_local = (int)nl; a.block_begin = 0; a.mode = kModeResidual; a.dt = 0;
  a.dt_dev = nullptr; a.grad = grad.data(); a.hn = hn.data(); a.h2s = h2s.data(); a.cell_nu = nullptr;
  a.vis.beta0 = betas[0]; a.vis.beta1 = betas[1]; a.vis.nu = 0; a.vis.cond = 0;
  const int n_blocks = (int)((nl + CPB - 1) / CPB);
  for (int col = 0; col < N; ++col) {   // mcfd_build_damping_matrices
    for (size_t i = 0; i < u.size(); ++i) u[i] = ((int)((i / kTile) % F) / 5 == col) ? 1.0 : 0.0;
    RunTraceOnHost<Q>(u.data(), geom.data(), grad.data(), hn.data(), (int)nc);
    for (int b = 0; b < n_blocks; ++b) RunStageBlockOnHost<Q, kHllc, CPB, NW, false, kDiffProbe>(a, b, smem.data());
    for (int64_t c = 0; c < nl; ++c) for (int r = 0; r < N; ++r) {
      double d[3], w;
      QuadraturePoint<Q>(r, geom.data() + (c >> 5) * kGeomRows * kTile + (c & 31), kTile, d, &w);
      damping[((c >> 5) * N * N + r * N + col) * kTile + (c & 31)] = res[((c >> 5) * F + r * 5) * kTile + (c & 31)] * w;
    }
  }
  for (int64_t c = 0; c < nl; ++c) for (int i = 0; i < N * N; ++i) damping_out[c * N * N + i] = damping[((c >> 5) * N * N + i) * kTile + (c & 31)];
  if (!column) return;
  for (int64_t c = 0; c < nc; ++c) for (int f = 0; f < F; ++f) u[((c >> 5) * F + f) * kTile + (c & 31)] = column[c * F + f];
  std::vector<double> length((size_t)nt * kTile, 1.0), nu((size_t)nt * 5 * kTile, 0.0);
  for (int64_t c = 0; c < nl; ++c) length[c] = mesh.length[c];
  std::vector<unsigned char> nn(6 * 8 * Q * Q);
  for (int i = 0; i < 6 * 8 * Q * Q; ++i) nn[i] = c_tab.nbr_node[i / (8 * Q * Q)][(i / (Q * Q)) % 8][i % (Q * Q)];
  double min_dt = 1e100;
  std::vector<double> shared(ViscosityShape<Q>::kRows);
  for (int64_t c = 0; c < nl; ++c) {
    ViscosityShared<Q>(u.data(), geom.data(), (int)(c >> 5), (int)(c & 31), 0, 1, shared.data(), 1);
    for (int k = 0; k < 5; ++k)
      min_dt = std::fmin(min_dt, ViscosityCell<Q>(u.data(), nbr.data(), code.data(), nn.data(), damping.data(),
          length.data(), (int)(c >> 5), (int)(c & 31), k, time_scale, shared.data(), 1, nu.data()));
  }
  for (int64_t c = 0; c < nl; ++c) for (int k = 0; k < 5; ++k) nu_out[c * 5 + k] = nu[((c >> 5) * 5 + k) * kTile + (c & 31)];
  *min_dt_out = min_dt;
}

extern "C" {

// EnergyBasedViscosity on the host: damping matrices [n_local][N][N] and, when a column is given, the viscosities
// [n_local][5] and min_dt_ of UpdateProperties.
int emu_energy_viscosity(int degree, int64_t n_nodes, const double *xyz, int64_t n_local, const int64_t *cell_nodes,
    int64_t n_bfaces, const int64_t *bface_nodes, const int32_t *bface_tags, const double *betas, const double *column,
    double time_scale, double *damping_out, double *nu_out, double *min_dt_out) {
  const int q = degree + 1;
  c_tab = BuildTables(degree, 0, (q == 3) ? kP2Warps : kP3Warps, kTile / ((q == 3) ? kP2Cells : kP3Cells));
  HostMesh mesh;
  if (!mesh.Build(degree, n_nodes, xyz, n_local, 0, cell_nodes, n_bfaces, bface_nodes, bface_tags, 0, nullptr)) return -1;
  if (q == 3) RunEnergy<3, kP2Cells, kP2Warps>(mesh, betas, column, time_scale, damping_out, nu_out, min_dt_out);
  else RunEnergy<4, kP3Cells, kP3Warps>(mesh, betas, column, time_scale, damping_out, nu_out, min_dt_out);
  return 0;
}

}  // extern "C"

template <int Q>
static void RunDgm(int riemann, DgmArgs a, const double *values, double *proj_out, int64_t nl, int64_t nc,
    std::vector<double> &u, std::vector<double> &tay, const double *geom, const double *bas) {
  if (values) {   // mcfd_project
    for (int64_t c = 0; c < nl; ++c) DgmProjectCell<Q>(values, geom, bas, proj_out, (int)c);
    return;
  }
  for (int64_t c = 0; c < nc; ++c) DgmTaylorCell<Q>(u.data(), bas, tay.data(), (int)(c >> 5), (int)(c & 31));
  for (int64_t c = 0; c < nl; ++c) {
    if (riemann == 0) DgmStageCell<Q, kExact>(a, (int)(c >> 5), (int)(c & 31));
    else if (riemann == 1) DgmStageCell<Q, kHllc>(a, (int)(c >> 5), (int)(c & 31));
    else DgmStageCell<Q, kAusm>(a, (int)(c >> 5), (int)(c & 31));
  }
}

extern "C" {

// dg::General on the modal polynomial, one stage on the host.  Columns hold 5 M coefficients per cell; with
// `values` ([n_local][Q^3][5]) the call is the L2 projection instead and column_out receives the coefficients.
// gauss_points_out (optional): [n_bfaces][Q*Q][3] boundary Gauss points in the library's boundary-face order.
int emu_dgm_stage(int degree, int riemann, int64_t n_nodes, const double *xyz, int64_t n_local, int64_t n_ghost,
    const int64_t *cell_nodes, int64_t n_bfaces, const int64_t *bface_nodes, const int32_t *bface_tags,
    const int32_t *tag_kinds, const double *bvals, const double *column_in, const double *column_u0, int mode, double dt,
    const double *values, double *column_out, double *gauss_points_out, double *quad_points_out) {
  const int q = degree + 1, M = degree == 2 ? 10 : 20, F = 5 * M;
  c_tab = BuildTables(degree, 0, (q == 3) ? kP2Warps : kP3Warps, kTile / ((q == 3) ? kP2Cells : kP3Cells));
  HostMesh mesh;
  if (!mesh.Build(degree, n_nodes, xyz, n_local, n_ghost, cell_nodes, n_bfaces, bface_nodes, bface_tags, 0, nullptr)) return -1;
  const int64_t nc = n_local + n_ghost, nt = (nc + kTile - 1) / kTile;
  if (gauss_points_out) {
    int last = -1;
    for (size_t b = 0; b < mesh.bound_faces.size(); ++b) {
      const int tg = mesh.bound_faces[b].tag;
      if (tg == last) continue;
      last = tg;
      mesh.BoundaryGaussPoints(c_tab, degree, tg, gauss_points_out + b * q * q * 3);
    }
  }
  if (quad_points_out) mesh.QuadraturePoints(c_tab, degree, quad_points_out);
  if (!column_out) return 0;
  auto geom = ToTiles(mesh.coef, kGeomRows, nc, nt);
  auto nbr = ToTiles(mesh.nbr, 6, nc, nt);
  auto code = ToTiles(mesh.code, 6, nc, nt);
  auto h2s = ToTiles(mesh.h2s, 18, nc, nt);
  auto bas = ToTiles(mesh.BuildModalBases(degree, true), q == 3 ? ModalShape<2>::kBasRows : ModalShape<3>::kBasRows, nc, nt);
  std::vector<unsigned char> bkind(mesh.bound_faces.size() + 1, 0);
  for (size_t b = 0; b < mesh.bound_faces.size(); ++b) bkind[b] = (unsigned char)tag_kinds[mesh.bound_faces[b].tag];
  std::vector<double> zero_vals((mesh.bound_faces.size() + 1) * q * q * 5, 0.0);
  std::vector<double> u((size_t)nt * F * kTile, 0.0), u0(u.size(), 0.0), out(u.size(), 0.0), tay(u.size(), 0.0);
  if (column_in) for (int64_t c = 0; c < nc; ++c) for (int f = 0; f < F; ++f) u[((c >> 5) * F + f) * kTile + (c & 31)] = column_in[c * F + f];
  if (column_u0) for (int64_t c = 0; c < n_local; ++c) for (int f = 0; f < F; ++f) u0[((c >> 5) * F + f) * kTile + (c & 31)] = column_u0[c * F + f];
  else u0 = u;
  DgmArgs a;
  a.u_in = u.data(); a.u0 = u0.data(); a.u_out = out.data(); a.tay = tay.data(); a.geom = geom.data(); a.bas = bas.data();
  a.h2s = h2s.data(); a.nbr = nbr.data(); a.code = code.data(); a.bkind = bkind.data();
  a.bvals = bvals ? bvals : zero_vals.data(); a.cell_begin = 0; a.cell_end = (int)n_local; a.mode = mode; a.dt = dt; a.dt_dev = nullptr;
  if (q == 3) RunDgm<3>(riemann, a, values, out.data(), n_local, nc, u, tay, geom.data(), bas.data());
  else RunDgm<4>(riemann, a, values, out.data(), n_local, nc, u, tay, geom.data(), bas.data());
  for (int64_t c = 0; c < n_local; ++c) for (int f = 0; f < F; ++f) column_out[c * F + f] = out[((c >> 5) * F + f) * kTile + (c & 31)];
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

// FormatG6 / FormatInteger (format.cuh) value by value: 16-byte token slots, lengths, and the "host must decide" flag
extern "C" {
void emu_format_g6(int64_t n, const double *values, char *slots, int *lens, unsigned char *unsure) {
  static std::vector<double> table;
  if (table.empty()) {
    table.resize(2 * (kPow10Max - kPow10Min + 1));
    BuildPow10Table(table.data());
  }
  for (int64_t i = 0; i < n; ++i) {
    const Token t = FormatG6(values[i], table.data());
    std::memcpy(slots + i * kTokenSlot, t.c, kTokenSlot);
    lens[i] = t.len;
    unsure[i] = t.unsure;
  }
}
void emu_format_integer(int64_t n, const int64_t *values, char *slots, int *lens) {
  for (int64_t i = 0; i < n; ++i) {
    const Token t = FormatInteger(values[i]);
    std::memcpy(slots + i * kTokenSlot, t.c, kTokenSlot);
    lens[i] = t.len;
  }
}
}  // extern "C"
