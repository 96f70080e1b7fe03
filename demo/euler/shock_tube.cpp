// demo/euler/shock_tube.cpp — Sod shock tube (BASELINE.json config 1) written against the same template
// API the reference's driver uses (demo/euler/sourceless.hpp:15-111, demo/euler/shock_tube.cpp:5-66,
// demo/euler/sourceless.cpp:107-221), with the GPU path selected by the include directory alone:
//     g++ -std=c++20 -I include demo/euler/shock_tube.cpp -L minicfd_b200 -lminicfd_b200
// Differences from the reference driver: the mesh is a structured box built in memory instead of
// gmsh -> CGNS -> METIS (out of scope), and there is no JSON / CGNS / VTK I/O.
// Prints the solution column's L1 norm per component and writes the final column to a binary file.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

using Scalar = double;
constexpr int kDimensions = 3;
#include "mini/riemann/euler/types.hpp"
#include "mini/riemann/euler/exact.hpp"
#include "mini/riemann/euler/hllc.hpp"
#include "mini/riemann/rotated/euler.hpp"
using Primitive = mini::riemann::euler::Primitives<Scalar, kDimensions>;
using Conservative = mini::riemann::euler::Conservatives<Scalar, kDimensions>;
using Gas = mini::riemann::euler::IdealGas<Scalar, 1.4>;
#ifdef USE_HLLC
using Unrotated = mini::riemann::euler::HartenLaxLeerContact<Gas, kDimensions>;
#else
using Unrotated = mini::riemann::euler::Exact<Gas, kDimensions>;
#endif
using Riemann = mini::riemann::rotated::Euler<Unrotated>;

constexpr int kComponents = 5;
#ifndef DEGREE
#define DEGREE 2   // demo/euler/sourceless.hpp:27; -DDEGREE=3 for the p = 3 schemes
#endif
constexpr int kDegrees = DEGREE;
#ifdef DGFEM   // demo/euler/sourceless.hpp:28-33
#include "mini/integrator/legendre.hpp"
using Gx = mini::integrator::Legendre<Scalar, kDegrees + 1>;
#include "mini/polynomial/projection.hpp"
using Polynomial = mini::polynomial::Projection<Scalar, kDimensions, kDegrees, kComponents>;
#else          // demo/euler/sourceless.hpp:35-49
#include "mini/integrator/lobatto.hpp"
using Gx = mini::integrator::Lobatto<Scalar, kDegrees + 1>;
#include "mini/polynomial/hexahedron.hpp"
using Interpolation = mini::polynomial::Hexahedron<Gx, Gx, Gx, kComponents, false>;
#ifdef LIMITER
#include "mini/polynomial/extrapolation.hpp"
using Polynomial = mini::polynomial::Extrapolation<Interpolation>;
#else
using Polynomial = Interpolation;
#endif
#endif
#include "mini/mesh/part.hpp"
using Part = mini::mesh::part::Part<int64_t, Polynomial>;
using Cell = typename Part::Cell;
using Global = typename Part::Global;
using Value = typename Part::Value;
#include "mini/spatial/fr/lobatto.hpp"
#include "mini/spatial/dg/general.hpp"
#ifdef DGFEM     // demo/euler/sourceless.hpp:62-74
using Unlimited = mini::spatial::dg::General<Part, Riemann>;
#else
using Unlimited = mini::spatial::fr::Lobatto<Part, Riemann>;
#endif
#ifdef LIMITER   // demo/euler/sourceless.hpp:76-80
using General = Unlimited;
#include "mini/limiter/weno.hpp"
#include "mini/limiter/reconstruct.hpp"
#include "mini/spatial/with_limiter.hpp"
#ifdef EIGEN_LIMITER   // demo/euler/rotorcraft.hpp:41
using Limiter = mini::limiter::weno::Eigen<Cell, Riemann>;
#else
using Limiter = mini::limiter::weno::Lazy<Cell>;
#endif
using Spatial = mini::spatial::WithLimiter<General, Limiter>;
#elif defined(VISCOSITY)   // demo/euler/sourceless.hpp:82-103
#include "mini/riemann/concept.hpp"
#include "mini/riemann/diffusive/linear.hpp"
#include "mini/riemann/diffusive/direct.hpp"
#include "mini/spatial/viscosity.hpp"
#include "mini/spatial/with_viscosity.hpp"
#include "mini/limiter/average.hpp"
#include "mini/limiter/majority.hpp"
using Diffusion = mini::riemann::diffusive::Direct<mini::riemann::diffusive::Isotropic<Scalar, kComponents>>;
using RiemannWithViscosity = mini::spatial::EnergyBasedViscosity<Part,
    mini::riemann::ConvectionDiffusion<Riemann, Diffusion>>;
static_assert(mini::riemann::ConvectiveDiffusive<RiemannWithViscosity>);
using General = mini::spatial::fr::Lobatto<Part, RiemannWithViscosity>;
using Spatial = mini::spatial::WithViscosity<General>;
#else
using Spatial = Unlimited;
#endif
#include "mini/mesh/vtk.hpp"
using VtkWriter = mini::mesh::vtk::Writer<Part>;   // demo/euler/sourceless.cpp:38
#include "mini/temporal/rk.hpp"
constexpr int kOrders = std::min(3, kDegrees + 1);   // demo/euler/sourceless.hpp:110
using Temporal = mini::temporal::RungeKutta<kOrders, Scalar>;

// demo/euler/shock_tube.cpp:5-26 of the reference
static Value given_value_left, given_value_right;
Value MyIC(const Global &xyz) { return xyz[0] < 2.5 ? given_value_left : given_value_right; }
Value GivenLeft(const Global &, double) { return given_value_left; }
Value GivenRight(const Global &, double) { return given_value_right; }

static void BuildBox(Part *part, int nx, int ny, int nz, double lx, double ly, double lz) {
  std::vector<double> nodes;
  for (int k = 0; k <= nz; ++k) for (int j = 0; j <= ny; ++j) for (int i = 0; i <= nx; ++i) {
    nodes.push_back(lx * i / nx); nodes.push_back(ly * j / ny); nodes.push_back(lz * k / nz);
  }
  auto nid = [&](int i, int j, int k) { return (int64_t)i + (nx + 1) * (j + (int64_t)(ny + 1) * k); };
  std::vector<int64_t> cells, bfaces;
  std::vector<std::string> names;
  static const int fn[6][4] = {{0, 3, 2, 1}, {0, 1, 5, 4}, {1, 2, 6, 5}, {2, 3, 7, 6}, {0, 4, 7, 3}, {4, 5, 6, 7}};
  for (int k = 0; k < nz; ++k) for (int j = 0; j < ny; ++j) for (int i = 0; i < nx; ++i) {
    int64_t c[8] = {nid(i, j, k), nid(i + 1, j, k), nid(i + 1, j + 1, k), nid(i, j + 1, k),
                    nid(i, j, k + 1), nid(i + 1, j, k + 1), nid(i + 1, j + 1, k + 1), nid(i, j + 1, k + 1)};
    cells.insert(cells.end(), c, c + 8);
    auto add = [&](int f, const char *name) { for (int m = 0; m < 4; ++m) bfaces.push_back(c[fn[f][m]]); names.push_back(name); };
    if (i == 0) add(4, "Left");
    if (i == nx - 1) add(2, "Right");
    if (j == 0) add(1, "Front");
    if (j == ny - 1) add(3, "Back");
    if (k == 0) add(0, "Bottom");
    if (k == nz - 1) add(5, "Top");
  }
  part->SetMesh(std::move(nodes), std::move(cells), (int64_t)nx * ny * nz, 0, std::move(bfaces), names);
}

int main(int argc, char *argv[]) {
  int n_steps = argc > 1 ? std::atoi(argv[1]) : 20;
  const char *out_name = argc > 2 ? argv[2] : nullptr;
  const char *frame_dir = argc > 3 ? argv[3] : nullptr;   // where Frame0 / Frame<n_steps> go (.pvtu + .vtu)
  // input/shock_tube.geo:4-31: [0,5] x [0,0.2] x [0,0.3], h = 0.1
  Part part(0, 1);
  BuildBox(&part, 50, 2, 3, 5.0, 0.2, 0.3);
  part.SetFieldNames({"Density", "MomentumX", "MomentumY", "MomentumZ", "EnergyStagnationDensity"});   // sourceless.cpp:109-110
  if (frame_dir) part.SetDirectoryName(frame_dir);
#ifdef LIMITER   // demo/euler/sourceless.cpp:114-120
  auto limiter = Limiter(/* w0 = */0.001, /* eps = */1e-6);
  Spatial spatial(&limiter, &part);
#ifdef EIGEN_LIMITER   // demo/euler/rotorcraft.cpp:93-96
  using Face = typename Part::Face;
  auto face_to_riemanns = [&spatial](Face const &face) -> auto const & { return spatial.GetRiemannSolvers(face); };
  spatial.limiter_ptr()->InstallRiemannSolvers(face_to_riemanns);
#endif
#elif defined(VISCOSITY)   // demo/euler/sourceless.cpp:117-126
  Diffusion::SetProperty(0.0);
#ifdef AS_SHIPPED   // demo/euler/shock_tube_sod.json:6, 19-21
  Diffusion::SetBetaValues(/* ddg_beta_0 = */2.0, /* ddg_beta_1 = */0.0833333333);
  Spatial spatial(&part);
  RiemannWithViscosity::SetTimeScale(/* time_scale = */5.0e-2);
#else
  Diffusion::SetBetaValues(/* ddg_beta_0 = */2.0, /* ddg_beta_1 = */1.0 / 12);
  Spatial spatial(&part);
  RiemannWithViscosity::SetTimeScale(/* time_scale = */1.0);
#endif
  for (int k = 0; k < kComponents; ++k) {   // sourceless.cpp:122-126
    VtkWriter::AddCellData("CellViscosity" + std::to_string(k + 1),
        [k](Cell const &cell) { return RiemannWithViscosity::GetPropertyOnCell(cell.id(), 0)[k]; });
  }
#else
  Spatial spatial(&part);
#endif
  Primitive left, right;
  left.rho() = 1.0; left.u() = 0; left.v() = 0; left.w() = 0; left.p() = 1.0;
  right.rho() = 0.125; right.u() = 0; right.v() = 0; right.w() = 0; right.p() = 0.1;
  auto cl = Gas::PrimitiveToConservative(left), cr = Gas::PrimitiveToConservative(right);
  for (int k = 0; k < 5; ++k) { given_value_left[k] = cl[k]; given_value_right[k] = cr[k]; }
  spatial.Approximate(MyIC);
#ifdef VISCOSITY   // demo/euler/sourceless.cpp:145-153; "initial_limiter": "majority" in shock_tube_sod.json
#ifdef AS_SHIPPED
  mini::limiter::majority::Reconstruct(spatial.part_ptr());
#else
  mini::limiter::average::Reconstruct(spatial.part_ptr());
#endif
  RiemannWithViscosity::Viscosity::UpdateProperties();
#endif
  // demo/euler/shock_tube.cpp:56-66
  spatial.SetSmartBoundary("Left", GivenLeft);
  spatial.SetSmartBoundary("Right", GivenRight);
  for (auto name : {"Top", "Bottom", "Front", "Back"}) spatial.SetInviscidWall(name);
  if (frame_dir) VtkWriter::WriteSolutions(part, "Frame0");   // sourceless.cpp:161-162
  Temporal temporal;
  double t = 0, t_stop = 0.2, dt_max = t_stop / 200;   // sourceless.cpp:68-74 with 10 frames x 20 steps
  for (int i = 0; i < n_steps; ++i) {                  // sourceless.cpp:187-215
    double dt = spatial.GetTimeStep(dt_max, kOrders);
    temporal.Update(&spatial, t, dt);
    t += dt;
  }
  if (frame_dir) VtkWriter::WriteSolutions(part, "Frame" + std::to_string(n_steps));   // sourceless.cpp:219-221
  auto column = spatial.GetSolutionColumn();
  double l1[5] = {0, 0, 0, 0, 0};
  for (size_t i = 0; i < column.size(); ++i) l1[i % 5] += std::abs(column[i]);
  std::printf("t = %.6f after %d steps; sum|u_k| = %.15e %.15e %.15e %.15e %.15e\n", t, n_steps, l1[0], l1[1], l1[2], l1[3], l1[4]);
#ifdef VISCOSITY   // what the reference writes as CellViscosity1..5 (demo/euler/sourceless.cpp:122-126)
  double nu_max = 0;
  for (int64_t c = 0; c < part.CountLocalCells(); ++c) {
    for (int k = 0; k < kComponents; ++k) nu_max = std::max(nu_max, RiemannWithViscosity::GetPropertyOnCell(c, 0)[k]);
  }
  std::printf("max cell viscosity %.15e\n", nu_max);
#endif
  if (out_name) {
    FILE *f = std::fopen(out_name, "wb");
    std::fwrite(column.data(), sizeof(double), column.size(), f);
    std::fclose(f);
  }
  return 0;
}
