// demo/euler/shock_tube.cpp — Sod shock tube (BASELINE.json config 1) written against the same template
// API the reference's driver uses (demo/euler/sourceless.hpp:15-111, demo/euler/shock_tube.cpp:5-66,
// demo/euler/sourceless.cpp:107-221), with the GPU path selected by the include directory alone:
//     g++ -std=c++20 -I include -I include/mini/b200/nompi demo/euler/shock_tube.cpp -L minicfd_b200 -lminicfd_b200
// (the second -I only where there is no MPI: it maps the driver's handful of MPI calls onto shared memory).
// Differences from the reference driver: the mesh is a structured box built in memory instead of
// gmsh -> CGNS -> METIS (out of scope), and there is no JSON / CGNS I/O.
// One process per part, like the reference:  MCFD_SESSION=s MCFD_WORLD_SIZE=2 MCFD_RANK=<r> ./shock_tube ...
// cuts the box into MCFD_WORLD_SIZE slabs along x; the ghost cells' coefficients travel between the GPUs behind
// Part::ShareGhostCellCoeffs / UpdateGhostCellCoeffs, the time step through MPI_Allreduce(MIN) as in sourceless.cpp:194.
// Prints the solution column's L1 norm per component and writes the final column to a binary file.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "mpi.h"
#include "../box_mesh.hpp"

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

static const char *SectionName(int face, int, int, int) {   // input/shock_tube.geo:33-38
  static const char *names[6] = {"Bottom", "Front", "Right", "Back", "Left", "Top"};
  return names[face];
}

int main(int argc, char *argv[]) {
  int n_steps = argc > 1 ? std::atoi(argv[1]) : 20;
  const char *out_name = argc > 2 ? argv[2] : nullptr;
  const char *frame_dir = argc > 3 ? argv[3] : nullptr;   // where Frame0 / Frame<n_steps> go (.pvtu + .vtu)
  MPI_Init(NULL, NULL);                       // sourceless.cpp:60-66
  int n_core, i_core;
  MPI_Comm_size(MPI_COMM_WORLD, &n_core);
  MPI_Comm_rank(MPI_COMM_WORLD, &i_core);
  // input/shock_tube.geo:4-31: [0,5] x [0,0.2] x [0,0.3], h = 0.1
  auto part = Part(i_core, n_core);           // sourceless.cpp:107, the mesh handed over in memory
  {
    auto xs = demo::Uniform(0, 5.0, 50), ys = demo::Uniform(0, 0.2, 2), zs = demo::Uniform(0, 0.3, 3);
    if (n_core == 1) demo::BuildBox(&part, xs, ys, zs, SectionName);
    else demo::BuildBoxPart(&part, xs, ys, zs, SectionName, n_core, 1, 1);
  }
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
    double dt_local = spatial.GetTimeStep(dt_max, kOrders);
    double dt;  // i.e. dt_global
    MPI_Allreduce(&dt_local, &dt, 1, MPI_DOUBLE, MPI_MIN, MPI_COMM_WORLD);
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
  if (out_name && n_core == 1) {
    FILE *f = std::fopen(out_name, "wb");
    std::fwrite(column.data(), sizeof(double), column.size(), f);
    std::fclose(f);
  } else if (out_name) {   // <out>.<rank>: the local cells' ids before partitioning, then their coefficients
    part.GatherSolutions();
    FILE *f = std::fopen((std::string(out_name) + "." + std::to_string(i_core)).c_str(), "wb");
    int64_t n = part.CountLocalCells();
    std::fwrite(&n, sizeof(n), 1, f);
    for (auto const &cell : part.GetLocalCells()) { int64_t g = cell.metis_id; std::fwrite(&g, sizeof(g), 1, f); }
    std::fwrite(part.cell_data().data(), sizeof(double), part.cell_data().size(), f);
    std::fclose(f);
  }
  MPI_Finalize();                             // sourceless.cpp:227
  return 0;
}
