// demo/laminar/cavity.cpp — lid-driven cavity written against the template API of the reference's laminar
// drivers (demo/laminar/shockless.hpp:20-107, demo/laminar/cavity.cpp:5-53, demo/laminar/shockless.cpp),
// with the GPU path selected by the include directory alone:
//     g++ -std=c++20 -I include demo/laminar/cavity.cpp -L minicfd_b200 -lminicfd_b200
// Differences from the reference driver: the mesh is a structured box built in memory (gmsh -> CGNS -> METIS is
// out of scope) and there is no JSON / CGNS / VTK I/O.  -DKLOCAL_FALSE selects the other expansion.
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

using Scalar = double;
constexpr int kDimensions = 3;
#include "mini/riemann/euler/types.hpp"
#include "mini/riemann/euler/hllc.hpp"
#include "mini/riemann/rotated/euler.hpp"
#include "mini/riemann/concept.hpp"
#include "mini/riemann/diffusive/navier_stokes.hpp"
#include "mini/riemann/diffusive/direct.hpp"
using Gas = mini::riemann::euler::IdealGas<Scalar, 1.4>;
using Unrotated = mini::riemann::euler::HartenLaxLeerContact<Gas, kDimensions>;
using Convection = mini::riemann::rotated::Euler<Unrotated>;
using NavierStokes = mini::riemann::diffusive::NavierStokes<Gas>;
using Diffusion = mini::riemann::diffusive::Direct<NavierStokes>;
using Riemann = mini::riemann::ConvectionDiffusion<Convection, Diffusion>;

constexpr int kDegrees = 2;
constexpr int kComponents = Riemann::kComponents;
#include "mini/integrator/lobatto.hpp"
using Gx = mini::integrator::Lobatto<Scalar, kDegrees + 1>;
#include "mini/polynomial/hexahedron.hpp"
#ifdef KLOCAL_FALSE
using Projection = mini::polynomial::Hexahedron<Gx, Gx, Gx, kComponents, false>;
#else   // demo/laminar/shockless.hpp:66-69: FR uses the kLocal == true expansion (coefficients = det(J) u)
using Projection = mini::polynomial::Hexahedron<Gx, Gx, Gx, kComponents, true>;
#endif
#include "mini/mesh/part.hpp"
using Part = mini::mesh::part::Part<int64_t, Projection>;
using Global = typename Part::Global;
using Value = typename Part::Value;
#include "mini/spatial/fr/lobatto.hpp"
using Spatial = mini::spatial::fr::Lobatto<Part, Riemann>;
#include "mini/temporal/rk.hpp"
constexpr int kOrders = 3;
using Temporal = mini::temporal::RungeKutta<kOrders, Scalar>;

// demo/laminar/cavity.cpp:7-38 of the reference (lid speed and pressure scaled to a Mach 0.2 lid)
static const Scalar rho_given = 1.0, u_given = 1.0, p_given = 1.0 / (1.4 * 0.2 * 0.2);
Value MyIC(const Global &) {
  Value value;
  value[0] = rho_given; value[1] = 0; value[2] = 0; value[3] = 0; value[4] = p_given / 0.4;
  return value;
}
Value Moving(const Global &, double) {
  Value value;
  value[0] = 0; value[1] = u_given; value[2] = 0; value[3] = 0; value[4] = 0;   // wall velocity, dT/dn
  return value;
}
Value Fixed(const Global &, double) {
  Value value;
  value[0] = 0; value[1] = 0; value[2] = 0; value[3] = 0; value[4] = 0;
  return value;
}

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
    if (j == 0) add(1, "Bottom");
    if (j == ny - 1) add(3, "Top");
    if (k == 0) add(0, "Back");
    if (k == nz - 1) add(5, "Front");
  }
  part->SetMesh(std::move(nodes), std::move(cells), (int64_t)nx * ny * nz, 0, std::move(bfaces), names);
}

int main(int argc, char *argv[]) {
  int n_steps = argc > 1 ? std::atoi(argv[1]) : 20;
  const char *out_name = argc > 2 ? argv[2] : nullptr;
  int n = argc > 3 ? std::atoi(argv[3]) : 8;
  // demo/laminar/cavity.cpp:49-52
  Riemann::Diffusion::SetBetaValues(2.0, 1.0 / 12);
  Riemann::Diffusion::SetProperty(/* nu = */0.01, /* prandtl = */0.708);
  Part part(0, 1);
  BuildBox(&part, n, n, 2, 1.0, 1.0, 0.25);
  Spatial spatial(&part);
  spatial.Approximate(MyIC);
  // demo/laminar/cavity.cpp:40-47
  spatial.SetNoSlipWall("Left", Fixed);
  spatial.SetNoSlipWall("Right", Fixed);
  spatial.SetNoSlipWall("Top", Moving);
  spatial.SetNoSlipWall("Bottom", Fixed);
  spatial.SetInviscidWall("Back");
  spatial.SetInviscidWall("Front");
  Temporal temporal;
  double t = 0, dt_max = 1e-3;
  for (int i = 0; i < n_steps; ++i) {   // demo/laminar/shockless.cpp time loop
    double dt = spatial.GetTimeStep(dt_max, kOrders);
    temporal.Update(&spatial, t, dt);
    t += dt;
  }
  auto column = spatial.GetSolutionColumn();
  double l1[5] = {0, 0, 0, 0, 0};
  for (size_t i = 0; i < column.size(); ++i) l1[i % 5] += std::abs(column[i]);
  std::printf("t = %.6f after %d steps; sum|u_k| = %.15e %.15e %.15e %.15e %.15e\n", t, n_steps, l1[0], l1[1], l1[2], l1[3], l1[4]);
  if (out_name) {
    FILE *f = std::fopen(out_name, "wb");
    std::fwrite(column.data(), sizeof(double), column.size(), f);
    std::fclose(f);
  }
  return 0;
}
