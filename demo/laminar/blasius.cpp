// demo/laminar/blasius.cpp — laminar boundary layer over a flat plate (BASELINE.json config 4) written against the
// template API of the reference's laminar drivers (demo/laminar/shockless.hpp:20-107 under FR,
// demo/laminar/blasius.cpp:7-74, demo/laminar/shockless.cpp), with the GPU path selected by the include directory:
//     g++ -std=c++20 -I include demo/laminar/blasius.cpp -L minicfd_b200 -lminicfd_b200
// Differences from the reference driver: the graded plate mesh of input/plate.py (0.2 x 0.02 x 0.1, 64 x 64 x 3 cells
// with first layers 1e-4 and 1e-5, coarsened by powers of two) is regenerated in memory from the same progression
// (input/thickness_ratio.py) instead of gmsh -> CGNS -> METIS; no JSON / CGNS input.
// Arguments: n_steps [column.bin [coarseness_x [coarseness_y [dt_max]]]].
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

using Scalar = double;
constexpr int kDimensions = 3;
#include "mini/constant/index.hpp"
#include "mini/riemann/euler/types.hpp"
#include "mini/riemann/euler/hllc.hpp"
#include "mini/riemann/rotated/euler.hpp"
#include "mini/riemann/concept.hpp"
#include "mini/riemann/diffusive/navier_stokes.hpp"
#include "mini/riemann/diffusive/direct.hpp"
using Primitive = mini::riemann::euler::Primitives<Scalar, kDimensions>;
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
using Projection = mini::polynomial::Hexahedron<Gx, Gx, Gx, kComponents, true>;   // shockless.hpp:66-69 (FR)
#include "mini/mesh/part.hpp"
using Part = mini::mesh::part::Part<int64_t, Projection>;
using Global = typename Part::Global;
using Value = typename Part::Value;
#include "mini/spatial/fr/lobatto.hpp"
using Spatial = mini::spatial::fr::Lobatto<Part, Riemann>;
#include "mini/temporal/rk.hpp"
constexpr int kOrders = 3;
using Temporal = mini::temporal::RungeKutta<kOrders, Scalar>;
#include "../box_mesh.hpp"

// demo/laminar/blasius.cpp:7-62 of the reference
Scalar mach_infty = 0.2;
Scalar total_pressure_infty = 104165;
Scalar total_temperature_infty = 290.304;
auto pressure_infty = Gas::TotalPressureToPressure(mach_infty, total_pressure_infty);  // 101300 Pa
auto temperature_infty = Gas::TotalTemperatureToTemperature(mach_infty, total_temperature_infty);
auto density_infty = pressure_infty / temperature_infty / Gas::R();
auto u_infty = mach_infty * Gas::GetSpeedOfSound(density_infty, pressure_infty);

Value MyIC(const Global &xyz) {
  auto x = xyz[mini::constant::index::X];
  auto y = xyz[mini::constant::index::Y];
  auto u_bottom = u_infty * (1 - x / 0.2);
  auto du_dy = (u_infty - u_bottom) / 0.02;
  auto u = du_dy * y + u_bottom;
  Primitive primitive;
  primitive.rho() = density_infty; primitive.u() = u; primitive.v() = 0; primitive.w() = 0; primitive.p() = pressure_infty;
  auto c = Gas::PrimitiveToConservative(primitive);
  Value value;
  for (int k = 0; k < 5; ++k) value[k] = c[k];
  return value;
}
auto outlet = [](const Global &, double) {
  Value value;
  for (int k = 0; k < 5; ++k) value[k] = 0;
  value[4] = pressure_infty;
  return value;
};
auto bottom = [](const Global &, double) {
  Value value;
  for (int k = 0; k < 5; ++k) value[k] = 0;   // wall velocity 0, value[4] = temperature gradient
  return value;
};
auto far_field = [](const Global &, double) {
  Primitive primitive;
  primitive.rho() = density_infty; primitive.u() = u_infty; primitive.v() = 0; primitive.w() = 0; primitive.p() = pressure_infty;
  auto c = Gas::PrimitiveToConservative(primitive);
  Value value;
  for (int k = 0; k < 5; ++k) value[k] = c[k];
  return value;
};

int main(int argc, char *argv[]) {
  int n_steps = argc > 1 ? std::atoi(argv[1]) : 20;
  const char *out_name = argc > 2 ? argv[2] : nullptr;
  int coarse_x = argc > 3 ? std::atoi(argv[3]) : 2, coarse_y = argc > 4 ? std::atoi(argv[4]) : 2;   // plate.py defaults
  double dt_max = argc > 5 ? std::atof(argv[5]) : 1e-7;
  // demo/laminar/blasius.cpp:70-72
  Riemann::Diffusion::SetBetaValues(2.0, 1.0 / 12);
  Riemann::Diffusion::SetProperty(/* nu = */1.57e-5, /* prandtl = */0.708);
  // input/plate.py:26-35: 64 layers with first thickness 1e-4 (x) and 1e-5 (y), coarsened by 2^c
  const double l_x = 0.2, l_y = 0.02, l_z = 0.1;
  const int n_x = 64 >> coarse_x, n_y = 64 >> coarse_y, n_z = 3;
  const double a_x = std::pow(demo::ThicknessRatio(0, l_x, 1e-4, 64), 1 << coarse_x);
  const double a_y = std::pow(demo::ThicknessRatio(0, l_y, 1e-5, 64), 1 << coarse_y);
  Part part(0, 1);
  demo::BuildBox(&part, demo::Layers(0, l_x, a_x, n_x), demo::Layers(0, l_y, a_y, n_y), demo::Uniform(0, l_z, n_z),
      [](int f, int, int, int) -> std::string {
        static const char *names[6] = {"Back", "Bottom", "Right", "Top", "Left", "Front"};
        return names[f];
      });
  Spatial spatial(&part);
  spatial.Approximate(MyIC);
  // demo/laminar/blasius.cpp:60-67
  spatial.SetSmartBoundary("Left", far_field);
  spatial.SetSubsonicOutlet("Right", outlet);
  spatial.SetSmartBoundary("Top", far_field);
  spatial.SetNoSlipWall("Bottom", bottom);
  spatial.SetInviscidWall("Back");
  spatial.SetInviscidWall("Front");
  Temporal temporal;
  double t = 0;
  for (int i = 0; i < n_steps; ++i) {   // demo/laminar/shockless.cpp time loop
    double dt = spatial.GetTimeStep(dt_max, kOrders);
    temporal.Update(&spatial, t, dt);
    t += dt;
  }
  auto column = spatial.GetSolutionColumn();
  double l1[5] = {0, 0, 0, 0, 0};
  for (size_t i = 0; i < column.size(); ++i) l1[i % 5] += std::abs(column[i]);
  std::printf("t = %.9e after %d steps; sum|u_k| = %.15e %.15e %.15e %.15e %.15e\n", t, n_steps, l1[0], l1[1], l1[2], l1[3], l1[4]);
  if (out_name && out_name[0]) {
    FILE *f = std::fopen(out_name, "wb");
    std::fwrite(column.data(), sizeof(double), column.size(), f);
    std::fclose(f);
  }
  return 0;
}
