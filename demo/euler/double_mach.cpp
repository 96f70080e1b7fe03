// demo/euler/double_mach.cpp — double Mach reflection (BASELINE.json config 3) written against the template API the
// reference's driver uses (demo/euler/sourceless.hpp:15-111 under DGSEM + LIMITER, demo/euler/double_mach.cpp:5-52,
// demo/euler/sourceless.cpp:107-221), with the GPU path selected by the include directory alone:
//     g++ -std=c++20 -I include demo/euler/double_mach.cpp -L minicfd_b200 -lminicfd_b200
// Differences from the reference driver: the mesh of input/double_mach.geo ([0,4] x [0,1] x [0,0.2], h = 0.1) is a
// structured box built in memory, whose "Gap" section is the part of the bottom left of the grid line nearest to
// x = 1/6; no JSON / CGNS input.  Arguments: n_steps [column.bin [frame directory [cells per unit length]]].
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

using Scalar = double;
constexpr int kDimensions = 3;
#include "mini/riemann/euler/types.hpp"
#include "mini/riemann/euler/hllc.hpp"
#include "mini/riemann/rotated/euler.hpp"
using Primitive = mini::riemann::euler::Primitives<Scalar, kDimensions>;
using Conservative = mini::riemann::euler::Conservatives<Scalar, kDimensions>;
using Gas = mini::riemann::euler::IdealGas<Scalar, 1.4>;
using Unrotated = mini::riemann::euler::HartenLaxLeerContact<Gas, kDimensions>;
using Riemann = mini::riemann::rotated::Euler<Unrotated>;

constexpr int kComponents = 5;
constexpr int kDegrees = 2;
#include "mini/integrator/lobatto.hpp"
using Gx = mini::integrator::Lobatto<Scalar, kDegrees + 1>;
#include "mini/polynomial/hexahedron.hpp"
using Interpolation = mini::polynomial::Hexahedron<Gx, Gx, Gx, kComponents, false>;
#include "mini/polynomial/extrapolation.hpp"
using Polynomial = mini::polynomial::Extrapolation<Interpolation>;
#include "mini/mesh/part.hpp"
using Part = mini::mesh::part::Part<int64_t, Polynomial>;
using Cell = typename Part::Cell;
using Global = typename Part::Global;
using Value = typename Part::Value;
#include "mini/spatial/dg/lobatto.hpp"
using General = mini::spatial::dg::Lobatto<Part, Riemann>;   // DGSEM, demo/euler/sourceless.hpp:67-69
#include "mini/limiter/weno.hpp"
#include "mini/limiter/reconstruct.hpp"
#include "mini/spatial/with_limiter.hpp"
using Limiter = mini::limiter::weno::Lazy<Cell>;
using Spatial = mini::spatial::WithLimiter<General, Limiter>;
#include "mini/mesh/vtk.hpp"
using VtkWriter = mini::mesh::vtk::Writer<Part>;
#include "mini/temporal/rk.hpp"
constexpr int kOrders = 3;
using Temporal = mini::temporal::RungeKutta<kOrders, Scalar>;
#include "../box_mesh.hpp"

// demo/euler/double_mach.cpp:5-33 of the reference
constexpr double kGammaPlus = 2.4, kGamma = 1.4, kGammaMinus = 0.4;
double rho_before = 1.4, p_before = 1.0;
double m_before = 10.0, a_before = 1.0, u_nu = m_before * a_before;
auto rho_after = rho_before * (m_before * m_before * kGammaPlus / 2.0)
    / (1.0 + m_before * m_before * kGammaMinus / 2.0);  // 8.0
auto p_after = p_before * (m_before * m_before * kGamma - kGammaMinus / 2.0)
    / (kGammaPlus / 2.0);  // 116.5
auto u_n_after = u_nu * (rho_after - rho_before) / rho_after;  // 8.25
auto tan_60 = std::sqrt(3.0), cos_30 = tan_60 * 0.5, sin_30 = 0.5;
auto u_after = u_n_after * cos_30, v_after = u_n_after * (-sin_30);
static Value value_after, value_before;
double x_gap = 1.0 / 6.0;

Value MyIC(const Global &xyz) {
  auto x = xyz[0], y = xyz[1];
  return ((x - x_gap) * tan_60 < y) ? value_after : value_before;
}
auto u_x = u_nu / cos_30;
auto moving_shock = [](const Global &xyz, double t) {
  auto x = xyz[0], y = xyz[1];
  return ((x - (x_gap + u_x * t)) * tan_60 < y) ? value_after : value_before;
};

int main(int argc, char *argv[]) {
  int n_steps = argc > 1 ? std::atoi(argv[1]) : 20;
  const char *out_name = argc > 2 ? argv[2] : nullptr;
  const char *frame_dir = argc > 3 && argv[3][0] ? argv[3] : nullptr;
  int per_unit = argc > 4 ? std::atoi(argv[4]) : 10;   // input/double_mach.geo: LC = 0.1
  Primitive after, before;
  after.rho() = rho_after; after.u() = u_after; after.v() = v_after; after.w() = 0; after.p() = p_after;
  before.rho() = rho_before; before.u() = 0; before.v() = 0; before.w() = 0; before.p() = p_before;
  auto ca = Gas::PrimitiveToConservative(after), cb = Gas::PrimitiveToConservative(before);
  for (int k = 0; k < 5; ++k) { value_after[k] = ca[k]; value_before[k] = cb[k]; }

  Part part(0, 1);
  const int nx = 4 * per_unit, ny = per_unit, nz = 2;
  const int gap_cells = (int)std::lround(x_gap * per_unit);
  demo::BuildBox(&part, demo::Uniform(0, 4, nx), demo::Uniform(0, 1, ny), demo::Uniform(0, 2.0 / per_unit, nz),
      [&](int f, int i, int, int) -> std::string {
        switch (f) {
        case 0: return "Back";
        case 5: return "Front";
        case 4: return "Left";
        case 2: return "Right";
        case 3: return "Top";
        default: return i < gap_cells ? "Gap" : "Wall";
        }
      });
  part.SetFieldNames({"Density", "MomentumX", "MomentumY", "MomentumZ", "EnergyStagnationDensity"});
  if (frame_dir) part.SetDirectoryName(frame_dir);
  auto limiter = Limiter(/* w0 = */0.001, /* eps = */1e-6);   // demo/euler/sourceless.cpp:114-116
  Spatial spatial(&limiter, &part);
  spatial.Approximate(MyIC);
  mini::limiter::Reconstruct(&part, &limiter);                 // sourceless.cpp:141
  // demo/euler/double_mach.cpp:35-52
  spatial.SetSupersonicInlet("Top", moving_shock);
  spatial.SetSupersonicInlet("Left", moving_shock);
  spatial.SetInviscidWall("Back");
  spatial.SetInviscidWall("Front");
  spatial.SetInviscidWall("Wall");
  spatial.SetSupersonicOutlet("Right");
  spatial.SetSupersonicOutlet("Gap");
  if (frame_dir) VtkWriter::WriteSolutions(part, "Frame0");
  Temporal temporal;
  double t = 0, t_stop = 0.2, dt_max = t_stop / 2000;
  for (int i = 0; i < n_steps; ++i) {                          // sourceless.cpp:187-215
    double dt = spatial.GetTimeStep(dt_max, kOrders);
    temporal.Update(&spatial, t, dt);
    t += dt;
  }
  if (frame_dir) VtkWriter::WriteSolutions(part, "Frame" + std::to_string(n_steps));
  auto column = spatial.GetSolutionColumn();
  double l1[5] = {0, 0, 0, 0, 0};
  for (size_t i = 0; i < column.size(); ++i) l1[i % 5] += std::abs(column[i]);
  std::printf("t = %.6f after %d steps; sum|u_k| = %.15e %.15e %.15e %.15e %.15e\n", t, n_steps, l1[0], l1[1], l1[2], l1[3], l1[4]);
  if (out_name && out_name[0]) {
    FILE *f = std::fopen(out_name, "wb");
    std::fwrite(column.data(), sizeof(double), column.size(), f);
    std::fclose(f);
  }
  return 0;
}
