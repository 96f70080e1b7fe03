// The part of mini::mesh::part::Part's surface a driver or a callback touches (part.hpp:186-291, 1052-1069,
// 1330-1409), printed for tests/test_gpu_cpp_shim.py::test_part_surface.
#include <cstdio>
#include <string>
#include <vector>

using Scalar = double;
#include "mini/integrator/lobatto.hpp"
#include "mini/polynomial/hexahedron.hpp"
#include "mini/mesh/part.hpp"
#include "mini/riemann/euler/types.hpp"
#include "mini/riemann/euler/hllc.hpp"
#include "mini/riemann/rotated/euler.hpp"
#include "mini/spatial/fr/lobatto.hpp"
#include "../../demo/box_mesh.hpp"

using Gx = mini::integrator::Lobatto<Scalar, 3>;
using Polynomial = mini::polynomial::Hexahedron<Gx, Gx, Gx, 5, false>;
using Part = mini::mesh::part::Part<int64_t, Polynomial>;
using Gas = mini::riemann::euler::IdealGas<Scalar, 1.4>;
using Riemann = mini::riemann::rotated::Euler<mini::riemann::euler::HartenLaxLeerContact<Gas, 3>>;
using Spatial = mini::spatial::fr::Lobatto<Part, Riemann>;

static const char *Name(int face, int, int, int) {
  static const char *names[6] = {"Bottom", "Front", "Right", "Back", "Left", "Top"};
  return names[face];
}

int main() {
  auto part = Part(0, 1);
  // a stretched box: cells of different size
  std::vector<double> xs = {0, 0.5, 1.5, 3.0, 3.5}, ys = {0, 1, 3}, zs = {0, 0.25, 0.5, 1.5};
  demo::BuildBox(&part, xs, ys, zs, Name);
  Spatial spatial(&part);
  spatial.Approximate([](Part::Global const &x) {
    Part::Value v; v[0] = 1 + x[0]; v[1] = x[1]; v[2] = x[2]; v[3] = 0; v[4] = 3 + x[0] * x[1]; return v; });
  std::printf("local %ld ghost %ld inner %zu inter %zu\n", (long)part.CountLocalCells(), (long)part.CountGhostCells(),
      part.GetInnerCellPointers().size(), part.GetInterCellPointers().size());
  for (auto const &cell : part.GetLocalCells()) {
    std::printf("cell %ld %ld %.17g %.17g %.17g %.17g %.17g %d %d\n", (long)cell.id(), (long)cell.metis_id, cell.center()[0],
        cell.center()[1], cell.center()[2], cell.volume(), cell.length(), (int)cell.inner(), cell.CountFields());
  }
  for (auto name : {"Left", "Right", "Front", "Back", "Bottom", "Top"}) {
    double area = 0, cx = 0;
    for (auto const &face : part.GetBoundaryFaces(name)) { area += face.area(); cx += face.center()[0] * face.area(); (void)face.holder().id(); }
    std::printf("section %s %zu %.17g %.17g\n", name, part.GetBoundaryFaces(name).size(), area, cx / area);
  }
  auto l1 = part.MeasureL1Error([](Part::Global const &x, double t) {   // part.hpp:1023-1037
    Part::Value v; v[0] = 1 + x[0] + 0.5; v[1] = x[1]; v[2] = x[2] - t; v[3] = 0; v[4] = 3 + x[0] * x[1]; return v; }, 0.25);
  std::printf("l1 %.17g %.17g %.17g %.17g %.17g\n", l1[0], l1[1], l1[2], l1[3], l1[4]);
  // Gather -> change on the host -> Scatter -> the column shows it (part.hpp:1052-1069)
  part.GatherSolutions();
  std::printf("coeff %.17g %.17g\n", part.GetLocalCells()[3].coeff()[0], part.cell_data()[part.GetCellDataOffset(3)]);
  for (auto &v : part.cell_data()) v *= 2;
  part.ScatterSolutions();
  auto column = spatial.GetSolutionColumn();
  std::printf("scattered %.17g %.17g\n", column[part.GetCellDataOffset(3)], 2 * part.GetLocalCells()[3].coeff()[0] / 2);
  return 0;
}
