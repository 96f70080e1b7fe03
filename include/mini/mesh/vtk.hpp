// Same include path as the reference header include/mini/mesh/vtk.hpp: `mini::mesh::vtk::Writer<Part>` with the
// reference's static interface (vtk.hpp:375-668) on the B200 build of Part.  The values at the 27 points of every
// cell's VTK tri-quadratic hexahedron are sampled on the GPU (mcfd_sample_vtk); this header lays them out as the
// reference's .vtu piece per part and the .pvtu index of rank 0: same arrays in the same order, same number format.
#ifndef MINI_B200_MESH_VTK_HPP_
#define MINI_B200_MESH_VTK_HPP_

#include <bit>
#include <cstdio>
#include <fstream>
#include <functional>
#include <string>
#include <utility>
#include <vector>

#include "mini/b200/shim.hpp"

namespace mini {
namespace mesh {
namespace vtk {

// RFC 4648 base64 of a byte range (the reference's EncodeBase64, vtk.hpp:61-98, produces the same text)
inline std::string EncodeBase64(const void *data, std::size_t n_byte) {
  static const char digits[] = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/";
  auto *in = static_cast<const unsigned char *>(data);
  std::string out;
  out.reserve((n_byte + 2) / 3 * 4);
  for (std::size_t i = 0; i < n_byte; i += 3) {
    const std::size_t left = n_byte - i;
    const unsigned b0 = in[i], b1 = left > 1 ? in[i + 1] : 0, b2 = left > 2 ? in[i + 2] : 0;
    out.push_back(digits[b0 >> 2]);
    out.push_back(digits[((b0 & 3) << 4) | (b1 >> 4)]);
    out.push_back(left > 1 ? digits[((b1 & 15) << 2) | (b2 >> 6)] : '=');
    out.push_back(left > 2 ? digits[b2 & 63] : '=');
  }
  return out;
}

enum class CellType { kHexahedron27 = 29 };   // vtk.hpp:127-139; the type WriteSolutions picks for 8-corner cells

template <typename Part>
class Writer {
 public:
  using Cell = typename Part::Cell;
  using Value = typename Part::Value;
  using Coord = typename Part::Global;
  using Scalar = typename Part::Scalar;
  using PointData = std::pair<std::string, std::function<Scalar(Cell const &, Coord const &, Value const &)>>;
  using CellData = std::pair<std::string, std::function<Scalar(Cell const &)>>;
  using ShiftByValue = std::function<void(Coord *, Value const &)>;

  template <typename Func>
  static void InstallShiftByValue(Func &&func) { shift_by_value_ = func; }          // vtk.hpp:390-393
  template <class F>
  static void AddPointData(std::string const &name, F &&f) { point_data_.emplace_back(name, std::forward<F>(f)); }
  template <class F>
  static void AddCellData(std::string const &name, F &&f) { cell_data_.emplace_back(name, std::forward<F>(f)); }
  static bool LittleEndian() { return std::endian::native == std::endian::little; }

  // vtk.hpp:526-668.  `modal` is what the reference gets from the polynomial type: Extrapolation and Projection
  // evaluate through their orthonormal projection, Hexahedron through its nodal basis.
  static void WriteSolutions(const Part &part, std::string const &soln_name) {
    mcfd_ctx *ctx = part.context_;
    if (!ctx) throw std::runtime_error("vtk::Writer: no spatial object has been built on this part");
    constexpr int K = Part::kComponents, kNodes = 27;
    const size_t n_cells = (size_t)part.CountLocalCells(), n_points = n_cells * kNodes;
    using Polynomial = typename Part::Polynomial;
    constexpr bool kModal = requires { Polynomial::kModal; } || requires { Polynomial::kModalOnly; };
    // the point values are only brought to the host when a callback wants to see them
    const bool need_values = !point_data_.empty() || static_cast<bool>(shift_by_value_);
    std::vector<double> coords(n_points * 3), values(need_values ? n_points * K : 0);
    b200::Check(mcfd_sample_vtk(ctx, kModal ? 1 : 0, coords.data(), need_values ? values.data() : nullptr), ctx,
        "mcfd_sample_vtk");
    std::vector<std::vector<Scalar>> point_data(point_data_.size()), cell_data(cell_data_.size());
    for (size_t c = 0; c < n_cells; ++c) {   // Prepare (vtk.hpp:421-494) on the sampled values
      Cell cell;
      cell.id_ = (typename Part::Index)c;
      for (int i = 0; need_values && i < kNodes; ++i) {
        const size_t at = c * kNodes + i;
        Coord global;
        Value value;
        for (int d = 0; d < 3; ++d) global[d] = coords[at * 3 + d];
        for (int k = 0; k < K; ++k) value[k] = values[at * K + k];
        if (shift_by_value_) {
          shift_by_value_(&global, value);
          for (int d = 0; d < 3; ++d) coords[at * 3 + d] = global[d];
        }
        for (size_t k = 0; k < point_data_.size(); ++k) point_data[k].push_back(point_data_[k].second(cell, global, value));
      }
      for (size_t k = 0; k < cell_data_.size(); ++k) cell_data[k].push_back(cell_data_[k].second(cell));
    }
    const char *endianness = LittleEndian() ? "\"LittleEndian\"" : "\"BigEndian\"";
    if (part.mpi_rank() == 0) {
      std::ofstream pvtu(part.GetDirectoryName() + "/" + soln_name + ".pvtu", std::ios::out);
      pvtu << "<VTKFile type=\"PUnstructuredGrid\" version=\"1.0\" byte_order=" << endianness << " header_type=\"UInt64\">\n";
      pvtu << "  <PUnstructuredGrid GhostLevel=\"1\">\n    <PPointData>\n";
      for (int k = 0; k < K; ++k) pvtu << "      <PDataArray type=\"Float64\" Name=\"" << part.GetFieldName(k) << "\"/>\n";
      for (auto &nf : point_data_) pvtu << "      <PDataArray type=\"Float64\" Name=\"" << nf.first << "\"/>\n";
      pvtu << "    </PPointData>\n    <PCellData>\n";
      for (auto &nf : cell_data_) pvtu << "      <PDataArray type=\"Float64\" Name=\"" << nf.first << "\"/>\n";
      pvtu << "    </PCellData>\n    <PPoints>\n";
      pvtu << "      <PDataArray type=\"Float64\" Name=\"Points\" NumberOfComponents=\"3\"/>\n    </PPoints>\n";
      for (int r = 0; r < part.mpi_size(); ++r) pvtu << "    <Piece Source=\"./" << soln_name << '/' << r << ".vtu\"/>\n";
      pvtu << "  </PUnstructuredGrid>\n</VTKFile>\n";
    }
    auto vtu = part.GetFileStream(soln_name, false, "vtu");
    // every number becomes text on the GPU (`vtu << v << ' '` = "%g "); the host only writes bytes
    std::string text;
    auto emit = [&](std::string const &type, std::string const &name, int64_t count, auto &&fill) {
      text.resize((size_t)count * 16 + 16);
      int64_t length = 0;
      b200::Check(fill(text.data(), (int64_t)text.size(), &length), ctx, "formatting a DataArray");
      vtu << "        <DataArray type=\"" << type << "\" Name=\"" << name << "\" format=\"ascii\">\n";
      vtu.write(text.data(), length);
      vtu << "\n        </DataArray>\n";
    };
    auto array = [&](std::string const &name, std::vector<Scalar> const &data) {
      emit("Float64", name, (int64_t)data.size(), [&](char *t, int64_t cap, int64_t *len) {
        return mcfd_format_doubles(ctx, (int64_t)data.size(), data.data(), t, cap, len); });
    };
    vtu << "<VTKFile type=\"UnstructuredGrid\" version=\"1.0\" byte_order=" << endianness << " header_type=\"UInt64\">\n";
    vtu << "  <UnstructuredGrid>\n";
    vtu << "    <Piece NumberOfPoints=\"" << n_points << "\" NumberOfCells=\"" << n_cells << "\">\n";
    vtu << "      <PointData>\n";
    for (int k = 0; k < K; ++k) {
      emit("Float64", part.GetFieldName(k), (int64_t)n_points, [&](char *t, int64_t cap, int64_t *len) {
        return mcfd_vtk_component_text(ctx, k, t, cap, len); });
    }
    for (size_t k = 0; k < point_data_.size(); ++k) array(point_data_[k].first, point_data[k]);
    vtu << "      </PointData>\n      <CellData>\n";
    for (size_t k = 0; k < cell_data_.size(); ++k) array(cell_data_[k].first, cell_data[k]);
    vtu << "      </CellData>\n      <Points>\n";
    vtu << "        <DataArray type=\"Float64\" Name=\"Points\" NumberOfComponents=\"3\" format=\"binary\">\n";
    const std::uint64_t n_byte = coords.size() * sizeof(double);
    vtu << EncodeBase64(&n_byte, sizeof n_byte) << EncodeBase64(coords.data(), n_byte);
    vtu << "\n        </DataArray>\n      </Points>\n      <Cells>\n";
    auto integers = [&](const char *type, const char *name, int64_t count, int64_t first, int64_t step) {
      emit(type, name, count, [&](char *t, int64_t cap, int64_t *len) {
        return mcfd_format_integers(ctx, count, first, step, t, cap, len); });
    };
    integers("Int32", "connectivity", (int64_t)n_points, 0, 1);
    integers("Int32", "offsets", (int64_t)n_cells, kNodes, kNodes);
    integers("UInt8", "types", (int64_t)n_cells, static_cast<int>(CellType::kHexahedron27), 0);
    vtu << "      </Cells>\n    </Piece>\n  </UnstructuredGrid>\n</VTKFile>\n";
  }

 private:
  inline static std::vector<PointData> point_data_;
  inline static std::vector<CellData> cell_data_;
  inline static ShiftByValue shift_by_value_;
};

}  // namespace vtk
}  // namespace mesh
}  // namespace mini

#endif  // MINI_B200_MESH_VTK_HPP_
