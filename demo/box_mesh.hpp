// In-memory stand-in for gmsh -> CGNS -> METIS (out of scope): a structured hexahedral box with given
// node coordinates along each axis, its boundary quads carrying the section names a callback chooses.
#ifndef DEMO_BOX_MESH_HPP_
#define DEMO_BOX_MESH_HPP_

#include <algorithm>
#include <cmath>
#include <map>
#include <cstdint>
#include <string>
#include <utility>
#include <vector>

namespace demo {

// local faces of a CGNS HEXA_8: 0 = z-, 1 = y-, 2 = x+, 3 = y+, 4 = x-, 5 = z+
// name_of(face, i, j, k) names the boundary section of that face of cell (i, j, k)
template <class Part, class NameOf>
void BuildBox(Part *part, std::vector<double> const &xs, std::vector<double> const &ys, std::vector<double> const &zs,
    NameOf &&name_of) {
  const int nx = (int)xs.size() - 1, ny = (int)ys.size() - 1, nz = (int)zs.size() - 1;
  std::vector<double> nodes;
  for (int k = 0; k <= nz; ++k) for (int j = 0; j <= ny; ++j) for (int i = 0; i <= nx; ++i) {
    nodes.push_back(xs[i]); nodes.push_back(ys[j]); nodes.push_back(zs[k]);
  }
  auto nid = [&](int i, int j, int k) { return (int64_t)i + (nx + 1) * (j + (int64_t)(ny + 1) * k); };
  std::vector<int64_t> cells, bfaces;
  std::vector<std::string> names;
  static const int fn[6][4] = {{0, 3, 2, 1}, {0, 1, 5, 4}, {1, 2, 6, 5}, {2, 3, 7, 6}, {0, 4, 7, 3}, {4, 5, 6, 7}};
  for (int k = 0; k < nz; ++k) for (int j = 0; j < ny; ++j) for (int i = 0; i < nx; ++i) {
    int64_t c[8] = {nid(i, j, k), nid(i + 1, j, k), nid(i + 1, j + 1, k), nid(i, j + 1, k),
                    nid(i, j, k + 1), nid(i + 1, j, k + 1), nid(i + 1, j + 1, k + 1), nid(i, j + 1, k + 1)};
    cells.insert(cells.end(), c, c + 8);
    auto add = [&](int f) {
      for (int m = 0; m < 4; ++m) bfaces.push_back(c[fn[f][m]]);
      names.push_back(name_of(f, i, j, k));
    };
    if (k == 0) add(0);
    if (j == 0) add(1);
    if (i == nx - 1) add(2);
    if (j == ny - 1) add(3);
    if (i == 0) add(4);
    if (k == nz - 1) add(5);
  }
  part->SetMesh(std::move(nodes), std::move(cells), (int64_t)nx * ny * nz, 0, std::move(bfaces), names);
}

// The same box cut into px x py x pz blocks, this rank's block as a part the way the reference's Part numbers one
// (part.hpp:746-775): local cells in the order inner (no neighbour in another part), then inter, each by ascending
// global id; then the ghost cells — the face neighbours owned by other parts — grouped by owner.  Send and receive
// lists name the same cells in the same order on both sides (ascending global id).  Stand-in for
// `mesh::Shuffler::PartitionAndShuffle` + `partition/<rank>.txt` (shuffler.hpp:436-506).
template <class Part, class NameOf>
void BuildBoxPart(Part *part, std::vector<double> const &xs, std::vector<double> const &ys, std::vector<double> const &zs,
    NameOf &&name_of, int px, int py, int pz) {
  const int nx = (int)xs.size() - 1, ny = (int)ys.size() - 1, nz = (int)zs.size() - 1;
  const int rank = part->mpi_rank();
  std::vector<double> nodes;
  for (int k = 0; k <= nz; ++k) for (int j = 0; j <= ny; ++j) for (int i = 0; i <= nx; ++i) {
    nodes.push_back(xs[i]); nodes.push_back(ys[j]); nodes.push_back(zs[k]);
  }
  auto nid = [&](int i, int j, int k) { return (int64_t)i + (nx + 1) * (j + (int64_t)(ny + 1) * k); };
  auto gid = [&](int i, int j, int k) { return (int64_t)i + nx * (j + (int64_t)ny * k); };
  auto owner = [&](int i, int j, int k) { return (i * px / nx) + px * ((j * py / ny) + py * (k * pz / nz)); };
  static const int step[6][3] = {{0, 0, -1}, {0, -1, 0}, {1, 0, 0}, {0, 1, 0}, {-1, 0, 0}, {0, 0, 1}};   // faces 0..5
  std::vector<int64_t> inner, inter;
  std::map<int, std::vector<int64_t>> send_to, recv_from;   // by part, ascending global id
  for (int k = 0; k < nz; ++k) for (int j = 0; j < ny; ++j) for (int i = 0; i < nx; ++i) {
    if (owner(i, j, k) != rank) continue;
    bool touches = false;
    for (int f = 0; f < 6; ++f) {
      const int a = i + step[f][0], b = j + step[f][1], c = k + step[f][2];
      if (a < 0 || a >= nx || b < 0 || b >= ny || c < 0 || c >= nz) continue;
      const int q = owner(a, b, c);
      if (q == rank) continue;
      touches = true;
      auto &s = send_to[q];
      if (s.empty() || s.back() != gid(i, j, k)) s.push_back(gid(i, j, k));
      recv_from[q].push_back(gid(a, b, c));
    }
    (touches ? inter : inner).push_back(gid(i, j, k));
  }
  std::vector<int64_t> order = inner;
  order.insert(order.end(), inter.begin(), inter.end());
  const int64_t n_local = (int64_t)order.size();
  typename Part::Halo halo;
  for (auto &kv : recv_from) {
    auto &r = kv.second;
    std::sort(r.begin(), r.end());
    r.erase(std::unique(r.begin(), r.end()), r.end());
    halo.parts.push_back(kv.first);
    halo.recv_counts.push_back((int64_t)r.size());
    for (int64_t g : r) { halo.recv_cells.push_back((int64_t)order.size()); order.push_back(g); }
  }
  std::map<int64_t, int64_t> local_of;
  for (int64_t c = 0; c < (int64_t)order.size(); ++c) local_of[order[c]] = c;
  for (auto &kv : send_to) {   // std::map: the same part order as recv_from (a face neighbour relation is symmetric)
    std::sort(kv.second.begin(), kv.second.end());
    halo.send_counts.push_back((int64_t)kv.second.size());
    for (int64_t g : kv.second) halo.send_cells.push_back(local_of[g]);
  }
  std::vector<int64_t> cells, bfaces;
  std::vector<std::string> names;
  static const int fn[6][4] = {{0, 3, 2, 1}, {0, 1, 5, 4}, {1, 2, 6, 5}, {2, 3, 7, 6}, {0, 4, 7, 3}, {4, 5, 6, 7}};
  for (int64_t idx = 0; idx < (int64_t)order.size(); ++idx) {
    const int64_t g = order[idx];
    const int i = (int)(g % nx), j = (int)((g / nx) % ny), k = (int)(g / ((int64_t)nx * ny));
    int64_t c[8] = {nid(i, j, k), nid(i + 1, j, k), nid(i + 1, j + 1, k), nid(i, j + 1, k),
                    nid(i, j, k + 1), nid(i + 1, j, k + 1), nid(i + 1, j + 1, k + 1), nid(i, j + 1, k + 1)};
    cells.insert(cells.end(), c, c + 8);
    if (idx >= n_local) continue;
    auto add = [&](int f) {
      for (int m = 0; m < 4; ++m) bfaces.push_back(c[fn[f][m]]);
      names.push_back(name_of(f, i, j, k));
    };
    if (k == 0) add(0);
    if (j == 0) add(1);
    if (i == nx - 1) add(2);
    if (j == ny - 1) add(3);
    if (i == 0) add(4);
    if (k == nz - 1) add(5);
  }
  std::vector<std::string> sections;   // every section of the box, also those this part has no face of
  for (int k = 0; k < nz; ++k) for (int j = 0; j < ny; ++j) for (int i = 0; i < nx; ++i) {
    const bool on[6] = {k == 0, j == 0, i == nx - 1, j == ny - 1, i == 0, k == nz - 1};
    for (int f = 0; f < 6; ++f) {
      if (!on[f]) continue;
      std::string name = name_of(f, i, j, k);
      if (std::find(sections.begin(), sections.end(), name) == sections.end()) sections.push_back(name);
    }
  }
  part->DeclareSections(sections);
  part->SetMesh(std::move(nodes), std::move(cells), n_local, (int64_t)order.size() - n_local, std::move(bfaces), names,
      std::move(halo));
  part->SetCellGlobalIds(order);
}

inline std::vector<double> Uniform(double lo, double hi, int n) {
  std::vector<double> x(n + 1);
  for (int i = 0; i <= n; ++i) x[i] = lo + (hi - lo) * i / n;
  return x;
}

// input/thickness_ratio.py:8-16: the ratio r with first layer d_0 and n layers spanning [y_0, y_n] (Newton, x0 = 1.01)
inline double ThicknessRatio(double y_0, double y_n, double d_0, int n) {
  double r = 1.01;
  for (int it = 0; it < 200; ++it) {
    const double f = (std::pow(r, n) - 1) / (r - 1) - (y_n - y_0) / d_0;
    const double df = n * std::pow(r, n - 1) / (r - 1) - (std::pow(r, n) - 1) / ((r - 1) * (r - 1));
    const double step = f / df;
    r -= step;
    if (std::abs(step) < 1e-15) break;
  }
  return r;
}
// input/thickness_ratio.py:66-76: node positions of n layers growing by the ratio r
inline std::vector<double> Layers(double y_0, double y_n, double r, int n) {
  const double d_0 = (y_n - y_0) * (r - 1) / (std::pow(r, n) - 1);
  std::vector<double> y(n + 1);
  for (int i = 0; i <= n; ++i) y[i] = y_0 + (std::pow(r, i) - 1) / (r - 1) * d_0;
  y[n] = y_n;
  return y;
}

}  // namespace demo
#endif  // DEMO_BOX_MESH_HPP_
