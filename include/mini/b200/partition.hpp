// partition.hpp — `partition/<rank>.txt`, the index file `mini::mesh::Shuffler::WritePartitionInfo` writes per part
// (include/mini/mesh/shuffler.hpp:436-506) and `mini::mesh::part::Part` reads in its constructor
// (include/mini/mesh/part.hpp:442-476 through BuildLocalNodes :535-576, ShareGhostNodes :577-613, BuildLocalCells
// :689-745, BuildAdj :806-827, BuildBoundaryFaces :1456-1549), parsed block by block, and a part built from it.
//
// What the file points into — `<case>/shuffled.cgns` — cannot be read here (no CGNS / HDF5 on the target image); the
// reordered mesh is handed over as arrays (`ShuffledMesh`: one zone, section 1 = the HEXA_8 cells, sections 2.. = one
// QUAD_4 section per boundary name), the stand-in for the cgp_*_read_data calls of those functions.  The Python host has
// the same reader and the writer (minicfd_b200/partition.py); tests/test_partition_txt.py holds both to each other.
#ifndef MINI_B200_PARTITION_HPP_
#define MINI_B200_PARTITION_HPP_

#include <algorithm>
#include <array>
#include <cstdint>
#include <fstream>
#include <map>
#include <set>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace mini {
namespace b200 {

struct PartitionInfo {
  std::vector<std::array<int64_t, 3>> node_ranges;   // i_zone i_node_head i_node_tail
  std::vector<std::array<int64_t, 2>> send_nodes;    // i_part i_node_metis
  std::vector<std::array<int64_t, 4>> adj_nodes;     // i_part i_node_metis i_zone i_node
  std::vector<std::array<int64_t, 4>> cell_ranges;   // i_zone i_sect i_cell_head i_cell_tail
  std::vector<std::array<int64_t, 2>> inner_adjs;    // i_cell_metis j_cell_metis
  std::vector<std::array<int64_t, 5>> inter_adjs;    // i_part i_cell_metis j_cell_metis i_node_npe j_node_npe
  std::vector<std::array<int64_t, 4>> face_ranges;   // i_zone i_sect i_face_head i_face_tail
};

// Seven blocks, each opened by a '#' line, closed by a lone '#'.
inline PartitionInfo ReadPartitionInfo(std::string const &path) {
  std::ifstream istrm(path);
  if (!istrm) throw std::runtime_error("cannot open " + path);
  PartitionInfo info;
  int block = -1;
  std::string line;
  auto take = [&](auto &rows) {
    typename std::remove_reference_t<decltype(rows)>::value_type row{};
    std::istringstream in(line);
    for (auto &x : row) if (!(in >> x)) throw std::runtime_error(path + ": short line in block " + std::to_string(block));
    rows.push_back(row);
  };
  while (std::getline(istrm, line)) {
    if (!line.empty() && line[0] == '#') { ++block; continue; }
    if (line.empty()) continue;
    switch (block) {
    case 0: take(info.node_ranges); break;
    case 1: take(info.send_nodes); break;
    case 2: take(info.adj_nodes); break;
    case 3: take(info.cell_ranges); break;
    case 4: take(info.inner_adjs); break;
    case 5: take(info.inter_adjs); break;
    case 6: take(info.face_ranges); break;
    default: throw std::runtime_error(path + ": data outside the seven blocks");
    }
  }
  if (block != 7) throw std::runtime_error(path + ": expected seven blocks closed by '#'");
  return info;
}

// The reordered mesh (what shuffled.cgns holds; ids 0-based here, 1-based in the file and in PartitionInfo).
struct ShuffledMesh {
  std::vector<double> nodes;          // [n_nodes][3]
  std::vector<int64_t> cells;         // [n_cells][8] node ids
  std::vector<int64_t> cell_metis;    // MetisIndex of every cell
  struct FaceSection { std::string name; std::vector<int64_t> quads; int64_t first; };   // [n][4] node ids; first CGNS element id
  std::vector<FaceSection> face_sections;   // section s + 2 of the zone
};

// Part::Part(case, rank, size) up to the geometry: local cells from the cell ranges, inner cells (no neighbour in
// another part) before inter cells (part.hpp:746-767), ghost cells = the far ends of the inter-part pairs in
// ascending id, send / receive lists per neighbouring part in ascending metis id (std::map, part.hpp:904-925),
// boundary faces from the face ranges.  `part` needs SetMesh, DeclareSections, SetCellGlobalIds and a Halo type.
template <class Part>
void BuildPartFromPartition(Part *part, PartitionInfo const &info, ShuffledMesh const &mesh) {
  const int64_t n_cells = (int64_t)mesh.cell_metis.size();
  std::map<int64_t, int64_t> new_of_metis;
  for (int64_t c = 0; c < n_cells; ++c) new_of_metis[mesh.cell_metis[c]] = c;
  std::vector<int64_t> local;
  for (auto const &r : info.cell_ranges) for (int64_t c = r[2]; c < r[3]; ++c) local.push_back(c - 1);
  std::map<int64_t, std::set<int64_t>> send, recv;   // part -> cells (new ids; ascending = ascending metis id within a part)
  std::set<int64_t> inter, ghosts;
  for (auto const &a : info.inter_adjs) {
    const int64_t i = new_of_metis.at(a[1]), j = new_of_metis.at(a[2]);
    send[a[0]].insert(i); recv[a[0]].insert(j);
    inter.insert(i); ghosts.insert(j);
  }
  std::vector<int64_t> order;
  for (int64_t c : local) if (!inter.count(c)) order.push_back(c);
  for (int64_t c : local) if (inter.count(c)) order.push_back(c);
  const int64_t n_local = (int64_t)order.size();
  order.insert(order.end(), ghosts.begin(), ghosts.end());
  std::map<int64_t, int64_t> id_of;
  for (int64_t k = 0; k < (int64_t)order.size(); ++k) id_of[order[k]] = k;
  typename Part::Halo halo;
  for (auto const &kv : send) {
    halo.parts.push_back(kv.first);
    halo.send_counts.push_back((int64_t)kv.second.size());
    for (int64_t c : kv.second) halo.send_cells.push_back(id_of.at(c));
    auto const &r = recv.at(kv.first);
    halo.recv_counts.push_back((int64_t)r.size());
    for (int64_t c : r) halo.recv_cells.push_back(id_of.at(c));
  }
  std::vector<int64_t> cell_nodes, ids;
  for (int64_t c : order) {
    cell_nodes.insert(cell_nodes.end(), mesh.cells.begin() + c * 8, mesh.cells.begin() + c * 8 + 8);
    ids.push_back(mesh.cell_metis[c]);
  }
  std::vector<int64_t> bfaces;
  std::vector<std::string> names, sections;
  for (auto const &s : mesh.face_sections) sections.push_back(s.name);
  for (auto const &r : info.face_ranges) {
    auto const &s = mesh.face_sections.at(r[1] - 2);
    for (int64_t f = r[2]; f < r[3]; ++f) {
      const int64_t k = f - s.first;
      bfaces.insert(bfaces.end(), s.quads.begin() + k * 4, s.quads.begin() + k * 4 + 4);
      names.push_back(s.name);
    }
  }
  part->DeclareSections(sections);
  part->SetMesh(mesh.nodes, std::move(cell_nodes), n_local, (int64_t)order.size() - n_local, std::move(bfaces), names,
      std::move(halo));
  part->SetCellGlobalIds(std::move(ids));
}

}  // namespace b200
}  // namespace mini
#endif  // MINI_B200_PARTITION_HPP_
