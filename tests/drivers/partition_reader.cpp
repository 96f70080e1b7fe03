// Reads partition/<rank>.txt with mini::b200::ReadPartitionInfo, builds the shim's Part from it and the reordered mesh
// (a binary dump written by tests/test_partition_txt.py), and prints the part's tables.  No GPU: no spatial object.
#include <cstdio>
#include <fstream>
#include <string>
#include <vector>

using Scalar = double;
#include "mini/integrator/lobatto.hpp"
#include "mini/polynomial/hexahedron.hpp"
#include "mini/mesh/part.hpp"
#include "mini/b200/partition.hpp"

using Gx = mini::integrator::Lobatto<Scalar, 3>;
using Polynomial = mini::polynomial::Hexahedron<Gx, Gx, Gx, 5, false>;
using Part = mini::mesh::part::Part<int64_t, Polynomial>;

template <class T>
static std::vector<T> Take(std::ifstream &in, size_t n) {
  std::vector<T> v(n);
  in.read(reinterpret_cast<char *>(v.data()), (std::streamsize)(n * sizeof(T)));
  return v;
}

int main(int argc, char *argv[]) {
  if (argc < 5) return 2;
  const std::string case_dir = argv[1], mesh_file = argv[4];
  const int rank = std::atoi(argv[2]), size = std::atoi(argv[3]);
  std::ifstream in(mesh_file, std::ios::binary);
  mini::b200::ShuffledMesh mesh;
  const int64_t n_nodes = Take<int64_t>(in, 1)[0];
  mesh.nodes = Take<double>(in, (size_t)n_nodes * 3);
  const int64_t n_cells = Take<int64_t>(in, 1)[0];
  mesh.cells = Take<int64_t>(in, (size_t)n_cells * 8);
  mesh.cell_metis = Take<int64_t>(in, (size_t)n_cells);
  const int64_t n_sections = Take<int64_t>(in, 1)[0];
  for (int64_t s = 0; s < n_sections; ++s) {
    mini::b200::ShuffledMesh::FaceSection sec;
    const int64_t len = Take<int64_t>(in, 1)[0];
    auto name = Take<char>(in, (size_t)len);
    sec.name.assign(name.begin(), name.end());
    sec.first = Take<int64_t>(in, 1)[0];
    const int64_t n_quads = Take<int64_t>(in, 1)[0];
    sec.quads = Take<int64_t>(in, (size_t)n_quads * 4);
    mesh.face_sections.push_back(std::move(sec));
  }
  auto info = mini::b200::ReadPartitionInfo(case_dir + "/partition/" + std::to_string(rank) + ".txt");
  auto part = Part(rank, size);
  mini::b200::BuildPartFromPartition(&part, info, mesh);
  std::printf("local %ld ghost %ld inner %zu\n", (long)part.CountLocalCells(), (long)part.CountGhostCells(), part.GetInnerCellPointers().size());
  std::printf("ids");
  for (auto g : part.global_ids_) std::printf(" %ld", (long)g);
  std::printf("\nneighbours");
  for (auto q : part.halo_.parts) std::printf(" %ld", (long)q);
  std::printf("\nsend_counts");
  for (auto q : part.halo_.send_counts) std::printf(" %ld", (long)q);
  std::printf("\nsend");
  for (auto q : part.halo_.send_cells) std::printf(" %ld", (long)q);
  std::printf("\nrecv");
  for (auto q : part.halo_.recv_cells) std::printf(" %ld", (long)q);
  std::printf("\nbfaces");
  for (auto q : part.bface_nodes_) std::printf(" %ld", (long)q);
  std::printf("\nbtags");
  for (auto q : part.bface_tags_) std::printf(" %d", (int)q);
  std::printf("\ncells");
  for (auto q : part.cell_nodes_) std::printf(" %ld", (long)q);
  std::printf("\n");
  return 0;
}
