// TEST INFRASTRUCTURE — NOT PART OF THE PRODUCT PATH.
//
// ref_harness.cpp: drives the UNMODIFIED miniCFD reference headers (included in place from
// /root/reference/include via -I; nothing is copied) through a small extern "C" interface so that
// tests/ and bench.py's cpu_baseline / --impl reference legs can obtain the reference's own
// residuals, RK updates and time steps on an in-memory hexahedral mesh.
//
// The reference's mesh container (mini/mesh/part.hpp) needs MPI + CGNS + METIS, which are not
// available here, so this file supplies `OraclePart`: a duck-typed stand-in exposing exactly the
// subset of the Part/Cell/Face interface that mini::spatial::* uses.  Its
// numbering rules restate the reference's:
//   * local cell ids: inner cells first, then inter cells (part.hpp:746-767),
//   * ghost cells appended after the locals (part.hpp:768-775; unordered_map order there,
//     canonicalised here to ascending global id),
//   * faces: local, then ghost, then boundary (part.hpp:115), local faces holder = smaller id
//     (shuffler.hpp:417-420), ghost faces grouped by neighbour part ascending (part.hpp:819-825),
//   * send/recv lists per neighbour part in ascending global id (part.hpp:800-805, 904-925),
//   * face nodes sorted so the normal points holder -> sharer (part.hpp:957-966,
//     coordinate/hexahedron.hpp:84-117, 378-389).
// P parts are emulated in one process, one std::thread per part, with a barrier-based copy standing
// in for MPI_Isend/Irecv/Waitall (part.hpp:1242-1314).
//
// Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load the resulting
// library (oracle/_ref/libminicfd_ref.so).
#include <algorithm>
#include <array>
#include <barrier>
#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sys/stat.h>
#include <functional>
#include <map>
#include <memory>
#include <ranges>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "mini/algebra/eigen.hpp"
#include "mini/coordinate/hexahedron.hpp"
#include "mini/coordinate/quadrangle.hpp"
#include "mini/integrator/hexahedron.hpp"
#include "mini/integrator/lobatto.hpp"
#include "mini/integrator/legendre.hpp"
#include "mini/polynomial/projection.hpp"
#include "mini/spatial/dg/general.hpp"
#include "mini/integrator/quadrangle.hpp"
#include "mini/limiter/average.hpp"
// limiter/majority.hpp:27 names `Scalar` unqualified; the drivers define that alias at global scope before
// including it (demo/euler/sourceless.hpp:11), and so do we.
using Scalar = double;
#include "mini/limiter/majority.hpp"
#include "mini/limiter/reconstruct.hpp"
#include "mini/limiter/weno.hpp"
#include "mini/polynomial/extrapolation.hpp"
#include "mini/spatial/with_limiter.hpp"
// spatial/viscosity.hpp names MPI_Request (its own buffers) without including an MPI header
struct ompi_request_t;
typedef ompi_request_t *MPI_Request;
#include "mini/spatial/viscosity.hpp"
#include "mini/spatial/with_viscosity.hpp"
#include "mini/polynomial/hexahedron.hpp"
#include "mini/riemann/euler/ausm.hpp"
#include "mini/riemann/euler/exact.hpp"
#include "mini/riemann/euler/hllc.hpp"
#include "mini/riemann/euler/types.hpp"
#include "mini/riemann/rotated/euler.hpp"
#include "mini/riemann/concept.hpp"
#include "mini/riemann/diffusive/direct.hpp"
#include "mini/riemann/diffusive/linear.hpp"
#include "mini/riemann/diffusive/navier_stokes.hpp"
#include "mini/spatial/dg/lobatto.hpp"
#include "mini/spatial/fr/lobatto.hpp"
#include "mini/temporal/rk.hpp"
#include "mini/mesh/vtk.hpp"

#ifndef REF_DEGREE
#define REF_DEGREE 2
#endif

namespace oracle {

using Scalar = double;
using Int = int64_t;

// ---------------------------------------------------------------------------------------------
// Global mesh shared by all emulated parts.
// ---------------------------------------------------------------------------------------------
struct GlobalMesh {
  std::vector<std::array<double, 3>> nodes;
  std::vector<std::array<Int, 8>> cells;   // CGNS HEXA_8 node order
  std::vector<int> cell_parts;
  int n_parts = 1;
  std::vector<std::array<Int, 4>> bfaces;  // boundary quads (any node order)
  std::vector<int> btags;
  // derived: for every cell, its neighbours (ascending global id)
  std::vector<std::vector<Int>> adj;
  // boundary face -> holder cell
  std::vector<Int> bholder;

  static std::array<Int, 4> Key(std::array<Int, 4> q) {
    std::sort(q.begin(), q.end());
    return q;
  }
  void Build() {
    // local face -> node slots, see coordinate/hexahedron.hpp:378-389
    static const int kFaceNodes[6][4] = {
      {0, 3, 2, 1}, {0, 1, 5, 4}, {1, 2, 6, 5}, {2, 3, 7, 6}, {0, 4, 7, 3}, {4, 5, 6, 7}};
    std::map<std::array<Int, 4>, std::vector<Int>> face_to_cells;
    for (Int c = 0; c < (Int)cells.size(); ++c) {
      for (int f = 0; f < 6; ++f) {
        std::array<Int, 4> q;
        for (int v = 0; v < 4; ++v) q[v] = cells[c][kFaceNodes[f][v]];
        face_to_cells[Key(q)].push_back(c);
      }
    }
    adj.assign(cells.size(), {});
    for (auto &[key, cs] : face_to_cells) {
      if (cs.size() == 2) {
        adj[cs[0]].push_back(cs[1]);
        adj[cs[1]].push_back(cs[0]);
      }
    }
    for (auto &a : adj) std::sort(a.begin(), a.end());
    bholder.resize(bfaces.size());
    for (size_t b = 0; b < bfaces.size(); ++b) {
      auto it = face_to_cells.find(Key(bfaces[b]));
      if (it == face_to_cells.end() || it->second.size() != 1) {
        std::fprintf(stderr, "ref_harness: boundary face %zu is not on the boundary\n", b);
        std::abort();
      }
      bholder[b] = it->second[0];
    }
  }
};

// ---------------------------------------------------------------------------------------------
// Duck-typed Cell / Face / Part (interface of mini/mesh/part.hpp:97-291, 404-445).
// ---------------------------------------------------------------------------------------------
template <class Poly>
struct OracleFace;

// What a polynomial type does not say itself: the concrete cell integrator and the 1-D rule of the face integrators.
// The nodal Hexahedron carries both; the modal Projection (polynomial/projection.hpp) only knows the abstract
// integrator::Cell, its rules come from the driver (demo/euler/sourceless.hpp:28-33: Legendre under DGFEM).
template <class Poly>
struct PolyTraits {
  using CellIntegrator = typename Poly::Integrator;
  using Gx = typename Poly::IntegratorX;
};
using ModalDgRule = mini::integrator::Legendre<double, REF_DEGREE + 1>;
using ModalDgPoly = mini::polynomial::Projection<double, 3, REF_DEGREE, 5>;
template <>
struct PolyTraits<ModalDgPoly> {
  using CellIntegrator = mini::integrator::Hexahedron<ModalDgRule, ModalDgRule, ModalDgRule>;
  using Gx = ModalDgRule;
};

template <class Poly>
struct OracleCell {
  using Polynomial = Poly;
  using Scalar = typename Polynomial::Scalar;
  using Integrator = mini::integrator::Cell<Scalar>;
  using Coordinate = mini::coordinate::Cell<Scalar>;
  using Local = typename Polynomial::Local;
  using Global = typename Polynomial::Global;
  using Value = typename Polynomial::Value;
  using Coeff = typename Polynomial::Coeff;
  static constexpr int K = Polynomial::K;
  static constexpr int N = Polynomial::N;
  static constexpr int P = Polynomial::P;
  static constexpr int D = Polynomial::D;
  static constexpr int kFields = K * N;
  using Face = OracleFace<Poly>;

  std::vector<OracleCell *> adj_cells_;
  std::vector<Face *> adj_faces_;
  std::vector<Face *> boundary_faces_;

  std::unique_ptr<mini::coordinate::Hexahedron8<Scalar>> coordinate_ptr_;
  std::unique_ptr<typename PolyTraits<Poly>::CellIntegrator> integrator_ptr_;
  std::unique_ptr<Polynomial> polynomial_ptr_;
  Scalar length_;
  Int metis_id{-1};
  Int id_{-1};
  bool inner_ = true;

  OracleCell(GlobalMesh const &mesh, Int m_cell) : metis_id(m_cell) {
    auto const &c = mesh.cells[m_cell];
    auto G = [&](int v) {
      auto const &x = mesh.nodes[c[v]];
      return Global(x[0], x[1], x[2]);
    };
    coordinate_ptr_.reset(new mini::coordinate::Hexahedron8<Scalar>{
        G(0), G(1), G(2), G(3), G(4), G(5), G(6), G(7)});
    integrator_ptr_ = std::make_unique<typename PolyTraits<Poly>::CellIntegrator>(*coordinate_ptr_);
    polynomial_ptr_ = std::make_unique<Polynomial>(*integrator_ptr_);
    length_ = std::cbrt(volume()) * 0.5;  // part.hpp:230
  }
  OracleCell(OracleCell const &) = delete;

  Scalar length() const { return length_; }
  Scalar volume() const { return integrator_ptr_->volume(); }
  Int id() const { return id_; }
  bool inner() const { return inner_; }
  bool inter() const { return !inner_; }
  Global const &center() const { return integrator().center(); }
  Integrator const &integrator() const { return *integrator_ptr_; }
  Coordinate const &coordinate() const { return *coordinate_ptr_; }
  Polynomial const &polynomial() const { return *polynomial_ptr_; }
  Polynomial &polynomial() { return *polynomial_ptr_; }
  Global LocalToGlobal(const Local &local) const { return coordinate().LocalToGlobal(local); }
  Value GlobalToValue(const Global &global) const { return polynomial().GlobalToValue(global); }
  int CountCorners() const { return 8; }
  int CountFields() const { return kFields; }
  auto GlobalToBasisValues(const Global &global) const {
    return polynomial().GlobalToBasisValues(global);
  }
  template <class Callable>
  void Approximate(Callable &&func) {
    polynomial().Approximate(std::forward<Callable>(func));
  }
};

template <class Poly>
struct OracleFace {
  using Scalar = typename Poly::Scalar;
  using Integrator = mini::integrator::Face<Scalar, 3>;
  using Coordinate = mini::coordinate::Face<Scalar, 3>;
  using Cell = OracleCell<Poly>;
  using Global = typename Cell::Global;
  using Gx = typename PolyTraits<Poly>::Gx;

  std::unique_ptr<mini::coordinate::Quadrangle4<Scalar, 3>> coordinate_ptr_;
  std::unique_ptr<mini::integrator::Quadrangle<3, Gx, Gx>> integrator_ptr_;
  Cell *holder_, *sharer_;
  Scalar holder_height_, sharer_height_;
  Global holder_to_sharer_;
  Int id_;
  int holder_face_;  // local face id on the holder (0..5)

  OracleFace(GlobalMesh const &mesh, std::array<Int, 4> const &sorted_nodes, int holder_face,
      Cell *holder, Cell *sharer, Int id)
      : holder_(holder), sharer_(sharer), id_(id), holder_face_(holder_face) {
    auto G = [&](int v) {
      auto const &x = mesh.nodes[sorted_nodes[v]];
      return Global(x[0], x[1], x[2]);
    };
    coordinate_ptr_.reset(new mini::coordinate::Quadrangle4<Scalar, 3>{G(0), G(1), G(2), G(3)});
    integrator_ptr_ = std::make_unique<mini::integrator::Quadrangle<3, Gx, Gx>>(*coordinate_ptr_);
    // part.hpp:118-133
    holder_height_ = holder->volume() / area();
    sharer_height_ = sharer ? sharer->volume() / area() : -1;
    holder_to_sharer_ = -holder->center();
    if (sharer) {
      holder_to_sharer_ += sharer->center();
    } else {
      holder_to_sharer_ += center();
      holder_to_sharer_ *= 2;
    }
  }
  OracleFace(OracleFace const &) = delete;

  Integrator const &integrator() const { return *integrator_ptr_; }
  Coordinate const &coordinate() const { return *coordinate_ptr_; }
  Global center() const { return integrator().center(); }
  Scalar area() const { return integrator().area(); }
  Scalar holder_height() const { return holder_height_; }
  Scalar sharer_height() const { return sharer_height_; }
  Int id() const { return id_; }
  Cell const &holder() const { return *holder_; }
  Cell *holder_ptr() { return holder_; }
  Cell const &sharer() const { return *sharer_; }
  Global const &HolderToSharer() const { return holder_to_sharer_; }
  Cell const *other(Cell const *cell) const { return cell == holder_ ? sharer_ : holder_; }
};

// Stand-in for MPI_COMM_WORLD among the emulated parts.
struct InProcessComm {
  int n_parts;
  std::barrier<> barrier;
  // send_buf[src][dst] -> packed doubles
  std::vector<std::map<int, std::vector<double>>> send_buf;
  explicit InProcessComm(int n) : n_parts(n), barrier(n), send_buf(n) {}
};

template <class Poly>
class OraclePart {
 public:
  using Index = Int;
  using Polynomial = Poly;
  using Face = OracleFace<Poly>;
  using Cell = OracleCell<Poly>;
  using Scalar = typename Cell::Scalar;
  using Global = typename Cell::Global;
  using Value = typename Cell::Value;
  static constexpr int kDegrees = Polynomial::P;
  static constexpr int kComponents = Polynomial::K;
  static constexpr int kPhysDim = 3;
  static constexpr int kFields = Cell::kFields;

  GlobalMesh const *mesh_;
  InProcessComm *comm_;
  int rank_, size_;
  std::vector<std::unique_ptr<Cell>> local_cells_;   // ordered by id (inner then inter)
  std::vector<std::unique_ptr<Cell>> ghost_cells_;   // ordered by id
  std::unordered_map<Int, Cell *> m_to_cell_;        // global id -> local or ghost cell
  std::array<std::vector<Cell *>, 2> inner_and_inter_cells_;
  std::vector<Cell *> local_cell_ptrs_;
  std::map<Int, std::vector<Cell *>> send_cell_ptrs_, recv_cell_ptrs_;
  std::vector<std::unique_ptr<Face>> local_faces_, ghost_faces_, bound_faces_;
  std::map<std::string, std::vector<Face *>> name_to_faces_;

  static const int (&FaceNodes())[6][4] {
    static const int t[6][4] = {
      {0, 3, 2, 1}, {0, 1, 5, 4}, {1, 2, 6, 5}, {2, 3, 7, 6}, {0, 4, 7, 3}, {4, 5, 6, 7}};
    return t;
  }
  // Which local face of `cell` has exactly the node set `key` (sorted)?
  int FindLocalFace(Int m_cell, std::array<Int, 4> const &key) const {
    for (int f = 0; f < 6; ++f) {
      std::array<Int, 4> q;
      for (int v = 0; v < 4; ++v) q[v] = mesh_->cells[m_cell][FaceNodes()[f][v]];
      if (GlobalMesh::Key(q) == key) return f;
    }
    return -1;
  }
  int FindSharedFace(Int m_holder, Int m_sharer) const {
    for (int f = 0; f < 6; ++f) {
      std::array<Int, 4> q;
      for (int v = 0; v < 4; ++v) q[v] = mesh_->cells[m_holder][FaceNodes()[f][v]];
      if (FindLocalFace(m_sharer, GlobalMesh::Key(q)) >= 0) return f;
    }
    return -1;
  }
  std::array<Int, 4> SortedFaceNodes(Int m_holder, int f) const {
    std::array<Int, 4> q;
    for (int v = 0; v < 4; ++v) q[v] = mesh_->cells[m_holder][FaceNodes()[f][v]];
    return q;
  }

  OraclePart(GlobalMesh const *mesh, InProcessComm *comm, int rank)
      : mesh_(mesh), comm_(comm), rank_(rank), size_(mesh->n_parts) {
    auto const &parts = mesh->cell_parts;
    Int n = mesh->cells.size();
    // classify local cells (ascending global id) into inner / inter
    std::vector<Int> inner, inter;
    std::map<Int, std::map<Int, int>> send, recv;  // [part][m_cell]
    std::map<Int, std::vector<std::pair<Int, Int>>> interpart;  // [part] -> (i, j)
    for (Int i = 0; i < n; ++i) {
      if (parts[i] != rank) continue;
      bool is_inter = false;
      for (Int j : mesh->adj[i]) {
        if (parts[j] != rank) {
          is_inter = true;
          send[parts[j]][i] = 8;
          recv[parts[j]][j] = 8;
          interpart[parts[j]].emplace_back(i, j);
        }
      }
      (is_inter ? inter : inner).push_back(i);
    }
    Int id = 0;
    for (auto *list : {&inner, &inter}) {
      for (Int m : *list) {
        auto cell = std::make_unique<Cell>(*mesh, m);
        cell->id_ = id++;
        cell->inner_ = (list == &inner);
        m_to_cell_[m] = cell.get();
        inner_and_inter_cells_[list == &inner ? 0 : 1].push_back(cell.get());
        local_cell_ptrs_.push_back(cell.get());
        local_cells_.emplace_back(std::move(cell));
      }
    }
    // ghost cells: ascending global id
    std::vector<Int> ghosts;
    for (auto &[p, cells] : recv) for (auto &[m, npe] : cells) ghosts.push_back(m);
    std::sort(ghosts.begin(), ghosts.end());
    ghosts.erase(std::unique(ghosts.begin(), ghosts.end()), ghosts.end());
    for (Int m : ghosts) {
      auto cell = std::make_unique<Cell>(*mesh, m);
      cell->id_ = id++;
      m_to_cell_[m] = cell.get();
      ghost_cells_.emplace_back(std::move(cell));
    }
    for (auto &[p, cells] : send)
      for (auto &[m, npe] : cells) send_cell_ptrs_[p].push_back(m_to_cell_.at(m));
    for (auto &[p, cells] : recv)
      for (auto &[m, npe] : cells) recv_cell_ptrs_[p].push_back(m_to_cell_.at(m));
    for (auto &[p, cells] : send_cell_ptrs_)
      comm_->send_buf[rank_][p].resize(cells.size() * kFields);
    // local faces: (i, j) with i < j, i ascending then j ascending
    for (Int i = 0; i < n; ++i) {
      if (parts[i] != rank) continue;
      for (Int j : mesh->adj[i]) {
        if (parts[j] != rank || j < i) continue;
        AddFace(&local_faces_, i, j, local_faces_.size());
      }
    }
    for (auto &[p, pairs] : interpart)
      for (auto [i, j] : pairs)
        AddFace(&ghost_faces_, i, j, local_faces_.size() + ghost_faces_.size());
    // boundary faces: grouped by tag ascending, input order within a tag
    std::map<int, std::vector<size_t>> by_tag;
    for (size_t b = 0; b < mesh->bfaces.size(); ++b)
      if (parts[mesh->bholder[b]] == rank) by_tag[mesh->btags[b]].push_back(b);
    for (auto &[tag, list] : by_tag) {
      auto &named = name_to_faces_[std::to_string(tag)];
      for (size_t b : list) {
        Int m = mesh->bholder[b];
        int f = FindLocalFace(m, GlobalMesh::Key(mesh->bfaces[b]));
        Cell *holder = m_to_cell_.at(m);
        Int fid = local_faces_.size() + ghost_faces_.size() + bound_faces_.size();
        auto face = std::make_unique<Face>(*mesh, SortedFaceNodes(m, f), f, holder, nullptr, fid);
        holder->boundary_faces_.push_back(face.get());
        named.push_back(face.get());
        bound_faces_.emplace_back(std::move(face));
      }
    }
  }
  void AddFace(std::vector<std::unique_ptr<Face>> *faces, Int i, Int j, Int fid) {
    int f = FindSharedFace(i, j);
    Cell *holder = m_to_cell_.at(i), *sharer = m_to_cell_.at(j);
    auto face = std::make_unique<Face>(*mesh_, SortedFaceNodes(i, f), f, holder, sharer, fid);
    holder->adj_cells_.push_back(sharer);
    holder->adj_faces_.push_back(face.get());
    if (faces == &local_faces_) {
      sharer->adj_cells_.push_back(holder);
      sharer->adj_faces_.push_back(face.get());
    }
    faces->emplace_back(std::move(face));
  }

  int mpi_rank() const { return rank_; }
  int mpi_size() const { return size_; }
  // what vtk::Writer<Part>::WriteSolutions asks of a part (part.hpp:477-485, 1565-1579), minus the MPI barrier
  std::string directory_ = ".";
  std::vector<std::string> field_names_{"Density", "MomentumX", "MomentumY", "MomentumZ", "EnergyStagnationDensity"};
  const std::string &GetFieldName(int i) const { return field_names_.at(i); }
  const std::string &GetDirectoryName() const { return directory_; }
  std::ofstream GetFileStream(std::string const &soln_name, bool binary, std::string const &suffix) const {
    ::mkdir(directory_.c_str(), 0777);   // the reference runs `mkdir -p` (part.hpp:1568-1573)
    ::mkdir((directory_ + "/" + soln_name).c_str(), 0777);
    return std::ofstream(directory_ + "/" + soln_name + "/" + std::to_string(rank_) + "." + suffix,
        std::ios::out | (binary ? (std::ios::binary) : std::ios::out));
  }
  Int CountLocalCells() const { return local_cells_.size(); }
  Int CountGhostCells() const { return ghost_cells_.size(); }
  Int GetCellDataOffset(Int i_cell) const { return i_cell * kFields; }
  Int GetCellDataSize() const { return CountLocalCells() * kFields; }
  bool IsLocal(Int i_cell) const { return 0 <= i_cell && i_cell < CountLocalCells(); }

  auto GetLocalCells() const {
    return local_cells_ | std::views::transform([](auto &u) -> Cell const & { return *u; });
  }
  auto GetGhostCells() const {
    return ghost_cells_ | std::views::transform([](auto &u) -> Cell const & { return *u; });
  }
  std::vector<Cell *> const &GetLocalCellPointers() { return local_cell_ptrs_; }
  std::vector<Cell *> const &GetInnerCellPointers() { return inner_and_inter_cells_[0]; }
  std::vector<Cell *> const &GetInterCellPointers() { return inner_and_inter_cells_[1]; }
  auto GetLocalFaces() const {
    return local_faces_ | std::views::transform([](auto &u) -> Face const & { return *u; });
  }
  auto GetGhostFaces() const {
    return ghost_faces_ | std::views::transform([](auto &u) -> Face const & { return *u; });
  }
  auto GetBoundaryFaces() const {
    return bound_faces_ | std::views::transform([](auto &u) -> Face const & { return *u; });
  }
  auto GetBoundaryFaces(std::string const &name) const {
    static const std::vector<Face *> empty;
    auto it = name_to_faces_.find(name);
    auto const &faces = (it == name_to_faces_.end()) ? empty : it->second;
    return faces | std::views::transform([](Face *p) -> Face const & { return *p; });
  }

  // part.hpp:1217-1231, 1242-1271, 1288-1308 as EnergyBasedViscosity uses them for its per-cell properties.  Its
  // state is class-static (one spatial object per process), so the emulated multi-part mode cannot host it: these
  // only serve a single part, where nothing is sent.
  void InitializeRequestsAndBuffers(Int n_scalar_per_cell, std::vector<MPI_Request> *requests_ptr,
      std::vector<std::vector<Scalar>> *send_data_ptr, std::vector<std::vector<Scalar>> *recv_data_ptr) const {
    (void)n_scalar_per_cell; (void)requests_ptr; (void)send_data_ptr; (void)recv_data_ptr;
    if (size_ != 1) std::abort();
  }
  template <class M>
  void ShareGhostCellData(std::vector<MPI_Request> *, std::vector<std::vector<Scalar>> *,
      std::vector<std::vector<Scalar>> *, M &&) {
    if (size_ != 1) std::abort();
  }
  template <class M>
  void UpdateGhostCellData(std::vector<MPI_Request> *, std::vector<std::vector<Scalar>> *, M &&) {
    if (size_ != 1) std::abort();
  }

  // part.hpp:1242-1278: pack the coefficients of the cells each neighbour needs.
  void ShareGhostCellCoeffs() {
    for (auto &[p, cells] : send_cell_ptrs_) {
      double *data = comm_->send_buf[rank_][p].data();
      for (Cell *c : cells) data = c->polynomial().WriteCoeffTo(data);
    }
  }
  // part.hpp:1288-1314: wait, then unpack into the ghost cells.
  void UpdateGhostCellCoeffs() {
    if (size_ == 1) return;
    comm_->barrier.arrive_and_wait();  // every part has packed
    for (auto &[p, cells] : recv_cell_ptrs_) {
      double const *data = comm_->send_buf[p].at(rank_).data();
      for (Cell *c : cells) data = c->polynomial().GetCoeffFrom(data);
    }
    comm_->barrier.arrive_and_wait();  // every part has unpacked; buffers may be reused
  }
};

// ---------------------------------------------------------------------------------------------
// Type zoo.
// ---------------------------------------------------------------------------------------------
using Gas = mini::riemann::euler::IdealGas<Scalar, 1.4>;
using Gx = mini::integrator::Lobatto<Scalar, REF_DEGREE + 1>;
using Poly = mini::polynomial::Hexahedron<Gx, Gx, Gx, 5, false>;
using Part = OraclePart<Poly>;
using Value = typename Poly::Value;
using Global = typename Poly::Global;
using BcFunc = std::function<Value(const Global &, double)>;
using Column = typename mini::temporal::System<Scalar>::Column;
// nodal expansion with its modal (orthonormal) extension, as demo/euler/sourceless.hpp:41-43 composes it under LIMITER
using ModalPoly = mini::polynomial::Extrapolation<Poly>;
using ModalPart = OraclePart<ModalPoly>;
// (w0, eps) of the next weno limiter to be created (demo/euler/sourceless.cpp:114); set by ref_pN_set_weno
static double g_weno_w0 = 0.001, g_weno_eps = 1e-6;

extern "C" typedef void (*ref_value_callback)(int tag, const double *xyz, double t, double *out,
    void *user);

struct FieldSpec {
  // kind: 0 = constant conservative state, 1 = isentropic vortex,
  //       2 = C callback, 3 = smooth trigonometric state (test pattern)
  int kind = 0;
  double params[16] = {0};
  ref_value_callback cb = nullptr;
  void *user = nullptr;
  int tag = 0;
};

static Value EvalField(FieldSpec const &s, Global const &x, double t) {
  Value v;
  switch (s.kind) {
  case 0:
    for (int k = 0; k < 5; ++k) v[k] = s.params[k];
    break;
  case 1: {
    // params: gamma-free; (x0, y0, u_inf, v_inf, w_inf, beta)
    double x0 = s.params[0], y0 = s.params[1];
    double ui = s.params[2], vi = s.params[3], wi = s.params[4], beta = s.params[5];
    const double g = 1.4, pi = 3.14159265358979323846;
    double dx = x[0] - x0 - ui * t, dy = x[1] - y0 - vi * t;
    double r2 = dx * dx + dy * dy;
    double e = std::exp(0.5 * (1 - r2));
    double du = -beta / (2 * pi) * e * dy, dv = beta / (2 * pi) * e * dx;
    double dT = -(g - 1) * beta * beta / (8 * g * pi * pi) * e * e;
    double T = 1 + dT;
    double rho = std::pow(T, 1 / (g - 1)), p = std::pow(rho, g);
    double u = ui + du, vv = vi + dv, w = wi;
    v[0] = rho; v[1] = rho * u; v[2] = rho * vv; v[3] = rho * w;
    v[4] = p / (g - 1) + 0.5 * rho * (u * u + vv * vv + w * w);
    break;
  }
  case 2:
    s.cb(s.tag, x.data(), t, v.data(), s.user);
    break;
  case 3: {
    // smooth, strictly admissible state with all components varying in all directions
    double a = s.params[0] ? s.params[0] : 1.0;
    double rho = 1.0 + 0.2 * std::sin(a * (x[0] + 2 * x[1] - x[2]) + 0.3 * t);
    double u = 0.5 + 0.3 * std::cos(a * (x[0] - x[1]) + 0.1);
    double vv = -0.2 + 0.25 * std::sin(a * (x[1] + x[2]) + 0.7);
    double w = 0.1 + 0.2 * std::cos(a * (x[2] - 2 * x[0]) - 0.4);
    double p = 1.0 + 0.3 * std::sin(a * (x[0] + x[1] + x[2]) + 1.1);
    v[0] = rho; v[1] = rho * u; v[2] = rho * vv; v[3] = rho * w;
    v[4] = p / 0.4 + 0.5 * rho * (u * u + vv * vv + w * w);
    break;
  }
  default:
    v.setZero();
  }
  return v;
}

struct SolverBase {
  virtual ~SolverBase() = default;
  virtual int n_parts() const = 0;
  virtual Int count_local(int r) const = 0;
  virtual Int count_ghost(int r) const = 0;
  virtual void cell_global_ids(int r, Int *out) const = 0;       // local then ghost
  virtual Int count_faces(int r, int kind) const = 0;             // 0 local, 1 ghost, 2 boundary
  virtual void face_table(int r, int kind, Int *holder, Int *sharer, int *holder_face) const = 0;
  virtual void halo_table(int r, int send, Int *n_parts_out, Int *parts, Int *counts, Int *cells) const = 0;
  virtual void set_bc(int tag, int kind, FieldSpec const &spec) = 0;
  virtual void approximate(FieldSpec const &spec) = 0;
  virtual void set_solution(int r, const double *col) = 0;
  virtual void set_solution_all(const double *cols) = 0;
  virtual int cell_viscosity(double *nu) const { (void)nu; return -1; }        // [n_local][5], scheme 7
  virtual int damping_matrices(double *d) { (void)d; return -1; }              // [n_local][N][N], scheme 7; zeroes the solution
  virtual void get_solution(int r, double *col) const = 0;
  virtual void residual_all(double t) = 0;
  virtual void get_residual(int r, double *col) const = 0;
  virtual void rk_update(int order, double t, double dt, int n_steps) = 0;
  virtual double time_step(double dt_guess, int rk_order) const = 0;
  virtual void cell_geometry(int r, double *length, double *volume) const = 0;
  virtual void limiter(int kind) = 0;
  virtual void write_vtu(const char *directory, const char *soln_name) = 0;
  virtual void sample_vtk(int r, double *points, double *values) const = 0;   // [n_local][27][3], [n_local][27][5]
};

struct NoLimiter { NoLimiter(double, double) {} };
template <class Spatial, bool kModal> struct LimiterOf { using type = NoLimiter; };
// WithLimiter<General, Limiter> keeps a Limiter *limiter_ptr(): recover the type from it
template <class Spatial> struct LimiterOf<Spatial, true> {
  using type = std::remove_pointer_t<decltype(std::declval<Spatial &>().limiter_ptr())>;
};

template <class Spatial>
struct SolverImpl : SolverBase {
  using Part = typename Spatial::Part;   // OraclePart<Poly>, or OraclePart<Extrapolation<Poly>> under a limiter
  static constexpr bool kLimited = requires(Spatial &sp) { sp.limiter_ptr(); };
  using Limiter = typename LimiterOf<Spatial, kLimited>::type;
  GlobalMesh mesh;
  std::unique_ptr<InProcessComm> comm;
  std::vector<std::unique_ptr<Limiter>> limiters;   // one per part: the limiter object is stateful
  std::vector<std::unique_ptr<Part>> parts;
  std::vector<std::unique_ptr<Spatial>> spatials;
  std::vector<Column> residuals;

  explicit SolverImpl(GlobalMesh &&m, double c_next = 0.0) : mesh(std::move(m)) {
    mesh.Build();
    comm = std::make_unique<InProcessComm>(mesh.n_parts);
    parts.resize(mesh.n_parts);
    spatials.resize(mesh.n_parts);
    residuals.resize(mesh.n_parts);
    ForEachPart([this](int r) {
      parts[r] = std::make_unique<Part>(&mesh, comm.get(), r);
    });
    limiters.resize(mesh.n_parts);
    ForEachPart([this, c_next](int r) {
      if constexpr (kLimited) {   // demo/euler/sourceless.cpp:114-120: Spatial(&limiter, &part)
        limiters[r] = std::make_unique<Limiter>(g_weno_w0, g_weno_eps);
        spatials[r] = std::make_unique<Spatial>(limiters[r].get(), parts[r].get());
        if constexpr (requires(Limiter &l) { l.InstallRiemannSolvers(0); }) {   // weno::Eigen, demo/euler/rotorcraft.cpp:93-96
          Spatial *sp = spatials[r].get();
          limiters[r]->InstallRiemannSolvers(
              [sp](typename Part::Face const &face) -> auto const & { return sp->GetRiemannSolvers(face); });
        }
      } else if constexpr (requires { typename Spatial::Riemann::Viscosity; }) {   // WithViscosity's constructor is variadic
        spatials[r] = std::make_unique<Spatial>(parts[r].get());
      } else if constexpr (std::is_constructible_v<Spatial, Part *, Scalar>) {
        spatials[r] = std::make_unique<Spatial>(parts[r].get(), c_next);   // fr::General(part, c_next)
      } else {
        spatials[r] = std::make_unique<Spatial>(parts[r].get());
      }
    });
  }
  template <class F>
  void ForEachPart(F &&f) {
    int n = mesh.n_parts;
    if (n == 1) { f(0); return; }
    std::vector<std::thread> threads;
    for (int r = 0; r < n; ++r) threads.emplace_back([&f, r] { f(r); });
    for (auto &t : threads) t.join();
  }
  int n_parts() const override { return mesh.n_parts; }
  Int count_local(int r) const override { return parts[r]->CountLocalCells(); }
  Int count_ghost(int r) const override { return parts[r]->CountGhostCells(); }
  void cell_global_ids(int r, Int *out) const override {
    for (auto &c : parts[r]->local_cells_) *out++ = c->metis_id;
    for (auto &c : parts[r]->ghost_cells_) *out++ = c->metis_id;
  }
  std::vector<std::unique_ptr<typename Part::Face>> const &faces(int r, int kind) const {
    return kind == 0 ? parts[r]->local_faces_ : kind == 1 ? parts[r]->ghost_faces_
                                                          : parts[r]->bound_faces_;
  }
  Int count_faces(int r, int kind) const override { return faces(r, kind).size(); }
  void face_table(int r, int kind, Int *holder, Int *sharer, int *holder_face) const override {
    for (auto &f : faces(r, kind)) {
      *holder++ = f->holder_->id();
      *sharer++ = f->sharer_ ? f->sharer_->id() : -1;
      *holder_face++ = f->holder_face_;
    }
  }
  void halo_table(int r, int send, Int *n_out, Int *plist, Int *counts, Int *cells) const override {
    auto const &table = send ? parts[r]->send_cell_ptrs_ : parts[r]->recv_cell_ptrs_;
    *n_out = table.size();
    for (auto &[p, list] : table) {
      if (plist) *plist++ = p;
      if (counts) *counts++ = list.size();
      if (cells) for (auto *c : list) *cells++ = c->id();
    }
  }
  void set_bc(int tag, int kind, FieldSpec const &spec) override {
    auto name = std::to_string(tag);
    FieldSpec s = spec;
    s.tag = tag;
    BcFunc func = [s](const Global &x, double t) { return EvalField(s, x, t); };
    for (auto &sp : spatials) {
      switch (kind) {  // kinds follow fem.hpp:196-222
      case 0: sp->SetInviscidWall(name); break;
      case 1: sp->SetSupersonicOutlet(name); break;
      case 2: sp->SetSupersonicInlet(name, func); break;
      case 3: sp->SetSubsonicInlet(name, func); break;
      case 4: sp->SetSubsonicOutlet(name, func); break;
      case 5: sp->SetSmartBoundary(name, func); break;
      case 6: sp->SetNoSlipWall(name, func); break;
      default: std::abort();
      }
    }
  }
  void approximate(FieldSpec const &spec) override {
    ForEachPart([&](int r) {
      spatials[r]->Approximate([&spec](const Global &x) { return EvalField(spec, x, 0.0); });
    });
  }
  void set_solution(int r, const double *col) override {
    Column c = Eigen::Map<const Column>(col, parts[r]->GetCellDataSize());
    spatials[r]->SetSolutionColumn(c);
  }
  void get_solution(int r, double *col) const override {
    Column c = spatials[r]->GetSolutionColumn();
    std::memcpy(col, c.data(), sizeof(double) * c.size());
  }
  void residual_all(double t) override {
    ForEachPart([&](int r) {
      spatials[r]->SetTime(t);
      residuals[r] = spatials[r]->GetResidualColumn();
    });
  }
  void get_residual(int r, double *col) const override {
    std::memcpy(col, residuals[r].data(), sizeof(double) * residuals[r].size());
  }
  void rk_update(int order, double t, double dt, int n_steps) override {
    ForEachPart([&](int r) {
      double tt = t;
      for (int s = 0; s < n_steps; ++s) {
        if (order == 1) {
          mini::temporal::RungeKutta<1, Scalar>().Update(spatials[r].get(), tt, dt);
        } else if (order == 2) {
          mini::temporal::RungeKutta<2, Scalar>().Update(spatials[r].get(), tt, dt);
        } else {
          mini::temporal::RungeKutta<3, Scalar>().Update(spatials[r].get(), tt, dt);
        }
        tt += dt;
      }
    });
  }
  double time_step(double dt_guess, int rk_order) const override {
    double dt = dt_guess;  // MPI_Allreduce(MIN), sourceless.cpp:191-194
    for (auto &sp : spatials) dt = std::min(dt, sp->GetTimeStep(dt_guess, rk_order));
    return dt;
  }
  void limiter(int kind) override {
    if constexpr (!mini::polynomial::Modal<typename Part::Polynomial>) {
      for (auto &p : parts) {
        if (kind == 0) mini::limiter::average::Reconstruct(p.get());
        else mini::limiter::majority::Reconstruct(p.get());
      }
    }
  }
  // mini::mesh::vtk::Writer<Part>::WriteSolutions(part, soln_name) as demo/euler/sourceless.cpp:161-162, 220-221 calls
  // it, part after part; with the CellViscosity1..5 cell data of sourceless.cpp:122-126 on the VISCOSITY scheme
  void write_vtu(const char *directory, const char *soln_name) override {
    using Writer = mini::mesh::vtk::Writer<Part>;
    if constexpr (requires { typename Spatial::Riemann::Viscosity; }) {
      static bool installed = false;
      if (!installed) {
        using V = typename Spatial::Riemann::Viscosity;
        for (int k = 0; k < 5; ++k) {
          Writer::AddCellData("CellViscosity" + std::to_string(k + 1),
              [k](typename Part::Cell const &cell) { return V::GetPropertyOnCell(cell.id(), 0)[k]; });
        }
        installed = true;
      }
    }
    for (auto &p : parts) {
      p->directory_ = directory;
      Writer::WriteSolutions(*p, soln_name);
    }
  }
  // what vtk::Writer<Part>::Prepare (vtk.hpp:421-494) collects per cell, at full precision
  void sample_vtk(int r, double *points, double *values) const override {
    using Cell = typename Part::Cell;
    using Coord = typename Cell::Global;
    for (Cell const &cell : parts[r]->GetLocalCells()) {
      for (int i = 0; i < 27; ++i) {
        Coord global;
        typename Cell::Value value;
        cell.polynomial().LocalToGlobalAndValue(mini::mesh::vtk::Hexahedron27<Coord>::locals[i], &global, &value);
        for (int d = 0; d < 3; ++d) *points++ = global[d];
        for (int k = 0; k < 5; ++k) *values++ = value[k];
      }
    }
  }
  int cell_viscosity(double *nu) const override {
    if constexpr (requires { typename Spatial::Riemann::Viscosity; }) {
      using V = typename Spatial::Riemann::Viscosity;
      for (Int c = 0; c < parts[0]->CountLocalCells(); ++c) {
        auto const &p = V::GetPropertyOnCell(c, 0);
        for (int k = 0; k < 5; ++k) *nu++ = p[k];
      }
      return 0;
    }
    return -1;
  }
  int damping_matrices(double *d) override {
    if constexpr (requires { typename Spatial::Riemann::Viscosity; }) {
      using V = typename Spatial::Riemann::Viscosity;
      auto mats = V::BuildDampingMatrices();   // viscosity.hpp:299-384; leaves the cells zeroed
      constexpr int N = Part::Cell::N;
      for (auto const &m : mats) for (int r = 0; r < N; ++r) for (int c = 0; c < N; ++c) *d++ = m(r, c);
      return 0;
    }
    return -1;
  }
  // all parts at once: under a limiter SetSolutionColumn exchanges ghost cells (with_limiter.hpp:58-61,
  // limiter/reconstruct.hpp:47-51), which needs every part to take part
  void set_solution_all(const double *cols) override {
    std::vector<const double *> begin(mesh.n_parts);
    for (int r = 0; r < mesh.n_parts; ++r) { begin[r] = cols; cols += parts[r]->GetCellDataSize(); }
    ForEachPart([&](int r) {
      Column c = Eigen::Map<const Column>(begin[r], parts[r]->GetCellDataSize());
      spatials[r]->SetSolutionColumn(c);
    });
  }
  void cell_geometry(int r, double *length, double *volume) const override {
    for (auto &c : parts[r]->local_cells_) {
      *length++ = c->length();
      *volume++ = c->volume();
    }
  }
};

template <class Unrotated>
using Rotated = mini::riemann::rotated::Euler<Unrotated>;
using RExact = Rotated<mini::riemann::euler::Exact<Gas, 3>>;
using RHllc = Rotated<mini::riemann::euler::HartenLaxLeerContact<Gas, 3>>;
using RAusm = Rotated<mini::riemann::euler::AdvectionUpstreamSplittingMethod<Gas, 3>>;

// Navier-Stokes with the direct-DG common gradient, as demo/laminar/shockless.hpp:43-57 composes it
using Diffusion = mini::riemann::diffusive::Direct<mini::riemann::diffusive::NavierStokes<Gas>>;
template <class Unrotated>
using Viscous = mini::riemann::ConvectionDiffusion<Rotated<Unrotated>, Diffusion>;
using VExact = Viscous<mini::riemann::euler::Exact<Gas, 3>>;
using VHllc = Viscous<mini::riemann::euler::HartenLaxLeerContact<Gas, 3>>;
using VAusm = Viscous<mini::riemann::euler::AdvectionUpstreamSplittingMethod<Gas, 3>>;

// Euler + constant isotropic diffusion of every component with the direct-DG face flux: the diffusive Riemann type
// that EnergyBasedViscosity wraps (demo/euler/sourceless.hpp:88-93), here with one uniform nu
using IsoDiffusion = mini::riemann::diffusive::Direct<mini::riemann::diffusive::Isotropic<Scalar, 5>>;
template <class Unrotated>
using IsoViscous = mini::riemann::ConvectionDiffusion<Rotated<Unrotated>, IsoDiffusion>;
using IExact = IsoViscous<mini::riemann::euler::Exact<Gas, 3>>;
using IHllc = IsoViscous<mini::riemann::euler::HartenLaxLeerContact<Gas, 3>>;

// Euler with the energy-based artificial viscosity (demo/euler/sourceless.hpp:82-103 under VISCOSITY)
template <class Unrotated>
using Energy = mini::spatial::EnergyBasedViscosity<Part, IsoViscous<Unrotated>>;
using EExact = Energy<mini::riemann::euler::Exact<Gas, 3>>;
using EHllc = Energy<mini::riemann::euler::HartenLaxLeerContact<Gas, 3>>;
// the same rule over a DIFFUSIVE base, Navier-Stokes + direct DG (demo/laminar/shocked.hpp:100-111): only to look at
// what the reference's own set-up makes of it (tests/test_oracle_pinned.py::test_energy_viscosity_over_navier_stokes_base)
using ENsHllc = mini::spatial::EnergyBasedViscosity<Part, VHllc>;

// the kLocal == true expansion (coefficients = det(J) u), as the reference's laminar drivers use it for FR
// (demo/laminar/shockless.hpp:62-69)
using PolyL = mini::polynomial::Hexahedron<Gx, Gx, Gx, 5, true>;
using PartL = OraclePart<PolyL>;
template <class Unrotated>
using ViscousL = Viscous<Unrotated>;

// modal DG on the stand-alone Projection polynomial with Gauss-Legendre rules (demo/euler/sourceless.hpp:28-33, 64-66)
using DgPart = OraclePart<ModalDgPoly>;

// c_next of the next fr::General to be created (fr/general.hpp:218-219); set by ref_pN_set_vincent
static double g_c_next = 0.0;

static SolverBase *Create(int scheme, int riemann, GlobalMesh &&mesh) {
  namespace sp = mini::spatial;
  if (scheme == 0) {
    switch (riemann) {
    case 0: return new SolverImpl<sp::fr::Lobatto<Part, RExact>>(std::move(mesh));
    case 1: return new SolverImpl<sp::fr::Lobatto<Part, RHllc>>(std::move(mesh));
    case 2: return new SolverImpl<sp::fr::Lobatto<Part, RAusm>>(std::move(mesh));
    }
  } else if (scheme == 1) {
    switch (riemann) {
    case 0: return new SolverImpl<sp::dg::Lobatto<Part, RExact>>(std::move(mesh));
    case 1: return new SolverImpl<sp::dg::Lobatto<Part, RHllc>>(std::move(mesh));
    case 2: return new SolverImpl<sp::dg::Lobatto<Part, RAusm>>(std::move(mesh));
    }
  } else if (scheme == 2) {
    switch (riemann) {
    case 0: return new SolverImpl<sp::fr::General<Part, RExact>>(std::move(mesh), g_c_next);
    case 1: return new SolverImpl<sp::fr::General<Part, RHllc>>(std::move(mesh), g_c_next);
    case 2: return new SolverImpl<sp::fr::General<Part, RAusm>>(std::move(mesh), g_c_next);
    }
  } else if (scheme == 4) {   // WithLimiter<fr::Lobatto, weno::Lazy> (demo/euler/sourceless.hpp:72-80 with FR + LIMITER)
    using Lazy = mini::limiter::weno::Lazy<typename ModalPart::Cell>;
    switch (riemann) {
    case 0: return new SolverImpl<sp::WithLimiter<sp::fr::Lobatto<ModalPart, RExact>, Lazy>>(std::move(mesh));
    case 1: return new SolverImpl<sp::WithLimiter<sp::fr::Lobatto<ModalPart, RHllc>, Lazy>>(std::move(mesh));
    }
  } else if (scheme == 13) {  // WithLimiter<dg::Lobatto, weno::Lazy> (demo/euler/sourceless.hpp:67-69, 76-80: DGSEM + LIMITER)
    using Lazy = mini::limiter::weno::Lazy<typename ModalPart::Cell>;
    switch (riemann) {
    case 0: return new SolverImpl<sp::WithLimiter<sp::dg::Lobatto<ModalPart, RExact>, Lazy>>(std::move(mesh));
    case 1: return new SolverImpl<sp::WithLimiter<sp::dg::Lobatto<ModalPart, RHllc>, Lazy>>(std::move(mesh));
    }
  } else if (scheme == 5) {   // WithLimiter<fr::Lobatto, weno::Eigen> (demo/euler/rotorcraft.hpp:41 names the limiter so)
    using Cell = typename ModalPart::Cell;
    switch (riemann) {
    case 0: return new SolverImpl<sp::WithLimiter<sp::fr::Lobatto<ModalPart, RExact>, mini::limiter::weno::Eigen<Cell, RExact>>>(std::move(mesh));
    case 1: return new SolverImpl<sp::WithLimiter<sp::fr::Lobatto<ModalPart, RHllc>, mini::limiter::weno::Eigen<Cell, RHllc>>>(std::move(mesh));
    }
  } else if (scheme == 11) {  // fr::Lobatto on Hexahedron<..., true>, Euler
    switch (riemann) {
    case 0: return new SolverImpl<sp::fr::Lobatto<PartL, RExact>>(std::move(mesh));
    case 1: return new SolverImpl<sp::fr::Lobatto<PartL, RHllc>>(std::move(mesh));
    case 2: return new SolverImpl<sp::fr::Lobatto<PartL, RAusm>>(std::move(mesh));
    }
  } else if (scheme == 12) {  // ... Navier-Stokes + direct DG: demo/laminar/shockless.hpp as shipped
    switch (riemann) {
    case 0: return new SolverImpl<sp::fr::Lobatto<PartL, VExact>>(std::move(mesh));
    case 1: return new SolverImpl<sp::fr::Lobatto<PartL, VHllc>>(std::move(mesh));
    case 2: return new SolverImpl<sp::fr::Lobatto<PartL, VAusm>>(std::move(mesh));
    }
  } else if (scheme == 10) {  // WithLimiter<dg::General, weno::Lazy> on Projection cells: the reference's DGFEM + LIMITER build
    using Lazy = mini::limiter::weno::Lazy<typename DgPart::Cell>;
    switch (riemann) {
    case 0: return new SolverImpl<sp::WithLimiter<sp::dg::General<DgPart, RExact>, Lazy>>(std::move(mesh));
    case 1: return new SolverImpl<sp::WithLimiter<sp::dg::General<DgPart, RHllc>, Lazy>>(std::move(mesh));
    }
  } else if (scheme == 9) {   // dg::General<Part<Projection>, Riemann>
    switch (riemann) {
    case 0: return new SolverImpl<sp::dg::General<DgPart, RExact>>(std::move(mesh));
    case 1: return new SolverImpl<sp::dg::General<DgPart, RHllc>>(std::move(mesh));
    case 2: return new SolverImpl<sp::dg::General<DgPart, RAusm>>(std::move(mesh));
    }
  } else if (scheme == 8) {   // fr::General with an arbitrary Vincent c on the Navier-Stokes Riemann type
    switch (riemann) {
    case 0: return new SolverImpl<sp::fr::General<Part, VExact>>(std::move(mesh), g_c_next);
    case 1: return new SolverImpl<sp::fr::General<Part, VHllc>>(std::move(mesh), g_c_next);
    }
  } else if (scheme == 7) {   // single part only: EnergyBasedViscosity keeps its state in class statics
    if (mesh.n_parts != 1) return nullptr;
    switch (riemann) {
    case 0: return new SolverImpl<sp::WithViscosity<sp::fr::Lobatto<Part, EExact>>>(std::move(mesh));
    case 1: return new SolverImpl<sp::WithViscosity<sp::fr::Lobatto<Part, EHllc>>>(std::move(mesh));
    }
  } else if (scheme == 20) {   // single part only, HLLC only
    if (mesh.n_parts != 1 || riemann != 1) return nullptr;
    return new SolverImpl<sp::WithViscosity<sp::fr::Lobatto<Part, ENsHllc>>>(std::move(mesh));
  } else if (scheme == 6) {
    switch (riemann) {
    case 0: return new SolverImpl<sp::fr::Lobatto<Part, IExact>>(std::move(mesh));
    case 1: return new SolverImpl<sp::fr::Lobatto<Part, IHllc>>(std::move(mesh));
    }
  } else if (scheme == 3) {
    switch (riemann) {
    case 0: return new SolverImpl<sp::fr::Lobatto<Part, VExact>>(std::move(mesh));
    case 1: return new SolverImpl<sp::fr::Lobatto<Part, VHllc>>(std::move(mesh));
    case 2: return new SolverImpl<sp::fr::Lobatto<Part, VAusm>>(std::move(mesh));
    }
  }
  return nullptr;
}

}  // namespace oracle

// -------------------------------------------------------------------------------------------------
// extern "C" surface; one library per polynomial degree (symbol prefix carries the degree).
// -------------------------------------------------------------------------------------------------
#define REF_CAT2(a, b) a##b
#define REF_CAT(a, b) REF_CAT2(a, b)
#define REF_FN(name) REF_CAT(REF_CAT(REF_CAT(ref_p, REF_DEGREE), _), name)

using oracle::FieldSpec;
using oracle::Int;
using oracle::SolverBase;

extern "C" {

int REF_FN(degree)() { return REF_DEGREE; }

void REF_FN(set_vincent)(double c_next) { oracle::g_c_next = c_next; }

// class-static configuration of the diffusive part (direct.hpp:80-83, navier_stokes.hpp:57-60)
void REF_FN(set_viscous)(double beta_0, double beta_1, double nu, double prandtl) {
  oracle::Diffusion::SetBetaValues(beta_0, beta_1);
  oracle::Diffusion::SetProperty(nu, prandtl);
  oracle::IsoDiffusion::SetBetaValues(beta_0, beta_1);   // linear.hpp:118-120
  oracle::IsoDiffusion::SetProperty(nu);
}

// EnergyBasedViscosity::SetTimeScale (viscosity.hpp:224-226, demo/euler/sourceless.cpp:121): class-static; call it after
// the solver exists (Initialize resets it to 1)
void REF_FN(set_time_scale)(double time_scale) {
  oracle::EExact::SetTimeScale(time_scale);
  oracle::EHllc::SetTimeScale(time_scale);
}

void *REF_FN(create)(int scheme, int riemann, Int n_nodes, const double *xyz, Int n_cells,
    const Int *cell_nodes, const int *cell_parts, int n_parts, Int n_bfaces,
    const Int *bface_nodes, const int *bface_tags) {
  oracle::GlobalMesh mesh;
  mesh.nodes.resize(n_nodes);
  for (Int i = 0; i < n_nodes; ++i) mesh.nodes[i] = {xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]};
  mesh.cells.resize(n_cells);
  for (Int i = 0; i < n_cells; ++i)
    for (int v = 0; v < 8; ++v) mesh.cells[i][v] = cell_nodes[8 * i + v];
  mesh.n_parts = std::max(1, n_parts);
  mesh.cell_parts.assign(n_cells, 0);
  if (cell_parts) mesh.cell_parts.assign(cell_parts, cell_parts + n_cells);
  mesh.bfaces.resize(n_bfaces);
  mesh.btags.resize(n_bfaces);
  for (Int i = 0; i < n_bfaces; ++i) {
    for (int v = 0; v < 4; ++v) mesh.bfaces[i][v] = bface_nodes[4 * i + v];
    mesh.btags[i] = bface_tags[i];
  }
  return oracle::Create(scheme, riemann, std::move(mesh));
}
void REF_FN(destroy)(void *h) { delete static_cast<SolverBase *>(h); }
int REF_FN(n_parts)(void *h) { return static_cast<SolverBase *>(h)->n_parts(); }
Int REF_FN(count_local)(void *h, int r) { return static_cast<SolverBase *>(h)->count_local(r); }
Int REF_FN(count_ghost)(void *h, int r) { return static_cast<SolverBase *>(h)->count_ghost(r); }
void REF_FN(cell_global_ids)(void *h, int r, Int *out) {
  static_cast<SolverBase *>(h)->cell_global_ids(r, out);
}
Int REF_FN(count_faces)(void *h, int r, int kind) {
  return static_cast<SolverBase *>(h)->count_faces(r, kind);
}
void REF_FN(face_table)(void *h, int r, int kind, Int *holder, Int *sharer, int *holder_face) {
  static_cast<SolverBase *>(h)->face_table(r, kind, holder, sharer, holder_face);
}
void REF_FN(halo_table)(void *h, int r, int send, Int *n_out, Int *parts, Int *counts, Int *cells) {
  static_cast<SolverBase *>(h)->halo_table(r, send, n_out, parts, counts, cells);
}
void REF_FN(set_bc)(void *h, int tag, int bc_kind, int field_kind, const double *params,
    int n_params, oracle::ref_value_callback cb, void *user) {
  FieldSpec s;
  s.kind = field_kind;
  for (int i = 0; i < n_params && i < 16; ++i) s.params[i] = params[i];
  s.cb = cb;
  s.user = user;
  static_cast<SolverBase *>(h)->set_bc(tag, bc_kind, s);
}
void REF_FN(approximate)(void *h, int field_kind, const double *params, int n_params,
    oracle::ref_value_callback cb, void *user) {
  FieldSpec s;
  s.kind = field_kind;
  for (int i = 0; i < n_params && i < 16; ++i) s.params[i] = params[i];
  s.cb = cb;
  s.user = user;
  static_cast<SolverBase *>(h)->approximate(s);
}
int REF_FN(cell_viscosity)(void *h, double *nu) { return static_cast<SolverBase *>(h)->cell_viscosity(nu); }
int REF_FN(damping_matrices)(void *h, double *d) { return static_cast<SolverBase *>(h)->damping_matrices(d); }
void REF_FN(set_weno)(double w0, double eps) { oracle::g_weno_w0 = w0; oracle::g_weno_eps = eps; }
void REF_FN(set_solution_all)(void *h, const double *cols) { static_cast<SolverBase *>(h)->set_solution_all(cols); }
void REF_FN(set_solution)(void *h, int r, const double *col) {
  static_cast<SolverBase *>(h)->set_solution(r, col);
}
void REF_FN(get_solution)(void *h, int r, double *col) {
  static_cast<SolverBase *>(h)->get_solution(r, col);
}
void REF_FN(residual_all)(void *h, double t) { static_cast<SolverBase *>(h)->residual_all(t); }
void REF_FN(get_residual)(void *h, int r, double *col) {
  static_cast<SolverBase *>(h)->get_residual(r, col);
}
void REF_FN(rk_update)(void *h, int order, double t, double dt, int n_steps) {
  static_cast<SolverBase *>(h)->rk_update(order, t, dt, n_steps);
}
double REF_FN(time_step)(void *h, double dt_guess, int rk_order) {
  return static_cast<SolverBase *>(h)->time_step(dt_guess, rk_order);
}
void REF_FN(limiter)(void *h, int kind) { static_cast<SolverBase *>(h)->limiter(kind); }
void REF_FN(sample_vtk)(void *h, int r, double *points, double *values) {
  static_cast<SolverBase *>(h)->sample_vtk(r, points, values);
}
void REF_FN(write_vtu)(void *h, const char *directory, const char *soln_name) {
  static_cast<SolverBase *>(h)->write_vtu(directory, soln_name);
}
void REF_FN(cell_geometry)(void *h, int r, double *length, double *volume) {
  static_cast<SolverBase *>(h)->cell_geometry(r, length, volume);
}
// One rotated Riemann problem through rotated::Euler<...>::GetFluxUpwind (rotated/euler.hpp:153-159):
// frame = {normal, tangent, bitangent} row-major, u_left/u_right conservative.  kind: 0 Exact,
// 1 HLLC, 2 AUSM; bc < 0 interior, bc == 0 inviscid wall (rotated/euler.hpp:160-174).
void REF_FN(riemann_flux)(int kind, int bc, const double *frame, const double *u_left,
    const double *u_right, double *flux) {
  using Vec = mini::algebra::Vector<double, 3>;
  std::array<Vec, 3> fr{Vec(frame[0], frame[1], frame[2]), Vec(frame[3], frame[4], frame[5]),
      Vec(frame[6], frame[7], frame[8])};
  auto run = [&](auto riemann) {
    using R = decltype(riemann);
    riemann.Rotate(fr);
    typename R::Conservative ul, ur;
    for (int k = 0; k < 5; ++k) { ul[k] = u_left[k]; ur[k] = u_right[k]; }
    typename R::Flux f = bc == 0 ? riemann.GetFluxOnInviscidWall(ul) : riemann.GetFluxUpwind(ul, ur);
    for (int k = 0; k < 5; ++k) flux[k] = f[k];
  };
  if (kind == 0) run(oracle::RExact());
  else if (kind == 1) run(oracle::RHllc());
  else run(oracle::RAusm());
}

}  // extern "C"
