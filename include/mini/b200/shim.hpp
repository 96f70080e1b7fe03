// mini/b200/shim.hpp — header-style C++ host side of the B200 path.
//
// Keeps the names, template parameters and member functions that a miniCFD driver uses for the hot
// path (demo/euler/sourceless.hpp:15-111, demo/euler/sourceless.cpp:107-221 of the reference) and forwards the
// work to the C ABI of include/minicfd_b200.h.  The classes are thin: geometry, connectivity and the
// solution live on the GPU; the host keeps the mesh description and the boundary functions.
//
//   reference type (include/mini/...)                          here
//   integrator::Lobatto<S,Q>          integrator/lobatto.hpp:20   tag carrying Q
//   polynomial::Hexahedron<Gx,Gy,Gz,K,kLocal>  polynomial/hexahedron.hpp:41  tag carrying P, K, N + value types
//   riemann::euler::IdealGas / Exact / HartenLaxLeerContact / AdvectionUpstreamSplittingMethod
//                                     riemann/euler/*.hpp         tags + host-side PrimitiveToConservative
//   riemann::rotated::Euler<U>        riemann/rotated/euler.hpp:23  tag
//   mesh::part::Part<Int,Poly>        mesh/part.hpp:404           in-memory part (no CGNS/MPI I/O)
//   spatial::fr::Lobatto / dg::Lobatto  spatial/fr/lobatto.hpp:36, spatial/dg/lobatto.hpp:34   GPU-backed System
//   temporal::System / Euler / RungeKutta   temporal/ode.hpp:17-134, temporal/rk.hpp:10-57
#ifndef MINI_B200_SHIM_HPP_
#define MINI_B200_SHIM_HPP_

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <fstream>
#include <sys/stat.h>
#include <functional>
#include <map>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include "minicfd_b200.h"
#include "mini/b200/comm.hpp"

namespace mini {

namespace algebra {
// Stand-in for Eigen's fixed-size vectors in driver code (value semantics, operator[]).
template <typename Scalar, int kSize>
struct Vector : std::array<Scalar, kSize> {
  Vector() { this->fill(Scalar(0)); }
  Vector(std::initializer_list<Scalar> il) { this->fill(Scalar(0)); std::copy(il.begin(), il.end(), this->begin()); }
  static Vector Zero() { return Vector(); }
};
// Stand-in for Eigen::Vector<Scalar, Dynamic>: the `Column` of temporal::System (temporal/ode.hpp:19).
template <typename Scalar>
struct DynamicVector : std::vector<Scalar> {
  using std::vector<Scalar>::vector;
  DynamicVector &operator+=(DynamicVector const &that) {
    for (size_t i = 0; i < this->size(); ++i) (*this)[i] += that[i];
    return *this;
  }
  DynamicVector &operator*=(Scalar s) { for (auto &x : *this) x *= s; return *this; }
  DynamicVector &operator/=(Scalar s) { for (auto &x : *this) x /= s; return *this; }
  Scalar squaredNorm() const { Scalar n = 0; for (auto x : *this) n += x * x; return n; }
  void setZero() { std::fill(this->begin(), this->end(), Scalar(0)); }
};
}  // namespace algebra

namespace integrator {
template <typename S, int N>
struct Lobatto {  // integrator/lobatto.hpp:20-93
  using Scalar = S;
  static constexpr int Q = N;
  static std::array<S, N> BuildPoints() {
    if constexpr (N == 3) return {-1.0, 0.0, 1.0};
    else return {-1.0, -std::sqrt(1.0 / 5), std::sqrt(1.0 / 5), 1.0};
  }
};
template <typename S, int N>
struct Legendre {  // integrator/legendre.hpp:20-80; the rule itself lives in the library (MCFD_DG_GENERAL)
  using Scalar = S;
  static constexpr int Q = N;
};
}  // namespace integrator

namespace polynomial {
template <class Gx, class Gy, class Gz, int kComponents, bool kL = false>
struct Hexahedron {  // polynomial/hexahedron.hpp:41-65
  static_assert(Gx::Q == Gy::Q && Gy::Q == Gz::Q, "the B200 path needs the same rule in all directions");
  static_assert(Gx::Q == 3 || Gx::Q == 4, "degree 2 or 3");
  using Scalar = typename Gx::Scalar;
  using IntegratorX = Gx;
  static constexpr bool kLocal = kL;
  static constexpr int P = Gx::Q - 1;
  static constexpr int N = Gx::Q * Gx::Q * Gx::Q;
  static constexpr int K = kComponents;
  static constexpr int D = 3;
  static constexpr int kFields = K * N;
  using Global = algebra::Vector<Scalar, 3>;
  using Local = Global;
  using Value = algebra::Vector<Scalar, K>;
};
// polynomial/projection.hpp:67-176: the stand-alone modal expansion on the cell's orthonormal basis, K x M coefficients
// per cell (M = 10 at P = 2, 20 at P = 3).  A tag here: the basis and the quadrature live in the library.
template <typename S, int kDimensions, int kDegrees, int kComponents>
struct Projection {
  static_assert(kDimensions == 3 && (kDegrees == 2 || kDegrees == 3), "three-dimensional, degree 2 or 3");
  using Scalar = S;
  static constexpr bool kLocal = false;
  static constexpr bool kModalOnly = true;
  static constexpr int P = kDegrees;
  static constexpr int Q = kDegrees + 1;                                  // Gauss-Legendre points per direction
  static constexpr int N = (kDegrees + 1) * (kDegrees + 2) * (kDegrees + 3) / 6;   // basis/taylor.hpp:27-31
  static constexpr int K = kComponents;
  static constexpr int D = 3;
  static constexpr int kFields = K * N;
  using Global = algebra::Vector<Scalar, 3>;
  using Local = Global;
  using Value = algebra::Vector<Scalar, K>;
};
// polynomial/extrapolation.hpp:20-188: the nodal expansion together with its projection on the cell's orthonormal
// basis.  On the GPU the modal side exists only inside the limiter kernels, so this is a tag.
template <class I>
struct Extrapolation : I {
  using Interpolation = I;
  static constexpr bool kModal = true;
};
}  // namespace polynomial

namespace riemann {
namespace euler {
template <typename S, int kDimensions>
struct Primitives : algebra::Vector<S, kDimensions + 2> {  // types.hpp:114-173
  using algebra::Vector<S, kDimensions + 2>::Vector;
  S &rho() { return (*this)[0]; }
  S &u() { return (*this)[1]; }
  S &v() { return (*this)[2]; }
  S &w() { return (*this)[3]; }
  S &p() { return (*this)[kDimensions + 1]; }
  S rho() const { return (*this)[0]; }
  S u() const { return (*this)[1]; }
  S v() const { return (*this)[2]; }
  S w() const { return (*this)[3]; }
  S p() const { return (*this)[kDimensions + 1]; }
};
template <typename S, int kDimensions>
struct Conservatives : algebra::Vector<S, kDimensions + 2> {  // types.hpp:175-184
  using algebra::Vector<S, kDimensions + 2>::Vector;
};
template <typename S, int kGammaTimes10>
struct IdealGasTag {
  using Scalar = S;
  static_assert(kGammaTimes10 == 14, "gamma = 1.4 (demo/euler/sourceless.hpp:21)");
  static constexpr S Gamma() { return 1.4; }
  static constexpr S R() { return 287.05; }
  // host-callable converters the drivers' initial and boundary conditions use (types.hpp:248-287)
  static S GetSpeedOfSound(S temperature) { return temperature > 0 ? std::sqrt(Gamma() * R() * temperature) : 0; }
  static S GetSpeedOfSound(S rho, S p) { return (rho > 0 && p > 0) ? std::sqrt(Gamma() * p / rho) : 0; }
  static S GetMachFactor(S mach) { return 1 + (Gamma() - 1) / 2 * mach * mach; }
  static S TotalTemperatureToTemperature(S mach, S total_temperature) { return total_temperature / GetMachFactor(mach); }
  static S TotalPressureToPressure(S mach, S total_pressure) {
    return total_pressure / std::pow(GetMachFactor(mach), Gamma() / (Gamma() - 1));
  }
  // types.hpp:302-311
  static Conservatives<S, 3> PrimitiveToConservative(Primitives<S, 3> const &q) {
    Conservatives<S, 3> c;
    c[0] = q.rho(); c[1] = q.rho() * q.u(); c[2] = q.rho() * q.v(); c[3] = q.rho() * q.w();
    c[4] = q.p() * (1 / (Gamma() - 1)) + q.rho() * (0.5 * (q.u() * q.u() + q.v() * q.v() + q.w() * q.w()));
    return c;
  }
};
// The reference spells IdealGas<Scalar, 1.4> with a floating-point non-type parameter (C++20).
template <typename S, double kGamma>
using IdealGas = IdealGasTag<S, static_cast<int>(kGamma * 10 + 0.5)>;

template <class G, int D> struct Exact { using Gas = G; static constexpr int kKind = MCFD_EXACT; };
template <class G, int D> struct HartenLaxLeerContact { using Gas = G; static constexpr int kKind = MCFD_HLLC; };
template <class G, int D> struct AdvectionUpstreamSplittingMethod { using Gas = G; static constexpr int kKind = MCFD_AUSM; };
}  // namespace euler

namespace rotated {
template <class Unrotated>
struct Euler {  // rotated/euler.hpp:23-40
  using Gas = typename Unrotated::Gas;
  using Scalar = typename Gas::Scalar;
  using Convection = Euler;
  static constexpr int kComponents = 5;
  static constexpr int kDimensions = 3;
  static constexpr int kKind = Unrotated::kKind;
  static constexpr int kDiffusion = MCFD_INVISCID;
  using Conservative = euler::Conservatives<Scalar, 3>;
  using Primitive = euler::Primitives<Scalar, 3>;
  using Value = algebra::Vector<Scalar, 5>;
};
}  // namespace rotated

namespace diffusive {
// riemann/diffusive/navier_stokes.hpp:20-228 as a tag: the class-static property (:54-63) is kept here and
// handed to the context before every residual evaluation (mcfd_set_diffusion).
template <class G>
struct NavierStokes {
  using Gas = G;
  using Scalar = typename Gas::Scalar;
  static constexpr int kComponents = 5;
  static constexpr int kDimensions = 3;
  static constexpr int kDiffusionTag = MCFD_NAVIER_STOKES_DDG;
  struct Property { Scalar nu, prandtl; };
  static void SetProperty(Scalar nu, Scalar prandtl) { property_ = Property{nu, prandtl}; }
  static Property const &GetProperty() { return property_; }
  inline static Property property_{0.0, 0.708};
};
// riemann/diffusive/linear.hpp:104-121: one viscosity for every component (F_k = -nu grad u_k), as a tag
template <typename S, int K>
struct Isotropic {
  using Scalar = S;
  static constexpr int kComponents = K;
  static constexpr int kDimensions = 3;
  static constexpr int kDiffusionTag = MCFD_CELL_VISCOSITY_DDG;
  struct Property { Scalar nu, prandtl; };
  static void SetProperty(Scalar nu) { property_ = Property{nu, 1.0}; }
  static Property const &GetProperty() { return property_; }
  inline static Property property_{0.0, 1.0};
};
// riemann/diffusive/direct.hpp:15-91
template <class DiffusionModel>
struct Direct : DiffusionModel {
  using Diffusion = Direct;
  using Scalar = typename DiffusionModel::Scalar;
  static void SetBetaValues(Scalar beta_0, Scalar beta_1) { beta_0_ = beta_0; beta_1_ = beta_1; }
  inline static Scalar beta_0_ = 2.0, beta_1_ = 1.0 / 12;
};
}  // namespace diffusive

// riemann/concept.hpp:41-91 on the tags: what a driver's static_assert can still check
template <typename R>
concept Convective = requires { typename R::Conservative; R::kKind; R::kComponents; R::kDimensions; };
template <typename R>
concept Diffusive = requires { typename R::Diffusion; } && (R::kDiffusion != MCFD_INVISCID);
template <typename R>
concept ConvectiveDiffusive = Convective<R> && Diffusive<R>;

// riemann/concept.hpp:93-121
template <class C, class D>
struct ConvectionDiffusion : C, D {
  using Convection = C;
  using Diffusion = D;
  using Scalar = typename C::Scalar;
  using Gas = typename C::Gas;
  using Conservative = typename C::Conservative;
  using Primitive = typename C::Primitive;
  using Value = typename C::Value;
  static constexpr int kComponents = C::kComponents;
  static constexpr int kDimensions = C::kDimensions;
  static constexpr int kKind = C::kKind;
  static constexpr int kDiffusion = D::kDiffusionTag;
};
}  // namespace riemann

namespace b200 {
inline void Check(int status, mcfd_ctx *ctx, const char *what) {
  if (status != 0) {
    throw std::runtime_error(std::string(what) + ": " + mcfd_last_error(ctx));
  }
}

// Wire the context of this rank's part to the contexts of the neighbouring parts, one process per part
// (minicfd_b200.h, "peer" section; the same hand-shake as minicfd_b200/peer.py): every rank publishes the CUDA IPC
// handles of its epoch words, time-step slots and receive stage together with its neighbour list, opens what its
// neighbours published, and connects.  Collective.  Halo: parts / send_counts / recv_counts as given to SetMesh.
template <class Halo>
void PeerConnect(mcfd_ctx *ctx, Halo const &halo, Comm &comm) {
  constexpr int kMaxNeighbours = 32;
  struct Card {
    mcfd_ipc_handle handles[3];
    int32_t n_neighbours;
    int32_t neighbours[kMaxNeighbours];
    int64_t recv_counts[kMaxNeighbours];
  };
  static_assert(sizeof(Card) <= Comm::kBlockBytes);
  const int world = comm.size(), rank = comm.rank(), n = (int)halo.parts.size();
  if (n > kMaxNeighbours) throw std::runtime_error("PeerConnect: more than 32 neighbouring parts");
  Card mine{};
  void *own[3];
  Check(mcfd_peer_export(ctx, world, mine.handles, own), ctx, "mcfd_peer_export");
  mine.n_neighbours = n;
  for (int i = 0; i < n; ++i) { mine.neighbours[i] = (int32_t)halo.parts[i]; mine.recv_counts[i] = halo.recv_counts[i]; }
  std::vector<Card> all(world);
  comm.Allgather(&mine, sizeof(Card), all.data());
  std::vector<std::array<void *, 3>> opened(world, {nullptr, nullptr, nullptr});
  auto pointer = [&](int r, int which) {
    if (r == rank) return own[which];
    if (!opened[r][which]) Check(mcfd_peer_open(ctx, &all[r].handles[which], &opened[r][which]), ctx, "mcfd_peer_open");
    return opened[r][which];
  };
  std::vector<void *> flags(n), recv(n), world_flags(world), world_slots(world);
  std::vector<int32_t> slot(n);
  std::vector<int64_t> items(n), offset(n);
  for (int i = 0; i < n; ++i) {
    const Card &q = all[halo.parts[i]];
    int there = -1;
    for (int j = 0; j < q.n_neighbours; ++j) if (q.neighbours[j] == rank) there = j;
    if (there < 0 || q.recv_counts[there] != halo.send_counts[i]) throw std::runtime_error("PeerConnect: the send / recv lists of two parts disagree");
    flags[i] = pointer((int)halo.parts[i], 0);
    recv[i] = pointer((int)halo.parts[i], 2);
    slot[i] = there;
    items[i] = 0;
    for (int j = 0; j < q.n_neighbours; ++j) { if (j == there) offset[i] = items[i]; items[i] += q.recv_counts[j]; }
  }
  for (int r = 0; r < world; ++r) { world_flags[r] = pointer(r, 0); world_slots[r] = pointer(r, 1); }
  Check(mcfd_peer_connect(ctx, rank, world, n, flags.data(), slot.data(), recv.data(), items.data(), offset.data(),
      world_flags.data(), world_slots.data()), ctx, "mcfd_peer_connect");
  comm.Barrier();   // nobody delivers before everybody is connected
}
}  // namespace b200

namespace mesh {
namespace part {

// In-memory part.  The reference constructs a Part from `<case>/shuffled.cgns` + `partition/<rank>.txt`
// (part.hpp:442-476); that I/O needs CGNS + MPI and stays out of scope, so the mesh is
// handed over as arrays, already numbered the way part.hpp:746-775 numbers it: local cells (inner
// first, then inter), then ghost cells.
template <typename Int, class Poly>
class Part {
 public:
  using Index = Int;
  using Polynomial = Poly;
  using Scalar = typename Poly::Scalar;
  using Global = typename Poly::Global;
  using Value = typename Poly::Value;
  static constexpr int kDegrees = Poly::P;
  static constexpr int kComponents = Poly::K;
  static constexpr int kFields = Poly::kFields;
  // Host-side views of the part's cells and boundary faces (part.hpp:186-291, 97-184): the geometry a driver or a
  // callback asks for.  The polynomial coefficients live on the device; GatherSolutions() brings them to the host
  // array that WriteSolutions / the cell views read (part.hpp:380-387, 1052-1069).
  class Cell {
   public:
    using Polynomial = Poly;
    using Scalar = typename Poly::Scalar;
    using Global = typename Poly::Global;
    using Value = typename Poly::Value;
    Int metis_id = -1;                                      // part.hpp:190: the cell's id before partitioning
    Int id() const { return id_; }                          // part.hpp:245
    bool inner() const { return inner_; }                   // part.hpp:248
    bool inter() const { return !inner_; }                  // part.hpp:251
    Global const &center() const { return center_; }        // part.hpp:254, coordinate/hexahedron.hpp center()
    Scalar volume() const { return volume_; }               // part.hpp:242
    Scalar length() const { return length_; }               // part.hpp:239: cbrt(volume) / 2
    int CountCorners() const { return 8; }                  // part.hpp:277
    int CountFields() const { return Poly::kFields; }       // part.hpp:280
    // the gathered coefficients of this cell, [N][K] (nodal schemes) or [K][M] (modal), after Part::GatherSolutions()
    Scalar const *coeff() const { return coeff_; }
    Int id_ = -1;
    bool inner_ = true;
    Global center_{};
    Scalar volume_ = 0, length_ = 0;
    Scalar const *coeff_ = nullptr;
  };
  class Face {   // boundary faces: holder only (part.hpp:97-184)
   public:
    Int id() const { return id_; }                          // part.hpp:162
    Global const &center() const { return center_; }        // part.hpp:148
    Scalar area() const { return area_; }                   // part.hpp:151
    Cell const &holder() const { return *holder_; }         // part.hpp:165
    Cell const *other(Cell const *cell) const { return cell == holder_ ? nullptr : holder_; }
    Int id_ = -1;
    Global center_{};
    Scalar area_ = 0;
    Cell const *holder_ = nullptr;
  };

  struct Halo {
    std::vector<int64_t> parts, send_counts, send_cells, recv_counts, recv_cells;
  };

  Part(int rank, int size) : rank_(rank), size_(size) {}

  // nodes [n_nodes][3]; cell_nodes [n_local + n_ghost][8] (CGNS HEXA_8); boundary quads with the
  // name of their section (the BC names of the drivers, e.g. "Left", demo/euler/shock_tube.cpp:56-66).
  void SetMesh(std::vector<Scalar> nodes, std::vector<int64_t> cell_nodes, Int n_local, Int n_ghost,
      std::vector<int64_t> bface_nodes, std::vector<std::string> const &bface_names, Halo halo = Halo()) {
    nodes_ = std::move(nodes); cell_nodes_ = std::move(cell_nodes);
    n_local_ = n_local; n_ghost_ = n_ghost; bface_nodes_ = std::move(bface_nodes); halo_ = std::move(halo);
    cells_built_ = false;
    bface_tags_.clear();
    for (auto const &name : bface_names) {
      auto it = name_to_tag_.find(name);
      if (it == name_to_tag_.end()) it = name_to_tag_.emplace(name, (int32_t)name_to_tag_.size()).first;
      bface_tags_.push_back(it->second);
    }
  }
  // The boundary sections of the whole mesh: a part knows every section of its zone, also those it has no face of
  // (part.hpp:1456-1549 reads all of them from the file).  Call before SetMesh.
  void DeclareSections(std::vector<std::string> const &names) {
    for (auto const &name : names) if (!name_to_tag_.count(name)) name_to_tag_.emplace(name, (int32_t)name_to_tag_.size());
  }
  int mpi_rank() const { return rank_; }
  int mpi_size() const { return size_; }
  // part.hpp:477-485, 1565-1579 (what the VTK writer asks of a part); the reference takes the directory from the
  // case name in its constructor
  void SetFieldNames(std::vector<std::string> names) { field_names_ = std::move(names); }
  const std::string &GetFieldName(int i) const { return field_names_.at(i); }
  void SetDirectoryName(std::string directory) { directory_ = std::move(directory); }
  const std::string &GetDirectoryName() const { return directory_; }
  std::ofstream GetFileStream(std::string const &soln_name, bool binary, std::string const &suffix) const {
    ::mkdir(directory_.c_str(), 0777);   // the reference runs `mkdir -p` (part.hpp:1568-1573)
    ::mkdir((directory_ + "/" + soln_name).c_str(), 0777);
    return std::ofstream(directory_ + "/" + soln_name + "/" + std::to_string(rank_) + "." + suffix,
        std::ios::out | (binary ? (std::ios::binary) : std::ios::out));
  }
  Int CountLocalCells() const { return n_local_; }                     // part.hpp:651-658
  Int CountGhostCells() const { return n_ghost_; }
  Int GetCellDataOffset(Int i_cell) const { return i_cell * kFields; }  // part.hpp:784-787
  Int GetCellDataSize() const { return n_local_ * kFields; }            // part.hpp:793-797
  bool IsLocal(Int i_cell) const { return 0 <= i_cell && i_cell < n_local_; }
  bool IsGhost(Int i_cell) const { return n_local_ <= i_cell && i_cell < n_local_ + n_ghost_; }   // part.hpp:1048
  // the cells' ids before partitioning (Cell::metis_id, part.hpp:190), local then ghost; optional
  void SetCellGlobalIds(std::vector<int64_t> ids) { global_ids_ = std::move(ids); cells_built_ = false; }

  // part.hpp:1330-1365.  Local cells in id order: inner cells (no ghost neighbour) first, then inter cells.
  std::vector<Cell> const &GetLocalCells() const { BuildCells(); return cells_; }
  std::vector<Cell *> const &GetInnerCellPointers() { BuildCells(); return inner_ptrs_; }
  std::vector<Cell *> const &GetInterCellPointers() { BuildCells(); return inter_ptrs_; }
  Cell const &GetCell(Int i_cell) const { BuildCells(); return cells_.at(i_cell); }
  // part.hpp:1399-1409: the faces of the boundary section `name`, in the order of the boundary face table
  std::vector<Face> const &GetBoundaryFaces(std::string const &name) const {
    BuildCells();
    auto it = bfaces_by_name_.find(name);
    if (it == bfaces_by_name_.end()) throw std::runtime_error("no boundary section named " + name);
    return it->second;
  }

  // part.hpp:1273-1314.  The coefficients of the send cells travel to the neighbouring parts' ghost cells: here
  // between the GPUs, enqueued on the context's stream (mcfd_peer_push / mcfd_peer_wait); every rank calls both.
  void ShareGhostCellCoeffs() {
    if (n_ghost_ == 0) return;
    RequireContext("ShareGhostCellCoeffs");
    b200::Check(mcfd_peer_push(context_), context_, "mcfd_peer_push");
  }
  void UpdateGhostCellCoeffs() {
    if (n_ghost_ == 0) return;
    RequireContext("UpdateGhostCellCoeffs");
    b200::Check(mcfd_peer_wait(context_), context_, "mcfd_peer_wait");
  }
  // part.hpp:1052-1069: polynomial coefficients -> the per-cell data array a frame is written from, and back
  void GatherSolutions() {
    RequireContext("GatherSolutions");
    cell_data_.resize((size_t)GetCellDataSize());
    b200::Check(mcfd_get_solution(context_, cell_data_.data()), context_, "mcfd_get_solution");
    BuildCells();
    for (Int c = 0; c < n_local_; ++c) cells_[c].coeff_ = cell_data_.data() + GetCellDataOffset(c);
  }
  void ScatterSolutions() {
    RequireContext("ScatterSolutions");
    if ((Int)cell_data_.size() != GetCellDataSize()) throw std::runtime_error("ScatterSolutions: nothing gathered or read");
    b200::Check(mcfd_set_solution(context_, cell_data_.data()), context_, "mcfd_set_solution");
    ShareGhostCellCoeffs();
    UpdateGhostCellCoeffs();
  }
  // part.hpp:1023-1037: sum over the local cells of the quadrature of |u - exact(x, t)|, per component (nodal schemes:
  // the cell's own Lobatto rule, whose points are the nodes).  Gathers the solution first.
  template <class Callable>
  Value MeasureL1Error(Callable &&exact_solution, Scalar t_curr) {
    static_assert(!requires { Poly::kModalOnly; }, "MeasureL1Error: nodal polynomials only");
    GatherSolutions();
    constexpr int Q = Poly::IntegratorX::Q, N = Poly::N, K = Poly::K;
    const std::array<Scalar, 4> w = Q == 3 ? std::array<Scalar, 4>{1.0 / 3, 4.0 / 3, 1.0 / 3, 0} : std::array<Scalar, 4>{1.0 / 6, 5.0 / 6, 5.0 / 6, 1.0 / 6};
    std::vector<double> det((size_t)n_local_ * N);
    b200::Check(mcfd_get_jacobian_determinants(context_, det.data()), context_, "mcfd_get_jacobian_determinants");
    Value l1_error{};
    for (Int c = 0; c < n_local_; ++c) {
      for (int q = 0; q < N; ++q) {
        const Scalar d = det[(size_t)c * N + q], weight = w[q / (Q * Q)] * w[(q / Q) % Q] * w[q % Q] * d;
        Value exact = exact_solution(NodeCoordinate(c, q), t_curr);
        for (int k = 0; k < K; ++k) {
          Scalar u = cell_data_[GetCellDataOffset(c) + q * K + k];
          if constexpr (Poly::kLocal) u /= d;   // the coefficient is det(J) u (hexahedron.hpp:344-349)
          l1_error[k] += std::abs(u - exact[k]) * weight;
        }
      }
    }
    return l1_error;
  }
  std::vector<Scalar> &cell_data() { return cell_data_; }
  std::vector<Scalar> const &cell_data() const { return cell_data_; }
  int32_t TagOf(std::string const &name) const {
    auto it = name_to_tag_.find(name);
    if (it == name_to_tag_.end()) throw std::runtime_error("no boundary section named " + name);
    return it->second;
  }
  // Global coordinates of the solution points of a local cell (integrator/hexahedron.hpp:66-80).
  Global NodeCoordinate(Int i_cell, int ijk) const {
    auto x = Poly::IntegratorX::BuildPoints();
    constexpr int Q = Poly::IntegratorX::Q;
    static const int v[8][3] = {{-1, -1, -1}, {1, -1, -1}, {1, 1, -1}, {-1, 1, -1},
                                {-1, -1, 1}, {1, -1, 1}, {1, 1, 1}, {-1, 1, 1}};
    Scalar loc[3] = {x[ijk / (Q * Q)], x[(ijk / Q) % Q], x[ijk % Q]};
    Global g;
    for (int c = 0; c < 8; ++c) {
      Scalar s = (1 + v[c][0] * loc[0]) * (1 + v[c][1] * loc[1]) * (1 + v[c][2] * loc[2]) / 8;
      for (int d = 0; d < 3; ++d) g[d] += s * nodes_[cell_nodes_[i_cell * 8 + c] * 3 + d];
    }
    return g;
  }

  void RequireContext(const char *what) const {
    if (!context_) throw std::runtime_error(std::string(what) + ": no spatial object has been built on this part yet");
  }
  // the host views: centres from the vertices, volume and length from the library (the reference's own quadrature,
  // part.hpp:230-232), inner / inter from the send lists (part.hpp:746-767)
  void BuildCells() const {
    if (cells_built_ && cells_context_ == context_) return;
    cells_.assign((size_t)n_local_, Cell());
    std::vector<double> length, volume;
    if (context_) {
      length.resize((size_t)n_local_); volume.resize((size_t)n_local_);
      b200::Check(mcfd_get_cell_geometry(context_, length.data(), volume.data()), context_, "mcfd_get_cell_geometry");
    }
    for (Int c = 0; c < n_local_; ++c) {
      Cell &cell = cells_[c];
      cell.id_ = c;
      cell.metis_id = global_ids_.empty() ? c : (Int)global_ids_[c];
      for (int v = 0; v < 8; ++v) for (int d = 0; d < 3; ++d) cell.center_[d] += nodes_[cell_nodes_[c * 8 + v] * 3 + d] / 8;
      if (context_) { cell.volume_ = volume[c]; cell.length_ = length[c]; }
    }
    for (int64_t s : halo_.send_cells) cells_[s].inner_ = false;
    inner_ptrs_.clear(); inter_ptrs_.clear();
    for (auto &cell : cells_) (cell.inner_ ? inner_ptrs_ : inter_ptrs_).push_back(&cell);
    bfaces_by_name_.clear();
    for (auto const &kv : name_to_tag_) bfaces_by_name_[kv.first];
    std::unordered_map<int32_t, std::string> name_of;
    for (auto const &kv : name_to_tag_) name_of[kv.second] = kv.first;
    // the holder of a boundary quad: the local cell that has its four nodes
    std::unordered_map<int64_t, std::vector<Int>> cells_of_node;
    for (Int c = 0; c < n_local_; ++c) for (int v = 0; v < 8; ++v) cells_of_node[cell_nodes_[c * 8 + v]].push_back(c);
    static const double gx[3] = {-0.7745966692414834, 0.0, 0.7745966692414834}, gw[3] = {5.0 / 9, 8.0 / 9, 5.0 / 9};
    for (size_t f = 0; f < bface_tags_.size(); ++f) {
      Face face;
      face.id_ = (Int)f;
      const int64_t *q = &bface_nodes_[f * 4];
      double x[4][3];
      for (int v = 0; v < 4; ++v) for (int d = 0; d < 3; ++d) { x[v][d] = nodes_[q[v] * 3 + d]; face.center_[d] += x[v][d] / 4; }
      for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) {   // |dx/da x dx/db| of the bilinear quad, 3 x 3 Gauss points
        double da[3], db[3];
        for (int d = 0; d < 3; ++d) {
          da[d] = ((1 - gx[b]) * (x[1][d] - x[0][d]) + (1 + gx[b]) * (x[2][d] - x[3][d])) / 4;
          db[d] = ((1 - gx[a]) * (x[3][d] - x[0][d]) + (1 + gx[a]) * (x[2][d] - x[1][d])) / 4;
        }
        const double n[3] = {da[1] * db[2] - da[2] * db[1], da[2] * db[0] - da[0] * db[2], da[0] * db[1] - da[1] * db[0]};
        face.area_ += gw[a] * gw[b] * std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
      }
      for (Int c : cells_of_node[q[0]]) {
        int hit = 0;
        for (int v = 0; v < 8; ++v) for (int m = 0; m < 4; ++m) hit += cell_nodes_[c * 8 + v] == q[m];
        if (hit == 4) { face.holder_ = &cells_[c]; break; }
      }
      bfaces_by_name_[name_of[bface_tags_[f]]].push_back(face);
    }
    cells_built_ = true;
    cells_context_ = context_;
  }

  std::vector<Scalar> nodes_;
  std::vector<int64_t> cell_nodes_, bface_nodes_, global_ids_;
  std::vector<Scalar> cell_data_;
  mutable std::vector<Cell> cells_;
  mutable std::vector<Cell *> inner_ptrs_, inter_ptrs_;
  mutable std::map<std::string, std::vector<Face>> bfaces_by_name_;
  mutable bool cells_built_ = false;
  mutable mcfd_ctx *cells_context_ = nullptr;
  std::vector<int32_t> bface_tags_;
  std::unordered_map<std::string, int32_t> name_to_tag_;
  Halo halo_;
  Int n_local_ = 0, n_ghost_ = 0;
  std::string directory_ = ".";
  std::vector<std::string> field_names_;
  // the context of the spatial object built on this part: what `mini::limiter::Reconstruct(&part, ...)` and the
  // like, which the reference hands a Part, act on
  mcfd_ctx *context_ = nullptr;

 private:
  int rank_, size_;
};

}  // namespace part
}  // namespace mesh

namespace temporal {

template <typename Scalar>
class System {  // temporal/ode.hpp:17-48
 public:
  using Column = algebra::DynamicVector<Scalar>;
  virtual ~System() = default;
  virtual void SetTime(double t_curr) = 0;
  virtual void SetSolutionColumn(Column const &column) = 0;
  virtual Column GetSolutionColumn() const = 0;
  virtual Column GetResidualColumn() const = 0;
  // GPU-backed systems advance a whole RungeKutta<order>::Update with fused stage kernels.
  virtual bool FusedUpdate(int order, double t_curr, double dt) { return false; }
};

template <typename Scalar>
class Solver {  // temporal/ode.hpp:92-105
 public:
  using Column = typename System<Scalar>::Column;
  virtual ~Solver() = default;
  virtual void Update(System<Scalar> *ode, double t_curr, double dt) = 0;
};

template <typename Scalar>
class Euler : public Solver<Scalar> {  // temporal/ode.hpp:112-134
 public:
  using Column = typename Solver<Scalar>::Column;
  static Column NextSolution(System<Scalar> *system, double t_curr, double dt) {
    system->SetTime(t_curr);
    auto u_next = system->GetSolutionColumn();
    auto residual = system->GetResidualColumn();
    residual *= dt;
    u_next += residual;
    return u_next;
  }
  void Update(System<Scalar> *system, double t_curr, double dt) final {
    if (system->FusedUpdate(1, t_curr, dt)) return;
    system->SetSolutionColumn(NextSolution(system, t_curr, dt));
  }
};

template <int kOrders, typename Scalar>
struct RungeKutta : public Solver<Scalar> {  // temporal/rk.hpp:10-57
  static_assert(1 <= kOrders && kOrders <= 3);
  using Column = typename Solver<Scalar>::Column;
  void Update(System<Scalar> *system, double t_curr, double dt) final {
    if (system->FusedUpdate(kOrders, t_curr, dt)) return;   // one fused kernel per stage
    UpdateGeneric(system, t_curr, dt);
  }
  // The reference's column algebra, through the System interface (every residual still on the GPU).
  void UpdateGeneric(System<Scalar> *system, double t_curr, double dt) {
    Euler<Scalar> euler;
    if constexpr (kOrders == 1) {
      system->SetSolutionColumn(euler.NextSolution(system, t_curr, dt));
    } else if constexpr (kOrders == 2) {
      auto u_curr = system->GetSolutionColumn();
      system->SetSolutionColumn(euler.NextSolution(system, t_curr, dt));
      auto u_next = euler.NextSolution(system, t_curr + dt, dt);
      u_next += u_curr;
      u_next *= 0.5;
      system->SetSolutionColumn(u_next);
    } else {
      auto u_curr = system->GetSolutionColumn();
      system->SetSolutionColumn(euler.NextSolution(system, t_curr, dt));
      auto u_next = u_curr;
      u_next *= 3;
      u_next += euler.NextSolution(system, t_curr + dt, dt);
      u_next /= 4;
      system->SetSolutionColumn(u_next);
      u_next = euler.NextSolution(system, t_curr + dt / 2, dt);
      u_next *= 2;
      u_next += u_curr;
      u_next /= 3;
      system->SetSolutionColumn(u_next);
    }
  }
};

}  // namespace temporal

namespace spatial {

// Common GPU-backed implementation of FiniteElement (spatial/fem.hpp:47-436) for the nodal schemes.
template <class P, class R, int kScheme>
class FiniteElementB200 : public temporal::System<typename P::Scalar> {
 public:
  using Part = P;
  using Riemann = R;
  using Scalar = typename Part::Scalar;
  using Global = typename Part::Global;
  using Value = typename Part::Value;
  using Column = typename temporal::System<Scalar>::Column;
  using Function = std::function<Value(const Global &, double)>;  // fem.hpp:72

  explicit FiniteElementB200(Part *part_ptr, int device = -1, double c_next = 0.0) : part_ptr_(part_ptr) {
    mcfd_config cfg{kScheme, Part::kDegrees, Riemann::kKind, device, c_next, Riemann::kDiffusion, 0};
    if constexpr (requires { Part::Polynomial::kLocal; }) cfg.k_local = Part::Polynomial::kLocal ? 1 : 0;
    if (mcfd_create(&cfg, &ctx_) != 0) throw std::runtime_error(std::string("mcfd_create: ") + mcfd_last_error(nullptr));
    auto &p = *part_ptr_;
    b200::Check(mcfd_set_mesh(ctx_, (int64_t)p.nodes_.size() / 3, p.nodes_.data(), p.n_local_, p.n_ghost_,
        p.cell_nodes_.data(), (int64_t)p.bface_tags_.size(), p.bface_nodes_.data(), p.bface_tags_.data(), 0, nullptr),
        ctx_, "mcfd_set_mesh");
    if (p.n_ghost_ > 0) {
      auto &h = p.halo_;
      b200::Check(mcfd_set_halo(ctx_, (int32_t)h.send_counts.size(), h.send_counts.data(), h.send_cells.data(),
          (int32_t)h.recv_counts.size(), h.recv_counts.data(), h.recv_cells.data()), ctx_, "mcfd_set_halo");
      if (p.mpi_size() > 1) {   // one process per part: the ghost exchange runs between the GPUs
        b200::PeerConnect(ctx_, h, b200::Comm::World());
        peer_connected_ = true;
      }
    }
    b200::Check(mcfd_set_time_callback(ctx_, &FiniteElementB200::OnTime, this), ctx_, "mcfd_set_time_callback");
    p.context_ = ctx_;
  }
  FiniteElementB200(const FiniteElementB200 &) = delete;
  FiniteElementB200 &operator=(const FiniteElementB200 &) = delete;
  ~FiniteElementB200() noexcept override {
    if (part_ptr_->context_ == ctx_) part_ptr_->context_ = nullptr;
    mcfd_destroy(ctx_);
  }

  Part *part_ptr() const { return part_ptr_; }
  Part const &part() const { return *part_ptr_; }
  // fem.hpp:113-116; the per-face solver objects do not exist on this side (nothing reads them)
  std::vector<Riemann> const &GetRiemannSolvers(typename Part::Face const &) const {
    static const std::vector<Riemann> none;
    return none;
  }
  mcfd_ctx *context() const { return ctx_; }
  bool peer_connected() const { return peer_connected_; }   // the ghost exchange runs between the GPUs (b200::PeerConnect)

  template <class Callable>
  void Approximate(Callable &&func) {  // fem.hpp:147-154
    if constexpr (requires { Part::Polynomial::kModalOnly; }) {   // Projection::Approximate: L2 projection
      constexpr int Q = Part::Polynomial::Q, n_points = Q * Q * Q, K = Part::kComponents;
      const size_t n_cells = (size_t)part().CountLocalCells();
      std::vector<double> xyz(n_cells * n_points * 3), values(n_cells * n_points * K);
      b200::Check(mcfd_get_quadrature_points(ctx_, xyz.data()), ctx_, "mcfd_get_quadrature_points");
      for (size_t i = 0; i < n_cells * n_points; ++i) {
        Global x;
        for (int d = 0; d < 3; ++d) x[d] = xyz[i * 3 + d];
        Value v = func(x);
        for (int k = 0; k < K; ++k) values[i * K + k] = v[k];
      }
      b200::Check(mcfd_project(ctx_, values.data()), ctx_, "mcfd_project");
    } else {
      Column column(part().GetCellDataSize());
      constexpr int N = Part::Polynomial::N, K = Part::kComponents;
      std::vector<double> det;
      if constexpr (Part::Polynomial::kLocal) {   // Hexahedron::SetValue: coeff = value * det J (hexahedron.hpp:344-349)
        det.resize((size_t)part().CountLocalCells() * N);
        b200::Check(mcfd_get_jacobian_determinants(ctx_, det.data()), ctx_, "mcfd_get_jacobian_determinants");
      }
      for (typename Part::Index c = 0; c < part().CountLocalCells(); ++c) {
        for (int q = 0; q < N; ++q) {
          Value v = func(part().NodeCoordinate(c, q));
          const double scale = det.empty() ? 1.0 : det[(size_t)c * N + q];
          for (int k = 0; k < K; ++k) column[(c * N + q) * K + k] = v[k] * scale;
        }
      }
      SetSolutionColumn(column);
    }
  }

  // fem.hpp:196-222
  template <typename F> void SetSmartBoundary(const std::string &name, F &&f) { SetBc(name, MCFD_BC_SMART, f); }
  template <typename F> void SetSupersonicInlet(const std::string &name, F &&f) { SetBc(name, MCFD_BC_SUPERSONIC_INLET, f); }
  template <typename F> void SetSubsonicInlet(const std::string &name, F &&f) { SetBc(name, MCFD_BC_SUBSONIC_INLET, f); }
  template <typename F> void SetSubsonicOutlet(const std::string &name, F &&f) { SetBc(name, MCFD_BC_SUBSONIC_OUTLET, f); }
  // ignored by inviscid Riemann types, like the reference (fr/general.hpp:455-459)
  template <typename F> void SetNoSlipWall(const std::string &name, F &&f) { SetBc(name, MCFD_BC_NO_SLIP_WALL, f); }
  void SetInviscidWall(const std::string &name) { SetBc(name, MCFD_BC_INVISCID_WALL, Function()); }
  void SetSupersonicOutlet(const std::string &name) { SetBc(name, MCFD_BC_SUPERSONIC_OUTLET, Function()); }

  // temporal::System (fem.hpp:225-258)
  void SetTime(double t_curr) override { t_curr_ = t_curr; }
  void SetSolutionColumn(Column const &column) override {
    b200::Check(mcfd_set_solution(ctx_, column.data()), ctx_, "mcfd_set_solution");
  }
  Column GetSolutionColumn() const override {
    Column column(part().GetCellDataSize());
    b200::Check(mcfd_get_solution(ctx_, column.data()), ctx_, "mcfd_get_solution");
    return column;
  }
  Column GetResidualColumn() const override {   // fem.hpp:248-258
    Column column(part().GetCellDataSize());
    PushDiffusion();
    if (peer_connected_) {
      part_ptr_->ShareGhostCellCoeffs();
      part_ptr_->UpdateGhostCellCoeffs();
    }
    b200::Check(mcfd_residual(ctx_, t_curr_, column.data(), nullptr), ctx_, "mcfd_residual");
    return column;
  }
  bool FusedUpdate(int order, double t_curr, double dt) override {
    PushDiffusion();
    if (part().CountGhostCells() == 0) {
      b200::Check(mcfd_rk_step(ctx_, order, t_curr, dt), ctx_, "mcfd_rk_step");
      return true;
    }
    if (!peer_connected_) return false;   // ghosts filled by the host: driven residual by residual
    // every stage: deliver the stage input's send cells, advance the inner cells meanwhile, then the inter cells
    // (fem.hpp:249-256 with the exchange of part.hpp:1273-1314 between the GPUs)
    b200::Check(mcfd_rk_begin(ctx_, order, t_curr, dt), ctx_, "mcfd_rk_begin");
    for (int stage = 0; stage < order; ++stage) {
      part_ptr_->ShareGhostCellCoeffs();
      b200::Check(mcfd_rk_stage_inner(ctx_, stage), ctx_, "mcfd_rk_stage_inner");
      part_ptr_->UpdateGhostCellCoeffs();
      b200::Check(mcfd_rk_stage_inter(ctx_, stage), ctx_, "mcfd_rk_stage_inter");
    }
    return true;
  }
  Scalar GetTimeStep(Scalar dt_guess, int rk_order) const {  // fem.hpp:417-435, fr/lobatto.hpp:220-222
    double dt = 0;
    b200::Check(mcfd_time_step(ctx_, dt_guess, rk_order, &dt), ctx_, "mcfd_time_step");
    return dt;
  }

 private:
  // the reference reads its class statics at every flux evaluation; hand their current values over
  void PushDiffusion() const {
    if constexpr (requires { typename Riemann::Viscosity; }) {
      // per-cell values owned by the viscosity rule; the betas went over when WithViscosity was constructed
    } else if constexpr (Riemann::kDiffusion != MCFD_INVISCID) {
      using D = typename Riemann::Diffusion;
      b200::Check(mcfd_set_diffusion(ctx_, D::beta_0_, D::beta_1_, D::GetProperty().nu, D::GetProperty().prandtl),
          ctx_, "mcfd_set_diffusion");
    }
  }
  struct Bc { Function func; std::vector<double> points, values; };
  void SetBc(const std::string &name, int kind, Function func) {
    int32_t tag = part().TagOf(name);
    b200::Check(mcfd_set_boundary(ctx_, tag, kind), ctx_, "mcfd_set_boundary");
    if (!func) return;
    Bc &bc = bcs_[tag];
    bc.func = std::move(func);
    constexpr int fq = (Part::kDegrees + 1) * (Part::kDegrees + 1);
    int64_t n = mcfd_count_boundary_faces(ctx_, tag);
    bc.points.resize(n * fq * 3);
    bc.values.resize(n * fq * 5);
    if (n) b200::Check(mcfd_get_boundary_points(ctx_, tag, bc.points.data()), ctx_, "mcfd_get_boundary_points");
    Sample(tag, t_curr_);
  }
  // the reference evaluates func(integrator.GetGlobal(f), t_curr_) inside the face loop
  // (fr/lobatto.hpp:350); here the values are sampled once per residual evaluation and uploaded
  void Sample(int32_t tag, double t) {
    Bc &bc = bcs_[tag];
    size_t n = bc.points.size() / 3;
    for (size_t i = 0; i < n; ++i) {
      Global x;
      for (int d = 0; d < 3; ++d) x[d] = bc.points[i * 3 + d];
      Value v = bc.func(x, t);
      for (int k = 0; k < 5; ++k) bc.values[i * 5 + k] = v[k];
    }
    if (n) b200::Check(mcfd_set_boundary_values(ctx_, tag, bc.values.data()), ctx_, "mcfd_set_boundary_values");
  }
  static void OnTime(void *self, double t) {
    auto *fe = static_cast<FiniteElementB200 *>(self);
    for (auto &kv : fe->bcs_) fe->Sample(kv.first, t);
  }

  Part *part_ptr_;
  mcfd_ctx *ctx_ = nullptr;
  bool peer_connected_ = false;
  double t_curr_ = 0;
  std::map<int32_t, Bc> bcs_;
};

}  // namespace spatial

namespace limiter {
namespace weno {
// limiter/weno.hpp:97-195; the object only carries (w0, eps), the work is mcfd_limiter_weno_lazy
template <typename Cell>
class Lazy {
 public:
  using Scalar = typename Cell::Scalar;
  static constexpr bool kEigen = false;
  Lazy(Scalar w0, Scalar eps, bool verbose = false) : w0_(w0), eps_(eps) { (void)verbose; }
  Scalar w0() const { return w0_; }
  Scalar eps() const { return eps_; }
 private:
  Scalar w0_, eps_;
};
// limiter/weno.hpp:197-349 -> mcfd_limiter_weno_eigen.  The face frames the reference takes from the installed
// Riemann solvers (rotated::Euler::a/b/c of the face's first quadrature point) are derived from the mesh by the
// library, so InstallRiemannSolvers has nothing left to do.
template <typename Cell, typename Riemann>
class Eigen {
 public:
  using Scalar = typename Cell::Scalar;
  static constexpr bool kEigen = true;
  Eigen(Scalar w0, Scalar eps) : w0_(w0), eps_(eps) {}
  template <typename FaceToRiemann> void InstallRiemannSolvers(FaceToRiemann &&) {}
  Scalar w0() const { return w0_; }
  Scalar eps() const { return eps_; }
 private:
  Scalar w0_, eps_;
};
}  // namespace weno

// limiter/reconstruct.hpp:21-54 on the part's context
template <typename Part, typename Limiter>
void Reconstruct(Part *part_ptr, Limiter *limiter_ptr) {
  mcfd_ctx *ctx = part_ptr->context_;
  if (!ctx) throw std::runtime_error("mini::limiter::Reconstruct: no spatial object has been built on this part");
  if constexpr (Limiter::kEigen) {
    b200::Check(mcfd_limiter_weno_eigen(ctx, limiter_ptr->w0(), limiter_ptr->eps(), -1), ctx, "mcfd_limiter_weno_eigen");
  } else {
    b200::Check(mcfd_limiter_weno_lazy(ctx, limiter_ptr->w0(), limiter_ptr->eps(), -1), ctx, "mcfd_limiter_weno_lazy");
  }
}
namespace average {
template <typename Part>
void Reconstruct(Part *part_ptr) {   // limiter/average.hpp:11-28
  mcfd_ctx *ctx = part_ptr->context_;
  if (!ctx) throw std::runtime_error("average::Reconstruct: no spatial object has been built on this part");
  b200::Check(mcfd_limiter_reconstruct(ctx, 0), ctx, "mcfd_limiter_reconstruct");
}
}  // namespace average
namespace majority {
template <typename Part>
void Reconstruct(Part *part_ptr) {   // limiter/majority.hpp:14-42
  mcfd_ctx *ctx = part_ptr->context_;
  if (!ctx) throw std::runtime_error("majority::Reconstruct: no spatial object has been built on this part");
  b200::Check(mcfd_limiter_reconstruct(ctx, 1), ctx, "mcfd_limiter_reconstruct");
}
}  // namespace majority
}  // namespace limiter

namespace spatial {

// spatial/with_limiter.hpp:21-68: the limiter follows every SetSolutionColumn / Approximate; a fused RK step
// runs it after every stage, which is what the column algebra of rk.hpp amounts to on such a system.
template <typename ConcreteFiniteElement, typename Limiter>
class WithLimiter : public ConcreteFiniteElement {
 public:
  using Base = ConcreteFiniteElement;
  using Column = typename Base::Column;
  template <class... Args>
  WithLimiter(Limiter *limiter_ptr, Args &&...args) : Base(std::forward<Args>(args)...), limiter_ptr_(limiter_ptr) {}
  Limiter *limiter_ptr() const { return limiter_ptr_; }
  void SetSolutionColumn(Column const &column) override {
    this->Base::SetSolutionColumn(column);
    Reconstruct();
  }
  template <class Callable>
  void Approximate(Callable &&func) {   // Base::Approximate ends in the virtual SetSolutionColumn above
    Base::Approximate(std::forward<Callable>(func));
  }
  bool FusedUpdate(int order, double t_curr, double dt) override {
    const bool ghosts = this->part().CountGhostCells() > 0;
    if (ghosts && !this->peer_connected()) return false;
    mcfd_ctx *ctx = this->context();
    b200::Check(mcfd_rk_begin(ctx, order, t_curr, dt), ctx, "mcfd_rk_begin");
    for (int stage = 0; stage < order; ++stage) {
      // the residual sees the neighbours' LIMITED coefficients (fem.hpp:249-256 after with_limiter.hpp:57-61)
      if (ghosts) this->part_ptr()->ShareGhostCellCoeffs();
      b200::Check(mcfd_rk_stage_inner(ctx, stage), ctx, "mcfd_rk_stage_inner");
      if (ghosts) this->part_ptr()->UpdateGhostCellCoeffs();
      b200::Check(mcfd_rk_stage_inter(ctx, stage), ctx, "mcfd_rk_stage_inter");
      Reconstruct();
    }
    return true;
  }
 private:
  void Limit(int phase) {
    mcfd_ctx *ctx = this->context();
    if constexpr (Limiter::kEigen) {
      b200::Check(mcfd_limiter_weno_eigen(ctx, limiter_ptr_->w0(), limiter_ptr_->eps(), phase), ctx, "mcfd_limiter_weno_eigen");
    } else {
      b200::Check(mcfd_limiter_weno_lazy(ctx, limiter_ptr_->w0(), limiter_ptr_->eps(), phase), ctx, "mcfd_limiter_weno_lazy");
    }
  }
  // mini::limiter::Reconstruct(part_ptr, limiter_ptr), limiter/reconstruct.hpp:21-54: the unlimited coefficients
  // travel while the inner cells are limited; the inter cells then borrow the limited inner cells and the
  // unlimited ghost cells
  void Reconstruct() {
    if (!limiter_ptr_) return;
    if (this->part().CountGhostCells() > 0 && this->peer_connected()) {
      this->part_ptr()->ShareGhostCellCoeffs();
      Limit(0);
      this->part_ptr()->UpdateGhostCellCoeffs();
      Limit(1);
    } else {
      Limit(-1);
    }
  }
  Limiter *limiter_ptr_;
};

// spatial/viscosity.hpp:35-632 as a tag with the class-static interface the drivers and WithViscosity use
// (demo/euler/sourceless.cpp:118-126, 153).  Like the reference's statics, one spatial object per process is
// served at a time: the context installed last.
template <typename P, typename R>
class EnergyBasedViscosity : public R {
 public:
  using Base = R;
  using Part = P;
  using Scalar = typename R::Scalar;
  using Viscosity = EnergyBasedViscosity;                   // viscosity.hpp:547
  using Property = algebra::Vector<Scalar, R::kComponents>;  // viscosity.hpp:54
  // over a diffusive base (Navier-Stokes, demo/laminar/shocked.hpp:100-111) the base's flux stays and the artificial
  // one comes on top (viscosity.hpp:76-139); over Direct<Isotropic> the rule owns the viscosities alone
  static constexpr bool kDiffusiveBase = R::kDiffusion == MCFD_NAVIER_STOKES_DDG;
  static constexpr int kDiffusion = kDiffusiveBase ? MCFD_NAVIER_STOKES_CELL_VISCOSITY_DDG : MCFD_CELL_VISCOSITY_DDG;

  template <class Spatial>
  static void Initialize(Spatial *spatial_ptr) {   // viscosity.hpp:549-554
    ctx_ = spatial_ptr->context();
    n_cells_ = spatial_ptr->part().CountLocalCells() + spatial_ptr->part().CountGhostCells();
    using D = typename R::Diffusion;
    if constexpr (kDiffusiveBase) {
      b200::Check(mcfd_set_diffusion(ctx_, D::beta_0_, D::beta_1_, D::GetProperty().nu, D::GetProperty().prandtl), ctx_, "mcfd_set_diffusion");
    } else {
      b200::Check(mcfd_set_diffusion(ctx_, D::beta_0_, D::beta_1_, 0.0, 1.0), ctx_, "mcfd_set_diffusion");
    }
    b200::Check(mcfd_build_damping_matrices(ctx_), ctx_, "mcfd_build_damping_matrices");
    SetTimeScale(1.0);
  }
  static Scalar GetTimeScale() { return time_scale_; }                     // viscosity.hpp:221-226
  static void SetTimeScale(Scalar time_scale) { time_scale_ = time_scale; }
  static void UpdateProperties() {                                          // viscosity.hpp:556-572
    b200::Check(mcfd_update_cell_viscosity(ctx_, time_scale_), ctx_, "mcfd_update_cell_viscosity");
    cache_valid_ = false;
  }
  // the ghost cells' values travel with mcfd_viscosity_halo_* where a host moves bytes between parts
  // viscosity.hpp:163-189.  With one part there is nothing to exchange.  With several parts driven from C++ the cell
  // viscosities would have to travel like the coefficients; that path exists on the Python host only
  // (mcfd_viscosity_halo_*, minicfd_b200/spatial.py), so a multi-part C++ run is refused when it is built
  // (WithViscosity's constructor) instead of running with stale ghost viscosities.
  static void ShareGhostCellProperties() {}
  static void UpdateGhostCellProperties() {}
  template <typename Int>
  static Property const &GetPropertyOnCell(Int i_cell, int /* i_node */) {  // viscosity.hpp:69-72: constant in a cell
    if (!cache_valid_) {
      flat_.resize((size_t)n_cells_ * R::kComponents);
      b200::Check(mcfd_get_cell_viscosity(ctx_, flat_.data()), ctx_, "mcfd_get_cell_viscosity");
      cache_.resize(n_cells_);
      for (int64_t c = 0; c < n_cells_; ++c) {
        for (int k = 0; k < R::kComponents; ++k) cache_[c][k] = flat_[c * R::kComponents + k];
      }
      cache_valid_ = true;
    }
    return cache_.at(i_cell);
  }
  static mcfd_ctx *context() { return ctx_; }

 private:
  inline static mcfd_ctx *ctx_ = nullptr;
  inline static int64_t n_cells_ = 0;
  inline static Scalar time_scale_ = 1.0;
  inline static bool cache_valid_ = false;
  inline static std::vector<Scalar> flat_;
  inline static std::vector<Property> cache_;
};

// spatial/with_viscosity.hpp:20-73
template <typename ConcreteFiniteElement>
class WithViscosity : public ConcreteFiniteElement {
 public:
  using Base = ConcreteFiniteElement;
  using Riemann = typename Base::Riemann;
  using Scalar = typename Base::Scalar;
  using Column = typename Base::Column;
  template <class... Args>
  WithViscosity(Args &&...args) : Base(std::forward<Args>(args)...) {
    if (this->part().CountGhostCells() > 0) {   // see EnergyBasedViscosity::ShareGhostCellProperties above
      throw std::runtime_error("WithViscosity: parts with ghost cells are not served from C++ (the exchange of the "
          "cell viscosities, with_viscosity.hpp:45-50, exists on the Python host only)");
    }
    Riemann::Viscosity::Initialize(this);
  }
  void SetSolutionColumn(Column const &column) override {
    this->Base::SetSolutionColumn(column);
    Riemann::Viscosity::UpdateProperties();
  }
  template <class Callable>
  void Approximate(Callable &&func) {   // Base::Approximate ends in the virtual SetSolutionColumn above
    Base::Approximate(std::forward<Callable>(func));
  }
  bool FusedUpdate(int order, double t_curr, double dt) override {
    if (this->part().CountGhostCells() > 0) return false;
    mcfd_ctx *ctx = this->context();
    b200::Check(mcfd_rk_begin(ctx, order, t_curr, dt), ctx, "mcfd_rk_begin");
    for (int stage = 0; stage < order; ++stage) {
      b200::Check(mcfd_rk_stage_inner(ctx, stage), ctx, "mcfd_rk_stage_inner");
      b200::Check(mcfd_rk_stage_inter(ctx, stage), ctx, "mcfd_rk_stage_inter");
      Riemann::Viscosity::UpdateProperties();
    }
    return true;
  }
  // with_viscosity.hpp:67-71: CFL * min_dt_ of the last UpdateProperties -- what mcfd_time_step returns on such a
  // context (Base::GetTimeStep)
};

namespace fr {
template <typename P, typename R>
class Lobatto : public FiniteElementB200<P, R, MCFD_FR_LOBATTO> {  // spatial/fr/lobatto.hpp:36-434
 public:
  using FiniteElementB200<P, R, MCFD_FR_LOBATTO>::FiniteElementB200;
  std::string name() const { return "FR::Lobatto"; }
};
template <typename P, typename R>
class General : public FiniteElementB200<P, R, MCFD_FR_GENERAL> {  // spatial/fr/general.hpp:34-641
 public:
  General(P *part_ptr, typename P::Scalar c_next) : FiniteElementB200<P, R, MCFD_FR_GENERAL>(part_ptr, -1, c_next) {}
  std::string name() const { return "FR::General"; }
};
}  // namespace fr
namespace dg {
template <typename P, typename R>
class General : public FiniteElementB200<P, R, MCFD_DG_GENERAL> {  // spatial/dg/general.hpp:20-200
 public:
  static_assert(P::Polynomial::kModalOnly, "dg::General is built for the modal Projection polynomial");
  using FiniteElementB200<P, R, MCFD_DG_GENERAL>::FiniteElementB200;
  std::string name() const { return "DG::General"; }
};
template <typename P, typename R>
class Lobatto : public FiniteElementB200<P, R, MCFD_DG_LOBATTO> {  // spatial/dg/lobatto.hpp:32-303
 public:
  using FiniteElementB200<P, R, MCFD_DG_LOBATTO>::FiniteElementB200;
  std::string name() const { return "DG::Lobatto"; }
};
}  // namespace dg

}  // namespace spatial
}  // namespace mini

#endif  // MINI_B200_SHIM_HPP_
