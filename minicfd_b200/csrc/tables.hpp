// tables.hpp — host-side construction of the per-degree constant tables and of the Part/Cell/Face
// connectivity (the "a1/a8" rows of SURVEY.md section 8): everything the reference finds at set-up
// by coordinate matching (fr/lobatto.hpp:72-163, polynomial/hexahedron.hpp:692-706, 821-928) is
// derived here from node ids instead.
#ifndef MINICFD_B200_TABLES_HPP_
#define MINICFD_B200_TABLES_HPP_

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <map>
#include <string>
#include <unordered_map>
#include <vector>

#include "kernels.cuh"

namespace mcfd {

// local face -> vertex slots, outward-pointing right-hand order
// (include/mini/coordinate/hexahedron.hpp:378-389)
static const int kFaceNodes[6][4] = {
  {0, 3, 2, 1}, {0, 1, 5, 4}, {1, 2, 6, 5}, {2, 3, 7, 6}, {0, 4, 7, 3}, {4, 5, 6, 7}};
// reference coordinates of the 8 vertex slots (coordinate/hexahedron.hpp:344-349)
static const int kVertex[8][3] = {
  {-1, -1, -1}, {+1, -1, -1}, {+1, +1, -1}, {-1, +1, -1},
  {-1, -1, +1}, {+1, -1, +1}, {+1, +1, +1}, {-1, +1, +1}};
// corners of a face in its own (a, b) parametrisation (coordinate/quadrangle.hpp:134-139)
static const int kCorner[4][2] = {{-1, -1}, {+1, -1}, {+1, +1}, {-1, +1}};

inline void LobattoRule(int q, double *x, double *w) {  // integrator/lobatto.hpp:38-57
  if (q == 3) {
    x[0] = -1; x[1] = 0; x[2] = 1;
    w[0] = 1.0 / 3; w[1] = 4.0 / 3; w[2] = 1.0 / 3;
  } else {  // q == 4
    x[0] = -1; x[1] = -std::sqrt(1.0 / 5); x[2] = +std::sqrt(1.0 / 5); x[3] = 1;
    w[0] = 1.0 / 6; w[1] = 5.0 / 6; w[2] = 5.0 / 6; w[3] = 1.0 / 6;
  }
}

inline void LegendreRule(int q, double *x, double *w) {  // integrator/legendre.hpp:50-73
  if (q == 3) {
    x[0] = -std::sqrt(0.6); x[1] = 0; x[2] = std::sqrt(0.6);
    w[0] = 5.0 / 9.0; w[1] = 8.0 / 9.0; w[2] = 5.0 / 9.0;
  } else {
    const double a = std::sqrt((3 + 2 * std::sqrt(1.2)) / 7), b = std::sqrt((3 - 2 * std::sqrt(1.2)) / 7);
    x[0] = -a; x[1] = -b; x[2] = b; x[3] = a;
    w[0] = w[3] = (18 - std::sqrt(30.0)) / 36; w[1] = w[2] = (18 + std::sqrt(30.0)) / 36;
  }
}

// l_m'(x_i): derivative of the m-th Lagrange polynomial at node i (basis/lagrange.hpp:72-133 obtains
// the same numbers through a Taylor-matrix inverse).
inline double LagrangeDerivative(int q, const double *x, int m, int i) {
  double sum = 0;
  for (int k = 0; k < q; ++k) {
    if (k == m) continue;
    double prod = 1.0 / (x[m] - x[k]);
    for (int l = 0; l < q; ++l) {
      if (l == m || l == k) continue;
      prod *= (x[i] - x[l]) / (x[m] - x[l]);
    }
    sum += prod;
  }
  return sum;
}

// Vincent::LocalToRightDerivative(1) with c = HuynhLumpingLobatto (basis/vincent.hpp:31-33, 75-91).
// l_m''(x_i) (basis/lagrange.hpp:130-133 with k = 2)
inline double LagrangeSecondDerivative(int q, const double *x, int m, int i) {
  double sum = 0;
  for (int k = 0; k < q; ++k) {
    if (k == m) continue;
    for (int j = 0; j < q; ++j) {
      if (j == m || j == k) continue;
      double prod = 1.0 / ((x[m] - x[k]) * (x[m] - x[j]));
      for (int l = 0; l < q; ++l) {
        if (l == m || l == k || l == j) continue;
        prod *= (x[i] - x[l]) / (x[m] - x[l]);
      }
      sum += prod;
    }
  }
  return sum;
}

inline double HuynhLumpedGPrimeAtOne(int degree) {
  double c_next = degree / (2.0 * degree + 1), c_prev = 1 - c_next;
  double prev = 0, curr = 0, next = 0;
  for (int k = 0; k <= degree; ++k) {
    if (k > 0) { prev = curr; curr = next; }
    next = (k + 1) * 1.0 /* P_k(1) */ + 1.0 * curr;
  }
  return 0.5 * (curr + c_prev * prev + c_next * next);
}

inline double LegendreP(int k, double x) {   // std::legendre
  double p0 = 1, p1 = x;
  if (k == 0) return p0;
  for (int n = 1; n < k; ++n) { double p2 = ((2 * n + 1) * x * p1 - n * p0) / (n + 1); p0 = p1; p1 = p2; }
  return p1;
}
// Vincent::LocalToRightDerivative (basis/vincent.hpp:75-91) for a given c_next
inline double VincentRightDerivative(int degree, double c_next, double local) {
  double c_prev = 1 - c_next, prev = 0, curr = 0, next = 0;
  for (int k = 0; k <= degree; ++k) {
    if (k > 0) { prev = curr; curr = next; }
    next = (k + 1) * LegendreP(k, local) + local * curr;
  }
  return 0.5 * (curr + c_prev * prev + c_next * next);
}

// Square symmetry o: my face corner m coincides with the neighbour's face corner pi(m);
// o = pi(0) + 4*reversed, reversed <=> pi(1) == pi(0) - 1 (mod 4).
inline int CornerImage(int o, int m) {
  int p0 = o & 3;
  return (o & 4) ? (p0 + 4 - m) % 4 : (p0 + m) % 4;
}

inline int NearestIndex(int q, const double *x, double v) {
  int best = 0;
  for (int i = 1; i < q; ++i) if (std::fabs(x[i] - v) < std::fabs(x[best] - v)) best = i;
  return best;
}

// node index on the cell of point (ia, ib) of face f's parametrisation
inline int FaceNode(int q, const double *x, int f, int ia, int ib) {
  double a = x[ia], b = x[ib];
  double w[4] = {(1 - a) * (1 - b) / 4, (1 + a) * (1 - b) / 4, (1 + a) * (1 + b) / 4, (1 - a) * (1 + b) / 4};
  double loc[3] = {0, 0, 0};
  for (int m = 0; m < 4; ++m)
    for (int d = 0; d < 3; ++d) loc[d] += w[m] * kVertex[kFaceNodes[f][m]][d];
  int i = NearestIndex(q, x, loc[0]), j = NearestIndex(q, x, loc[1]), k = NearestIndex(q, x, loc[2]);
  return (i * q + j) * q + k;
}

inline Tables BuildTables(int degree, int scheme, int n_warps, int sub, double c_next = 0.0) {
  Tables t{};
  const int q = degree + 1, n = q * q * q, fq = q * q;
  double x[kMaxQ], w[kMaxQ];
  LobattoRule(q, x, w);
  for (int i = 0; i < q; ++i) { t.xi[i] = x[i]; t.w[i] = w[i]; }
  LegendreRule(q, t.gx, t.gw);
  for (int i = 0; i < q; ++i)
    for (int m = 0; m < q; ++m) {
      t.d1[i][m] = LagrangeDerivative(q, x, m, i);
      t.d2[i][m] = LagrangeSecondDerivative(q, x, m, i);
    }
  const double g_prime = HuynhLumpedGPrimeAtOne(degree);
  for (int i = 0; i < q; ++i) {
    for (int m = 0; m < q; ++m) {
      t.a[i][m] = (scheme == 0) ? -LagrangeDerivative(q, x, m, i)
                                : w[m] * LagrangeDerivative(q, x, i, m) / w[i];
    }
  }
  t.coef_f = (scheme == 0) ? g_prime : 1.0 / w[q - 1];
  t.coef_F = (scheme == 0) ? g_prime : 0.0;
  t.general = scheme == 2;
  if (scheme == 2) {   // fr::General: g'_left(xi) = -g'_right(-xi) (basis/vincent.hpp:99-101)
    for (int i = 0; i < q; ++i) {
      t.g[1][i] = VincentRightDerivative(degree, c_next, x[i]);
      t.g[0][i] = -VincentRightDerivative(degree, c_next, -x[i]);
      for (int m = 0; m < q; ++m) t.a[i][m] = -LagrangeDerivative(q, x, m, i);
    }
  }
  // face-point index of every cell node on each of its faces
  int face_node[6][kMaxFaceQ];
  for (int f = 0; f < 6; ++f)
    for (int ia = 0; ia < q; ++ia)
      for (int ib = 0; ib < q; ++ib) {
        face_node[f][ia * q + ib] = FaceNode(q, x, f, ia, ib);
        t.face_node[f][ia * q + ib] = (unsigned char)face_node[f][ia * q + ib];
      }
  for (int c = 0; c < n; ++c) {
    int ijk[3] = {c / fq, (c / q) % q, c % q};
    for (int d = 0; d < 3; ++d) {
      t.node_q[c][d] = -1;
      if (ijk[d] != 0 && ijk[d] != q - 1) continue;
      bool plus = ijk[d] != 0;
      int f = (d == 0) ? (plus ? 2 : 4) : (d == 1) ? (plus ? 3 : 1) : (plus ? 5 : 0);
      for (int p = 0; p < fq; ++p) if (face_node[f][p] == c) t.node_q[c][d] = (signed char)p;
    }
  }
  // neighbour node for (sharer face, orientation, my face point)
  for (int f2 = 0; f2 < 6; ++f2) {
    for (int o = 0; o < 8; ++o) {
      for (int ia = 0; ia < q; ++ia) {
        for (int ib = 0; ib < q; ++ib) {
          double a = x[ia], b = x[ib];
          double wm[4] = {(1 - a) * (1 - b) / 4, (1 + a) * (1 - b) / 4, (1 + a) * (1 + b) / 4,
                          (1 - a) * (1 + b) / 4};
          double a2 = 0, b2 = 0;
          for (int m = 0; m < 4; ++m) {
            a2 += wm[m] * kCorner[CornerImage(o, m)][0];
            b2 += wm[m] * kCorner[CornerImage(o, m)][1];
          }
          t.nbr_node[f2][o][ia * q + ib] =
              (unsigned char)face_node[f2][NearestIndex(q, x, a2) * q + NearestIndex(q, x, b2)];
        }
      }
    }
  }
  // deal nodes to (warp, slot): longest-processing-time first on the number of face points
  const int slots = (n + n_warps * sub - 1) / (n_warps * sub);
  for (auto &v : t.slot_node) v = -1;
  std::vector<std::pair<int, int>> weight;  // (-faces, node)
  for (int c = 0; c < n; ++c) {
    int nf = 0;
    for (int d = 0; d < 3; ++d) nf += t.node_q[c][d] >= 0;
    weight.emplace_back(-nf, c);
  }
  std::sort(weight.begin(), weight.end());
  std::vector<int> load(n_warps, 0), used(n_warps, 0);
  for (auto [neg, c] : weight) {
    int best = -1;
    for (int wp = 0; wp < n_warps; ++wp) {
      if (used[wp] >= slots * sub) continue;
      if (best < 0 || load[wp] < load[best]) best = wp;
    }
    // position inside the warp: slot-major, sub-minor
    t.slot_node[best * slots * sub + used[best]] = c;
    used[best]++;
    load[best] += -neg;
  }
  return t;
}

// det J at local coordinates (x, e, z) from one cell's 21 coefficients (host copy of ComputeMetric's determinant)
struct HostMetric { double det; };
inline HostMetric ComputeMetricHost(const double *g, double x, double e, double z) {
  double r[3][3];
  for (int d = 0; d < 3; ++d) {
    r[0][d] = g[0 + d] + g[9 + d] * e + g[15 + d] * z + g[18 + d] * e * z;
    r[1][d] = g[3 + d] + g[9 + d] * x + g[12 + d] * z + g[18 + d] * x * z;
    r[2][d] = g[6 + d] + g[12 + d] * e + g[15 + d] * x + g[18 + d] * x * e;
  }
  return HostMetric{r[0][0] * (r[1][1] * r[2][2] - r[1][2] * r[2][1]) - r[0][1] * (r[1][0] * r[2][2] - r[1][2] * r[2][0])
                    + r[0][2] * (r[1][0] * r[2][1] - r[1][1] * r[2][0])};
}

// ------------------------------------------------------------------------------------------------
// Connectivity of one part.
// ------------------------------------------------------------------------------------------------
struct FaceRec {
  int64_t holder, sharer;       // cell ids; sharer < 0 on a boundary
  int holder_face, sharer_face; // local face ids
  int orientation;              // holder corner m <-> sharer corner CornerImage(o, m)
  int tag;                      // boundary section
  bool periodic = false;        // the two sides were identified through a periodic pair
};

struct HostMesh {
  int64_t n_local = 0, n_ghost = 0;
  std::vector<FaceRec> local_faces, ghost_faces, bound_faces;
  std::vector<int> nbr;              // [cell][6]
  std::vector<unsigned char> code;   // [cell][6]
  std::vector<double> coef;          // [cell][21]
  std::vector<double> c0;            // [cell][3] constant term of the map = x(0,0,0)
  std::vector<double> h2s;           // [cell][6][3] Face::HolderToSharer of the face behind local face f
  std::vector<double> frames;        // [cell][6][9] normal frame (normal, tangent, bitangent) of that face at its
                                     // quadrature point 0, as rotated::Euler::Rotate receives it (fem.hpp:103-106)
  std::vector<double> length, volume;  // local cells
  bool all_affine = false;             // every local cell is a parallelepiped (c4..c7 below 64 ulp of c1..c3): J constant per cell
  // optional: the cells' ids in the whole mesh.  The reference decides holder / sharer of a face between two
  // local cells by those (holder = smaller id, mesh/shuffler.hpp:417-420), not by the part-local numbering.
  std::vector<int64_t> global_ids;
  std::string error;

  using Key = std::array<int64_t, 4>;
  static Key Sorted(Key k) { std::sort(k.begin(), k.end()); return k; }

  // Flux-point coordinates [face][Q*Q][3] of the boundary faces carrying `tag`: face point p of
  // local face f is the cell node n with node_q[n][d] == p, and its position is the trilinear map
  // at the node's Lobatto coordinates (= face.integrator().GetGlobal(p) of the reference up to
  // rounding; fr/lobatto.hpp:350).
  void BoundaryPoints(Tables const &t, int degree, int tag, double *xyz) const {
    const int q = degree + 1, fq = q * q, n_nodes = q * fq;
    for (auto const &fr : bound_faces) {
      if (fr.tag != tag) continue;
      const int f = fr.holder_face;
      const int d = (f == 2 || f == 4) ? 0 : (f == 1 || f == 3) ? 1 : 2;
      const double *g = &coef[fr.holder * kGeomRows];
      const double *o = &c0[fr.holder * 3];
      for (int n = 0; n < n_nodes; ++n) {
        int ijk[3] = {n / fq, (n / q) % q, n % q};
        const bool plus = ijk[d] != 0;
        const int fn = (d == 0) ? (plus ? 2 : 4) : (d == 1) ? (plus ? 3 : 1) : (plus ? 5 : 0);
        int p = t.node_q[n][d];
        if (p < 0 || fn != f) continue;
        double x = t.xi[ijk[0]], e = t.xi[ijk[1]], z = t.xi[ijk[2]];
        for (int k = 0; k < 3; ++k) {
          xyz[p * 3 + k] = o[k] + g[0 + k] * x + g[3 + k] * e + g[6 + k] * z + g[9 + k] * x * e
              + g[12 + k] * e * z + g[15 + k] * z * x + g[18 + k] * x * e * z;
        }
      }
      xyz += fq * 3;
    }
  }

  // Per-cell orthonormal basis of the modal extension (basis::OrthoNormal, basis/linear.hpp:87-141):
  // Gram-Schmidt on the Taylor monomials about the cell centre with the cell's own quadrature
  // (integrator/function.hpp:138-172).  Row layout per cell: ModalShape<P> in kernels.cuh.
  std::vector<double> BuildModalBases(int degree, bool legendre = false) const {
    const int q = degree + 1, m = degree == 2 ? 10 : 20, tri = m * (m + 1) / 2, rows = 3 + tri + 3;
    const int64_t nc = n_local + n_ghost;
    double x[kMaxQ], w[kMaxQ];
    if (legendre) LegendreRule(q, x, w);   // the stand-alone Projection of dg::General integrates with Gauss-Legendre
    else LobattoRule(q, x, w);
    std::vector<double> bas((size_t)nc * rows, 0.0);
    std::vector<double> A(m * m), S(m * m);
    for (int64_t c = 0; c < nc; ++c) {
      const double *g = &coef[c * kGeomRows];
      std::fill(A.begin(), A.end(), 0.0);
      for (int i = 0; i < q; ++i) for (int j = 0; j < q; ++j) for (int k = 0; k < q; ++k) {
        const double xi = x[i], et = x[j], ze = x[k];
        double d[3], r[3][3];
        for (int a = 0; a < 3; ++a) {
          d[a] = g[0 + a] * xi + g[3 + a] * et + g[6 + a] * ze + g[9 + a] * (xi * et) + g[12 + a] * (et * ze)
               + g[15 + a] * (ze * xi) + g[18 + a] * (xi * et * ze);
          r[0][a] = g[0 + a] + g[9 + a] * et + g[15 + a] * ze + g[18 + a] * et * ze;
          r[1][a] = g[3 + a] + g[9 + a] * xi + g[12 + a] * ze + g[18 + a] * xi * ze;
          r[2][a] = g[6 + a] + g[12 + a] * et + g[15 + a] * xi + g[18 + a] * xi * et;
        }
        const double det = r[0][0] * (r[1][1] * r[2][2] - r[1][2] * r[2][1])
                         - r[0][1] * (r[1][0] * r[2][2] - r[1][2] * r[2][0])
                         + r[0][2] * (r[1][0] * r[2][1] - r[1][1] * r[2][0]);
        const double wq = w[i] * w[j] * w[k] * det;
        double t[20];
        t[0] = 1; t[1] = d[0]; t[2] = d[1]; t[3] = d[2];
        t[4] = d[0] * d[0]; t[5] = d[0] * d[1]; t[6] = d[0] * d[2]; t[7] = d[1] * d[1]; t[8] = d[1] * d[2]; t[9] = d[2] * d[2];
        if (m == 20) {
          t[10] = d[0] * t[4]; t[11] = d[0] * t[5]; t[12] = d[0] * t[6]; t[13] = d[0] * t[7]; t[14] = d[0] * t[8];
          t[15] = d[0] * t[9]; t[16] = d[1] * t[7]; t[17] = d[1] * t[8]; t[18] = d[1] * t[9]; t[19] = d[2] * t[9];
        }
        for (int a = 0; a < m; ++a) for (int b = 0; b < m; ++b) A[a * m + b] += t[a] * t[b] * wq;
      }
      std::fill(S.begin(), S.end(), 0.0);
      for (int i = 0; i < m; ++i) S[i * m + i] = 1;
      S[0] = 1 / std::sqrt(A[0]);
      for (int i = 1; i < m; ++i) {
        for (int j = 0; j < i; ++j) {
          double temp = 0;
          for (int k = 0; k <= j; ++k) temp += S[j * m + k] * A[k * m + i];
          for (int l = 0; l <= j; ++l) S[i * m + l] -= temp * S[j * m + l];
        }
        double norm_sq = 0;
        for (int j = 0; j <= i; ++j) {
          double sum = 0, sij = S[i * m + j];
          for (int k = 0; k < j; ++k) sum += 2 * S[i * m + k] * A[k * m + j];
          norm_sq += sij * (sij * A[j * m + j] + sum);
        }
        const double nrm = std::sqrt(norm_sq);
        for (int j = 0; j < m; ++j) S[i * m + j] /= nrm;
      }
      double *out = &bas[(size_t)c * rows];
      for (int a = 0; a < 3; ++a) out[a] = c0[c * 3 + a];
      int r = 3;
      for (int i = 0; i < m; ++i) for (int j = 0; j <= i; ++j) out[r++] = S[i * m + j];
      const double volume = 1 / (S[0] * S[0]);                      // OrthoNormal::Measure, basis/linear.hpp:127-130
      for (int k = 1; k <= 3; ++k) {                                // Smoothness::GetWeight, limiter/weno.hpp:50-53
        const double kf = std::tgamma(k + 1.0);
        out[r++] = std::pow(volume, k * 2. / 3 - 1) / (kf * kf);
      }
    }
    return bas;
  }

  // The same for dg::General: the Gauss-Legendre points of the boundary faces carrying `tag`, point (ia, ib) of a
  // face at index ia*Q + ib with ia, ib running along the face's two local axes in increasing axis order (the order
  // DgmStageCell reads its boundary values in; the reference evaluates the boundary function at
  // face.integrator().GetGlobal(q), dg/general.hpp:131-133).
  void BoundaryGaussPoints(Tables const &t, int degree, int tag, double *xyz) const {
    const int q = degree + 1;
    for (auto const &fr : bound_faces) {
      if (fr.tag != tag) continue;
      const int f = fr.holder_face;
      const int d = (f == 2 || f == 4) ? 0 : (f == 1 || f == 3) ? 1 : 2;
      const double sgn = (f == 2 || f == 3 || f == 5) ? 1.0 : -1.0;
      const double *g = &coef[fr.holder * kGeomRows];
      const double *o = &c0[fr.holder * 3];
      for (int ia = 0; ia < q; ++ia) for (int ib = 0; ib < q; ++ib) {
        const double a = t.gx[ia], b = t.gx[ib];
        const double x = d == 0 ? sgn : a, e = d == 1 ? sgn : (d == 0 ? a : b), z = d == 2 ? sgn : b;
        for (int k = 0; k < 3; ++k)
          xyz[(ia * q + ib) * 3 + k] = o[k] + g[0 + k] * x + g[3 + k] * e + g[6 + k] * z + g[9 + k] * x * e
              + g[12 + k] * e * z + g[15 + k] * z * x + g[18 + k] * x * e * z;
      }
      xyz += q * q * 3;
    }
  }
  // Gauss-Legendre points of the local cells, [cell][Q^3][3] in the integrator's order (integrator/hexahedron.hpp:66-80)
  void QuadraturePoints(Tables const &t, int degree, double *xyz) const {
    const int q = degree + 1;
    for (int64_t c = 0; c < n_local; ++c) {
      const double *g = &coef[c * kGeomRows];
      const double *o = &c0[c * 3];
      for (int i = 0; i < q; ++i) for (int j = 0; j < q; ++j) for (int k2 = 0; k2 < q; ++k2) {
        const double x = t.gx[i], e = t.gx[j], z = t.gx[k2];
        for (int k = 0; k < 3; ++k)
          *xyz++ = o[k] + g[0 + k] * x + g[3 + k] * e + g[6 + k] * z + g[9 + k] * x * e + g[12 + k] * e * z
              + g[15 + k] * z * x + g[18 + k] * x * e * z;
      }
    }
  }

  bool Build(int degree, int64_t n_nodes, const double *xyz, int64_t nl, int64_t ng,
      const int64_t *cn, int64_t n_bfaces, const int64_t *bf, const int32_t *btag,
      int64_t n_pairs, const int64_t *pairs) {
    n_local = nl; n_ghost = ng;
    const int64_t nc = nl + ng;
    // every node id must name a row of xyz: a malformed mesh is an error, not an out-of-bounds read
    auto bad = [n_nodes](int64_t id) { return id < 0 || id >= n_nodes; };
    for (int64_t k = 0; k < nc * 8; ++k) if (bad(cn[k])) { error = "cell " + std::to_string(k / 8) + " names node " + std::to_string(cn[k]) + " of " + std::to_string(n_nodes); return false; }
    for (int64_t k = 0; k < n_bfaces * 4; ++k) if (bad(bf[k])) { error = "boundary face " + std::to_string(k / 4) + " names node " + std::to_string(bf[k]) + " of " + std::to_string(n_nodes); return false; }
    for (int64_t k = 0; k < n_pairs * 8; ++k) if (bad(pairs[k])) { error = "periodic pair " + std::to_string(k / 8) + " names node " + std::to_string(pairs[k]) + " of " + std::to_string(n_nodes); return false; }
    // periodic identification: node of the second quad -> node of the first quad, per pair, applied
    // only to the face that carries it
    std::map<Key, std::array<int64_t, 8>> periodic;  // sorted key of B -> (B nodes, A nodes)
    for (int64_t p = 0; p < n_pairs; ++p) {
      Key a, b;
      for (int m = 0; m < 4; ++m) { a[m] = pairs[p * 8 + m]; b[m] = pairs[p * 8 + 4 + m]; }
      std::array<int64_t, 8> rec;
      for (int m = 0; m < 4; ++m) { rec[m] = b[m]; rec[4 + m] = a[m]; }
      periodic[Sorted(b)] = rec;
    }
    struct Side { int64_t cell; int face; Key nodes; bool aliased; };  // nodes in the cell's face order, aliased
    std::map<Key, std::vector<Side>> sides;
    for (int64_t c = 0; c < nc; ++c) {
      for (int f = 0; f < 6; ++f) {
        Key k;
        for (int m = 0; m < 4; ++m) k[m] = cn[c * 8 + kFaceNodes[f][m]];
        auto it = periodic.find(Sorted(k));
        const bool aliased = it != periodic.end();
        if (aliased) {
          for (int m = 0; m < 4; ++m)
            for (int j = 0; j < 4; ++j)
              if (k[m] == it->second[j]) { k[m] = it->second[4 + j]; break; }
        }
        sides[Sorted(k)].push_back(Side{c, f, k, aliased});
      }
    }
    nbr.assign(nc * 6, INT32_MIN);
    code.assign(nc * 6, 0);
    // interior faces in the reference's order: holder = smaller id, ascending (holder, sharer)
    std::vector<FaceRec> all;
    for (auto &[key, list] : sides) {
      if (list.size() == 2) {
        Side h = list[0], s = list[1];
        const bool both_local = h.cell < nl && s.cell < nl && (int64_t)global_ids.size() == nc;
        if (both_local ? global_ids[s.cell] < global_ids[h.cell] : s.cell < h.cell) std::swap(h, s);
        if (h.cell >= nl) continue;  // ghost-ghost
        int pi[4];
        for (int m = 0; m < 4; ++m) {
          pi[m] = -1;
          for (int j = 0; j < 4; ++j) if (s.nodes[j] == h.nodes[m]) pi[m] = j;
          if (pi[m] < 0) { error = "face vertex mismatch"; return false; }
        }
        int o = pi[0] + ((pi[1] == (pi[0] + 3) % 4) ? 4 : 0);
        for (int m = 0; m < 4; ++m)
          if (CornerImage(o, m) != pi[m]) { error = "face is not a square symmetry"; return false; }
        all.push_back(FaceRec{h.cell, s.cell, h.face, s.face, o, -1, h.aliased || s.aliased});
      } else if (list.size() > 2) {
        error = "more than two cells share a face";
        return false;
      }
    }
    const bool have_ids = (int64_t)global_ids.size() == nc;
    std::sort(all.begin(), all.end(), [&](FaceRec const &a, FaceRec const &b) {
      const bool ga = a.sharer >= nl, gb = b.sharer >= nl;   // ghost faces keep the part-local order
      if (ga != gb) return gb;
      if (have_ids && !ga) {
        const int64_t ah = global_ids[a.holder], bh = global_ids[b.holder];
        return ah != bh ? ah < bh : global_ids[a.sharer] < global_ids[b.sharer];
      }
      return a.holder != b.holder ? a.holder < b.holder : a.sharer < b.sharer;
    });
    for (auto &fr : all) {
      (fr.sharer < nl ? local_faces : ghost_faces).push_back(fr);
      // the sharer sees the inverse correspondence
      int inv = -1;
      for (int o2 = 0; o2 < 8 && inv < 0; ++o2) {
        bool ok = true;
        for (int m = 0; m < 4; ++m) ok = ok && CornerImage(o2, CornerImage(fr.orientation, m)) == m;
        if (ok) inv = o2;
      }
      nbr[fr.holder * 6 + fr.holder_face] = (int)fr.sharer;
      code[fr.holder * 6 + fr.holder_face] = (unsigned char)(fr.sharer_face | fr.orientation << 3 | 64);
      nbr[fr.sharer * 6 + fr.sharer_face] = (int)fr.holder;
      code[fr.sharer * 6 + fr.sharer_face] = (unsigned char)(fr.holder_face | inv << 3);
    }
    // boundary faces: grouped by tag ascending, input order inside a tag (part.hpp:1476-1541)
    std::vector<int64_t> order(n_bfaces);
    for (int64_t b = 0; b < n_bfaces; ++b) order[b] = b;
    std::stable_sort(order.begin(), order.end(), [&](int64_t a, int64_t b) { return btag[a] < btag[b]; });
    for (int64_t b : order) {
      Key k;
      for (int m = 0; m < 4; ++m) k[m] = bf[b * 4 + m];
      auto it = sides.find(Sorted(k));
      if (it == sides.end() || it->second.size() != 1) { error = "boundary quad is not a boundary face"; return false; }
      Side h = it->second[0];
      if (h.cell >= nl) continue;  // belongs to a ghost cell: another part's face
      nbr[h.cell * 6 + h.face] = -1 - (int)bound_faces.size();
      code[h.cell * 6 + h.face] = 64;
      bound_faces.push_back(FaceRec{h.cell, -1, h.face, -1, 0, btag[b]});
    }
    for (int64_t c = 0; c < nl; ++c)
      for (int f = 0; f < 6; ++f)
        if (nbr[c * 6 + f] == INT32_MIN) { error = "a local cell face has neither neighbour nor boundary"; return false; }
    for (int64_t c = nl; c < nc; ++c)
      for (int f = 0; f < 6; ++f) if (nbr[c * 6 + f] == INT32_MIN) nbr[c * 6 + f] = 0;
    // geometry: trilinear coefficients, volume (integrator/element.hpp:101-114) and length (part.hpp:230)
    const int q = degree + 1;
    double x[kMaxQ], w[kMaxQ];
    LobattoRule(q, x, w);
    coef.assign(nc * kGeomRows, 0.0);
    c0.assign(nc * 3, 0.0);
    all_affine = true;
    length.resize(nl);
    volume.resize(nl);
    static const int kTerm[7][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {1, 1, 0}, {0, 1, 1}, {1, 0, 1}, {1, 1, 1}};
    for (int64_t c = 0; c < nc; ++c) {
      for (int d = 0; d < 3; ++d) {
        double sum = 0;
        for (int v = 0; v < 8; ++v) sum += xyz[cn[c * 8 + v] * 3 + d];
        c0[c * 3 + d] = sum / 8;
      }
      for (int t = 0; t < 7; ++t) {
        for (int d = 0; d < 3; ++d) {
          double sum = 0;
          for (int v = 0; v < 8; ++v) {
            double s = 1;
            for (int e = 0; e < 3; ++e) if (kTerm[t][e]) s *= kVertex[v][e];
            sum += s * xyz[cn[c * 8 + v] * 3 + d];
          }
          coef[c * kGeomRows + t * 3 + d] = sum / 8;
        }
      }
      if (c >= nl) continue;
      {   // a parallelepiped up to the rounding of its vertex coordinates: the bilinear and trilinear terms vanish
        double lin = 0, rest = 0;
        for (int k = 0; k < 9; ++k) lin = std::max(lin, std::fabs(coef[c * kGeomRows + k]));
        for (int k = 9; k < kGeomRows; ++k) rest = std::max(rest, std::fabs(coef[c * kGeomRows + k]));
        if (rest > 64 * 2.220446049250313e-16 * lin) all_affine = false;
      }
      double vol = 0;
      for (int i = 0; i < q; ++i) for (int j = 0; j < q; ++j) for (int k = 0; k < q; ++k) {
        double r[3][3];
        const double *g = &coef[c * kGeomRows];
        for (int d = 0; d < 3; ++d) {
          r[0][d] = g[0 + d] + g[9 + d] * x[j] + g[15 + d] * x[k] + g[18 + d] * x[j] * x[k];
          r[1][d] = g[3 + d] + g[9 + d] * x[i] + g[12 + d] * x[k] + g[18 + d] * x[i] * x[k];
          r[2][d] = g[6 + d] + g[12 + d] * x[j] + g[15 + d] * x[i] + g[18 + d] * x[i] * x[j];
        }
        double det = r[0][0] * (r[1][1] * r[2][2] - r[1][2] * r[2][1])
                   - r[0][1] * (r[1][0] * r[2][2] - r[1][2] * r[2][0])
                   + r[0][2] * (r[1][0] * r[2][1] - r[1][1] * r[2][0]);
        if (!(det > 0)) {   // an inverted or degenerate cell; the reference divides by this determinant (fr/general.hpp:256-266)
          error = "cell " + std::to_string(c) + " has a non-positive Jacobian determinant at a solution point "
                  "(node order must be CGNS HEXA_8, right-handed)";
          return false;
        }
        vol += w[i] * w[j] * w[k] * det;
      }
      volume[c] = vol;
      length[c] = std::cbrt(vol) * 0.5;
    }
    // Face::HolderToSharer (mesh/part.hpp:118-133): between the cell centres x(0,0,0), or twice
    // centre -> face centre on a boundary; across a periodic pair each side contributes its own
    // centre -> face-centre vector.
    h2s.assign(nc * 18, 0.0);
    auto face_center = [&](int64_t c, int f, double *out) {
      for (int d = 0; d < 3; ++d) {
        double sum = 0;
        for (int m = 0; m < 4; ++m) sum += xyz[cn[c * 8 + kFaceNodes[f][m]] * 3 + d];
        out[d] = sum / 4;
      }
    };
    auto set_faces = [&](std::vector<FaceRec> const &faces) {
      for (auto const &fr : faces) {
        double v[3];
        if (fr.sharer < 0) {
          face_center(fr.holder, fr.holder_face, v);
          for (int d = 0; d < 3; ++d) v[d] = (v[d] - c0[fr.holder * 3 + d]) * 2;
        } else if (fr.periodic) {
          double a[3], b[3];
          face_center(fr.holder, fr.holder_face, a);
          face_center(fr.sharer, fr.sharer_face, b);
          for (int d = 0; d < 3; ++d) v[d] = (a[d] - c0[fr.holder * 3 + d]) + (c0[fr.sharer * 3 + d] - b[d]);
        } else {
          for (int d = 0; d < 3; ++d) v[d] = c0[fr.sharer * 3 + d] - c0[fr.holder * 3 + d];
        }
        for (int d = 0; d < 3; ++d) {
          h2s[(fr.holder * 6 + fr.holder_face) * 3 + d] = v[d];
          if (fr.sharer >= 0) h2s[(fr.sharer * 6 + fr.sharer_face) * 3 + d] = v[d];
        }
      }
    };
    set_faces(local_faces); set_faces(ghost_faces); set_faces(bound_faces);
    // Face frames at face point 0 = the holder's first face corner (Lobatto points include the ends):
    // tangent along corner 0 -> 1, normal = d/da x d/db (coordinate/face.hpp:54-62, quadrangle.hpp:77-105)
    frames.assign(nc * 54, 0.0);
    auto set_frames = [&](std::vector<FaceRec> const &faces) {
      for (auto const &fr : faces) {
        const double *P[4];
        for (int m = 0; m < 4; ++m) P[m] = &xyz[cn[fr.holder * 8 + kFaceNodes[fr.holder_face][m]] * 3];
        const double a = x[0], b = x[0];
        double ja[3] = {0, 0, 0}, jb[3] = {0, 0, 0};
        for (int m = 0; m < 4; ++m) {
          const double fa = 1 + kCorner[m][0] * a, fb = 1 + kCorner[m][1] * b;
          for (int k = 0; k < 3; ++k) {
            ja[k] += kCorner[m][0] * (fb / 4) * P[m][k];
            jb[k] += kCorner[m][1] * (fa / 4) * P[m][k];
          }
        }
        double f9[9];
        double nn[3] = {ja[1] * jb[2] - ja[2] * jb[1], ja[2] * jb[0] - ja[0] * jb[2], ja[0] * jb[1] - ja[1] * jb[0]};
        const double ln = std::sqrt(nn[0] * nn[0] + nn[1] * nn[1] + nn[2] * nn[2]);
        const double la = std::sqrt(ja[0] * ja[0] + ja[1] * ja[1] + ja[2] * ja[2]);
        for (int k = 0; k < 3; ++k) { f9[k] = nn[k] / ln; f9[3 + k] = ja[k] / la; }
        f9[6] = f9[1] * f9[5] - f9[2] * f9[4];
        f9[7] = f9[2] * f9[3] - f9[0] * f9[5];
        f9[8] = f9[0] * f9[4] - f9[1] * f9[3];
        for (int k = 0; k < 9; ++k) {
          frames[(fr.holder * 6 + fr.holder_face) * 9 + k] = f9[k];
          if (fr.sharer >= 0) frames[(fr.sharer * 6 + fr.sharer_face) * 9 + k] = f9[k];
        }
      }
    };
    set_frames(local_faces); set_frames(ghost_faces); set_frames(bound_faces);
    return true;
  }
};

}  // namespace mcfd
#endif  // MINICFD_B200_TABLES_HPP_
