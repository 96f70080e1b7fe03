// kernels.cuh — device data layout and kernels of libminicfd_b200 (sm_100a).
//
// Layout ("tiles"): cells are grouped 32 at a time; within a tile every scalar is stored as a row of
// 32 doubles, one per cell, so a warp whose lanes are the 32 cells of a tile reads and writes whole
// 256-byte rows:
//     U   [tile][node*5 + component][lane]     solution coefficients (kFields rows per tile)
//     G   [tile][21][lane]                     trilinear map coefficients c1..c7 (3 doubles each)
//     NBR [tile][6][lane]  int32               neighbour cell id across local face f, or -1-b for
//                                              boundary face b
//     CODE[tile][6][lane]  uint8               sharer face | orientation<<3 | i_am_holder<<6
// Ghost cells are ordinary cells appended after the local ones.
#ifndef MINICFD_B200_KERNELS_CUH_
#define MINICFD_B200_KERNELS_CUH_

#include <cstdint>

#include "physics.cuh"

namespace mcfd {

constexpr int kTile = 32;       // cells per tile
constexpr int kGeomRows = 21;   // c1..c7, xyz each
constexpr int kMaxQ = 4;
constexpr int kMaxN = kMaxQ * kMaxQ * kMaxQ;
constexpr int kMaxFaceQ = kMaxQ * kMaxQ;
constexpr int kMaxSlots = 8;    // node slots per warp
constexpr int kMaxWarps = 16;
// launch shape per degree: cells per CTA and warps per CTA (see StageKernel)
// p = 2, inviscid: 16-cell CTAs of 7 warps, four per SM (55 KB of shared memory each, 72 registers): 14 node / face-point
// items per pass leave 1 of 28 node slots and 2 of 56 point slots idle; 6 % faster than 32-cell CTAs of 10 warps, two
// per SM (tools/run_stage_variants.sh).  The viscous kernels keep 32-cell CTAs.
#ifndef MCFD_P2_WARPS
#define MCFD_P2_WARPS 7
#endif
#ifndef MCFD_P2_CELLS
#define MCFD_P2_CELLS 16
#endif
#ifndef MCFD_P2_MINB
#define MCFD_P2_MINB 4
#endif
// the viscous instantiation carries more live state per face point: its own launch shape
#ifndef MCFD_P2_VWARPS
#define MCFD_P2_VWARPS 12
#endif
#ifndef MCFD_P2_VMINB
#define MCFD_P2_VMINB 1
#endif
#ifndef MCFD_TRACE_MINB
#define MCFD_TRACE_MINB 2
#endif
// resident CTAs per SM the limiter kernels (160 threads, 43-101 KB of shared memory) are compiled for
#ifndef MCFD_WENO_MINB
#define MCFD_WENO_MINB 4   // p = 2: 102 registers, 20 warps per SM (1 CTA per SM ran the limiter 1.7x slower)
#endif
// resident one-warp CTAs per SM the modal DG stage kernel is compiled for
#ifndef MCFD_WENO_P3_MINB
#define MCFD_WENO_P3_MINB 3
#endif
#ifndef MCFD_WENO_P3_BAS_SHARED
#define MCFD_WENO_P3_BAS_SHARED 0
#endif
#ifndef MCFD_DGM_MINB
#define MCFD_DGM_MINB 1
#endif
// p = 3, inviscid: 8-cell CTAs of 8 warps, three per SM (61 KB of shared memory each, 85 registers); measured 16 %
// faster than one 16-cell CTA of 16 warps per SM (tools/run_p3_variants.sh).  The viscous kernels keep the latter.
#ifndef MCFD_P3_CELLS
#define MCFD_P3_CELLS 8
#endif
#ifndef MCFD_P3_WARPS
#define MCFD_P3_WARPS 8
#endif
#ifndef MCFD_P3_MINB
#define MCFD_P3_MINB 3
#endif
constexpr int kP2Cells = MCFD_P2_CELLS, kP2Warps = MCFD_P2_WARPS, kP3Cells = MCFD_P3_CELLS, kP3Warps = MCFD_P3_WARPS;
#ifndef MCFD_P2_VCELLS
#define MCFD_P2_VCELLS 32
#endif
constexpr int kP3ViscousCells = 16, kP3ViscousWarps = 16, kP2ViscousCells = MCFD_P2_VCELLS;
constexpr int kP2ViscousWarps = MCFD_P2_VWARPS;

// Per-degree tables, filled by the host (tables.cpp) and copied to constant memory.
struct Tables {
  double xi[kMaxQ];                    // Lobatto points (integrator/lobatto.hpp:38-57)
  double w[kMaxQ];                     // Lobatto weights
  double a[kMaxQ][kMaxQ];              // R[c] += a[i][m] * Fhat_d(node m on the d-line)[c]
  double coef_f;                       // R -= coef_f * |n| * f_upwind
  double coef_F;                       // R += coef_F * sgn * Fhat_d   (0 for DG)
  // general Vincent correction (fr/general.hpp:85-211): g'(xi_i) of the correction function of the
  // - and + end of a node line; the lumped Lobatto scheme only has g[0][0] and g[1][Q-1]
  int general;
  double g[2][kMaxQ];
  double gx[kMaxQ], gw[kMaxQ];         // Gauss-Legendre rule of dg::General (integrator/legendre.hpp:50-73)
  double d1[kMaxQ][kMaxQ];             // l_m'(x_i)  (viscous path: local gradients)
  double d2[kMaxQ][kMaxQ];             // l_m''(x_i) (viscous path: local Hessians)
  int slot_node[kMaxWarps * kMaxSlots];  // node handled by (warp, slot), -1 = idle
  // face-point index q of node n on its face in direction d (or -1 when not on a face)
  signed char node_q[kMaxN][3];
  // neighbour node for (sharer face f', orientation o, my face-point q)
  unsigned char nbr_node[6][8][kMaxFaceQ];
  // cell node of face point q of local face f
  unsigned char face_node[6][kMaxFaceQ];
};

struct StageArgs {
  const double *u_in;     // stage state (tiles, incl. ghosts)
  const double *u0;       // state at the beginning of the step (tiles)
  double *u_out;          // next stage state, or the residual when mode == kModeResidual
  const double *geom;
  const int *nbr;
  const unsigned char *code;
  const unsigned char *bkind;   // [n_bfaces] boundary kind
  const double *bvals;          // [n_bfaces][Q*Q][5] pre-sampled boundary function
  int n_local;                  // cells [0, n_local) are advanced
  int block_begin;              // first CTA-sized cell block of this launch
  int mode;
  double dt;
  const double *dt_dev;         // when not null the time step is read from device memory instead
  // viscous path only (ConvectionDiffusion<Euler, Direct<NavierStokes>>): traces written by TraceKernel
  const double *grad;           // [tile][node*15 + x*5 + c][lane]  global gradient at every node
  const double *hn;             // [tile][point*15 + x*5 + c][lane] Hessian . own outward unit normal
  const double *h2s;            // [tile][face*3 + x][lane]         Face::HolderToSharer of my face f
  const double *cell_nu;        // [tile][comp][lane] kDiffIsotropic: viscosity of every component, per cell
  Viscosity vis;
  const unsigned char *line_bytes;   // LineStageKernel only: LineTableBytes() in device memory
  // LineStageKernel as ONE launch per stage on a part with neighbours (peer.cuh; all null / 0 otherwise).
  // Receiving: the blocks with inter cells, which come last (part.hpp:746-767 numbers those cells last), first wait until
  // this part's epoch words carry the count of the delivery of the stage input, and read the ghost states from the
  // receive stage [row][ghost item] instead of ghost rows.  Sending: they store a second copy of their send cells'
  // new values into the send stage [row][send item], which the copy engines take to the neighbours after the launch.
  struct PeerStage {
    int on, first_block, first_cell, n_neighbours, wait_add, n_send, n_recv;
    const unsigned long long *flags;    // my epoch words
    const unsigned long long *counts;   // peer.cuh: kPeerPlain, kPeerStageBase; expected = plain + stage_base + wait_add
    const double *recv;                 // receive stage of the stage INPUT
    const int *recv_index;              // [n_ghost] ghost cell n_local + i -> its item
    double *send;                       // send stage of the stage OUTPUT
    const int *tgt_offset;              // [n_local - first_cell + 1] send items of local cell first_cell + i ...
    const int *tgt_item;                // ... listed here
    int *error;
  } peer;
};

// diffusive part of the Riemann type, a template parameter of the stage kernel
enum : int {
  kDiffNone = 0,
  kDiffNavierStokes = 1,   // diffusive::Direct<diffusive::NavierStokes<Gas>>
  kDiffIsotropic = 2,      // diffusive::Direct<diffusive::Isotropic<Scalar, 5>> / EnergyBasedViscosity: nu per component and cell
  // EnergyBasedViscosity::BuildDampingMatrices (spatial/viscosity.hpp:299-384): the viscous operator alone, nu = 1,
  // every neighbour and every boundary side zero (AddFluxToHolder / AddFluxToSharer, fr/general.hpp:356-408)
  kDiffProbe = 3,
  kDiffNsIsotropic = 4     // EnergyBasedViscosity over the Navier-Stokes base: both of the above (viscosity.hpp:76-129)
};

enum : int {
  kModeResidual = 0,  // out = R
  kModeEuler = 1,     // out = u + R*dt                        (ode.hpp:121-133)
  kModeRk3b = 2,      // out = (u0*3 + (u + R*dt)) / 4         (rk.hpp:36-40)
  kModeRk3c = 3,      // out = ((u + R*dt)*2 + u0) / 3         (rk.hpp:42-47)
  kModeRk2b = 4,      // out = ((u + R*dt) + u0) * 0.5         (rk.hpp:24-29)
  kModeProbe = 5      // host-side only: a residual with the kDiffProbe instantiation (damping matrices)
};

MCFD_CONSTANT Tables c_tab;

// Rows r0,r1,r2 of the Jacobian J = d(x,y,z)/d(xi,eta,zeta) of the trilinear map at (x,e,z)
// (coordinate/hexahedron.hpp:431-456, element.hpp:44-47) from the coefficient rows in `g`
// (stride `gs` doubles between rows), then J* = det(J) J^-1 by cofactors
// (polynomial/hexahedron.hpp:221-231): column d of J* is the cross product of the other two rows.
struct Metric {
  double s[3][3];  // s[d][x] = J*(x, d): the (unnormalised) normal of the d-faces
  double det;
};

MCFD_HD Metric ComputeMetric(const double *g, int gs, double x, double e, double z) {
  double r[3][3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    double c1 = g[(0 + k) * gs], c2 = g[(3 + k) * gs], c3 = g[(6 + k) * gs];
    double c4 = g[(9 + k) * gs], c5 = g[(12 + k) * gs], c6 = g[(15 + k) * gs], c7 = g[(18 + k) * gs];
    r[0][k] = fma(z, fma(e, c7, c6), fma(e, c4, c1));
    r[1][k] = fma(z, fma(x, c7, c5), fma(x, c4, c2));
    r[2][k] = fma(e, fma(x, c7, c5), fma(x, c6, c3));
  }
  Metric m;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    const double *a = r[(d + 1) % 3], *b = r[(d + 2) % 3];
    m.s[d][0] = a[1] * b[2] - a[2] * b[1];
    m.s[d][1] = a[2] * b[0] - a[0] * b[2];
    m.s[d][2] = a[0] * b[1] - a[1] * b[0];
  }
  m.det = r[0][0] * m.s[0][0] + r[0][1] * m.s[0][1] + r[0][2] * m.s[0][2];
  return m;
}

template <class T>
MCFD_HD T Sel3(int d, T a0, T a1, T a2) { return d == 0 ? a0 : (d == 1 ? a1 : a2); }

// Column d of J* only (the un-normalised normal of the d-faces at this node).
MCFD_HD void ComputeNormal(const double *g, int gs, double x, double e, double z, int d, double *n) {
  double r[3][3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    double c1 = g[(0 + k) * gs], c2 = g[(3 + k) * gs], c3 = g[(6 + k) * gs];
    double c4 = g[(9 + k) * gs], c5 = g[(12 + k) * gs], c6 = g[(15 + k) * gs], c7 = g[(18 + k) * gs];
    r[0][k] = fma(z, fma(e, c7, c6), fma(e, c4, c1));
    r[1][k] = fma(z, fma(x, c7, c5), fma(x, c4, c2));
    r[2][k] = fma(e, fma(x, c7, c5), fma(x, c6, c3));
  }
  double a[3], b[3];   // rows (d+1)%3 and (d+2)%3
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    a[k] = Sel3(d, r[1][k], r[2][k], r[0][k]);
    b[k] = Sel3(d, r[2][k], r[0][k], r[1][k]);
  }
  n[0] = a[1] * b[2] - a[2] * b[1];
  n[1] = a[2] * b[0] - a[0] * b[2];
  n[2] = a[0] * b[1] - a[1] * b[0];
}

// One RK stage (or a bare residual) for the cells of a CTA-sized block.
//   CPB  cells per CTA: 32 (a whole tile: a warp handles one node or face point of 32 cells at a
//        time) or 16 (half a tile, two per warp) — bounded by the shared-memory flux cache of
//        N*15*CPB doubles.
//   NW   warps per CTA.
// Phases, separated by CTA barriers:
//   A   node-centric     Fhat = F(u) J* at every node -> shared memory sF[N][15][CPB]
//   B1  node-centric     R = (sum_lines a * Fhat  +  own-face Fhat terms) / det      (registers)
//   --  sF is dead from here on and is reused: sR[N][5][CPB] then sC[6*Q*Q][5][CPB]
//   B2  face-point-centric  one Riemann problem per face point -> lifted flux term -> sC
//   C   node-centric     R + face terms, RK update, store
// The body is written as per-thread phase functions of (thread id, block id, shared memory) so that
// tests/host_emulation can run the very same source on the host to debug index logic without a GPU;
// the __global__ kernel strings them together with __syncthreads().
template <int Q, int CPB, int NW>
struct StageGeometry {
  static constexpr int N = Q * Q * Q;
  static constexpr int F = 5 * N;
  static constexpr int FQ = Q * Q;
  static constexpr int SUB = kTile / CPB;                 // nodes / face points handled at once by a warp
  static constexpr int SLOTS = (N + NW * SUB - 1) / (NW * SUB);
  static constexpr int POINTS = 6 * FQ;                   // face points per cell
  static constexpr int ITEMS = (POINTS + NW * SUB - 1) / (NW * SUB);
  static constexpr size_t kSharedBytes =
      (size_t)(N * 15 * CPB + kGeomRows * CPB) * sizeof(double) + 6 * 8 * FQ + 6 * CPB * (sizeof(int) + 1);
  static_assert(POINTS * 5 <= N * 10, "face terms must fit behind the residual rows");
  int lane, warp, lc, sub, block, tile, tl, cell;
  bool active;
  double *sF, *sR, *sC, *sG;
  unsigned char *sNbrNode, *sCode;
  int *sNbr;
  MCFD_HD StageGeometry(const StageArgs &p, int tid, int bid, double *smem) {
    sF = smem;                                      // [N][15][CPB]
    sR = smem;                                      // [N][5][CPB]        (after B1)
    sC = smem + N * 5 * CPB;                        // [POINTS][5][CPB]   (after B1)
    sG = smem + N * 15 * CPB;                       // [21][CPB]
    sNbrNode = reinterpret_cast<unsigned char *>(sG + kGeomRows * CPB);  // [6][8][FQ]
    sNbr = reinterpret_cast<int *>(sNbrNode + 6 * 8 * FQ);               // [6][CPB] neighbour of the block's cells
    sCode = reinterpret_cast<unsigned char *>(sNbr + 6 * CPB);           // [6][CPB]
    lane = tid & 31; warp = tid >> 5;
    lc = lane % CPB;                                // cell within the block
    sub = lane / CPB;
    block = p.block_begin + bid;
    tile = block / SUB;
    tl = (block % SUB) * CPB + lc;                  // lane within the tile
    cell = tile * kTile + tl;
    active = cell < p.n_local;
  }
  MCFD_HD int Node(int s) const {
    int slot = (warp * SLOTS + s) * SUB + sub;
    return (slot < kMaxWarps * kMaxSlots) ? c_tab.slot_node[slot] : -1;
  }
  // face point handled in item k: points are dealt round-robin so that all warps get the same count
  MCFD_HD int Point(int k) const {
    int pt = (k * NW + warp) * SUB + sub;
    return pt < POINTS ? pt : -1;
  }
};

MCFD_HD int FaceOf(int d, bool plus) {  // local face id of the -/+ side in direction d
  return (d == 0) ? (plus ? 2 : 4) : (d == 1) ? (plus ? 3 : 1) : (plus ? 5 : 0);
}

// ---- viscous traces -------------------------------------------------------------------------------
// What a face point of the direct-DG scheme needs from a cell besides the value (fr/general.hpp:316-350):
// the gradient and the Hessian of the nodal polynomial in physical space
// (Hexahedron::GetGlobalValueGradientHessian with kLocal == false, polynomial/hexahedron.hpp:586-594,
// 496-524).  The Hessian only enters through  (H_sharer - H_holder) . normal  (direct.hpp:66-76), so each
// cell stores  H . n_own  per face point, n_own its own outward unit normal there; the neighbour's
// outward normal is the opposite one, hence  jump . n_holder = -(Hn_own + Hn_neighbour)  on both sides.
//
//   grad out: rows (n*15 + x*5 + c),  hn out: rows ((f*Q*Q + q)*15 + x*5 + c),  `os` doubles apart.
//
// With J = d(x,y,z)/d(xi,eta,zeta) (rows = local directions), inv = J^-1 and T_a = d(J^-T)/d(xi_a) =
// -(inv dJ_a inv)^T (hexahedron.hpp:233-236), the reference evaluates per component
//     S(a, j) = sum_b Hl(a, b) inv(j, b) + sum_b gl(b) T_a(b, j),     H = inv S.
// Contracting with n first:  w = inv^T n,  v_a = w^T dJ_a,  m_a(b) = -sum_s v_a(s) inv(s, b),
//     t(a) = sum_b Hl(a, b) w(b) + sum_b gl(b) m_a(b),     (H n)(i) = sum_a inv(i, a) t(a).
// dJ_a of the trilinear map has three distinct rows: A = d2x/dxi deta, B = d2x/dzeta dxi, C = d2x/deta dzeta.
// kLocal == true (polynomial/hexahedron.hpp:171-219, 397-408, 526-562): `u` holds U = det(J) u, the interpolated
// quantity; with dg = grad_xi det, dh = Hessian_xi det (coordinate/cell.hpp:157-199) the same contraction gives
//   grad u = inv (gl - dg u) / det,
//   t(a)  = [sum_b Hl(a,b) w_b]/det + sum_b gl_b (m_a(b)/det - w_b dg_a/det^2) - (dg.w)/det^2 gl_a
//           - [(sum_b dh(a,b) w_b + sum_b dg_b m_a(b))/det^2 - 2 dg_a (dg.w)/det^3] U,        (H n)(i) = sum_a inv(i,a) t(a)
// so that the outputs are again the gradient and Hessian.normal of u and the stage kernel does not change.
template <int Q, bool kLocal = false>
MCFD_HD void TraceCellNode(int n, const double *u, int us, const double *geo, int gs, double *grad,
    double *hn, int os) {
  constexpr int FQ = Q * Q;
  const int idx[3] = {n / FQ, (n / Q) % Q, n % Q};
  const int str[3] = {FQ, Q, 1};
  double gl[3][5];
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    const int base = n - idx[d] * str[d];
#pragma unroll
    for (int c = 0; c < 5; ++c) gl[d][c] = 0;
#pragma unroll
    for (int m = 0; m < Q; ++m) {
      const double coef = c_tab.d1[idx[d]][m];
      const double *src = u + (size_t)(base + m * str[d]) * 5 * us;
#pragma unroll
      for (int c = 0; c < 5; ++c) gl[d][c] = fma(coef, src[c * us], gl[d][c]);
    }
  }
  const double x = c_tab.xi[idx[0]], e = c_tab.xi[idx[1]], z = c_tab.xi[idx[2]];
  const Metric m = ComputeMetric(geo, gs, x, e, z);
  const double inv_det = Rcp(m.det);
  double inv[3][3];   // inv[xyz][local] = J*(xyz, local) / det
#pragma unroll
  for (int k = 0; k < 3; ++k) {
#pragma unroll
    for (int d = 0; d < 3; ++d) inv[k][d] = m.s[d][k] * inv_det;
  }
  double A[3], B[3], C[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const double c4 = geo[(9 + k) * gs], c5 = geo[(12 + k) * gs], c6 = geo[(15 + k) * gs], c7 = geo[(18 + k) * gs];
    A[k] = fma(z, c7, c4);
    B[k] = fma(e, c7, c6);
    C[k] = fma(x, c7, c5);
  }
  double dg[3] = {0, 0, 0}, dh[3][3], big_u[5];
  if (kLocal) {
    // P_a = inv dJ_a (xyz x xyz); dJ_0 has rows (0, A, B), dJ_1 (A, 0, C), dJ_2 (B, C, 0)
    double Pm[3][3][3], tr[3];
#pragma unroll
    for (int xx = 0; xx < 3; ++xx) {
#pragma unroll
      for (int y = 0; y < 3; ++y) {
        Pm[0][xx][y] = inv[xx][1] * A[y] + inv[xx][2] * B[y];
        Pm[1][xx][y] = inv[xx][0] * A[y] + inv[xx][2] * C[y];
        Pm[2][xx][y] = inv[xx][0] * B[y] + inv[xx][1] * C[y];
      }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      tr[a] = Pm[a][0][0] + Pm[a][1][1] + Pm[a][2][2];
      dg[a] = m.det * tr[a];
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
      for (int b = 0; b < 3; ++b) {
        double t2 = 0, tpp = 0;
        if (a != b) {
          const int p3 = 3 - a - b;
#pragma unroll
          for (int xx = 0; xx < 3; ++xx) t2 += inv[xx][p3] * geo[(18 + xx) * gs];
        }
#pragma unroll
        for (int xx = 0; xx < 3; ++xx) {
#pragma unroll
          for (int y = 0; y < 3; ++y) tpp += Pm[a][xx][y] * Pm[b][y][xx];
        }
        dh[a][b] = (t2 + tr[a] * tr[b] - tpp) * m.det;
      }
    }
#pragma unroll
    for (int c = 0; c < 5; ++c) big_u[c] = u[(size_t)(n * 5 + c) * us];
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
#pragma unroll
    for (int c = 0; c < 5; ++c) {
      if (kLocal) {
        const double uc = big_u[c] * inv_det;
        grad[(size_t)(n * 15 + k * 5 + c) * os] = (inv[k][0] * (gl[0][c] - dg[0] * uc) + inv[k][1] * (gl[1][c] - dg[1] * uc)
            + inv[k][2] * (gl[2][c] - dg[2] * uc)) * inv_det;
      } else {
        grad[(size_t)(n * 15 + k * 5 + c) * os] = inv[k][0] * gl[0][c] + inv[k][1] * gl[1][c] + inv[k][2] * gl[2][c];
      }
    }
  }
  if (idx[0] != 0 && idx[0] != Q - 1 && idx[1] != 0 && idx[1] != Q - 1 && idx[2] != 0 && idx[2] != Q - 1) return;
  // local Hessian XX XY XZ YY YZ ZZ (hexahedron.hpp:155-170, 439-447)
  double hl[6][5];
#pragma unroll
  for (int h = 0; h < 6; ++h) {
#pragma unroll
    for (int c = 0; c < 5; ++c) hl[h][c] = 0;
  }
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    const int base = n - idx[d] * str[d];
    const int h = d == 0 ? 0 : (d == 1 ? 3 : 5);
#pragma unroll
    for (int mm = 0; mm < Q; ++mm) {
      const double coef = c_tab.d2[idx[d]][mm];
      const double *src = u + (size_t)(base + mm * str[d]) * 5 * us;
#pragma unroll
      for (int c = 0; c < 5; ++c) hl[h][c] = fma(coef, src[c * us], hl[h][c]);
    }
  }
#pragma unroll
  for (int pair = 0; pair < 3; ++pair) {
    const int d0 = pair == 2 ? 1 : 0, d1 = pair == 0 ? 1 : 2;     // (x,y) (x,z) (y,z)
    const int h = pair == 0 ? 1 : (pair == 1 ? 2 : 4);
    const int base = n - idx[d0] * str[d0] - idx[d1] * str[d1];
#pragma unroll
    for (int a = 0; a < Q; ++a) {
      const double ca = c_tab.d1[idx[d0]][a];
#pragma unroll
      for (int b = 0; b < Q; ++b) {
        const double coef = ca * c_tab.d1[idx[d1]][b];
        const double *src = u + (size_t)(base + a * str[d0] + b * str[d1]) * 5 * us;
#pragma unroll
        for (int c = 0; c < 5; ++c) hl[h][c] = fma(coef, src[c * us], hl[h][c]);
      }
    }
  }
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    if (idx[d] != 0 && idx[d] != Q - 1) continue;
    const bool plus = idx[d] != 0;
    const int pt = FaceOf(d, plus) * FQ + c_tab.node_q[n][d];
    const double n2 = m.s[d][0] * m.s[d][0] + m.s[d][1] * m.s[d][1] + m.s[d][2] * m.s[d][2];
    const double sc = (plus ? 1.0 : -1.0) * Rsqrt(n2);
    const double nrm[3] = {sc * m.s[d][0], sc * m.s[d][1], sc * m.s[d][2]};
    double w[3], v[3][3], mm[3][3];
#pragma unroll
    for (int b = 0; b < 3; ++b) w[b] = nrm[0] * inv[0][b] + nrm[1] * inv[1][b] + nrm[2] * inv[2][b];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      v[0][k] = w[1] * A[k] + w[2] * B[k];
      v[1][k] = w[0] * A[k] + w[2] * C[k];
      v[2][k] = w[0] * B[k] + w[1] * C[k];
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
      for (int b = 0; b < 3; ++b) mm[a][b] = -(v[a][0] * inv[0][b] + v[a][1] * inv[1][b] + v[a][2] * inv[2][b]);
    }
    double dgw = 0, mbu[3] = {0, 0, 0};
    if (kLocal) {
      dgw = dg[0] * w[0] + dg[1] * w[1] + dg[2] * w[2];
      const double id2 = inv_det * inv_det;
#pragma unroll
      for (int a = 0; a < 3; ++a)
        mbu[a] = ((dh[a][0] * w[0] + dh[a][1] * w[1] + dh[a][2] * w[2]) + (dg[0] * mm[a][0] + dg[1] * mm[a][1] + dg[2] * mm[a][2])) * id2
               - 2 * dg[a] * dgw * id2 * inv_det;
    }
#pragma unroll
    for (int c = 0; c < 5; ++c) {
      double t[3];
      t[0] = hl[0][c] * w[0] + hl[1][c] * w[1] + hl[2][c] * w[2];
      t[1] = hl[1][c] * w[0] + hl[3][c] * w[1] + hl[4][c] * w[2];
      t[2] = hl[2][c] * w[0] + hl[4][c] * w[1] + hl[5][c] * w[2];
#pragma unroll
      for (int a = 0; a < 3; ++a) t[a] += gl[0][c] * mm[a][0] + gl[1][c] * mm[a][1] + gl[2][c] * mm[a][2];
      if (kLocal) {
        const double glw = gl[0][c] * w[0] + gl[1][c] * w[1] + gl[2][c] * w[2];
        const double id2 = inv_det * inv_det;
#pragma unroll
        for (int a = 0; a < 3; ++a)
          t[a] = t[a] * inv_det - glw * dg[a] * id2 - dgw * id2 * gl[a][c] - mbu[a] * big_u[c];
      }
#pragma unroll
      for (int k = 0; k < 3; ++k)
        hn[(size_t)(pt * 15 + k * 5 + c) * os] = inv[k][0] * t[0] + inv[k][1] * t[1] + inv[k][2] * t[2];
    }
  }
}

template <int Q, int CPB, int NW>
struct SlotStates {
  double u[StageGeometry<Q, CPB, NW>::SLOTS][5];
};
template <int Q, int CPB, int NW>
struct SlotResiduals {
  double r[StageGeometry<Q, CPB, NW>::SLOTS][5];
  double inv_det[StageGeometry<Q, CPB, NW>::SLOTS];
};

// All slots' loads are issued before the first use (and, on the GPU, before the prologue's barrier):
// one memory latency per phase instead of one per slot.
template <int Q, int CPB, int NW>
MCFD_HD void StageLoadSlots(const StageArgs &p, int tid, int bid, double *smem, const double *src,
    SlotStates<Q, CPB, NW> *st) {
  using G = StageGeometry<Q, CPB, NW>;
  G g(p, tid, bid, smem);
  const double *uin = src + (size_t)g.tile * G::F * kTile + g.tl;
#pragma unroll
  for (int s = 0; s < G::SLOTS; ++s) {
    const int n = g.Node(s);
#pragma unroll
    for (int c = 0; c < 5; ++c) st->u[s][c] = n >= 0 ? uin[(n * 5 + c) * kTile] : 0.0;
  }
}

template <int Q, int CPB, int NW>
MCFD_HD void StagePrologue(const StageArgs &p, int tid, int bid, double *smem) {
  StageGeometry<Q, CPB, NW> g(p, tid, bid, smem);
  constexpr int FQ = Q * Q;
  for (int i = tid; i < 6 * 8 * FQ; i += NW * 32) {
    g.sNbrNode[i] = c_tab.nbr_node[i / (8 * FQ)][(i / FQ) % 8][i % FQ];
  }
  for (int i = tid; i < kGeomRows * CPB; i += NW * 32) {
    int row = i / CPB, c = i % CPB;
    g.sG[i] = p.geom[((size_t)g.tile * kGeomRows + row) * kTile + (g.block % g.SUB) * CPB + c];
  }
  // the face tables too: a face point's neighbour state is then one global latency away, not two
  for (int i = tid; i < 6 * CPB; i += NW * 32) {
    const size_t at = ((size_t)g.tile * 6 + i / CPB) * kTile + (g.block % g.SUB) * CPB + i % CPB;
    g.sNbr[i] = p.nbr[at];
    g.sCode[i] = p.code[at];
  }
}

// ---- phase A: contravariant fluxes Fhat = F(u) J* at every node -> shared memory --------------
// (fr/general.hpp:276-279; fem.hpp:392-395; types.hpp:322-333; polynomial/hexahedron.hpp:607-611)
template <int Q, int CPB, int NW, int kViscous = kDiffNone>
MCFD_HD void StagePhaseA(const StageArgs &p, int tid, int bid, double *smem,
    const SlotStates<Q, CPB, NW> &st, SlotResiduals<Q, CPB, NW> *res) {
  using G = StageGeometry<Q, CPB, NW>;
  G g(p, tid, bid, smem);
  constexpr int FQ = G::FQ;
#pragma unroll
  for (int s = 0; s < G::SLOTS; ++s) {
    int n = g.Node(s);
    res->inv_det[s] = 0;
    if (n < 0) continue;
    int i = n / FQ, j = (n / Q) % Q, k = n % Q;
    double u[5];
#pragma unroll
    for (int c = 0; c < 5; ++c) u[c] = st.u[s][c];
    Prim q = ConsToPrim(u[0], u[1], u[2], u[3], u[4]);
    Metric m = ComputeMetric(g.sG + g.lc, CPB, c_tab.xi[i], c_tab.xi[j], c_tab.xi[k]);
    res->inv_det[s] = Rcp(m.det);                  // fr/general.hpp:256-266; dg/lobatto.hpp:143-158
    double h0 = u[4] + q.p;
    double vel[3] = {q.u, q.v, q.w};
    double fx[3][5];  // flux matrix columns (x, y, z)
#pragma unroll
    for (int x = 0; x < 3; ++x) {
      fx[x][0] = u[1 + x];
      fx[x][1] = u[1] * vel[x];
      fx[x][2] = u[2] * vel[x];
      fx[x][3] = u[3] * vel[x];
      fx[x][1 + x] += q.p;
      fx[x][4] = h0 * vel[x];
    }
    if (kViscous) {   // fem.hpp:397-404: F -= viscous flux of (u, grad u) at the node
      double gr[3][5];
      const double *src = p.grad + ((size_t)g.tile * G::N * 15 + n * 15) * kTile + g.tl;
#pragma unroll
      for (int x = 0; x < 3; ++x) {
#pragma unroll
        for (int c = 0; c < 5; ++c) gr[x][c] = src[(x * 5 + c) * kTile];
      }
      if (kViscous == kDiffNavierStokes || kViscous == kDiffNsIsotropic) {
        NsMinusFluxMatrix(p.vis, u, gr, fx);
      }
      if (kViscous == kDiffNavierStokes) {
      } else if (kViscous == kDiffProbe) {
        const double one[5] = {1, 1, 1, 1, 1};
#pragma unroll
        for (int x = 0; x < 3; ++x) {
#pragma unroll
          for (int c = 0; c < 5; ++c) fx[x][c] = 0;
        }
        IsoMinusFluxMatrix(one, gr, fx);
      } else {
        double nu[5];
#pragma unroll
        for (int c = 0; c < 5; ++c) nu[c] = p.cell_nu[((size_t)g.tile * 5 + c) * kTile + g.tl];
        IsoMinusFluxMatrix(nu, gr, fx);
      }
    }
    double *dst = g.sF + (size_t)n * 15 * CPB + g.lc;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
#pragma unroll
      for (int c = 0; c < 5; ++c) {
        dst[(d * 5 + c) * CPB] = fx[0][c] * m.s[d][0] + fx[1][c] * m.s[d][1] + fx[2][c] * m.s[d][2];
      }
    }
  }
}

// ---- phase B1: flux divergence along the three node lines + the node's own face-flux terms ------
// (fr/general.hpp:280-286, 290-294, 305-313; dg/lobatto.hpp:161-170)
template <int Q, int CPB, int NW, bool kGeneral>
MCFD_HD void StagePhaseB1(const StageArgs &p, int tid, int bid, double *smem, SlotResiduals<Q, CPB, NW> *res) {
  using G = StageGeometry<Q, CPB, NW>;
  G g(p, tid, bid, smem);
  constexpr int FQ = G::FQ;
  const int lc = g.lc;
#pragma unroll
  for (int s = 0; s < G::SLOTS; ++s) {
    const int n = g.Node(s);
    double r[5] = {0, 0, 0, 0, 0};
    if (n >= 0) {
      const int i0 = n / FQ, i1 = (n / Q) % Q, i2 = n % Q;
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        const int id = d == 0 ? i0 : (d == 1 ? i1 : i2);
        const int st = d == 0 ? FQ : (d == 1 ? Q : 1);
        const int base = n - id * st;
#pragma unroll
        for (int mm = 0; mm < Q; ++mm) {
          const double a = c_tab.a[id][mm];
          const double *src = g.sF + ((size_t)(base + mm * st) * 15 + d * 5) * CPB + lc;
#pragma unroll
          for (int c = 0; c < 5; ++c) r[c] = fma(a, src[c * CPB], r[c]);
        }
        if (kGeneral) {                             // both line ends correct every node of the line
#pragma unroll
          for (int side = 0; side < 2; ++side) {
            const double gg = c_tab.g[side][id];
            const double *end = g.sF + ((size_t)(base + (side ? (Q - 1) * st : 0)) * 15 + d * 5) * CPB + lc;
#pragma unroll
            for (int c = 0; c < 5; ++c) r[c] = fma(gg, end[c * CPB], r[c]);
          }
        } else if (id == 0 || id == Q - 1) {        // the node lies on a face in direction d
          const double cF = id != 0 ? c_tab.coef_F : -c_tab.coef_F;
          const double *own = g.sF + ((size_t)n * 15 + d * 5) * CPB + lc;
#pragma unroll
          for (int c = 0; c < 5; ++c) r[c] = fma(cF, own[c * CPB], r[c]);
        }
      }
    }
#pragma unroll
    for (int c = 0; c < 5; ++c) res->r[s][c] = r[c];
  }
}

// after the barrier that retires sF: park the node residuals in shared memory
template <int Q, int CPB, int NW>
MCFD_HD void StageStoreResiduals(const StageArgs &p, int tid, int bid, double *smem,
    const SlotResiduals<Q, CPB, NW> &res) {
  using G = StageGeometry<Q, CPB, NW>;
  G g(p, tid, bid, smem);
#pragma unroll
  for (int s = 0; s < G::SLOTS; ++s) {
    const int n = g.Node(s);
    if (n < 0) continue;
#pragma unroll
    for (int c = 0; c < 5; ++c) g.sR[(n * 5 + c) * CPB + g.lc] = res.r[s][c];
  }
}

// ---- phase B2: one Riemann problem per face point ------------------------------------------------
// (fr/lobatto.hpp:232-251, 278-428; fr/general.hpp:296-315; dg/lobatto.hpp:171-191)
struct PointLoads {   // everything a face point needs from global memory
  double u[5], v[5];
  int nb, code;
};

template <int Q, int CPB, int NW>
MCFD_HD void LoadPoint(const StageArgs &p, const StageGeometry<Q, CPB, NW> &g, int pt, PointLoads *ld) {
  constexpr int F = StageGeometry<Q, CPB, NW>::F, FQ = Q * Q;
  ld->nb = -1; ld->code = 0;
  if (pt < 0 || !g.active) return;      // the other fields are never read for such points
  const int f = pt / FQ, q = pt % FQ;
  const int n = c_tab.face_node[f][q];
  const double *uin = p.u_in + (size_t)g.tile * F * kTile + g.tl;
#pragma unroll
  for (int c = 0; c < 5; ++c) ld->u[c] = uin[(n * 5 + c) * kTile];
  ld->nb = g.sNbr[f * CPB + g.lc];
  ld->code = g.sCode[f * CPB + g.lc];
  if (ld->nb >= 0) {
    const int n2 = g.sNbrNode[((ld->code & 7) * 8 + ((ld->code >> 3) & 7)) * FQ + q];
    const double *src = p.u_in + ((size_t)(ld->nb >> 5) * F + n2 * 5) * kTile + (ld->nb & 31);
#pragma unroll
    for (int c = 0; c < 5; ++c) ld->v[c] = src[c * kTile];
  }
}

template <int Q, int kRiemann, int CPB, int NW, bool kGeneral, int kViscous = kDiffNone>
MCFD_HD void StagePhaseB2(const StageArgs &p, int tid, int bid, double *smem) {
  using G = StageGeometry<Q, CPB, NW>;
  G g(p, tid, bid, smem);
  constexpr int FQ = G::FQ;
  const int lc = g.lc;
  PointLoads cur, next;
  LoadPoint<Q, CPB, NW>(p, g, g.Point(0), &cur);
#pragma unroll 1
  for (int k = 0; k < G::ITEMS; ++k) {
    const int pt = g.Point(k);
    // request the next point's data before solving this one
    LoadPoint<Q, CPB, NW>(p, g, (k + 1 < G::ITEMS) ? g.Point(k + 1) : -1, &next);
    if (pt >= 0 && g.active) {
      const int f = pt / FQ, q = pt % FQ;
      const int n = c_tab.face_node[f][q];
      const int d = (f == 2 || f == 4) ? 0 : ((f == 1 || f == 3) ? 1 : 2);
      const double sgn = (f == 2 || f == 3 || f == 5) ? 1.0 : -1.0;
      double nrm[3];
      ComputeNormal(g.sG + lc, CPB, c_tab.xi[n / FQ], c_tab.xi[(n / Q) % Q], c_tab.xi[n % Q], d, nrm);
      const double nx = nrm[0], ny = nrm[1], nz = nrm[2];
      const double n2 = nx * nx + ny * ny + nz * nz;
      const double inv_area = Rsqrt(n2), area = n2 * inv_area;
      double fl[5];
      // lumped: the lifted term itself; general: |n| f* along my outward normal, lifted in phase C
      double scale = kGeneral ? area : -c_tab.coef_f * area;
      if (kViscous == kDiffProbe) {
        // zero on the other side of EVERY face, boundary ones included; no Riemann flux; nu = 1
        const bool holder = (cur.code & 64) != 0;
        const double hs = (holder ? sgn : -sgn) * inv_area;
        const double os_ = sgn * inv_area;
        const double o[3] = {os_ * nx, os_ * ny, os_ * nz};
        const double *hv = p.h2s + ((size_t)g.tile * 18 + f * 3) * kTile + g.tl;
        const double dh = (hs * nx) * hv[0] + (hs * ny) * hv[kTile] + (hs * nz) * hv[2 * kTile];
        const double pen0 = p.vis.beta0 * Rcp(dh), pen1 = p.vis.beta1 * dh;
        const double *go = p.grad + ((size_t)g.tile * G::N * 15 + n * 15) * kTile + g.tl;
        const double *ho = p.hn + ((size_t)g.tile * G::POINTS * 15 + pt * 15) * kTile + g.tl;
        double gc[3][5];
#pragma unroll
        for (int x = 0; x < 3; ++x) {
#pragma unroll
          for (int c = 0; c < 5; ++c) {
            const int r = (x * 5 + c) * kTile;
            gc[x][c] = go[r] * 0.5 - (pen0 * o[x]) * cur.u[c] - pen1 * ho[r];
          }
        }
        const double one[5] = {1, 1, 1, 1, 1};
#pragma unroll
        for (int c = 0; c < 5; ++c) fl[c] = 0;
        IsoMinusFluxNormal(one, gc, hs * nx, hs * ny, hs * nz, fl);
        if (!holder) scale = -scale;
      } else if (cur.nb >= 0) {
        // The Riemann problem is posed in the holder's frame, as the reference does: when the
        // neighbour holds the face, swap the states, flip the normal and negate the result.
        const bool holder = (cur.code & 64) != 0;
        const double hs = (holder ? sgn : -sgn) * inv_area;
        double ul[5], ur[5];
#pragma unroll
        for (int c = 0; c < 5; ++c) {
          ul[c] = holder ? cur.u[c] : cur.v[c];
          ur[c] = holder ? cur.v[c] : cur.u[c];
        }
        FluxUpwind<kRiemann>(ul, ur, hs * nx, hs * ny, hs * nz, fl);
        if (kViscous) {
          // direct-DG common gradient (direct.hpp:45-78) and the viscous flux of the averaged state
          // along the holder's normal (fr/general.hpp:328-337).  In terms of MY outward unit normal o
          // and (v - u) = neighbour minus own value the expression is the same on both sides:
          //   G = (g_own + g_nbr)/2 + beta0/dist * o (v - u)^T - beta1*dist * (Hn_own + Hn_nbr)
          const double os_ = sgn * inv_area;
          const double o[3] = {os_ * nx, os_ * ny, os_ * nz};
          const double *hv = p.h2s + ((size_t)g.tile * 18 + f * 3) * kTile + g.tl;
          const double dh = (hs * nx) * hv[0] + (hs * ny) * hv[kTile] + (hs * nz) * hv[2 * kTile];
          const double pen0 = p.vis.beta0 * Rcp(dh), pen1 = p.vis.beta1 * dh;
          const int code = cur.code;
          const int f2 = code & 7;
          const int node2 = g.sNbrNode[(f2 * 8 + ((code >> 3) & 7)) * FQ + q];
          const int d2 = (f2 == 2 || f2 == 4) ? 0 : ((f2 == 1 || f2 == 3) ? 1 : 2);
          const int pt2 = f2 * FQ + c_tab.node_q[node2][d2];
          const double *go = p.grad + ((size_t)g.tile * G::N * 15 + n * 15) * kTile + g.tl;
          const double *ho = p.hn + ((size_t)g.tile * G::POINTS * 15 + pt * 15) * kTile + g.tl;
          const double *gn = p.grad + ((size_t)(cur.nb >> 5) * G::N * 15 + node2 * 15) * kTile + (cur.nb & 31);
          const double *hb = p.hn + ((size_t)(cur.nb >> 5) * G::POINTS * 15 + pt2 * 15) * kTile + (cur.nb & 31);
          double gc[3][5], uc[5];
#pragma unroll
          for (int c = 0; c < 5; ++c) uc[c] = (ul[c] + ur[c]) * 0.5;
#pragma unroll
          for (int x = 0; x < 3; ++x) {
#pragma unroll
            for (int c = 0; c < 5; ++c) {
              const int r = (x * 5 + c) * kTile;
              gc[x][c] = (go[r] + gn[r]) * 0.5 + (pen0 * o[x]) * (cur.v[c] - cur.u[c]) - pen1 * (ho[r] + hb[r]);
            }
          }
          if (kViscous == kDiffNavierStokes || kViscous == kDiffNsIsotropic) {
            NsMinusFluxNormal(p.vis, uc, gc, hs * nx, hs * ny, hs * nz, fl);
          }
          if (kViscous != kDiffNavierStokes) {   // property = (mine + neighbour's) / 2, fr/general.hpp:331-335
            double nu[5];
#pragma unroll
            for (int c = 0; c < 5; ++c)
              nu[c] = (p.cell_nu[((size_t)g.tile * 5 + c) * kTile + g.tl]
                     + p.cell_nu[((size_t)(cur.nb >> 5) * 5 + c) * kTile + (cur.nb & 31)]) * 0.5;
            IsoMinusFluxNormal(nu, gc, hs * nx, hs * ny, hs * nz, fl);
          }
        }
        if (!holder) scale = -scale;
      } else {
        const int b = -1 - cur.nb;
        const int kind = p.bkind[b];
        const double inv = sgn * inv_area;
        if (kViscous && !kGeneral && kind == kBcNoSlip) {   // fr::General leaves AddFluxOnNoSlipWalls empty (fem.hpp:382-383)
          // fr/lobatto.hpp:301-319, fr/general.hpp:460-479: inviscid-wall flux minus the viscous flux
          // with the velocity penalty beta0/dist; `given` = (-, wall velocity, n . grad T)
          const double *gv = p.bvals + ((size_t)b * FQ + q) * 5;
          const double wall[5] = {gv[0], gv[1], gv[2], gv[3], gv[4]};
          BoundaryFlux<kRiemann>(kBcWall, cur.u, wall, inv * nx, inv * ny, inv * nz, fl);
          const double *hv = p.h2s + ((size_t)g.tile * 18 + f * 3) * kTile + g.tl;
          const double dh = (inv * nx) * hv[0] + (inv * ny) * hv[kTile] + (inv * nz) * hv[2 * kTile];
          const double *go = p.grad + ((size_t)g.tile * G::N * 15 + n * 15) * kTile + g.tl;
          double gr[3][5];
#pragma unroll
          for (int x = 0; x < 3; ++x) {
#pragma unroll
            for (int c = 0; c < 5; ++c) gr[x][c] = go[(x * 5 + c) * kTile];
          }
          if (kViscous == kDiffNavierStokes || kViscous == kDiffNsIsotropic) {
            NsMinusFluxNoSlip(p.vis, wall, cur.u, gr, inv * nx, inv * ny, inv * nz, p.vis.beta0 * Rcp(dh), fl);
          }
          // over a diffusive base EnergyBasedViscosity::MinusViscousFluxOnNoSlipWall goes on with its own
          // MinusViscousFlux(flux, nu, value, gradient, normal), which subtracts the BASE's normal flux of the
          // unpenalised gradient once more before the artificial one (viscosity.hpp:114-139)
          if (kViscous == kDiffNsIsotropic) NsMinusFluxNormal(p.vis, cur.u, gr, inv * nx, inv * ny, inv * nz, fl);
          if (kViscous != kDiffNavierStokes) {   // linear.hpp:85-91: the cell's own gradient, no penalty
            double nu[5];
#pragma unroll
            for (int c = 0; c < 5; ++c) nu[c] = p.cell_nu[((size_t)g.tile * 5 + c) * kTile + g.tl];
            IsoMinusFluxNormal(nu, gr, inv * nx, inv * ny, inv * nz, fl);
          }
        } else if (kind == kBcNoSlip) {
          // no face term at all for inviscid systems (fem.hpp:382-383): cancel the node's own
          // face-flux term that phase B1 has already added, -coef_F * sgn * (F(u) . J*col_d) / det
          double fm[5][3];
          Prim pr = ConsToPrim(cur.u[0], cur.u[1], cur.u[2], cur.u[3], cur.u[4]);
          const double vel[3] = {pr.u, pr.v, pr.w};
          const double h0 = cur.u[4] + pr.p;
#pragma unroll
          for (int x = 0; x < 3; ++x) {
            fm[0][x] = cur.u[1 + x];
            fm[1][x] = cur.u[1] * vel[x]; fm[2][x] = cur.u[2] * vel[x]; fm[3][x] = cur.u[3] * vel[x];
            fm[1 + x][x] += pr.p;
            fm[4][x] = h0 * vel[x];
          }
          if (kViscous == kDiffNavierStokes || kViscous == kDiffIsotropic || kViscous == kDiffNsIsotropic) {   // the cached flux carried the viscous part too
            const double *go = p.grad + ((size_t)g.tile * G::N * 15 + n * 15) * kTile + g.tl;
            double gr[3][5], fx[3][5];
#pragma unroll
            for (int x = 0; x < 3; ++x) {
#pragma unroll
              for (int c = 0; c < 5; ++c) { gr[x][c] = go[(x * 5 + c) * kTile]; fx[x][c] = fm[c][x]; }
            }
            if (kViscous == kDiffNavierStokes || kViscous == kDiffNsIsotropic) {
              NsMinusFluxMatrix(p.vis, cur.u, gr, fx);
            }
            if (kViscous != kDiffNavierStokes) {
              double nu[5];
#pragma unroll
              for (int c = 0; c < 5; ++c) nu[c] = p.cell_nu[((size_t)g.tile * 5 + c) * kTile + g.tl];
              IsoMinusFluxMatrix(nu, gr, fx);
            }
#pragma unroll
            for (int x = 0; x < 3; ++x) {
#pragma unroll
              for (int c = 0; c < 5; ++c) fm[c][x] = fx[x][c];
            }
          }
#pragma unroll
          for (int c = 0; c < 5; ++c) fl[c] = fm[c][0] * nx + fm[c][1] * ny + fm[c][2] * nz;
          scale = kGeneral ? sgn : -c_tab.coef_F * sgn;
        } else {
          State5 ui, given;
          const double *gv = p.bvals + ((size_t)b * FQ + q) * 5;
#pragma unroll
          for (int c = 0; c < 5; ++c) { ui.v[c] = cur.u[c]; given.v[c] = gv[c]; }
          const State5 bf = BoundaryFluxOutOfLine<kRiemann>(kind, ui, given, inv * nx, inv * ny, inv * nz);
#pragma unroll
          for (int c = 0; c < 5; ++c) fl[c] = bf.v[c];
        }
      }
      double *dst = g.sC + (size_t)pt * 5 * CPB + lc;
#pragma unroll
      for (int c = 0; c < 5; ++c) dst[c * CPB] = scale * fl[c];
    }
    cur = next;
  }
}

// ---- phase C: gather the face terms, RK update, store ---------------------------------------------
template <int Q, int CPB, int NW, bool kGeneral>
MCFD_HD void StagePhaseC(const StageArgs &p, int tid, int bid, double *smem,
    const SlotStates<Q, CPB, NW> &su, const SlotStates<Q, CPB, NW> &su0, const SlotResiduals<Q, CPB, NW> &sres) {
  using G = StageGeometry<Q, CPB, NW>;
  G g(p, tid, bid, smem);
  constexpr int F = G::F, FQ = G::FQ;
  if (!g.active) return;
  const int lc = g.lc;
  double *out = p.u_out + (size_t)g.tile * F * kTile + g.tl;
  const double dt = p.dt_dev ? *p.dt_dev : p.dt;
#pragma unroll
  for (int s = 0; s < G::SLOTS; ++s) {
    const int n = g.Node(s);
    if (n < 0) continue;
    const int i0 = n / FQ, i1 = (n / Q) % Q, i2 = n % Q;
    double r[5];
#pragma unroll
    for (int c = 0; c < 5; ++c) r[c] = g.sR[(n * 5 + c) * CPB + lc];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      const int id = d == 0 ? i0 : (d == 1 ? i1 : i2);
      if (kGeneral) {
        const int st = d == 0 ? FQ : (d == 1 ? Q : 1);
        const int base = n - id * st;
#pragma unroll
        for (int side = 0; side < 2; ++side) {
          const int e = base + (side ? (Q - 1) * st : 0);            // end node of my line
          const int pt = FaceOf(d, side != 0) * FQ + c_tab.node_q[e][d];
          const double gg = side ? -c_tab.g[1][id] : c_tab.g[0][id];   // -g' * sgn of that face
#pragma unroll
          for (int c = 0; c < 5; ++c) r[c] = fma(gg, g.sC[((size_t)pt * 5 + c) * CPB + lc], r[c]);
        }
        continue;
      }
      const int q = c_tab.node_q[n][d];
      if (q < 0) continue;
      const int pt = FaceOf(d, id != 0) * FQ + q;
#pragma unroll
      for (int c = 0; c < 5; ++c) r[c] += g.sC[((size_t)pt * 5 + c) * CPB + lc];
    }
#pragma unroll
    for (int c = 0; c < 5; ++c) {
      const double u = su.u[s][c], u0 = su0.u[s][c], res = r[c] * sres.inv_det[s];
      double val;
      switch (p.mode) {
      case kModeResidual: val = res; break;
      case kModeEuler: val = u + res * dt; break;
      case kModeRk3b: val = (u0 * 3 + (u + res * dt)) * 0.25; break;
      case kModeRk3c: val = ((u + res * dt) * 2 + u0) / 3; break;
      default: val = ((u + res * dt) + u0) * 0.5; break;
      }
      out[(n * 5 + c) * kTile] = val;
    }
  }
}

// One CTA-sized block, all phases; `sync` is __syncthreads() on the device and a no-op marker on the
// host, where RunStageBlockOnHost() below replays the phases thread by thread.
#ifdef __CUDACC__
template <int Q, int kRiemann, int CPB, int NW, bool kGeneral, int kViscous = kDiffNone>
__global__ void __launch_bounds__(NW * 32, (Q == 3 ? (kViscous ? MCFD_P2_VMINB : MCFD_P2_MINB) : (kViscous ? 1 : MCFD_P3_MINB)))
StageKernel(StageArgs p) {
  extern __shared__ double smem[];
  const int tid = threadIdx.x, bid = blockIdx.x;
  SlotStates<Q, CPB, NW> su;
  SlotResiduals<Q, CPB, NW> res;
  StageLoadSlots<Q, CPB, NW>(p, tid, bid, smem, p.u_in, &su);
  StagePrologue<Q, CPB, NW>(p, tid, bid, smem);
  __syncthreads();
  StagePhaseA<Q, CPB, NW, kViscous>(p, tid, bid, smem, su, &res);
  __syncthreads();
  StagePhaseB1<Q, CPB, NW, kGeneral>(p, tid, bid, smem, &res);
  __syncthreads();
  StageStoreResiduals<Q, CPB, NW>(p, tid, bid, smem, res);
  StagePhaseB2<Q, kRiemann, CPB, NW, kGeneral, kViscous>(p, tid, bid, smem);
  // the states needed by the update are requested before the last barrier
  SlotStates<Q, CPB, NW> su0;
  StageLoadSlots<Q, CPB, NW>(p, tid, bid, smem, p.u_in, &su);
  StageLoadSlots<Q, CPB, NW>(p, tid, bid, smem, p.mode >= kModeRk3b ? p.u0 : p.u_in, &su0);
  __syncthreads();
  StagePhaseC<Q, CPB, NW, kGeneral>(p, tid, bid, smem, su, su0, res);
}
#else
template <int Q, int kRiemann, int CPB, int NW, bool kGeneral, int kViscous = kDiffNone>
void RunStageBlockOnHost(const StageArgs &p, int bid, double *smem) {
  constexpr int T = NW * 32;
  static SlotStates<Q, CPB, NW> su[T], su0[T];
  static SlotResiduals<Q, CPB, NW> res[T];
  for (int t = 0; t < T; ++t) {
    StageLoadSlots<Q, CPB, NW>(p, t, bid, smem, p.u_in, &su[t]);
    StagePrologue<Q, CPB, NW>(p, t, bid, smem);
  }
  for (int t = 0; t < T; ++t) StagePhaseA<Q, CPB, NW, kViscous>(p, t, bid, smem, su[t], &res[t]);
  for (int t = 0; t < T; ++t) StagePhaseB1<Q, CPB, NW, kGeneral>(p, t, bid, smem, &res[t]);
  for (int t = 0; t < T; ++t) StageStoreResiduals<Q, CPB, NW>(p, t, bid, smem, res[t]);
  for (int t = 0; t < T; ++t) {
    StagePhaseB2<Q, kRiemann, CPB, NW, kGeneral, kViscous>(p, t, bid, smem);
    StageLoadSlots<Q, CPB, NW>(p, t, bid, smem, p.mode >= kModeRk3b ? p.u0 : p.u_in, &su0[t]);
  }
  for (int t = 0; t < T; ++t) StagePhaseC<Q, CPB, NW, kGeneral>(p, t, bid, smem, su[t], su0[t], res[t]);
}
#endif

// ---- WENO limiter on the modal extension of the nodal expansion -----------------
// mini::polynomial::Extrapolation<Hexahedron> keeps, next to the nodal values, their L2 projection on a
// per-cell orthonormal basis phi = S * taylor(x - center) (polynomial/extrapolation.hpp:60-90,
// basis/linear.hpp:87-141, integrator/function.hpp:138-172); weno::Lazy (limiter/weno.hpp:97-195) blends
// that projection with the neighbours' polynomials re-projected on the cell's own basis, weighted by
// Taylor-derivative smoothness indicators (weno.hpp:25-95), and the nodal values are re-sampled from the
// blend.  Per cell the host stores BAS = centre (3), S packed row-wise (M(M+1)/2), the smoothness weights
// volume^(2k/3-1)/(k!)^2 for k = 1..3; the device works per (cell, component) because the components never
// mix in weno::Lazy.  All sums over quadrature points use moments against the *Taylor* monomials:
//     g_k = sum_q W_q taylor_k(x_q - c) f(x_q),     projection on phi:  a = S g,    Taylor form:  t = S^T a.
template <int P>
struct ModalShape {
  static constexpr int M = (P == 2) ? 10 : 20;            // basis/taylor.hpp:27-31
  static constexpr int kTri = M * (M + 1) / 2;
  static constexpr int kBasRows = 3 + kTri + 3;
  static constexpr int kSOffset = 3, kWOffset = 3 + kTri;
};

template <int P>
MCFD_HD void TaylorValues(const double *d, double *t) {   // basis/taylor.hpp:204-218
  const double x = d[0], y = d[1], z = d[2];
  t[0] = 1; t[1] = x; t[2] = y; t[3] = z;
  t[4] = x * x; t[5] = x * y; t[6] = x * z; t[7] = y * y; t[8] = y * z; t[9] = z * z;
  if (P == 3) {
    t[10] = x * t[4]; t[11] = x * t[5]; t[12] = x * t[6]; t[13] = x * t[7]; t[14] = x * t[8]; t[15] = x * t[9];
    t[16] = y * t[7]; t[17] = y * t[8]; t[18] = y * t[9]; t[19] = z * t[9];
  }
}

// Sum of the squared partial derivatives of each order at one point (Taylor::GetPartialDerivatives,
// basis/taylor.hpp:241-372, squared and grouped as weno.hpp:55-75 groups them): sq[k-1] for order k.
template <int P>
MCFD_HD void TaylorDerivativeSquares(const double *d, const double *c, double *sq) {
  const double x = d[0], y = d[1], z = d[2];
  double r1 = c[1] + c[4] * (2 * x) + c[5] * y + c[6] * z;
  double r2 = c[2] + c[5] * x + c[7] * (2 * y) + c[8] * z;
  double r3 = c[3] + c[6] * x + c[8] * y + c[9] * (2 * z);
  double r4 = c[4], r5 = c[5], r6 = c[6], r7 = c[7], r8 = c[8], r9 = c[9];
  if (P == 3) {
    const double xx = x * x, xy = x * y, xz = x * z, yy = y * y, yz = y * z, zz = z * z;
    r1 += c[10] * (3 * xx) + c[11] * (2 * xy) + c[12] * (2 * xz) + c[13] * yy + c[14] * yz + c[15] * zz;
    r2 += c[11] * xx + c[14] * xz + c[13] * (2 * xy) + c[16] * (3 * yy) + c[17] * (2 * yz) + c[18] * zz;
    r3 += c[12] * xx + c[14] * xy + c[15] * (2 * xz) + c[17] * yy + c[18] * (2 * yz) + c[19] * (3 * zz);
    r4 += c[11] * y + c[12] * z + c[10] * x * 3;
    r5 += c[11] * x * 2 + c[13] * y * 2 + c[14] * z;
    r6 += c[12] * x * 2 + c[15] * z * 2 + c[14] * y;
    r7 += c[13] * x + c[17] * z + c[16] * y * 3;
    r8 += c[14] * x + c[17] * y * 2 + c[18] * z * 2;
    r9 += c[15] * x + c[18] * y + c[19] * z * 3;
  }
  r4 += r4; r7 += r7; r9 += r9;
  sq[0] = r1 * r1 + r2 * r2 + r3 * r3;
  sq[1] = r4 * r4 + r5 * r5 + r6 * r6 + r7 * r7 + r8 * r8 + r9 * r9;
  if (P == 3) {
    const double a = c[10] * 6, b = c[11] * 2, e = c[12] * 2, f = c[13] * 2, g = c[14], h = c[15] * 2;
    const double i = c[16] * 6, j = c[17] * 2, k = c[18] * 2, l = c[19] * 6;
    sq[2] = a * a + b * b + e * e + f * f + g * g + h * h + i * i + j * j + k * k + l * l;
  }
}

// offset of quadrature point n from the cell centre, and its global weight (local weight x det J)
template <int Q>
MCFD_HD void QuadraturePoint(int n, const double *geo, int gs, double *d, double *w, bool legendre = false) {
  constexpr int FQ = Q * Q;
  const int i = n / FQ, j = (n / Q) % Q, k = n % Q;
  const double *px = legendre ? c_tab.gx : c_tab.xi, *pw = legendre ? c_tab.gw : c_tab.w;
  const double x = px[i], e = px[j], z = px[k];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    d[a] = geo[(0 + a) * gs] * x + geo[(3 + a) * gs] * e + geo[(6 + a) * gs] * z + geo[(9 + a) * gs] * (x * e)
         + geo[(12 + a) * gs] * (e * z) + geo[(15 + a) * gs] * (z * x) + geo[(18 + a) * gs] * (x * e * z);
  }
  const Metric m = ComputeMetric(geo, gs, x, e, z);
  *w = pw[i] * pw[j] * pw[k] * m.det;
}

// a = S g (orthonormal coefficients from Taylor moments) and t = S^T a (Taylor coefficients);
// S packed row-wise, rows `ss` doubles apart
template <int M>
MCFD_HD void MomentsToModal(const double *S, int ss, const double *g, double *a) {
  int r = 0;
  for (int i = 0; i < M; ++i) {
    double v = 0;
    for (int j = 0; j <= i; ++j) v = fma(S[(size_t)(r + j) * ss], g[j], v);
    a[i] = v;
    r += i + 1;
  }
}
template <int M>
MCFD_HD void ModalToTaylor(const double *S, int ss, const double *a, double *t) {
  for (int j = 0; j < M; ++j) t[j] = 0;
  int r = 0;
  for (int i = 0; i < M; ++i) {
    for (int j = 0; j <= i; ++j) t[j] = fma(a[i], S[(size_t)(r + j) * ss], t[j]);
    r += i + 1;
  }
}

// Extrapolation::UpdateModalCoeff for one (cell, component): u rows `us` apart (node-major, component
// already selected), quadrature table quad[n*4 + {dx,dy,dz,W}] rows `qs` apart, outputs rows `os` apart.
template <int Q>
MCFD_HD void ModalProjectCell(const double *u, int us, const double *quad, int qs, const double *bas, int bs,
    double *mod, double *tay, int os) {
  constexpr int P = Q - 1, N = Q * Q * Q, M = ModalShape<P>::M;
  double g[M], a[M], t[M];
  for (int k = 0; k < M; ++k) g[k] = 0;
  for (int n = 0; n < N; ++n) {
    const double d[3] = {quad[(size_t)(n * 4) * qs], quad[(size_t)(n * 4 + 1) * qs], quad[(size_t)(n * 4 + 2) * qs]};
    const double wf = quad[(size_t)(n * 4 + 3) * qs] * u[(size_t)(n * 5) * us];
    double tv[M];
    TaylorValues<P>(d, tv);
    for (int k = 0; k < M; ++k) g[k] = fma(tv[k], wf, g[k]);
  }
  const double *S = bas + (size_t)ModalShape<P>::kSOffset * bs;
  MomentsToModal<M>(S, bs, g, a);
  ModalToTaylor<M>(S, bs, a, t);
  for (int k = 0; k < M; ++k) { mod[(size_t)k * os] = a[k]; tay[(size_t)k * os] = t[k]; }
}

// weno::GetSmoothness of one component given on my orthonormal basis (limiter/weno.hpp:78-95, 37-75)
template <int Q>
MCFD_HD double WenoSmoothness(const double *a, const double *quad, int qs, const double *bas, int bs) {
  constexpr int P = Q - 1, N = Q * Q * Q, M = ModalShape<P>::M;
  double t[M], sum[3] = {0, 0, 0};
  ModalToTaylor<M>(bas + (size_t)ModalShape<P>::kSOffset * bs, bs, a, t);
  for (int n = 0; n < N; ++n) {
    const double d[3] = {quad[(size_t)(n * 4) * qs], quad[(size_t)(n * 4 + 1) * qs], quad[(size_t)(n * 4 + 2) * qs]};
    const double w = quad[(size_t)(n * 4 + 3) * qs];
    double sq[3];
    TaylorDerivativeSquares<P>(d, t, sq);
    for (int k = 0; k < P; ++k) sum[k] = fma(sq[k], w, sum[k]);
  }
  double beta = 0;
  for (int k = 0; k < P; ++k) beta += sum[k] * bas[(size_t)(ModalShape<P>::kWOffset + k) * bs];
  return beta;
}

struct WenoArgs {
  double *u;              // state tiles (nodal values); cells of the batch are overwritten
  const double *geom, *bas, *h2s;
  const int *nbr;
  const unsigned char *code;
  double *mod, *tay;      // [tile][comp*M + k][lane] orthonormal / Taylor coefficients of every cell
  double *mod_new, *tay_new;  // the limited cells' new coefficients (copied over after the batch)
  int cell_begin, cell_end;
  double w0, eps;
  // coefficient (comp, k) of a cell sits at row comp*comp_stride + k*mode_stride of its 5 M rows: (M, 1) for the
  // limiter's own tables, (1, 5) when the tables ARE a modal state (dg::General on Projection cells: modal_state != 0,
  // the blend is written straight into `u`, nothing is resampled, the rule is Gauss-Legendre)
  int comp_stride, mode_stride, modal_state, legendre;
};

// weno::Lazy::Reconstruct for one (cell, component) (limiter/weno.hpp:127-194) + Extrapolation::SetCoeff
// (polynomial/extrapolation.hpp:119-122, 66-71).  `tile`, `tl`: my cell; quad/bas as above.
template <int Q>
MCFD_HD void WenoLazyCell(const WenoArgs &p, int tile, int tl, int comp, const double *quad, int qs,
    const double *bas, int bs) {
  constexpr int P = Q - 1, N = Q * Q * Q, F = 5 * N, M = ModalShape<P>::M;
  const double *S = bas + (size_t)ModalShape<P>::kSOffset * bs;
  const double s00 = S[0];
  double mine[M], fresh[M];
  const size_t ms = (size_t)p.mode_stride * kTile;
  const double *my_mod = p.mod + ((size_t)tile * 5 * M + comp * p.comp_stride) * kTile + tl;
  for (int k = 0; k < M; ++k) { mine[k] = my_mod[k * ms]; fresh[k] = 0; }
  const double my_average = mine[0] * s00;                  // projection.hpp:46-51
  int adj_cnt = 0;
  double sum_w = 0;
  for (int f = 0; f < 6; ++f) {
    const int nb = p.nbr[((size_t)tile * 6 + f) * kTile + tl];
    if (nb < 0) continue;
    ++adj_cnt;
    // my centre -> neighbour's centre: +-HolderToSharer (through the face, so periodic pairs work too)
    const bool holder = (p.code[((size_t)tile * 6 + f) * kTile + tl] & 64) != 0;
    const double *hv = p.h2s + ((size_t)tile * 18 + f * 3) * kTile + tl;
    const double sg = holder ? 1.0 : -1.0;
    const double off[3] = {sg * hv[0], sg * hv[kTile], sg * hv[2 * kTile]};
    const double *nt = p.tay + ((size_t)(nb >> 5) * 5 * M + comp * p.comp_stride) * kTile + (nb & 31);
    double tb[M], g[M], cand[M];
    for (int k = 0; k < M; ++k) { tb[k] = nt[k * ms]; g[k] = 0; }
    for (int n = 0; n < N; ++n) {   // Borrow: neighbour's polynomial at my quadrature points, projected on my basis
      const double d[3] = {quad[(size_t)(n * 4) * qs], quad[(size_t)(n * 4 + 1) * qs], quad[(size_t)(n * 4 + 2) * qs]};
      const double dn[3] = {d[0] - off[0], d[1] - off[1], d[2] - off[2]};
      double tv[M];
      TaylorValues<P>(dn, tv);
      double v = 0;
      for (int k = 0; k < M; ++k) v = fma(tb[k], tv[k], v);
      v *= quad[(size_t)(n * 4 + 3) * qs];
      TaylorValues<P>(d, tv);
      for (int k = 0; k < M; ++k) g[k] = fma(tv[k], v, g[k]);
    }
    MomentsToModal<M>(S, bs, g, cand);
    cand[0] += (my_average - cand[0] * s00) / s00;           // projection.hpp:257-261
    double beta = WenoSmoothness<Q>(cand, quad, qs, bas, bs) + p.eps;
    const double w = p.w0 / (beta * beta);
    sum_w += w;
    for (int k = 0; k < M; ++k) fresh[k] = fma(w, cand[k], fresh[k]);
  }
  {
    double beta = WenoSmoothness<Q>(mine, quad, qs, bas, bs) + p.eps;
    const double w = (p.w0 * (-adj_cnt) + 1.0) / (beta * beta);
    sum_w += w;
    for (int k = 0; k < M; ++k) fresh[k] = fma(w, mine[k], fresh[k]);
  }
  const double inv = 1.0 / sum_w;
  for (int k = 0; k < M; ++k) fresh[k] *= inv;
  if (p.modal_state) {   // Projection::SetCoeff: the blend is the new state
    double *state = p.u + ((size_t)tile * 5 * M + comp * p.comp_stride) * kTile + tl;
    for (int k = 0; k < M; ++k) state[k * ms] = fresh[k];
    return;
  }
  double t[M];
  ModalToTaylor<M>(S, bs, fresh, t);
  double *new_mod = p.mod_new + ((size_t)tile * 5 * M + comp * M) * kTile + tl;
  double *new_tay = p.tay_new + ((size_t)tile * 5 * M + comp * M) * kTile + tl;
  for (int k = 0; k < M; ++k) { new_mod[(size_t)k * kTile] = fresh[k]; new_tay[(size_t)k * kTile] = t[k]; }
  double *dst = p.u + (size_t)tile * F * kTile + tl;
  for (int n = 0; n < N; ++n) {   // UpdateNodalCoeff: the blend sampled at my nodes
    const double d[3] = {quad[(size_t)(n * 4) * qs], quad[(size_t)(n * 4 + 1) * qs], quad[(size_t)(n * 4 + 2) * qs]};
    double tv[M];
    TaylorValues<P>(d, tv);
    double v = 0;
    for (int k = 0; k < M; ++k) v = fma(t[k], tv[k], v);
    dst[(size_t)(n * 5 + comp) * kTile] = v;
  }
}

// ---- weno::Eigen (limiter/weno.hpp:197-349): the characteristic variant ---------------------------------
// Same borrowing as weno::Lazy, but only for the cells its jump detector flags, and the blend is done once
// per adjacent face in the characteristic variables of that face's frame (EigenMatrices of the cell mean),
// the per-face results averaged with the neighbours' volumes as weights.  The components mix through L and R,
// so the (cell, component) threads of a tile meet twice per face in scratch memory:
//   cand  [(slot*5 + comp)*M + k][lane]   slot 0..5 = face, 6 = the cell's own projection
//   blend [comp*M + k][lane]              the blended characteristic coefficients of the current face
struct WenoEigenArgs {
  WenoArgs w;
  const double *frames;    // [tile][face*9 + i][lane]
  double *cand, *blend;    // scratch of this CTA
};
template <int P>
struct EigenState {
  double total[ModalShape<P>::M];
  double total_volume;
  int adj_cnt;
  bool troubled;
};

// Eigen::IsNotSmooth (weno.hpp:230-258) for my cell; identical in all five component threads.
template <int Q>
MCFD_HD bool WenoEigenDetect(const WenoArgs &p, int tile, int tl, const double *quad, int qs, const double *bas, int bs,
    int *adj_cnt_out) {
  constexpr int P = Q - 1, N = Q * Q * Q, M = ModalShape<P>::M, kBasRows = ModalShape<P>::kBasRows;
  const double s00 = bas[(size_t)ModalShape<P>::kSOffset * bs];
  double volume = 0;
  for (int n = 0; n < N; ++n) volume += quad[(size_t)(n * 4 + 3) * qs];
  double mine[2], max_avg[2], sum_diff[2] = {0, 0};
  for (int t = 0; t < 2; ++t) {
    const int k = t * 4;
    const size_t row = ((size_t)tile * 5 * M + k * M) * kTile + tl;
    mine[t] = p.tay[row];                                   // my polynomial at my centre
    max_avg[t] = fmax(1e-9, fabs(p.mod[row] * s00));
  }
  int adj_cnt = 0;
  for (int f = 0; f < 6; ++f) {
    const int nb = p.nbr[((size_t)tile * 6 + f) * kTile + tl];
    if (nb < 0) continue;
    ++adj_cnt;
    const bool holder = (p.code[((size_t)tile * 6 + f) * kTile + tl] & 64) != 0;
    const double *hv = p.h2s + ((size_t)tile * 18 + f * 3) * kTile + tl;
    const double sg = holder ? -1.0 : 1.0;                  // my centre seen from the neighbour's
    const double dn[3] = {sg * hv[0], sg * hv[kTile], sg * hv[2 * kTile]};
    double tv[M];
    TaylorValues<P>(dn, tv);
    const double s00n = p.bas[((size_t)(nb >> 5) * kBasRows + ModalShape<P>::kSOffset) * kTile + (nb & 31)];
    for (int t = 0; t < 2; ++t) {
      const int k = t * 4;
      const size_t row = ((size_t)(nb >> 5) * 5 * M + k * M) * kTile + (nb & 31);
      double v = 0;
      for (int i = 0; i < M; ++i) v = fma(p.tay[row + (size_t)i * kTile], tv[i], v);
      sum_diff[t] += fabs(mine[t] - v);
      max_avg[t] = fmax(max_avg[t], fabs(p.mod[row] * s00n));
    }
  }
  *adj_cnt_out = adj_cnt;
  const double divisor = Pow(volume, (P + 1.0) / 2.0 / 3) * adj_cnt;
  const double reference = P < 3 ? 1.0 : 3.0;
  return (sum_diff[0] / max_avg[0] / divisor > reference) || (sum_diff[1] / max_avg[1] / divisor > reference);
}

// Borrow (weno.hpp:269-283): my component of every candidate -> cand scratch
template <int Q>
MCFD_HD void WenoEigenBorrow(const WenoEigenArgs &e, int tile, int tl, int comp, const double *quad, int qs,
    const double *bas, int bs) {
  constexpr int P = Q - 1, N = Q * Q * Q, M = ModalShape<P>::M;
  const WenoArgs &p = e.w;
  const double *S = bas + (size_t)ModalShape<P>::kSOffset * bs;
  const double s00 = S[0];
  const double *my_mod = p.mod + ((size_t)tile * 5 * M + comp * M) * kTile + tl;
  const double my_average = my_mod[0] * s00;
  for (int f = 0; f < 6; ++f) {
    const int nb = p.nbr[((size_t)tile * 6 + f) * kTile + tl];
    if (nb < 0) continue;
    const bool holder = (p.code[((size_t)tile * 6 + f) * kTile + tl] & 64) != 0;
    const double *hv = p.h2s + ((size_t)tile * 18 + f * 3) * kTile + tl;
    const double sg = holder ? 1.0 : -1.0;
    const double off[3] = {sg * hv[0], sg * hv[kTile], sg * hv[2 * kTile]};
    const double *nt = p.tay + ((size_t)(nb >> 5) * 5 * M + comp * M) * kTile + (nb & 31);
    double tb[M], g[M], cand[M];
    for (int k = 0; k < M; ++k) { tb[k] = nt[(size_t)k * kTile]; g[k] = 0; }
    for (int n = 0; n < N; ++n) {
      const double d[3] = {quad[(size_t)(n * 4) * qs], quad[(size_t)(n * 4 + 1) * qs], quad[(size_t)(n * 4 + 2) * qs]};
      const double dn[3] = {d[0] - off[0], d[1] - off[1], d[2] - off[2]};
      double tv[M];
      TaylorValues<P>(dn, tv);
      double v = 0;
      for (int k = 0; k < M; ++k) v = fma(tb[k], tv[k], v);
      v *= quad[(size_t)(n * 4 + 3) * qs];
      TaylorValues<P>(d, tv);
      for (int k = 0; k < M; ++k) g[k] = fma(tv[k], v, g[k]);
    }
    MomentsToModal<M>(S, bs, g, cand);
    cand[0] += (my_average - cand[0] * s00) / s00;
    double *dst = e.cand + ((size_t)(f * 5 + comp) * M) * kTile + tl;
    for (int k = 0; k < M; ++k) dst[(size_t)k * kTile] = cand[k];
  }
  double *dst = e.cand + ((size_t)(6 * 5 + comp) * M) * kTile + tl;
  for (int k = 0; k < M; ++k) dst[(size_t)k * kTile] = my_mod[(size_t)k * kTile];
}

// ReconstructOnFace, first half (weno.hpp:290-330): characteristic component `comp` of every candidate, its
// smoothness, the weights, the blend -> blend scratch
template <int Q>
MCFD_HD void WenoEigenFaceBlend(const WenoEigenArgs &e, int tile, int tl, int comp, int face, int adj_cnt,
    const double *quad, int qs, const double *bas, int bs) {
  constexpr int P = Q - 1, M = ModalShape<P>::M;
  const WenoArgs &p = e.w;
  const double s00 = bas[(size_t)ModalShape<P>::kSOffset * bs];
  double big_u[5], frame[9], row[5];
  for (int c = 0; c < 5; ++c) big_u[c] = p.mod[((size_t)tile * 5 * M + c * M) * kTile + tl] * s00;
  for (int i = 0; i < 9; ++i) frame[i] = e.frames[((size_t)tile * 54 + face * 9 + i) * kTile + tl];
  const EigenCommon ec = EigenPrepare(big_u, frame);
  EigenRowL(ec, frame, comp, row);
  double blend[M], sum_w = 0;
  for (int k = 0; k < M; ++k) blend[k] = 0;
  for (int slot = 0; slot < 7; ++slot) {
    if (slot < 6 && p.nbr[((size_t)tile * 6 + slot) * kTile + tl] < 0) continue;
    double rot[M];
    for (int k = 0; k < M; ++k) {
      double v = 0;
      for (int c = 0; c < 5; ++c) v = fma(row[c], e.cand[((size_t)(slot * 5 + c) * M + k) * kTile + tl], v);
      rot[k] = v;
    }
    const double beta = WenoSmoothness<Q>(rot, quad, qs, bas, bs) + p.eps;
    const double w = (slot < 6 ? p.w0 : p.w0 * (-adj_cnt) + 1.0) / (beta * beta);
    sum_w += w;
    for (int k = 0; k < M; ++k) blend[k] = fma(w, rot[k], blend[k]);
  }
  const double inv = 1.0 / sum_w;
  double *dst = e.blend + ((size_t)comp * M) * kTile + tl;
  for (int k = 0; k < M; ++k) dst[(size_t)k * kTile] = blend[k] * inv;
}

// ReconstructOnFace, second half (weno.hpp:331-337): back to conservative variables, weighted by the
// neighbour's volume
template <int Q>
MCFD_HD void WenoEigenFaceAccumulate(const WenoEigenArgs &e, int tile, int tl, int comp, int face, const double *bas,
    int bs, EigenState<Q - 1> *st) {
  constexpr int P = Q - 1, M = ModalShape<P>::M, kBasRows = ModalShape<P>::kBasRows;
  const WenoArgs &p = e.w;
  const double s00 = bas[(size_t)ModalShape<P>::kSOffset * bs];
  double big_u[5], frame[9], row[5];
  for (int c = 0; c < 5; ++c) big_u[c] = p.mod[((size_t)tile * 5 * M + c * M) * kTile + tl] * s00;
  for (int i = 0; i < 9; ++i) frame[i] = e.frames[((size_t)tile * 54 + face * 9 + i) * kTile + tl];
  const EigenCommon ec = EigenPrepare(big_u, frame);
  EigenRowR(ec, frame, comp, row);
  const int nb = p.nbr[((size_t)tile * 6 + face) * kTile + tl];
  const double s00n = p.bas[((size_t)(nb >> 5) * kBasRows + ModalShape<P>::kSOffset) * kTile + (nb & 31)];
  const double adj_volume = 1.0 / (s00n * s00n);
  for (int k = 0; k < M; ++k) {
    double v = 0;
    for (int c = 0; c < 5; ++c) v = fma(row[c], e.blend[((size_t)c * M + k) * kTile + tl], v);
    st->total[k] = fma(v, adj_volume, st->total[k]);
  }
  st->total_volume += adj_volume;
}

// weno.hpp:342-348 + Extrapolation::SetCoeff: normalise, publish the new coefficients, resample the nodes;
// cells the detector left alone only pass their coefficients on
template <int Q>
MCFD_HD void WenoEigenFinish(const WenoEigenArgs &e, int tile, int tl, int comp, const double *quad, int qs,
    const double *bas, int bs, const EigenState<Q - 1> &st) {
  constexpr int P = Q - 1, N = Q * Q * Q, F = 5 * N, M = ModalShape<P>::M;
  const WenoArgs &p = e.w;
  const size_t row = ((size_t)tile * 5 * M + comp * M) * kTile + tl;
  if (!st.troubled) {
    for (int k = 0; k < M; ++k) {
      p.mod_new[row + (size_t)k * kTile] = p.mod[row + (size_t)k * kTile];
      p.tay_new[row + (size_t)k * kTile] = p.tay[row + (size_t)k * kTile];
    }
    return;
  }
  double fresh[M], t[M];
  const double inv = 1.0 / st.total_volume;
  for (int k = 0; k < M; ++k) fresh[k] = st.total[k] * inv;
  ModalToTaylor<M>(bas + (size_t)ModalShape<P>::kSOffset * bs, bs, fresh, t);
  for (int k = 0; k < M; ++k) { p.mod_new[row + (size_t)k * kTile] = fresh[k]; p.tay_new[row + (size_t)k * kTile] = t[k]; }
  double *dst = p.u + (size_t)tile * F * kTile + tl;
  for (int n = 0; n < N; ++n) {
    const double d[3] = {quad[(size_t)(n * 4) * qs], quad[(size_t)(n * 4 + 1) * qs], quad[(size_t)(n * 4 + 2) * qs]};
    double tv[M];
    TaylorValues<P>(d, tv);
    double v = 0;
    for (int k = 0; k < M; ++k) v = fma(t[k], tv[k], v);
    dst[(size_t)(n * 5 + comp) * kTile] = v;
  }
}

#ifdef __CUDACC__
// One CTA per tile, five warps: warp = component, lanes = cells.  Quadrature offsets/weights and the
// cell's basis are staged in shared memory by all warps first.
template <int Q>
struct WenoShape {
  static constexpr int P = Q - 1, N = Q * Q * Q;
  static constexpr size_t kSharedBytes = (size_t)(N * 4 + ModalShape<P>::kBasRows) * kTile * sizeof(double);
  // ModalKernel / WenoLazyKernel at p = 3 leave the basis table in global memory (L1/L2 serve it): 65 KB instead of
  // 121 KB per CTA, so that three CTAs share an SM instead of one
  static constexpr bool kBasInShared = MCFD_WENO_P3_BAS_SHARED || Q == 3;
  static constexpr size_t kLazySharedBytes = kBasInShared ? kSharedBytes : (size_t)(N * 4) * kTile * sizeof(double);
};
template <int Q>
__device__ __forceinline__ void WenoStage(const double *geom, const double *bas, int tile, double *sQuad, double *sBas,
    bool legendre = false) {
  constexpr int N = Q * Q * Q, kBasRows = ModalShape<Q - 1>::kBasRows;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int n = warp; n < N; n += 5) {
    double d[3], w;
    QuadraturePoint<Q>(n, geom + (size_t)tile * kGeomRows * kTile + lane, kTile, d, &w, legendre);
    sQuad[(n * 4 + 0) * kTile + lane] = d[0]; sQuad[(n * 4 + 1) * kTile + lane] = d[1];
    sQuad[(n * 4 + 2) * kTile + lane] = d[2]; sQuad[(n * 4 + 3) * kTile + lane] = w;
  }
  if (sBas) for (int i = threadIdx.x; i < kBasRows * kTile; i += 160) sBas[i] = bas[(size_t)tile * kBasRows * kTile + i];
  __syncthreads();
}
template <int Q>
__global__ void __launch_bounds__(160, (Q == 3 ? MCFD_WENO_MINB : MCFD_WENO_P3_MINB)) ModalKernel(const double *u, const double *geom, const double *bas,
    double *mod, double *tay, int tile_begin, int cell_begin, int cell_end) {
  constexpr int P = Q - 1, N = Q * Q * Q, F = 5 * N, M = ModalShape<P>::M;
  extern __shared__ double smem[];
  const int tile = tile_begin + blockIdx.x, lane = threadIdx.x & 31, comp = threadIdx.x >> 5;
  double *sQuad = smem, *sBas = WenoShape<Q>::kBasInShared ? smem + N * 4 * kTile : nullptr;
  WenoStage<Q>(geom, bas, tile, sQuad, sBas);
  const double *tBas = sBas ? sBas : bas + (size_t)tile * ModalShape<P>::kBasRows * kTile;
  const int cell = tile * kTile + lane;
  if (cell < cell_begin || cell >= cell_end) return;
  ModalProjectCell<Q>(u + (size_t)tile * F * kTile + comp * kTile + lane, kTile, sQuad + lane, kTile, tBas + lane, kTile,
      mod + ((size_t)tile * 5 * M + comp * M) * kTile + lane, tay + ((size_t)tile * 5 * M + comp * M) * kTile + lane, kTile);
}
template <int Q>
__global__ void __launch_bounds__(160, (Q == 3 ? MCFD_WENO_MINB : MCFD_WENO_P3_MINB)) WenoLazyKernel(WenoArgs p, int tile_begin) {
  constexpr int N = Q * Q * Q;
  extern __shared__ double smem[];
  const int tile = tile_begin + blockIdx.x, lane = threadIdx.x & 31, comp = threadIdx.x >> 5;
  double *sQuad = smem, *sBas = WenoShape<Q>::kBasInShared ? smem + N * 4 * kTile : nullptr;
  WenoStage<Q>(p.geom, p.bas, tile, sQuad, sBas, p.legendre != 0);
  const double *tBas = sBas ? sBas : p.bas + (size_t)tile * ModalShape<Q - 1>::kBasRows * kTile;
  const int cell = tile * kTile + lane;
  if (cell < p.cell_begin || cell >= p.cell_end) return;
  WenoLazyCell<Q>(p, tile, lane, comp, sQuad + lane, kTile, tBas + lane, kTile);
}
// weno::Eigen: persistent CTAs (each owns one scratch slot) walking over the tiles of the batch
template <int Q>
struct WenoEigenShape {
  static constexpr int M = ModalShape<Q - 1>::M;
  static constexpr size_t kScratchDoubles = (size_t)(7 * 5 * M + 5 * M) * kTile;   // per CTA
};
template <int Q>
__global__ void __launch_bounds__(160) WenoEigenKernel(WenoEigenArgs e, double *scratch, int tile_begin, int tile_end) {
  constexpr int P = Q - 1, N = Q * Q * Q, M = ModalShape<P>::M;
  extern __shared__ double smem[];
  double *sQuad = smem, *sBas = smem + N * 4 * kTile;
  const int lane = threadIdx.x & 31, comp = threadIdx.x >> 5;
  e.cand = scratch + (size_t)blockIdx.x * WenoEigenShape<Q>::kScratchDoubles;
  e.blend = e.cand + (size_t)7 * 5 * M * kTile;
  for (int tile = tile_begin + blockIdx.x; tile < tile_end; tile += gridDim.x) {
    __syncthreads();                       // the previous tile's shared tables are no longer read
    WenoStage<Q>(e.w.geom, e.w.bas, tile, sQuad, sBas);
    const int cell = tile * kTile + lane;
    const bool active = cell >= e.w.cell_begin && cell < e.w.cell_end;
    EigenState<P> st;
    st.total_volume = 0; st.adj_cnt = 0; st.troubled = false;
    for (int k = 0; k < M; ++k) st.total[k] = 0;
    if (active) st.troubled = WenoEigenDetect<Q>(e.w, tile, lane, sQuad + lane, kTile, sBas + lane, kTile, &st.adj_cnt);
    if (active && st.troubled) WenoEigenBorrow<Q>(e, tile, lane, comp, sQuad + lane, kTile, sBas + lane, kTile);
    __syncthreads();
    for (int f = 0; f < 6; ++f) {
      const bool work = active && st.troubled && e.w.nbr[((size_t)tile * 6 + f) * kTile + lane] >= 0;
      if (work) WenoEigenFaceBlend<Q>(e, tile, lane, comp, f, st.adj_cnt, sQuad + lane, kTile, sBas + lane, kTile);
      __syncthreads();
      if (work) WenoEigenFaceAccumulate<Q>(e, tile, lane, comp, f, sBas + lane, kTile, &st);
      __syncthreads();
    }
    if (active) WenoEigenFinish<Q>(e, tile, lane, comp, sQuad + lane, kTile, sBas + lane, kTile, st);
  }
}
// EigenMatrices of n states (test hook for SURVEY row a28): L and R row-major [n][5][5]
__global__ void EigenMatricesKernel(int n, const double *cons, const double *frames, double *L, double *R) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const EigenCommon ec = EigenPrepare(cons + i * 5, frames + i * 9);
  for (int k = 0; k < 5; ++k) {
    EigenRowL(ec, frames + i * 9, k, L + (size_t)i * 25 + k * 5);
    EigenRowR(ec, frames + i * 9, k, R + (size_t)i * 25 + k * 5);
  }
}
// new coefficients of the cells of a finished batch -> the tables the next batch borrows from
__global__ void CopyCellRowsKernel(const double *src, double *dst, int rows, int cell_begin, int cell_end) {
  const int cell = cell_begin + blockIdx.x * blockDim.x + threadIdx.x;
  if (cell >= cell_end) return;
  const size_t base = (size_t)(cell >> 5) * rows * kTile + (cell & 31);
  for (int r = 0; r < rows; ++r) dst[base + (size_t)r * kTile] = src[base + (size_t)r * kTile];
}
#endif

// ---- EnergyBasedViscosity::UpdateProperties (spatial/viscosity.hpp:556-572) --------------------------------
// What the five components of a cell share, computed once per cell (rows `ss` doubles apart):
//   [0, N)            quadrature weight of node n (local weight x det J): their sum is the volume
//   [N, 2N)           lambda_max at node n (rotated/euler.hpp:340-355)
//   [2N, 2N + 6 FQ)   weight of face point (f, q): w_a w_b |J* column| from the cell's own metric
// `part` of `n_parts` callers split the items among themselves (the kernel's five warps; one caller on the host).
template <int Q>
struct ViscosityShape {
  static constexpr int N = Q * Q * Q, FQ = Q * Q, kRows = 2 * N + 6 * FQ;
  static constexpr size_t kSharedBytes = (size_t)kRows * kTile * sizeof(double) + 6 * 8 * FQ;
};
template <int Q>
MCFD_HD void ViscosityShared(const double *u, const double *geom, int tile, int tl, int part, int n_parts,
    double *sh, int ss) {
  constexpr int N = Q * Q * Q, F = 5 * N, FQ = Q * Q;
  const double *mine = u + (size_t)tile * F * kTile + tl;
  const double *geo = geom + (size_t)tile * kGeomRows * kTile + tl;
  for (int n = part; n < N; n += n_parts) {
    double d[3], w;
    QuadraturePoint<Q>(n, geo, kTile, d, &w);
    sh[(size_t)n * ss] = w;
    const double *s = mine + (size_t)n * 5 * kTile;
    const Prim pr = ConsToPrim(s[0], s[kTile], s[2 * kTile], s[3 * kTile], s[4 * kTile]);
    const double cs = SoundSpeed(pr.rho, pr.p), vel = Sqrt(pr.u * pr.u + pr.v * pr.v + pr.w * pr.w + 1e-300);
    sh[(size_t)(N + n) * ss] = fmax(cs + vel, cs - vel);
  }
  for (int pt = part; pt < 6 * FQ; pt += n_parts) {
    const int f = pt / FQ, q = pt % FQ;
    const int dir = (f == 2 || f == 4) ? 0 : ((f == 1 || f == 3) ? 1 : 2);
    const int n = c_tab.face_node[f][q];
    const int i = n / FQ, j = (n / Q) % Q, k = n % Q;
    double nrm[3];
    ComputeNormal(geo, kTile, c_tab.xi[i], c_tab.xi[j], c_tab.xi[k], dir, nrm);
    const int ia = dir == 0 ? j : i, ib = dir == 2 ? j : k;     // the two in-face indices
    sh[(size_t)(2 * N + pt) * ss] = c_tab.w[ia] * c_tab.w[ib] * Sqrt(nrm[0] * nrm[0] + nrm[1] * nrm[1] + nrm[2] * nrm[2]);
  }
}

// One (cell, component): the jump integral over the cell's interior faces (IntegrateJumpOnFaces, :446-475; the
// height is volume / area), the reference square (rotated/euler.hpp:340-355), the damping rate -u^T D u and the
// clamped viscosity (:477-544).  Returns this thread's contribution to min_dt_.
// D: [tile][r*N + c][lane], nu out: [tile][comp][lane]; sh: ViscosityShared's rows of this cell.
template <int Q>
MCFD_HD double ViscosityCell(const double *u, const int *nbr, const unsigned char *code,
    const unsigned char *nbr_node, const double *damping, const double *length, int tile, int tl, int comp,
    double time_scale, const double *sh, int ss, double *nu_out) {
  constexpr int P = Q - 1, N = Q * Q * Q, F = 5 * N, FQ = Q * Q;
  const double *mine = u + (size_t)tile * F * kTile + tl;
  double volume = 0;
  for (int n = 0; n < N; ++n) volume += sh[(size_t)n * ss];
  double jump = 0;
  for (int f = 0; f < 6; ++f) {
    const int nb = nbr[((size_t)tile * 6 + f) * kTile + tl];
    if (nb < 0) continue;
    const int cd = code[((size_t)tile * 6 + f) * kTile + tl];
    double area = 0, sum = 0;
    for (int q = 0; q < FQ; ++q) {
      const int n = c_tab.face_node[f][q];
      const double w = sh[(size_t)(2 * N + f * FQ + q) * ss];
      const int n2 = nbr_node[((cd & 7) * 8 + ((cd >> 3) & 7)) * FQ + q];
      const double v = u[((size_t)(nb >> 5) * F + n2 * 5 + comp) * kTile + (nb & 31)];
      const double dj = mine[(size_t)(n * 5 + comp) * kTile] - v;
      area += w;
      sum = fma(dj * dj, w, sum);
    }
    jump = fma(sum, volume / area, jump);
  }
  double max_speed = 1e-5, ref = 1e-5;
  for (int n = 0; n < N; ++n) {
    const double lam = sh[(size_t)(N + n) * ss];
    max_speed = fmax(max_speed, lam);
    const double r0 = mine[(size_t)n * 5 * kTile];
    double sq = r0 * r0;
    if (comp == 4) { const double r4 = mine[(size_t)(n * 5 + 4) * kTile]; sq = r4 * r4; }
    else if (comp != 0) sq = r0 * r0 * lam * lam;
    ref = fmax(ref, sq);
  }
  const double len = length[(size_t)tile * kTile + tl];
  double min_dt = len / max_speed;
  const double max_visc = max_speed * len / P, time_base = len / max_speed;
  double scale = len * len * len * len * len * len;            // h^(P+1) * h^3, P = 2
  if (P == 3) scale *= len;
  double nu = 0;
  if (!(jump < ref * scale)) {
    const double *D = damping + (size_t)tile * N * N * kTile + tl;
    double rate = 0;
    for (int r = 0; r < N; ++r) {
      double du = 0;
      for (int c = 0; c < N; ++c) du = fma(D[(size_t)(r * N + c) * kTile], mine[(size_t)(c * 5 + comp) * kTile], du);
      rate = fma(mine[(size_t)(r * 5 + comp) * kTile], du, rate);
    }
    rate = -rate;
    nu = fmin(max_visc, fmax(0.0, jump / (rate * (time_base * time_scale))));
    const double factor = P == 2 ? 10.0 : 40.0;                // GetSpatialFactor, viscosity.hpp:579-600
    min_dt = fmin(min_dt, len / (max_speed + factor * nu / len));
  }
  nu_out[((size_t)tile * 5 + comp) * kTile + tl] = nu;
  return min_dt;
}

#ifdef __CUDACC__
// thread = (cell, component): lanes = cells of the tile, warp = component; the warps first split the per-cell
// quantities all components share (metric, wave speeds, face weights) and hand them over in shared memory
template <int Q>
__global__ void __launch_bounds__(160) ViscosityKernel(const double *u, const double *geom, const int *nbr,
    const unsigned char *code, const double *damping, const double *length, int n_local, double time_scale,
    double *nu_out, unsigned long long *min_dt_bits) {
  extern __shared__ double smem[];
  constexpr int kRows = ViscosityShape<Q>::kRows;
  double *sh = smem;
  unsigned char *sNbrNode = reinterpret_cast<unsigned char *>(smem + (size_t)kRows * kTile);
  for (int i = threadIdx.x; i < 6 * 8 * Q * Q; i += 160) sNbrNode[i] = c_tab.nbr_node[i / (8 * Q * Q)][(i / (Q * Q)) % 8][i % (Q * Q)];
  const int tile = blockIdx.x, lane = threadIdx.x & 31, comp = threadIdx.x >> 5;
  const bool live = tile * kTile + lane < n_local;
  if (live) ViscosityShared<Q>(u, geom, tile, lane, comp, 5, sh + lane, kTile);
  __syncthreads();
  double dt = 1e100;
  if (live)
    dt = ViscosityCell<Q>(u, nbr, code, sNbrNode, damping, length, tile, lane, comp, time_scale, sh + lane, kTile, nu_out);
  for (int o = 16; o > 0; o >>= 1) dt = fmin(dt, __shfl_xor_sync(0xffffffffu, dt, o));
  if (lane == 0) atomicMin(min_dt_bits, (unsigned long long)__double_as_longlong(dt));
}
// column `col` of every cell's damping matrix from the probe residual (component 0), scaled by the row's
// quadrature weight (viscosity.hpp:335-337, 371-382)
template <int Q>
__global__ void DampingColumnKernel(const double *res, const double *geom, double *damping, int col, int n_local) {
  constexpr int N = Q * Q * Q, F = 5 * N;
  const int cell = blockIdx.x * blockDim.x + threadIdx.x;
  if (cell >= n_local) return;
  const int tile = cell >> 5, tl = cell & 31;
  for (int r = 0; r < N; ++r) {
    double d[3], w;
    QuadraturePoint<Q>(r, geom + (size_t)tile * kGeomRows * kTile + tl, kTile, d, &w);
    damping[((size_t)tile * N * N + r * N + col) * kTile + tl] = res[((size_t)tile * F + r * 5) * kTile + tl] * w;
  }
}
// every cell's nodal values := 1 at node `col`, 0 elsewhere, in all components (viscosity.hpp:341-346)
__global__ void ProbeStateKernel(double *u, int rows_per_tile, int n_tiles, int col) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)n_tiles * rows_per_tile * kTile) return;
  const int row = (int)((i / kTile) % rows_per_tile);
  u[i] = (row / 5 == col) ? 1.0 : 0.0;
}
#endif

// ---- dg::General on the modal Projection polynomial --------------------------------------------
// spatial/dg/general.hpp:58-200 with polynomial/projection.hpp:67-176 on Gauss-Legendre rules (demo/euler/
// sourceless.hpp:28-33, 64-66).  State: [tile][mode*5 + comp][lane], M modes per component on the cell's orthonormal
// basis phi = S taylor(x - centre) (BAS, built with the Legendre rule).  One thread per cell: values come from the
// Taylor form t = S^T a (TAY, written for every cell, ghosts included, by DgmTaylorCell), the residual is gathered
// as Taylor moments G_j = sum_q (...) taylor_j / grad taylor_j and turned into modal coefficients by S once.
struct DgmArgs {
  const double *u_in, *u0;      // modal coefficients (tiles)
  double *u_out;
  const double *tay;            // Taylor coefficients of u_in for every cell
  const double *geom, *bas, *h2s;
  const int *nbr;
  const unsigned char *code, *bkind;
  const double *bvals;          // [n_bfaces][Q*Q][5] boundary function at the face's Gauss points
  int cell_begin, cell_end, mode;
  double dt;
  const double *dt_dev;
};

template <int P>
MCFD_HD void TaylorGradients(const double *d, double (*tg)[3]) {   // basis/taylor.hpp:374-470, per monomial
  constexpr int M = ModalShape<P>::M;
  const double x = d[0], y = d[1], z = d[2];
  for (int j = 0; j < M; ++j) { tg[j][0] = 0; tg[j][1] = 0; tg[j][2] = 0; }
  tg[1][0] = 1; tg[2][1] = 1; tg[3][2] = 1;
  tg[4][0] = 2 * x; tg[5][0] = y; tg[5][1] = x; tg[6][0] = z; tg[6][2] = x;
  tg[7][1] = 2 * y; tg[8][1] = z; tg[8][2] = y; tg[9][2] = 2 * z;
  if (P == 3) {
    const double xx = x * x, xy = x * y, xz = x * z, yy = y * y, yz = y * z, zz = z * z;
    tg[10][0] = 3 * xx;
    tg[11][0] = 2 * xy; tg[11][1] = xx;
    tg[12][0] = 2 * xz; tg[12][2] = xx;
    tg[13][0] = yy; tg[13][1] = 2 * xy;
    tg[14][0] = yz; tg[14][1] = xz; tg[14][2] = xy;
    tg[15][0] = zz; tg[15][2] = 2 * xz;
    tg[16][1] = 3 * yy;
    tg[17][1] = 2 * yz; tg[17][2] = yy;
    tg[18][1] = zz; tg[18][2] = 2 * yz;
    tg[19][2] = 3 * zz;
  }
}

// offset from the cell centre of the point with local coordinates (x, e, z), and the metric there
MCFD_HD Metric DgmPoint(const double *geo, int gs, double x, double e, double z, double *d) {
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    d[a] = geo[(0 + a) * gs] * x + geo[(3 + a) * gs] * e + geo[(6 + a) * gs] * z + geo[(9 + a) * gs] * (x * e)
         + geo[(12 + a) * gs] * (e * z) + geo[(15 + a) * gs] * (z * x) + geo[(18 + a) * gs] * (x * e * z);
  }
  return ComputeMetric(geo, gs, x, e, z);
}

// t = S^T a for the five components of one cell (Projection::GetCoeffOnTaylorBasis, projection.hpp:117-119)
template <int Q>
MCFD_HD void DgmTaylorCell(const double *u, const double *bas, double *tay, int tile, int tl) {
  constexpr int P = Q - 1, M = ModalShape<P>::M, kBasRows = ModalShape<P>::kBasRows;
  const double *S = bas + ((size_t)tile * kBasRows + ModalShape<P>::kSOffset) * kTile + tl;
  for (int k = 0; k < 5; ++k) {
    double a[M], t[M];
    for (int i = 0; i < M; ++i) a[i] = u[((size_t)tile * 5 * M + i * 5 + k) * kTile + tl];
    ModalToTaylor<M>(S, kTile, a, t);
    for (int i = 0; i < M; ++i) tay[((size_t)tile * 5 * M + i * 5 + k) * kTile + tl] = t[i];
  }
}

template <int P>
MCFD_HD void DgmValue(const double *t, size_t stride, const double *d, double *val) {   // sum_j t_j taylor_j(d)
  constexpr int M = ModalShape<P>::M;
  double tv[M];
  TaylorValues<P>(d, tv);
#pragma unroll
  for (int k = 0; k < 5; ++k) val[k] = 0;
  for (int j = 0; j < M; ++j) {
#pragma unroll
    for (int k = 0; k < 5; ++k) val[k] = fma(t[(size_t)(j * 5 + k) * stride], tv[j], val[k]);
  }
}

// One cell: residual (dg/general.hpp:58-89 + the boundary loops) and RK update of its modal coefficients.
template <int Q, int kRiemann>
MCFD_HD void DgmStageCell(const DgmArgs &p, int tile, int tl) {
  constexpr int P = Q - 1, M = ModalShape<P>::M, kBasRows = ModalShape<P>::kBasRows, FQ = Q * Q;
  const double *geo = p.geom + (size_t)tile * kGeomRows * kTile + tl;
  const double *mine = p.tay + (size_t)tile * 5 * M * kTile + tl;
  double G[M][5];
  for (int j = 0; j < M; ++j) for (int k = 0; k < 5; ++k) G[j][k] = 0;
  for (int i = 0; i < Q; ++i) for (int j2 = 0; j2 < Q; ++j2) for (int k2 = 0; k2 < Q; ++k2) {   // volume
    double d[3], u[5];
    const Metric mt = DgmPoint(geo, kTile, c_tab.gx[i], c_tab.gx[j2], c_tab.gx[k2], d);
    const double w = c_tab.gw[i] * c_tab.gw[j2] * c_tab.gw[k2] * mt.det;
    DgmValue<P>(mine, kTile, d, u);
    const Prim q = ConsToPrim(u[0], u[1], u[2], u[3], u[4]);
    const double h0 = u[4] + q.p, vel[3] = {q.u, q.v, q.w};
    double fx[3][5];
#pragma unroll
    for (int x = 0; x < 3; ++x) {
      fx[x][0] = u[1 + x];
      fx[x][1] = u[1] * vel[x]; fx[x][2] = u[2] * vel[x]; fx[x][3] = u[3] * vel[x];
      fx[x][1 + x] += q.p;
      fx[x][4] = h0 * vel[x];
#pragma unroll
      for (int c = 0; c < 5; ++c) fx[x][c] *= w;
    }
    double tg[M][3];
    TaylorGradients<P>(d, tg);
    for (int j = 1; j < M; ++j) {
#pragma unroll
      for (int c = 0; c < 5; ++c) G[j][c] += fx[0][c] * tg[j][0] + fx[1][c] * tg[j][1] + fx[2][c] * tg[j][2];
    }
  }
  for (int f = 0; f < 6; ++f) {   // faces
    const int nb = p.nbr[((size_t)tile * 6 + f) * kTile + tl];
    const int cd = p.code[((size_t)tile * 6 + f) * kTile + tl];
    const int dir = (f == 2 || f == 4) ? 0 : ((f == 1 || f == 3) ? 1 : 2);
    const double sgn = (f == 2 || f == 3 || f == 5) ? 1.0 : -1.0;
    const bool holder = (cd & 64) != 0;
    double off[3] = {0, 0, 0};
    const double *theirs = mine;
    if (nb >= 0) {
      const double *hv = p.h2s + ((size_t)tile * 18 + f * 3) * kTile + tl;
      const double sg = holder ? 1.0 : -1.0;
      off[0] = sg * hv[0]; off[1] = sg * hv[kTile]; off[2] = sg * hv[2 * kTile];
      theirs = p.tay + (size_t)(nb >> 5) * 5 * M * kTile + (nb & 31);
    }
    for (int ia = 0; ia < Q; ++ia) for (int ib = 0; ib < Q; ++ib) {
      // local coordinates: +-1 in `dir`, the Gauss points in the other two (in increasing axis order)
      const double a = c_tab.gx[ia], b = c_tab.gx[ib];
      const double x = dir == 0 ? sgn : a, e = dir == 1 ? sgn : (dir == 0 ? a : b), z = dir == 2 ? sgn : b;
      double d[3], u[5], fl[5];
      const Metric mt = DgmPoint(geo, kTile, x, e, z, d);
      const double nx = mt.s[dir][0], ny = mt.s[dir][1], nz = mt.s[dir][2];
      const double n2 = nx * nx + ny * ny + nz * nz;
      const double inv_area = Rsqrt(n2), area = n2 * inv_area;
      const double w = c_tab.gw[ia] * c_tab.gw[ib] * area;
      DgmValue<P>(mine, kTile, d, u);
      double out_sign = 1.0;       // flux along MY outward normal = out_sign * (flux in the holder's frame)
      if (nb >= 0) {
        const double dn[3] = {d[0] - off[0], d[1] - off[1], d[2] - off[2]};
        double v[5], ul[5], ur[5];
        DgmValue<P>(theirs, kTile, dn, v);
        const double hs = (holder ? sgn : -sgn) * inv_area;
#pragma unroll
        for (int c = 0; c < 5; ++c) { ul[c] = holder ? u[c] : v[c]; ur[c] = holder ? v[c] : u[c]; }
        FluxUpwind<kRiemann>(ul, ur, hs * nx, hs * ny, hs * nz, fl);
        if (!holder) out_sign = -1.0;
      } else {
        const int bf = -1 - nb;
        const int kind = p.bkind[bf];
        if (kind == kBcNoSlip) continue;     // dg::General does not override AddFluxOnNoSlipWalls (fem.hpp:382-383)
        const double inv = sgn * inv_area;
        State5 ui, given;
        const double *gv = p.bvals + ((size_t)bf * FQ + ia * Q + ib) * 5;
#pragma unroll
        for (int c = 0; c < 5; ++c) { ui.v[c] = u[c]; given.v[c] = gv[c]; }
        const State5 r = BoundaryFluxOutOfLine<kRiemann>(kind, ui, given, inv * nx, inv * ny, inv * nz);
#pragma unroll
        for (int c = 0; c < 5; ++c) fl[c] = r.v[c];
      }
      double tv[M];
      TaylorValues<P>(d, tv);
      const double ws = -w * out_sign;
      for (int j = 0; j < M; ++j) {
#pragma unroll
        for (int c = 0; c < 5; ++c) G[j][c] = fma(ws * fl[c], tv[j], G[j][c]);
      }
    }
  }
  // residual on the orthonormal basis: R_m = sum_{j <= m} S[m][j] G_j; then the RK update (rk.hpp:16-49)
  const double *S = p.bas + ((size_t)tile * kBasRows + ModalShape<P>::kSOffset) * kTile + tl;
  const double dt = p.dt_dev ? *p.dt_dev : p.dt;
  int row = 0;
  for (int m = 0; m < M; ++m) {
    double r[5] = {0, 0, 0, 0, 0};
    for (int j = 0; j <= m; ++j) {
      const double s = S[(size_t)(row + j) * kTile];
#pragma unroll
      for (int c = 0; c < 5; ++c) r[c] = fma(s, G[j][c], r[c]);
    }
    row += m + 1;
#pragma unroll
    for (int c = 0; c < 5; ++c) {
      const size_t at = ((size_t)tile * 5 * M + m * 5 + c) * kTile + tl;
      const double u = p.u_in[at], u0 = p.mode >= kModeRk3b ? p.u0[at] : u, res = r[c];
      double val;
      switch (p.mode) {
      case kModeResidual: val = res; break;
      case kModeEuler: val = u + res * dt; break;
      case kModeRk3b: val = (u0 * 3 + (u + res * dt)) * 0.25; break;
      case kModeRk3c: val = ((u + res * dt) * 2 + u0) / 3; break;
      default: val = ((u + res * dt) + u0) * 0.5; break;
      }
      p.u_out[at] = val;
    }
  }
}

// FiniteElement::GetTimeStep on a modal cell: Cell::N = M, so the first M Gauss points are sampled (fem.hpp:419-426)
template <int Q>
MCFD_HD double DgmCellSpeed(const double *tay, const double *geom, int tile, int tl) {
  constexpr int P = Q - 1, M = ModalShape<P>::M;
  const double *geo = geom + (size_t)tile * kGeomRows * kTile + tl;
  const double *mine = tay + (size_t)tile * 5 * M * kTile + tl;
  double vmax = 1e-5;
  for (int n = 0; n < M; ++n) {
    const int i = n / (Q * Q), j = (n / Q) % Q, k = n % Q;
    double d[3], u[5];
    DgmPoint(geo, kTile, c_tab.gx[i], c_tab.gx[j], c_tab.gx[k], d);
    DgmValue<P>(mine, kTile, d, u);
    const Prim q = ConsToPrim(u[0], u[1], u[2], u[3], u[4]);
    const double c = SoundSpeed(q.rho, q.p), v = sqrt(q.u * q.u + q.v * q.v + q.w * q.w);
    vmax = fmax(vmax, fmax(c + v, c - v));
  }
  return vmax;
}

// Projection::Approximate (projection.hpp:28-43) from the function's values at the cell's Gauss points,
// values [cell][Q^3][5] in the reference's column order; writes the modal coefficients of one cell
template <int Q>
MCFD_HD void DgmProjectCell(const double *values, const double *geom, const double *bas, double *u, int cell) {
  constexpr int P = Q - 1, M = ModalShape<P>::M, N = Q * Q * Q, kBasRows = ModalShape<P>::kBasRows;
  const int tile = cell >> 5, tl = cell & 31;
  const double *geo = geom + (size_t)tile * kGeomRows * kTile + tl;
  const double *S = bas + ((size_t)tile * kBasRows + ModalShape<P>::kSOffset) * kTile + tl;
  for (int k = 0; k < 5; ++k) {
    double g[M], a[M];
    for (int j = 0; j < M; ++j) g[j] = 0;
    for (int n = 0; n < N; ++n) {
      const int i = n / (Q * Q), j2 = (n / Q) % Q, k2 = n % Q;
      double d[3], tv[M];
      const Metric mt = DgmPoint(geo, kTile, c_tab.gx[i], c_tab.gx[j2], c_tab.gx[k2], d);
      const double wf = c_tab.gw[i] * c_tab.gw[j2] * c_tab.gw[k2] * mt.det * values[((size_t)cell * N + n) * 5 + k];
      TaylorValues<P>(d, tv);
      for (int j = 0; j < M; ++j) g[j] = fma(tv[j], wf, g[j]);
    }
    MomentsToModal<M>(S, kTile, g, a);
    for (int j = 0; j < M; ++j) u[((size_t)tile * 5 * M + j * 5 + k) * kTile + tl] = a[j];
  }
}

#ifdef __CUDACC__
template <int Q, int kRiemann>
__global__ void __launch_bounds__(kTile, MCFD_DGM_MINB) DgmStageKernel(DgmArgs p, int tile_begin) {
  const int tile = tile_begin + blockIdx.x, cell = tile * kTile + threadIdx.x;
  if (cell < p.cell_begin || cell >= p.cell_end) return;
  DgmStageCell<Q, kRiemann>(p, tile, threadIdx.x);
}
template <int Q>
__global__ void __launch_bounds__(kTile) DgmTaylorKernel(const double *u, const double *bas, double *tay, int tile_begin,
    int cell_begin, int cell_end) {
  const int tile = tile_begin + blockIdx.x, cell = tile * kTile + threadIdx.x;
  if (cell < cell_begin || cell >= cell_end) return;
  DgmTaylorCell<Q>(u, bas, tay, tile, threadIdx.x);
}
template <int Q>
__global__ void __launch_bounds__(kTile) DgmTimeStepKernel(const double *tay, const double *geom, const double *length,
    int n_local, unsigned long long *result) {
  const int tile = blockIdx.x, cell = tile * kTile + threadIdx.x;
  double dt = 1e100;
  if (cell < n_local) dt = length[cell] / DgmCellSpeed<Q>(tay, geom, tile, threadIdx.x);
  for (int o = 16; o > 0; o >>= 1) dt = fmin(dt, __shfl_xor_sync(0xffffffffu, dt, o));
  if (threadIdx.x == 0) atomicMin(result, (unsigned long long)__double_as_longlong(dt));
}
template <int Q>
__global__ void DgmProjectKernel(const double *values, const double *geom, const double *bas, double *u, int n_local) {
  const int cell = blockIdx.x * blockDim.x + threadIdx.x;
  if (cell < n_local) DgmProjectCell<Q>(values, geom, bas, u, cell);
}
#endif

// Traces of cells [cell_begin, cell_end) (local and ghost alike): one CTA per tile, the tile's state and
// geometry staged in shared memory, lanes = cells, nodes dealt to the warps.
template <int Q>
struct TraceShape {
  static constexpr int N = Q * Q * Q, F = 5 * N;
#ifndef MCFD_TRACE_WARPS
#define MCFD_TRACE_WARPS 4
#endif
  static constexpr int NW = MCFD_TRACE_WARPS;     // small CTAs, several per SM: one stages while another computes
  static constexpr size_t kSharedBytes = (size_t)(F + kGeomRows) * kTile * sizeof(double);
};
#ifdef __CUDACC__
template <int Q, bool kLocal = false>
__global__ void __launch_bounds__(TraceShape<Q>::NW * 32, MCFD_TRACE_MINB) TraceKernel(const double *u,
    const double *geom, double *grad, double *hn, int tile_begin, int cell_begin, int cell_end) {
  using S = TraceShape<Q>;
  constexpr int T = S::NW * 32;
  extern __shared__ double smem[];
  double *sU = smem, *sG = smem + S::F * kTile;
  const int tile = tile_begin + blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  {  // stage the tile: batches of independent 16-byte loads, then the stores
    const double2 *src = reinterpret_cast<const double2 *>(u + (size_t)tile * S::F * kTile);
    double2 *dst = reinterpret_cast<double2 *>(sU);
    constexpr int kVec = S::F * kTile / 2, kBatch = 6;
    for (int base = threadIdx.x; base < kVec; base += T * kBatch) {
      double2 r[kBatch];
#pragma unroll
      for (int j = 0; j < kBatch; ++j) if (base + j * T < kVec) r[j] = src[base + j * T];
#pragma unroll
      for (int j = 0; j < kBatch; ++j) if (base + j * T < kVec) dst[base + j * T] = r[j];
    }
    for (int i = threadIdx.x; i < kGeomRows * kTile; i += T) sG[i] = geom[(size_t)tile * kGeomRows * kTile + i];
  }
  __syncthreads();
  if (kLocal) {   // the interpolated quantity is U = det(J) u (the state holds u): scale the staged tile
    for (int n = warp; n < S::N; n += S::NW) {
      const Metric mt = ComputeMetric(sG + lane, kTile, c_tab.xi[n / (Q * Q)], c_tab.xi[(n / Q) % Q], c_tab.xi[n % Q]);
#pragma unroll
      for (int c = 0; c < 5; ++c) sU[(n * 5 + c) * kTile + lane] *= mt.det;
    }
    __syncthreads();
  }
  const int cell = tile * kTile + lane;
  if (cell < cell_begin || cell >= cell_end) return;
  double *gdst = grad + (size_t)tile * S::N * 15 * kTile + lane;
  double *hdst = hn + (size_t)tile * 6 * Q * Q * 15 * kTile + lane;
  for (int n = warp; n < S::N; n += S::NW)
    TraceCellNode<Q, kLocal>(n, sU + lane, kTile, sG + lane, kTile, gdst, hdst, kTile);
}
// kLocal == true contexts keep u on the device; columns hold U = det(J) u (polynomial/hexahedron.hpp:338-343).
// In-place conversion of a tile range (invert: U -> u) or of a column [cell][node][comp] (u -> U, or residuals).
template <int Q>
__global__ void ScaleTilesKernel(double *tiles, const double *geom, int cell_begin, int cell_end, int invert) {
  constexpr int N = Q * Q * Q, F = 5 * N;
  const int cell = cell_begin + blockIdx.x * blockDim.x + threadIdx.x;
  if (cell >= cell_end) return;
  const int tile = cell >> 5, tl = cell & 31;
  for (int n = 0; n < N; ++n) {
    const Metric mt = ComputeMetric(geom + (size_t)tile * kGeomRows * kTile + tl, kTile, c_tab.xi[n / (Q * Q)],
        c_tab.xi[(n / Q) % Q], c_tab.xi[n % Q]);
    const double s = invert ? 1.0 / mt.det : mt.det;
    for (int c = 0; c < 5; ++c) tiles[((size_t)tile * F + n * 5 + c) * kTile + tl] *= s;
  }
}
template <int Q>
__global__ void ScaleColumnKernel(double *column, const double *geom, int n_cells) {
  constexpr int N = Q * Q * Q;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;     // (cell, node)
  if (i >= n_cells * N) return;
  const int cell = i / N, n = i % N;
  const Metric mt = ComputeMetric(geom + (size_t)(cell >> 5) * kGeomRows * kTile + (cell & 31), kTile, c_tab.xi[n / (Q * Q)],
      c_tab.xi[(n / Q) % Q], c_tab.xi[n % Q]);
  for (int c = 0; c < 5; ++c) column[(size_t)i * 5 + c] *= mt.det;
}
#else
template <int Q, bool kLocal = false>
void RunTraceOnHost(const double *u, const double *geom, double *grad, double *hn, int n_cells) {
  constexpr int N = Q * Q * Q, F = 5 * N;
  static double big[F];   // U = det u of the current cell when kLocal
  for (int cell = 0; cell < n_cells; ++cell) {
    const int tile = cell >> 5, lane = cell & 31;
    const double *src = u + (size_t)tile * F * kTile + lane;
    const double *geo = geom + (size_t)tile * kGeomRows * kTile + lane;
    int us = kTile;
    if (kLocal) {
      for (int n = 0; n < N; ++n) {
        const Metric mt = ComputeMetric(geo, kTile, c_tab.xi[n / (Q * Q)], c_tab.xi[(n / Q) % Q], c_tab.xi[n % Q]);
        for (int c = 0; c < 5; ++c) big[n * 5 + c] = src[(size_t)(n * 5 + c) * kTile] * mt.det;
      }
      src = big; us = 1;
    }
    for (int n = 0; n < N; ++n)
      TraceCellNode<Q, kLocal>(n, src, us, geo, kTile, grad + (size_t)tile * N * 15 * kTile + lane,
          hn + (size_t)tile * 6 * Q * Q * 15 * kTile + lane, kTile);
  }
}
#endif

#ifdef __CUDACC__
// FiniteElement::GetTimeStep's reduction (fem.hpp:417-432; rotated/euler.hpp:340-345): min over cells of
// length / max(1e-5, max_nodes(c + |v|)).  One CTA per tile, lanes = cells, the nodes dealt to the warps
// (whole 256-byte rows per load); per-cell maxima meet in shared memory, the CTA's minimum is folded
// into *result by atomicMin on the bit pattern (valid for non-negative doubles).
template <int Q>
__global__ void __launch_bounds__(256) TimeStepKernel(const double *u, const double *length, int n_local,
    unsigned long long *result) {
  constexpr int N = Q * Q * Q, F = 5 * N, NW = 8;
  __shared__ double s_max[NW][kTile];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cell = blockIdx.x * kTile + lane;
  const double *src = u + (size_t)blockIdx.x * F * kTile + lane;
  double vmax = 1e-5;
  if (cell < n_local) {
    for (int n = warp; n < N; n += NW) {
      Prim q = ConsToPrim(src[(n * 5) * kTile], src[(n * 5 + 1) * kTile], src[(n * 5 + 2) * kTile],
                          src[(n * 5 + 3) * kTile], src[(n * 5 + 4) * kTile]);
      double c = SoundSpeed(q.rho, q.p);
      double v = sqrt(q.u * q.u + q.v * q.v + q.w * q.w);
      vmax = fmax(vmax, fmax(c + v, c - v));
    }
  }
  s_max[warp][lane] = vmax;
  __syncthreads();
  if (warp == 0) {
    double dt = 1e100;
    if (cell < n_local) {
#pragma unroll
      for (int w = 1; w < NW; ++w) vmax = fmax(vmax, s_max[w][lane]);
      dt = length[cell] / vmax;
    }
    for (int o = 16; o > 0; o >>= 1) dt = fmin(dt, __shfl_xor_sync(0xffffffffu, dt, o));
    if (lane == 0) atomicMin(result, (unsigned long long)__double_as_longlong(dt));
  }
}

// Initial-condition clean-up limiters (limiter/average.hpp:11-28, limiter/majority.hpp:14-42): every node
// of a cell gets the cell average (sum_q u_q w_q detJ_q / volume, polynomial/hexahedron.hpp:626-632) or,
// per component, the value that more than half of the nodes carry.  One thread per cell.
template <int Q>
__global__ void CellConstantKernel(double *u, const double *geom, int n_local, int majority) {
  constexpr int N = Q * Q * Q, F = 5 * N;
  const int cell = blockIdx.x * blockDim.x + threadIdx.x;
  if (cell >= n_local) return;
  double *dst = u + (size_t)(cell >> 5) * F * kTile + (cell & 31);
  double value[5] = {0, 0, 0, 0, 0};
  if (majority) {
    for (int k = 0; k < 5; ++k) {
      bool found = false;
      for (int q = 0; q < N && !found; ++q) {
        const double val = dst[(q * 5 + k) * kTile];
        int count = 0;
        for (int r = 0; r < N; ++r) count += dst[(r * 5 + k) * kTile] == val;
        if (count * 2 > N) { value[k] = val; found = true; }
      }
      // the reference asserts that a majority exists; without one it leaves value[k] unset
    }
  } else {
    const double *g = geom + (size_t)(cell >> 5) * kGeomRows * kTile + (cell & 31);
    double volume = 0;
    for (int q = 0; q < N; ++q) {
      const int i = q / (Q * Q), j = (q / Q) % Q, k = q % Q;
      const Metric m = ComputeMetric(g, kTile, c_tab.xi[i], c_tab.xi[j], c_tab.xi[k]);
      const double w = c_tab.w[i] * c_tab.w[j] * c_tab.w[k] * m.det;   // global weight
      volume += w;
      for (int c = 0; c < 5; ++c) value[c] += dst[(q * 5 + c) * kTile] * w;
    }
    for (int c = 0; c < 5; ++c) value[c] /= volume;
  }
  for (int q = 0; q < N; ++q)
    for (int c = 0; c < 5; ++c) dst[(q * 5 + c) * kTile] = value[c];
}

// dt = min(cfl * min_dt, dt_guess), doubled and capped again for FR-Lobatto (fem.hpp:433-434,
// fr/lobatto.hpp:220-222), entirely on the device so that a time loop never has to wait for the host.
__global__ void TimeStepFinishKernel(const unsigned long long *min_bits, double cfl, double dt_guess,
    int fr_doubling, double *dt_out) {
  double dt = fmin(cfl * __longlong_as_double((long long)*min_bits), dt_guess);
  if (fr_doubling) dt = fmin(dt_guess, 2.0 * dt);
  *dt_out = dt;
}

__global__ void TimeStepResetKernel(unsigned long long *min_bits) {
  *min_bits = (unsigned long long)__double_as_longlong(1e100);
}

// column[cell*F + f]  <->  tiles[(tile*F + f)*32 + lane]; one CTA per tile, staged through shared
// memory so both sides are coalesced.  `cells` (optional) lists the cells to move (halo pack/unpack:
// item i of the column is cell cells[i]).
template <int F, bool kToTiles>
__global__ void ConvertKernel(double *tiles, double *column, const int *cells, int n_items) {
  constexpr int FP = F + 1;  // padded: conflict-free transposed access
  extern __shared__ double sbuf[];  // [32][FP]
  int first = blockIdx.x * kTile;
  int count = min(kTile, n_items - first);
  if (kToTiles) {
    for (int i = threadIdx.x; i < count * F; i += blockDim.x)
      sbuf[(i / F) * FP + i % F] = column[(size_t)first * F + i];
    __syncthreads();
    for (int i = threadIdx.x; i < F * kTile; i += blockDim.x) {
      int f = i / kTile, l = i % kTile;
      if (l < count) {
        int cell = cells ? cells[first + l] : first + l;
        tiles[((size_t)(cell >> 5) * F + f) * kTile + (cell & 31)] = sbuf[l * FP + f];
      }
    }
  } else {
    for (int i = threadIdx.x; i < F * kTile; i += blockDim.x) {
      int f = i / kTile, l = i % kTile;
      if (l < count) {
        int cell = cells ? cells[first + l] : first + l;
        sbuf[l * FP + f] = tiles[((size_t)(cell >> 5) * F + f) * kTile + (cell & 31)];
      }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < count * F; i += blockDim.x)
      column[(size_t)first * F + i] = sbuf[(i / F) * FP + i % F];
  }
}

// Ghost exchange of the per-cell viscosities (EnergyBasedViscosity::ShareGhostCellProperties /
// UpdateGhostCellProperties, spatial/viscosity.hpp:163-189): 5 doubles per cell, buffer[i][5] <-> nu[tile][comp][lane]
__global__ void ViscosityHaloKernel(double *nu, double *buffer, const int *cells, int n_items, int unpack) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_items * 5) return;
  const int cell = cells[i / 5], k = i % 5;
  double *at = nu + ((size_t)(cell >> 5) * 5 + k) * kTile + (cell & 31);
  if (unpack) *at = buffer[i];
  else buffer[i] = *at;
}

// ---- frame output: mini::mesh::vtk::Writer<Part>::Prepare (mesh/vtk.hpp:421-494) ------------------------------
// Every hexahedral cell becomes one VTK tri-quadratic hexahedron (vtk.hpp:400-409) whose 27 points carry the
// solution evaluated at their local coordinates -1/0/+1 (VTK's node order, vtk.hpp:305-325: corners, edges,
// faces, centre).  values[cell][point][component], the reference's memory order of its `values` vector.
//   nodal:  Hexahedron::LocalToValue (polynomial/hexahedron.hpp:256-264): sum_k coeff_k phi_k(local), for
//           kLocal == true divided by det J at the point (the coefficients are det(J_k) u_k there);
//   modal:  Projection::LocalToGlobalAndValue (projection.hpp:128-132; Extrapolation forwards to its projection,
//           extrapolation.hpp:180-183): the Taylor form at LocalToGlobal(local) - centre.
struct VtkSampleArgs {
  const double *u, *geom, *tay;   // tay: Taylor coefficients, row (mode * mode_stride + comp * comp_stride)
  double *values;
  int n_local, modal, k_local, comp_stride, mode_stride;
  double lag[3][4];               // Lagrange basis on the Q solution nodes evaluated at -1, 0, +1
};
__constant__ signed char kVtkHex27[27][3] = {
  {-1, -1, -1}, {1, -1, -1}, {1, 1, -1}, {-1, 1, -1}, {-1, -1, 1}, {1, -1, 1}, {1, 1, 1}, {-1, 1, 1},
  {0, -1, -1}, {1, 0, -1}, {0, 1, -1}, {-1, 0, -1}, {0, -1, 1}, {1, 0, 1}, {0, 1, 1}, {-1, 0, 1},
  {-1, -1, 0}, {1, -1, 0}, {1, 1, 0}, {-1, 1, 0},
  {-1, 0, 0}, {1, 0, 0}, {0, -1, 0}, {0, 1, 0}, {0, 0, -1}, {0, 0, 1}, {0, 0, 0}};

template <int Q>
__global__ void __launch_bounds__(256) VtkSampleKernel(VtkSampleArgs p) {
  constexpr int P = Q - 1, N = Q * Q * Q, F = 5 * N, M = ModalShape<P>::M, kRow = 27 * 5;
  __shared__ double sOut[kTile * kRow];
  const int tile = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cell = tile * kTile + lane;
  const double *geo = p.geom + (size_t)tile * kGeomRows * kTile + lane;
  if (cell < p.n_local) {
    for (int pt = warp; pt < 27; pt += 8) {
      const int a = kVtkHex27[pt][0], b = kVtkHex27[pt][1], c = kVtkHex27[pt][2];
      double v[5] = {0, 0, 0, 0, 0};
      if (p.modal) {
        double d[3], t[M];
        DgmPoint(geo, kTile, (double)a, (double)b, (double)c, d);
        TaylorValues<P>(d, t);
        const double *src = p.tay + (size_t)tile * 5 * M * kTile + lane;
        for (int j = 0; j < M; ++j) {
#pragma unroll
          for (int k = 0; k < 5; ++k) v[k] = fma(src[(size_t)(j * p.mode_stride + k * p.comp_stride) * kTile], t[j], v[k]);
        }
      } else {
        const double *src = p.u + (size_t)tile * F * kTile + lane;
        for (int i = 0; i < Q; ++i) {
          const double wi = p.lag[a + 1][i];
          if (wi == 0) continue;
          for (int j = 0; j < Q; ++j) {
            const double wj = wi * p.lag[b + 1][j];
            if (wj == 0) continue;
            for (int k = 0; k < Q; ++k) {
              double w = wj * p.lag[c + 1][k];
              if (w == 0) continue;
              if (p.k_local) w *= ComputeMetric(geo, kTile, c_tab.xi[i], c_tab.xi[j], c_tab.xi[k]).det;
              const int n = (i * Q + j) * Q + k;
#pragma unroll
              for (int m = 0; m < 5; ++m) v[m] = fma(w, src[(n * 5 + m) * kTile], v[m]);
            }
          }
        }
        if (p.k_local) {
          const double inv = 1.0 / ComputeMetric(geo, kTile, (double)a, (double)b, (double)c).det;
#pragma unroll
          for (int m = 0; m < 5; ++m) v[m] *= inv;
        }
      }
#pragma unroll
      for (int m = 0; m < 5; ++m) sOut[lane * kRow + pt * 5 + m] = v[m];
    }
  }
  __syncthreads();
  const int valid = min(kTile, p.n_local - tile * kTile) * kRow;   // the tile's rows are contiguous in the output
  double *dst = p.values + (size_t)tile * kTile * kRow;
  for (int i = threadIdx.x; i < valid; i += 256) dst[i] = sOut[i];
}

// Test hook: n independent rotated Riemann problems.
template <int kRiemann>
__global__ void RiemannKernel(int bc_kind, int n, const double *normals, const double *ul,
    const double *ur, double *flux) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double a[5], b[5], f[5];
  for (int c = 0; c < 5; ++c) { a[c] = ul[i * 5 + c]; b[c] = ur[i * 5 + c]; }
  if (bc_kind < 0) FluxUpwind<kRiemann>(a, b, normals[3 * i], normals[3 * i + 1], normals[3 * i + 2], f);
  else BoundaryFlux<kRiemann>(bc_kind, a, b, normals[3 * i], normals[3 * i + 1], normals[3 * i + 2], f);
  for (int c = 0; c < 5; ++c) flux[i * 5 + c] = f[c];
}

#endif  // __CUDACC__

}  // namespace mcfd

#include "peer.cuh"
#include "stage_lines.cuh"
#include "stage_nodes.cuh"

#endif  // MINICFD_B200_KERNELS_CUH_
