// stage_lines.cuh — the inviscid stage kernel of the lumped Lobatto schemes (fr::Lobatto, dg::Lobatto) at p = 2,
// "node line" formulation.  Included at the end of kernels.cuh.
//
// Same mathematics and the same HBM layout as StageKernel (kernels.cuh), different mapping onto the machine:
//   * lanes are cells of a tile: 16 cells per CTA and two node lines per warp, so a global access is a 128-byte half
//     row per line;
//   * a thread owns one LINE of Q nodes (i, j, 0..Q-1) of its cell, k compile time; (i, j) comes from a table indexed
//     by (warp, lane / 16) whose lines are sorted by CLASS (how many of i, j lie on a face), so a warp never mixes
//     classes and every branch on "my line has face points in direction xi / eta" is warp uniform;
//   * phase A: per node the primitive state, the three contravariant fluxes, the zeta part of the divergence (which
//     never leaves the thread).  Shared memory per node and cell: F^_xi (5), F^_eta (5), the residual accumulator
//     (5), the primitive state and sound speed (6) — 21 doubles;
//   * phase F (no barrier between A and F: a thread touches only what it wrote itself): every face point is solved
//     by the thread that owns its node, two points per pass through a branch-free HLLC (two independent
//     instruction streams per thread); boundary kinds, clipped states and the other solvers go through one
//     out-of-line copy per point.  The neighbours' states at the face points were requested in phase A and arrive
//     in shared memory without passing through registers (LDGSTS), `sN[6 Q^2][5][cells]`;
//   * ONE barrier, then phase N: the xi and eta parts of the divergence from the other lines' F^, 1/det J, the RK
//     update, the store — and, when the part has neighbours on other GPUs, the packed copy of a send cell's row
//     into the send stage (peer.cuh);
//   * the Jacobian rows of the trilinear map split into a part that is constant along the line and a part
//     linear in zeta, so the metric of a node costs 6 FMA + 3 cross products, and the face normals of a line
//     come from the same pieces; a mesh whose cells are all parallelepipeds (up to rounding) takes an instantiation
//     with one metric per cell (kAffine: 155 instead of 164 us per stage; such meshes go to stage_nodes.cuh by default);
//   * reciprocals and square roots: one MUFU seed + one cubic step; HLLC with the two square roots per side
//     merged (a_K q_K is one root) and one reciprocal per star region;
//   * the per-degree tables travel as a kernel parameter (constant bank), the byte tables are staged in shared
//     memory by the prologue together with the geometry rows and the face tables of the CTA's cells.
// Shape at p = 2: 5 warps per CTA, 111 KB of shared memory, 166 registers, two CTAs per SM (the register file is
// 16 K registers per SM sub-partition; the shared-memory carve-out has to be raised to hold two CTAs).
// Measured history of this kernel and of the variants that did not pay off: profiles/r2_kernel_history.md.
// Reference members restated: fr::General::AddFluxDivergence spatial/fr/general.hpp:271-288,
// fr::Lobatto::AddFluxToHolderAndSharer spatial/fr/lobatto.hpp:232-251, fr::General::GetFluxForHolderAndSharer
// fr/general.hpp:296-315, dg::Lobatto dg/lobatto.hpp:143-191, boundary loops fr/lobatto.hpp:278-428,
// IdealGas::GetFluxMatrix riemann/euler/types.hpp:322-333, HartenLaxLeerContact riemann/euler/hllc.hpp:30-109,
// Hexahedron::GlobalFluxToLocalFlux polynomial/hexahedron.hpp:606-624, RungeKutta temporal/rk.hpp:16-56.
#ifndef MINICFD_B200_STAGE_LINES_CUH_
#define MINICFD_B200_STAGE_LINES_CUH_

#include <vector>

namespace mcfd {

// The part of `Tables` this kernel reads.  It travels as a kernel parameter (constant bank 0), so every
// context launches with its own copy and contexts of different degree or scheme may run concurrently.
struct LineTables {
  double xi[kMaxQ];
  double a[kMaxQ][kMaxQ];
  double coef_f, coef_F;
  signed char node_q[kMaxN][3];
  unsigned char nbr_node[6][8][kMaxFaceQ];
  // node line (i, j, .) of slot warp * SUB + sub (-1 = idle) and its class: 0 (i end, j end), 1 (i end, j inner),
  // 2 (i inner, j end), 3 (inner, inner)
  signed char line_i[kMaxFaceQ], line_j[kMaxFaceQ], line_cls[kMaxFaceQ];
  // face points of a line of each class, two per pass: k | direction << 2 | padding << 4
  unsigned char cls_pairs[4];
  unsigned char cls_pts[4][12];
};

// q = degree + 1, sub = node lines per warp (32 / cells per CTA)
inline LineTables MakeLineTables(const Tables &t, int q, int sub) {
  LineTables l{};
  int slot = 0;
  for (int cls = 0; cls < 4; ++cls) {
    for (int i = 0; i < q; ++i) for (int j = 0; j < q; ++j) {
      const bool ie = i == 0 || i == q - 1, je = j == 0 || j == q - 1;
      if ((ie ? 0 : 2) + (je ? 0 : 1) != cls) continue;
      l.line_i[slot] = (signed char)i; l.line_j[slot] = (signed char)j; l.line_cls[slot] = (signed char)cls; ++slot;
    }
    while (slot % sub) { l.line_i[slot] = l.line_j[slot] = -1; l.line_cls[slot] = (signed char)cls; ++slot; }   // a warp never mixes classes
    // the class's face points: both ends of the line (direction 2), then every node on the i-face, on the j-face
    int n = 0;
    l.cls_pts[cls][n++] = (unsigned char)(0 | 2 << 2);
    l.cls_pts[cls][n++] = (unsigned char)((q - 1) | 2 << 2);
    if (cls == 0 || cls == 1) for (int k = 0; k < q; ++k) l.cls_pts[cls][n++] = (unsigned char)(k | 0 << 2);
    if (cls == 0 || cls == 2) for (int k = 0; k < q; ++k) l.cls_pts[cls][n++] = (unsigned char)(k | 1 << 2);
    if (n % 2) { l.cls_pts[cls][n] = (unsigned char)(l.cls_pts[cls][n - 1] | 16); ++n; }   // odd count: the last point twice, once for nothing
    l.cls_pairs[cls] = (unsigned char)(n / 2);
  }
  for (; slot < kMaxFaceQ; ++slot) { l.line_i[slot] = l.line_j[slot] = -1; l.line_cls[slot] = 3; }
  for (int i = 0; i < kMaxQ; ++i) {
    l.xi[i] = t.xi[i];
    for (int m = 0; m < kMaxQ; ++m) l.a[i][m] = t.a[i][m];
  }
  l.coef_f = t.coef_f; l.coef_F = t.coef_F;
  for (int n = 0; n < kMaxN; ++n) for (int d = 0; d < 3; ++d) l.node_q[n][d] = t.node_q[n][d];
  for (int f = 0; f < 6; ++f) for (int o = 0; o < 8; ++o) for (int q2 = 0; q2 < kMaxFaceQ; ++q2) l.nbr_node[f][o][q2] = t.nbr_node[f][o][q2];
  return l;
}

// the byte tables the kernel stages in shared memory, as they lie there: nbr_node[6][8][q*q], node_q[q^3][3] (padded to 4)
inline std::vector<unsigned char> LineTableBytes(const LineTables &l, int q) {
  const int fq = q * q, n = q * fq;
  std::vector<unsigned char> b;
  for (int f = 0; f < 6; ++f) for (int o = 0; o < 8; ++o) for (int k = 0; k < fq; ++k) b.push_back(l.nbr_node[f][o][k]);
  for (int k = 0; k < n; ++k) for (int d = 0; d < 3; ++d) b.push_back((unsigned char)l.node_q[k][d]);
  while (b.size() % 4) b.push_back(0);
  return b;
}

// cells per CTA and resident CTAs per SM the kernel is compiled for
#ifndef MCFD_LINE_P2_CELLS
#define MCFD_LINE_P2_CELLS 16
#endif
#ifndef MCFD_LINE_P3_CELLS
#define MCFD_LINE_P3_CELLS 16
#endif
#ifndef MCFD_LINE_P2_MINB
#define MCFD_LINE_P2_MINB 2
#endif
#ifndef MCFD_LINE_P3_MINB
#define MCFD_LINE_P3_MINB 1
#endif
constexpr int kLineP2Cells = MCFD_LINE_P2_CELLS, kLineP3Cells = MCFD_LINE_P3_CELLS;

#ifdef __CUDACC__
// 1/x and 1/sqrt(x) from the MUFU seed (relative error < 2^-19) and one third-order step: error ~ e^3 plus
// two or three roundings.  Same caveats as Rcp / Rsqrt in physics.cuh (normal, non-zero arguments).
MCFD_HD double RcpC(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  const double e = fma(-x, r, 1.0);
  return fma(r, fma(e, e, e), r);
}
MCFD_HD double RsqrtC(double x) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double e = fma(-x * y, y, 1.0);
  return fma(y, e * fma(0.375, e, 0.5), y);
}
#else
MCFD_HD double RcpC(double x) { return 1.0 / x; }
MCFD_HD double RsqrtC(double x) { return 1.0 / sqrt(x); }
#endif
MCFD_HD double SqrtC(double x) { return x * RsqrtC(x); }

// ConservativeToPrimitive + GetSpeedOfSound of one state (types.hpp:194-202, 252-254, 289-301).  a == 0 marks a
// clipped or vacuum state (the reference's SetZeroIfNegative / rho == 0 branches).
struct PrimA {
  double rho, u, v, w, p, a;
};
MCFD_HD PrimA ConsToPrimA(const double *c) {
  PrimA q;
  const double r = c[0], e = c[4];
  const double inv = RcpC(r);
  q.rho = r;
  q.u = c[1] * inv; q.v = c[2] * inv; q.w = c[3] * inv;
  const double ke = 0.5 * (q.u * q.u + q.v * q.v + q.w * q.w);
  q.p = (e - r * ke) * kGm1;
  if (r > 0 && e >= 0 && q.p > 0) {                 // everywhere but in clipped or vacuum states
    q.a = SqrtC((kGamma * q.p) * inv);
  } else {
    if (r == 0 && e >= 0) { q.u = c[1]; q.v = c[2]; q.w = c[3]; q.p = e; }   // returned unchanged by the reference
    else if (r < 0 || e < 0 || q.p < 0) { q.rho = q.u = q.v = q.w = q.p = 0; }
    q.a = 0.0;
  }
  return q;
}
// the same without any branch, for states that are expected to be ordinary; *ok tells whether it was
MCFD_HD PrimA ConsToPrimFlat(const double *c, bool *ok) {
  PrimA q;
  const double r = c[0], e = c[4];
  const double inv = RcpC(r);
  q.rho = r;
  q.u = c[1] * inv; q.v = c[2] * inv; q.w = c[3] * inv;
  const double ke = 0.5 * (q.u * q.u + q.v * q.v + q.w * q.w);
  q.p = (e - r * ke) * kGm1;
  q.a = SqrtC((kGamma * q.p) * inv);
  *ok = r > 0 && e >= 0 && q.p > 0;
  return q;
}

// HLLC flux (hllc.hpp:30-109) of the Riemann problem (left = L, right = R) along the unit normal n pointing from L
// to R, in the global frame, for two ordinary states (rho, p > 0), without a branch: two of these in a row are two
// independent instruction streams for the scheduler.  The reference rotates into the face frame and poses the
// problem holder -> sharer; the solver is mirror symmetric, so posing it own -> neighbour along the own outward
// normal gives the negated flux up to rounding.  With p >= 0, max(max(p_pvrs, 0) - p_K, 0) = max(p_pvrs - p_K, 0), and
// q_K = sqrt(1 + (gamma+1)/(2 gamma) max(p_pvrs - p_K, 0) / p_K) takes one reciprocal of p_L p_R for both sides.
MCFD_HD void HllcFlux(const PrimA &L, const PrimA &R, double nx, double ny, double nz, double *flux) {
#ifdef MCFD_EXPERIMENT_CHEAP_FLUX   // timing experiment only: how much of the stage is the solver itself
  {   // central flux 0.5 (F(L) + F(R)) . n: about a quarter of the solver's arithmetic, stable for the few steps timed
    const double ul = L.u * nx + L.v * ny + L.w * nz, ur = R.u * nx + R.v * ny + R.w * nz;
    const double ml = L.rho * ul, mr = R.rho * ur, pm = 0.5 * (L.p + R.p);
    const double el = L.p * kOneOverGm1 + 0.5 * L.rho * (L.u * L.u + L.v * L.v + L.w * L.w);
    const double er = R.p * kOneOverGm1 + 0.5 * R.rho * (R.u * R.u + R.v * R.v + R.w * R.w);
    flux[0] = 0.5 * (ml + mr);
    flux[1] = 0.5 * (ml * L.u + mr * R.u) + pm * nx;
    flux[2] = 0.5 * (ml * L.v + mr * R.v) + pm * ny;
    flux[3] = 0.5 * (ml * L.w + mr * R.w) + pm * nz;
    flux[4] = 0.5 * ((el + L.p) * ul + (er + R.p) * ur);
    return;
  }
#endif
  const double unl = L.u * nx + L.v * ny + L.w * nz;
  const double unr = R.u * nx + R.v * ny + R.w * nz;
  const double rho_avg = (L.rho + R.rho) * 0.5, a_avg = (L.a + R.a) * 0.5;
  const double p_pvrs = (L.p + R.p - (unr - unl) * rho_avg * a_avg) * 0.5;
  constexpr double kQ = kGp1Over2 / kGamma;
  double dl = p_pvrs - L.p, dr = p_pvrs - R.p;
  dl = dl > 0 ? dl : 0.0; dr = dr > 0 ? dr : 0.0;
  const double ip = RcpC(L.p * R.p);
  const double aql = L.a * SqrtC(fma(kQ * dl, R.p * ip, 1.0)), aqr = R.a * SqrtC(fma(kQ * dr, L.p * ip, 1.0));
  const double wl = unl - aql, wr = unr + aqr;
  const double ml = L.rho * (wl - unl), mr = R.rho * (wr - unr);
  const double ws = (R.p - L.p + ml * unl - mr * unr) * RcpC(ml - mr);
  const bool left = (0.0 <= wl) || (!(wr <= 0.0) && (0.0 <= ws));
  const bool star = !(0.0 <= wl) && !(wr <= 0.0);
  const double rho = left ? L.rho : R.rho, un = left ? unl : unr, p = left ? L.p : R.p;
  const double vx = left ? L.u : R.u, vy = left ? L.v : R.v, vz = left ? L.w : R.w;
  const double wk = left ? wl : wr, mk = left ? ml : mr;
  const double ke = 0.5 * (vx * vx + vy * vy + vz * vz);
  const double energy = p * kOneOverGm1 + rho * ke;
  const double mass0 = rho * un;
  const double mom0 = mass0 * un + p;
  const double en0 = energy * un + p * un;
  const double inv = RcpC(wk - ws);
  const double temp = mk * inv;                                   // rho (S_K - u_K) / (S_K - S*)
  const double e_star = (energy * (wk - un) + (ws - un) * fma(mk, ws, p)) * inv;
  const double mass = star ? fma(temp - rho, wk, mass0) : mass0;
  const double mom_n = star ? fma(temp * ws - mass0, wk, mom0) : mom0;
  const double en = star ? fma(e_star - energy, wk, en0) : en0;
  const double tx = vx - un * nx, ty = vy - un * ny, tz = vz - un * nz;
  flux[0] = mass;
  flux[1] = mom_n * nx + mass * tx;
  flux[2] = mom_n * ny + mass * ty;
  flux[3] = mom_n * nz + mass * tz;
  flux[4] = en;
}

// Launch shape.  CPB cells of a tile per CTA, SUB = 32 / CPB node lines per warp.  The Q*Q lines of a cell are dealt
// to (warp, sub) in CLASS order, so that the lines of one warp agree on "my line lies on an i-face / a j-face" and
// every branch on that is warp uniform (LineTables::line_i / line_j / line_cls, filled by MakeLineTables).
template <int Q, int CPB>
struct LineShape {
  static constexpr int N = Q * Q * Q, F = 5 * N, FQ = Q * Q;
  static constexpr int SUB = kTile / CPB;
  static constexpr int kSlots = (Q == 3 && SUB == 2) ? 10 : FQ;   // Q = 3: classes of 4, 2, 2, 1 lines
  static constexpr int NW = kSlots / SUB;
  static constexpr int kThreads = NW * 32;
  static constexpr int kMinBlocks = (Q == 3) ? MCFD_LINE_P2_MINB : MCFD_LINE_P3_MINB;
  // registers per thread for kMinBlocks resident CTAs: the register file is 16 K registers per SM sub-partition and
  // the warps of the resident CTAs are dealt round-robin to the four sub-partitions
  static constexpr int kWarpsPerPartition = (kMinBlocks * NW + 3) / 4;
  static constexpr int kRegsRaw = 16384 / (32 * kWarpsPerPartition) / 8 * 8;
  static constexpr int kMaxRegs = kRegsRaw > 255 ? 255 : kRegsRaw;
  // shared memory per node and cell: F^_xi (5), F^_eta (5), residual accumulator (5), primitive state + sound speed (6)
  static constexpr int kNodeDoubles = 21;
  static constexpr int kStageDoubles = 6 * FQ * 5 * CPB;   // neighbour states at the face points
  static constexpr int kNodeQBytes = (N * 3 + 3) / 4 * 4;
  // affine cells (parallelepipeds): per cell and direction the column of J* (3), 1 / its length, its length; det J
  static constexpr int kAffineRows = 16;
  static constexpr size_t kSharedBytes = (size_t)(N * kNodeDoubles * CPB + kStageDoubles + (kGeomRows + kAffineRows) * CPB) * sizeof(double)
      + 6 * 8 * FQ + kNodeQBytes + 6 * CPB * (sizeof(int) + 1);
  static_assert(SUB == 1 || SUB == 2 || (SUB == 4 && Q == 4), "lines per warp");
};

// what a thread keeps between the phases
template <int Q>
struct LineRegs {
  double u[Q][5];    // own state at the nodes of the line
};

template <int Q, int CPB>
struct LineThread {
  using S = LineShape<Q, CPB>;
  int lane, warp, lc, sub, block, tile, tl, cell, i, j, cls;
  bool active, valid;
  double *sF, *sN, *sG, *sA;
  unsigned char *sNbrNode, *sNodeQ, *sCode;
  int *sNbr;
  MCFD_HD LineThread(const StageArgs &p, const LineTables &tab, int tid, int bid, double *smem) {
    sF = smem;                                                           // [N][21][CPB]
    sN = smem + S::N * S::kNodeDoubles * CPB;                            // [6 FQ][5][CPB]
    sG = sN + S::kStageDoubles;                                          // [21][CPB]
    sA = sG + kGeomRows * CPB;                                           // [16][CPB]
    sNbrNode = reinterpret_cast<unsigned char *>(sA + S::kAffineRows * CPB);   // [6][8][FQ]
    sNodeQ = sNbrNode + 6 * 8 * S::FQ;                                   // [N][3]
    sNbr = reinterpret_cast<int *>(sNodeQ + S::kNodeQBytes);             // [6][CPB]
    sCode = reinterpret_cast<unsigned char *>(sNbr + 6 * CPB);           // [6][CPB]
    lane = tid & 31; warp = tid >> 5;
    lc = lane % CPB; sub = lane / CPB;
    block = p.block_begin + bid;
    tile = block / S::SUB;
    tl = (block % S::SUB) * CPB + lc;
    cell = tile * kTile + tl;
    active = cell < p.n_local;
    i = tab.line_i[warp * S::SUB + sub];
    j = tab.line_j[warp * S::SUB + sub];
    cls = tab.line_cls[warp * S::SUB];
    valid = i >= 0;
    if (!valid) { i = 1; j = 1; }
  }
};

#ifdef __CUDA_ARCH__
#define MCFD_PREFETCH_L2(ptr) asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr))
// 8 bytes global -> shared without a register in between (LDGSTS); completion through the thread's async group
MCFD_HD void AsyncCopy8(double *dst_shared, const double *src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(dst_shared)), "l"(src) : "memory");
}
MCFD_HD void AsyncCopy4(void *dst_shared, const void *src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(dst_shared)), "l"(src) : "memory");
}
MCFD_HD void AsyncCopyWait() { asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory"); }
#else
#define MCFD_PREFETCH_L2(ptr) ((void)(ptr))
MCFD_HD void AsyncCopy8(double *dst, const double *src) { *dst = *src; }
MCFD_HD void AsyncCopy4(void *dst, const void *src) { *static_cast<int *>(dst) = *static_cast<const int *>(src); }
MCFD_HD void AsyncCopyWait() {}
#endif

template <int Q, int CPB>
MCFD_HD void LineLoad(const StageArgs &p, const LineTables &tab, int tid, int bid, double *smem, LineRegs<Q> *t) {
  using S = LineShape<Q, CPB>;
  LineThread<Q, CPB> g(p, tab, tid, bid, smem);
  const size_t row0 = (size_t)g.tile * S::F * kTile + (size_t)((g.i * Q + g.j) * Q) * 5 * kTile + g.tl;
  const double *uin = p.u_in + row0;
#pragma unroll
  for (int k = 0; k < Q; ++k) {
#pragma unroll
    for (int c = 0; c < 5; ++c) t->u[k][c] = uin[(k * 5 + c) * kTile];
  }
  if (p.mode >= kModeRk3b && g.valid) {   // the second stream of the update: bring it towards the SM early
    const double *u0 = p.u0 + row0;
#pragma unroll
    for (int k = 0; k < Q; ++k) {
#pragma unroll
      for (int c = 0; c < 5; ++c) MCFD_PREFETCH_L2(u0 + (k * 5 + c) * kTile);
    }
  }
}

// The Jacobian rows of the trilinear map along the line (xi_i, eta_j, .): rows 0 and 1 are linear in zeta,
// row 2 does not depend on it (coordinate/hexahedron.hpp:431-456; the same FMA nesting as ComputeMetric).
struct LineJacobian {
  double A0[3], B0[3], A1[3], B1[3], R2[3];   // r0 = A0 + zeta B0,  r1 = A1 + zeta B1,  r2 = R2
};
MCFD_HD LineJacobian ComputeLineJacobian(const double *g, int gs, double x, double e) {
  LineJacobian J;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const double c1 = g[(0 + k) * gs], c2 = g[(3 + k) * gs], c3 = g[(6 + k) * gs];
    const double c4 = g[(9 + k) * gs], c5 = g[(12 + k) * gs], c6 = g[(15 + k) * gs], c7 = g[(18 + k) * gs];
    J.A0[k] = fma(e, c4, c1); J.B0[k] = fma(e, c7, c6);
    J.A1[k] = fma(x, c4, c2); J.B1[k] = fma(x, c7, c5);
    J.R2[k] = fma(e, J.B1[k], fma(x, c6, c3));
  }
  return J;
}
// the same for a parallelepiped: r0 = c1, r1 = c2, r2 = c3
MCFD_HD LineJacobian AffineLineJacobian(const double *g, int gs) {
  LineJacobian J;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    J.A0[k] = g[(0 + k) * gs]; J.A1[k] = g[(3 + k) * gs]; J.R2[k] = g[(6 + k) * gs];
    J.B0[k] = 0; J.B1[k] = 0;
  }
  return J;
}
MCFD_HD void Cross(const double *a, const double *b, double *c) {
  c[0] = a[1] * b[2] - a[2] * b[1];
  c[1] = a[2] * b[0] - a[0] * b[2];
  c[2] = a[0] * b[1] - a[1] * b[0];
}

// Everything the CTA shares goes global -> shared asynchronously (no register, no stall at a store): the byte tables
// (p.line_bytes = nbr_node[6][8][FQ] then node_q[N][3], a device copy of LineTables' members), the geometry rows
// and the face tables of the block's cells.  LinePrologueWait() + a barrier make them visible.
template <int Q, int CPB, bool kAffine>
MCFD_HD void LinePrologue(const StageArgs &p, const LineTables &tab, int tid, int bid, double *smem) {
  using S = LineShape<Q, CPB>;
  LineThread<Q, CPB> g(p, tab, tid, bid, smem);
  if (kAffine && tid < CPB) {
    // Every cell of the mesh is a parallelepiped (c4..c7 = 0): J is constant in the cell.  One thread per cell of the
    // block takes c1, c2, c3 straight from global memory and leaves the three columns of J*, their lengths and det J
    // in shared memory — the same operations on the same numbers as the general path performs per node and per face
    // point, so the results are bit-identical.
    const double *gm = p.geom + (size_t)g.tile * kGeomRows * kTile + (g.block % S::SUB) * CPB + tid;
    double c[3][3], nrm[3][3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
      for (int x = 0; x < 3; ++x) c[r][x] = gm[(r * 3 + x) * kTile];
    }
    Cross(c[1], c[2], nrm[0]); Cross(c[2], c[0], nrm[1]); Cross(c[0], c[1], nrm[2]);
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      const double n2 = nrm[d][0] * nrm[d][0] + nrm[d][1] * nrm[d][1] + nrm[d][2] * nrm[d][2];
      const double inv_area = RsqrtC(n2);
#pragma unroll
      for (int x = 0; x < 3; ++x) g.sA[(d * 5 + x) * CPB + tid] = nrm[d][x];
      g.sA[(d * 5 + 3) * CPB + tid] = inv_area;
      g.sA[(d * 5 + 4) * CPB + tid] = n2 * inv_area;
    }
    g.sA[15 * CPB + tid] = c[0][0] * nrm[0][0] + c[0][1] * nrm[0][1] + c[0][2] * nrm[0][2];
  }
  constexpr int kTableWords = (6 * 8 * S::FQ + S::kNodeQBytes) / 4;
  for (int x = tid; x < kTableWords; x += S::kThreads) AsyncCopy4(g.sNbrNode + 4 * x, p.line_bytes + 4 * x);
  const int off = (g.block % S::SUB) * CPB;
  for (int x = tid; x < kGeomRows * CPB; x += S::kThreads)
    AsyncCopy8(g.sG + x, p.geom + ((size_t)g.tile * kGeomRows + x / CPB) * kTile + off + x % CPB);
  for (int x = tid; x < 6 * CPB; x += S::kThreads)
    AsyncCopy4(g.sNbr + x, p.nbr + ((size_t)g.tile * 6 + x / CPB) * kTile + off + x % CPB);
  for (int x = tid; x < 6 * CPB / 4; x += S::kThreads)
    AsyncCopy4(g.sCode + 4 * x, p.code + ((size_t)g.tile * 6 + (4 * x) / CPB) * kTile + off + (4 * x) % CPB);
}

// F^_d = F(u) . s for one column s of J* (types.hpp:322-333 with the primitive state q of u; hexahedron.hpp:607-611)
MCFD_HD void ContravariantFlux(const double *u, const PrimA &q, const double *s, double *f) {
  const double un = q.u * s[0] + q.v * s[1] + q.w * s[2];
  f[0] = u[1] * s[0] + u[2] * s[1] + u[3] * s[2];
  f[1] = fma(u[1], un, q.p * s[0]);
  f[2] = fma(u[2], un, q.p * s[1]);
  f[3] = fma(u[3], un, q.p * s[2]);
  f[4] = (u[4] + q.p) * un;
}

// The neighbour's state at the node matching face point q of local face f goes to the thread's own slot of sN;
// nothing moves for a boundary face.
template <int Q, int CPB>
MCFD_HD void LineStageNeighbour(const StageArgs &p, const LineThread<Q, CPB> &g, int f, int q) {
  using S = LineShape<Q, CPB>;
  const int nb = g.sNbr[f * CPB + g.lc], code = g.sCode[f * CPB + g.lc];
  if (nb < 0) return;
  const int n2node = g.sNbrNode[((code & 7) * 8 + ((code >> 3) & 7)) * S::FQ + q];
  double *dst = g.sN + (size_t)((f * S::FQ + q) * 5) * CPB + g.lc;
  if (p.peer.on && nb >= p.n_local) {   // a ghost cell whose state lies in the receive stage [row][item] (peer.cuh)
    const double *src = p.peer.recv + (size_t)(n2node * 5) * p.peer.n_recv + p.peer.recv_index[nb - p.n_local];
#pragma unroll
    for (int c = 0; c < 5; ++c) AsyncCopy8(dst + c * CPB, src + (size_t)c * p.peer.n_recv);
    return;
  }
  const double *src = p.u_in + ((size_t)(nb >> 5) * S::F + n2node * 5) * kTile + (nb & 31);
#pragma unroll
  for (int c = 0; c < 5; ++c) AsyncCopy8(dst + c * CPB, src + c * kTile);
}

// ---- phase A: per node of the line: primitive state, contravariant fluxes, the zeta part of the divergence ----
// Shared memory receives, per node of the cell, F^_xi and F^_eta (read by the other lines in phase B), and for this
// thread's later use the finished zeta-direction sum and the primitive state.  The node's own face-flux terms
// +-coef_F F^_d(own) (fr/general.hpp:305-313, "- F^_h[:, d]") ride on the diagonal of the divergence matrix.
// The neighbour states of the line's face points are requested first and travel meanwhile.
template <int Q, int CPB, bool kAffine>
MCFD_HD void LinePhaseA(const StageArgs &p, const LineTables &tab, int tid, int bid, double *smem, LineRegs<Q> *t) {
  using S = LineShape<Q, CPB>;
  constexpr int ND = S::kNodeDoubles;
  LineThread<Q, CPB> g(p, tab, tid, bid, smem);
  if (!g.valid) return;
  const int n0 = (g.i * Q + g.j) * Q;
  if (g.active) {
    const bool i_end = g.i == 0 || g.i == Q - 1, j_end = g.j == 0 || g.j == Q - 1;
#pragma unroll
    for (int k = 0; k < Q; ++k) {
      if (i_end) LineStageNeighbour<Q, CPB>(p, g, g.i == 0 ? 4 : 2, g.sNodeQ[(n0 + k) * 3 + 0]);
      if (j_end) LineStageNeighbour<Q, CPB>(p, g, g.j == 0 ? 1 : 3, g.sNodeQ[(n0 + k) * 3 + 1]);
      if (k == 0 || k == Q - 1) LineStageNeighbour<Q, CPB>(p, g, k == 0 ? 0 : 5, g.sNodeQ[(n0 + k) * 3 + 2]);
    }
  }
  const LineJacobian J = kAffine ? AffineLineJacobian(g.sG + g.lc, CPB)
                                 : ComputeLineJacobian(g.sG + g.lc, CPB, tab.xi[g.i], tab.xi[g.j]);
  double *dst = g.sF + (size_t)n0 * ND * CPB + g.lc;
  double f2[Q][5];
#pragma unroll
  for (int k = 0; k < Q; ++k) {
    const double z = tab.xi[k];
    double r0[3], r1[3], s[3];
#pragma unroll
    for (int x = 0; x < 3; ++x) {   // (affine: the same rows at every node, the products below are computed once)
      r0[x] = kAffine ? J.A0[x] : fma(z, J.B0[x], J.A0[x]);
      r1[x] = kAffine ? J.A1[x] : fma(z, J.B1[x], J.A1[x]);
    }
    bool ok;
    PrimA q = ConsToPrimFlat(t->u[k], &ok);       // branch-free: the Q nodes of the line are Q independent streams
    if (!ok) q = ConsToPrimA(t->u[k]);             // clipped or vacuum state
    double *dk = dst + k * ND * CPB;
    dk[15 * CPB] = q.rho; dk[16 * CPB] = q.u; dk[17 * CPB] = q.v; dk[18 * CPB] = q.w; dk[19 * CPB] = q.p; dk[20 * CPB] = q.a;
    double f[5];
    Cross(r1, J.R2, s);
    ContravariantFlux(t->u[k], q, s, f);
#pragma unroll
    for (int c = 0; c < 5; ++c) dk[c * CPB] = f[c];
    Cross(J.R2, r0, s);
    ContravariantFlux(t->u[k], q, s, f);
#pragma unroll
    for (int c = 0; c < 5; ++c) dk[(5 + c) * CPB] = f[c];
    Cross(r0, r1, s);
    ContravariantFlux(t->u[k], q, s, f2[k]);
  }
  const double cF = tab.coef_F;
#pragma unroll
  for (int k = 0; k < Q; ++k) {
    double a2[Q];
#pragma unroll
    for (int m = 0; m < Q; ++m) a2[m] = tab.a[k][m];
    if (k == 0) a2[0] -= cF;
    if (k == Q - 1) a2[Q - 1] += cF;
#pragma unroll
    for (int c = 0; c < 5; ++c) {
      double acc = a2[0] * f2[0][c];
#pragma unroll
      for (int m = 1; m < Q; ++m) acc = fma(a2[m], f2[m][c], acc);
      dst[(k * ND + 10 + c) * CPB] = acc;
    }
  }
  AsyncCopyWait();   // own slots of sN are complete (read back by this thread only)
}

// One face point of node k of the line, any kind: the boundary fluxes, clipped states, and every point of the
// solvers that are not written branch-free (exact, AUSM).  fl: flux along the own outward normal times *scale.
//   nr: column dd of J* at the node (the un-normalised normal of the dd-faces); (ox, oy, oz), area: its unit
//   outward form and length.
template <int Q, int kRiemann, int CPB>
MCFD_HD void LinePointGeneric(const StageArgs &p, const LineTables &tab, const LineThread<Q, CPB> &g, int k, int f,
    int q, const double *nr, double ox, double oy, double oz, double area, double *fl, double *scale) {
  using S = LineShape<Q, CPB>;
  constexpr int FQ = S::FQ;
  const int nb = g.sNbr[f * CPB + g.lc], code = g.sCode[f * CPB + g.lc];
  const double sgn = (f == 2 || f == 3 || f == 5) ? 1.0 : -1.0;
  double uo[5];
  const double *own = p.u_in + (size_t)g.tile * S::F * kTile + (size_t)(((g.i * Q + g.j) * Q + k) * 5) * kTile + g.tl;
#pragma unroll
  for (int c = 0; c < 5; ++c) uo[c] = own[c * kTile];
  *scale = -tab.coef_f * area;     // the lifted term  -g' |n| f*  (FR) or  -|n| f* / w_end  (DG)
  if (nb >= 0) {
    // posed in the holder's frame like the reference (the exact solver's Newton iteration starts from
    // the left pressure): swap the states, flip the normal and negate when the neighbour holds the face
    const double *src = g.sN + (size_t)((f * FQ + q) * 5) * CPB + g.lc;
    const bool holder = (code & 64) != 0;
    const double hs = holder ? 1.0 : -1.0;
    double ul[5], ur[5];
#pragma unroll
    for (int c = 0; c < 5; ++c) { const double v = src[c * CPB]; ul[c] = holder ? uo[c] : v; ur[c] = holder ? v : uo[c]; }
    FluxUpwind<kRiemann>(ul, ur, hs * ox, hs * oy, hs * oz, fl);
    *scale *= hs;
  } else {
    const int bf = -1 - nb;
    const int kind = p.bkind[bf];
    if (kind == kBcNoSlip) {
      // no face term at all for inviscid systems (fem.hpp:382-383): take the node's own face-flux term back
      const PrimA L = ConsToPrimA(uo);
      ContravariantFlux(uo, L, nr, fl);
      *scale = -tab.coef_F * sgn;
    } else {
      State5 ui, given;
      const double *gv = p.bvals + ((size_t)bf * FQ + q) * 5;
#pragma unroll
      for (int c = 0; c < 5; ++c) { ui.v[c] = uo[c]; given.v[c] = gv[c]; }
      const State5 bfl = BoundaryFluxOutOfLine<kRiemann>(kind, ui, given, ox, oy, oz);
#pragma unroll
      for (int c = 0; c < 5; ++c) fl[c] = bfl.v[c];
    }
  }
}
// the rare points of the branch-free HLLC path: one shared out-of-line copy
struct PointResult { double fl[5], scale; };
template <int Q, int CPB>
MCFD_NOINLINE PointResult LinePointRare(const StageArgs &p, const LineTables &tab, const LineThread<Q, CPB> &g, int k,
    int f, int q, double n0, double n1, double n2, double ox, double oy, double oz, double area) {
  PointResult out;
  const double nr[3] = {n0, n1, n2};
  LinePointGeneric<Q, kHllc, CPB>(p, tab, g, k, f, q, nr, ox, oy, oz, area, out.fl, &out.scale);
  return out;
}

// geometry and addresses of face point (k, dd) of the line
struct PointSetup {
  int k, f, q;
  bool pad;
  double nr[3], ox, oy, oz, area;
};
template <int Q, int CPB, bool kAffine>
MCFD_HD PointSetup LinePointSetup(const LineTables &tab, const LineThread<Q, CPB> &g, const LineJacobian &J, int desc) {
  PointSetup s;
  s.k = desc & 3;
  const int dd = (desc >> 2) & 3;
  s.pad = (desc & 16) != 0;
  if (kAffine) {   // the cell's column dd of J*, its length and the reciprocal are waiting in shared memory
    const double *a = g.sA + dd * 5 * CPB + g.lc;
    s.f = dd == 0 ? (g.i == 0 ? 4 : 2) : (dd == 1 ? (g.j == 0 ? 1 : 3) : (s.k == 0 ? 0 : 5));
    s.q = g.sNodeQ[(((g.i * Q + g.j) * Q) + s.k) * 3 + dd];
    s.nr[0] = a[0]; s.nr[1] = a[CPB]; s.nr[2] = a[2 * CPB];
    const double os_ = ((s.f == 2 || s.f == 3 || s.f == 5) ? 1.0 : -1.0) * a[3 * CPB];
    s.area = a[4 * CPB];
    s.ox = os_ * s.nr[0]; s.oy = os_ * s.nr[1]; s.oz = os_ * s.nr[2];
    return s;
  }
  const double z = tab.xi[s.k];
  double r0[3], r1[3];
#pragma unroll
  for (int x = 0; x < 3; ++x) { r0[x] = fma(z, J.B0[x], J.A0[x]); r1[x] = fma(z, J.B1[x], J.A1[x]); }
  if (dd == 0) { Cross(r1, J.R2, s.nr); s.f = g.i == 0 ? 4 : 2; }
  else if (dd == 1) { Cross(J.R2, r0, s.nr); s.f = g.j == 0 ? 1 : 3; }
  else { Cross(r0, r1, s.nr); s.f = s.k == 0 ? 0 : 5; }
  s.q = g.sNodeQ[(((g.i * Q + g.j) * Q) + s.k) * 3 + dd];
  const double sgn = (s.f == 2 || s.f == 3 || s.f == 5) ? 1.0 : -1.0;
  const double n2 = s.nr[0] * s.nr[0] + s.nr[1] * s.nr[1] + s.nr[2] * s.nr[2];
  const double inv_area = RsqrtC(n2);
  s.area = n2 * inv_area;
  const double os_ = sgn * inv_area;
  s.ox = os_ * s.nr[0]; s.oy = os_ * s.nr[1]; s.oz = os_ * s.nr[2];   // own outward unit normal
  return s;
}

// ---- phase F: the face points of the line, two at a time.  Touches only what this thread wrote in phase A and
// its own staged neighbour states, so no barrier separates the two.
template <int Q, int kRiemann, int CPB, bool kAffine>
MCFD_HD void LinePhaseF(const StageArgs &p, const LineTables &tab, int tid, int bid, double *smem) {
  using S = LineShape<Q, CPB>;
  constexpr int FQ = S::FQ, ND = S::kNodeDoubles;
  LineThread<Q, CPB> g(p, tab, tid, bid, smem);
  if (!g.valid || !g.active) return;
  const int i = g.i, j = g.j, lc = g.lc;
  const int n0 = (i * Q + j) * Q;
  double *own = g.sF + (size_t)n0 * ND * CPB + lc;       // this line's nodes
  LineJacobian J;
  if (!kAffine) J = ComputeLineJacobian(g.sG + lc, CPB, tab.xi[i], tab.xi[j]);
  // Each pass solves two of them in straight-line code — two independent dependency chains per
  // warp — and adds the lifted terms to the accumulators of their nodes in shared memory.
  const int n_pairs = tab.cls_pairs[g.cls];
#pragma unroll 1
  for (int t = 0; t < n_pairs; ++t) {
    const PointSetup sa = LinePointSetup<Q, CPB, kAffine>(tab, g, J, tab.cls_pts[g.cls][2 * t]);
    const PointSetup sb = LinePointSetup<Q, CPB, kAffine>(tab, g, J, tab.cls_pts[g.cls][2 * t + 1]);
    double fa[5], fb[5], sca, scb;
    if (kRiemann == kHllc) {
      const double *la = own + (sa.k * ND + 15) * CPB, *lb = own + (sb.k * ND + 15) * CPB;
      const PrimA La = {la[0], la[CPB], la[2 * CPB], la[3 * CPB], la[4 * CPB], la[5 * CPB]};
      const PrimA Lb = {lb[0], lb[CPB], lb[2 * CPB], lb[3 * CPB], lb[4 * CPB], lb[5 * CPB]};
      const int nba = g.sNbr[sa.f * CPB + lc], nbb = g.sNbr[sb.f * CPB + lc];
      const double *va = g.sN + (size_t)((sa.f * FQ + sa.q) * 5) * CPB + lc, *vb = g.sN + (size_t)((sb.f * FQ + sb.q) * 5) * CPB + lc;
      double ra[5], rb[5];
#pragma unroll
      for (int c = 0; c < 5; ++c) { ra[c] = va[c * CPB]; rb[c] = vb[c * CPB]; }
      bool oka, okb;
      const PrimA Ra = ConsToPrimFlat(ra, &oka), Rb = ConsToPrimFlat(rb, &okb);
      HllcFlux(La, Ra, sa.ox, sa.oy, sa.oz, fa);
      HllcFlux(Lb, Rb, sb.ox, sb.oy, sb.oz, fb);
      sca = -tab.coef_f * sa.area; scb = -tab.coef_f * sb.area;
      if (!(oka && nba >= 0 && La.a > 0)) {   // boundary face, clipped or vacuum state
        const PointResult pr = LinePointRare<Q, CPB>(p, tab, g, sa.k, sa.f, sa.q, sa.nr[0], sa.nr[1], sa.nr[2], sa.ox, sa.oy, sa.oz, sa.area);
#pragma unroll
        for (int c = 0; c < 5; ++c) fa[c] = pr.fl[c];
        sca = pr.scale;
      }
      if (!(okb && nbb >= 0 && Lb.a > 0)) {
        const PointResult pr = LinePointRare<Q, CPB>(p, tab, g, sb.k, sb.f, sb.q, sb.nr[0], sb.nr[1], sb.nr[2], sb.ox, sb.oy, sb.oz, sb.area);
#pragma unroll
        for (int c = 0; c < 5; ++c) fb[c] = pr.fl[c];
        scb = pr.scale;
      }
    } else {
      LinePointGeneric<Q, kRiemann, CPB>(p, tab, g, sa.k, sa.f, sa.q, sa.nr, sa.ox, sa.oy, sa.oz, sa.area, fa, &sca);
      LinePointGeneric<Q, kRiemann, CPB>(p, tab, g, sb.k, sb.f, sb.q, sb.nr, sb.ox, sb.oy, sb.oz, sb.area, fb, &scb);
    }
    if (sb.pad) scb = 0.0;
    double *acc_a = own + (sa.k * ND + 10) * CPB, *acc_b = own + (sb.k * ND + 10) * CPB;
#pragma unroll
    for (int c = 0; c < 5; ++c) acc_a[c * CPB] = fma(sca, fa[c], acc_a[c * CPB]);
#pragma unroll
    for (int c = 0; c < 5; ++c) acc_b[c * CPB] = fma(scb, fb[c], acc_b[c * CPB]);
  }
}

// ---- phase N, behind the barrier that publishes F^_xi and F^_eta: per node the xi and eta parts of the divergence,
// 1/det (fr/general.hpp:256-266; dg/lobatto.hpp:143-158), RK update (rk.hpp:16-56), store.  Its global loads are
// issued before the barrier.
template <int Q, int CPB>
struct LineUpdateLoads {
  double uo[Q][5], u0[Q][5];
};
template <int Q, int CPB>
MCFD_HD void LineLoadUpdate(const StageArgs &p, const LineTables &tab, int tid, int bid, double *smem, LineUpdateLoads<Q, CPB> *ld) {
  using S = LineShape<Q, CPB>;
  LineThread<Q, CPB> g(p, tab, tid, bid, smem);
  if (!g.valid || !g.active) return;
  const size_t row0 = (size_t)g.tile * S::F * kTile + (size_t)((g.i * Q + g.j) * Q) * 5 * kTile + g.tl;
#pragma unroll
  for (int k = 0; k < Q; ++k) {
#pragma unroll
    for (int c = 0; c < 5; ++c) {
      ld->uo[k][c] = p.u_in[row0 + (k * 5 + c) * kTile];
      ld->u0[k][c] = p.mode >= kModeRk3b ? p.u0[row0 + (k * 5 + c) * kTile] : 0.0;
    }
  }
}
template <int Q, int CPB, bool kAffine>
MCFD_HD void LinePhaseN(const StageArgs &p, const LineTables &tab, int tid, int bid, double *smem, const LineUpdateLoads<Q, CPB> &ld) {
  using S = LineShape<Q, CPB>;
  constexpr int FQ = S::FQ, ND = S::kNodeDoubles;
  LineThread<Q, CPB> g(p, tab, tid, bid, smem);
  if (!g.valid || !g.active) return;
  const int i = g.i, j = g.j, lc = g.lc;
  const bool i_end = i == 0 || i == Q - 1, j_end = j == 0 || j == Q - 1;
  const int n0 = (i * Q + j) * Q;
  const double *own = g.sF + (size_t)n0 * ND * CPB + lc;
  LineJacobian J;
  if (!kAffine) J = ComputeLineJacobian(g.sG + lc, CPB, tab.xi[i], tab.xi[j]);
  const double cF = tab.coef_F;
  double a0[Q], a1[Q], out[Q][5];
#pragma unroll
  for (int m = 0; m < Q; ++m) {
    a0[m] = tab.a[i][m] + ((m == i && i_end) ? (i == 0 ? -cF : cF) : 0.0);
    a1[m] = tab.a[j][m] + ((m == j && j_end) ? (j == 0 ? -cF : cF) : 0.0);
  }
  const double *s0 = g.sF + (size_t)(j * Q) * ND * CPB + lc;              // nodes (m, j, k), direction 0
  const double *s1 = g.sF + (size_t)(i * FQ) * ND * CPB + 5 * CPB + lc;    // nodes (i, m, k), direction 1
  const size_t row0 = (size_t)g.tile * S::F * kTile + (size_t)n0 * 5 * kTile + g.tl;
  const double dt = p.dt_dev ? *p.dt_dev : p.dt;
  const int mode = p.mode;
  // kModeEuler: w;  kModeRk3b: (u0 3 + w) / 4;  kModeRk3c: (w 2 + u0) / 3;  kModeRk2b: (w + u0) / 2
  const double cw = mode == kModeRk3c ? 2.0 : 1.0;
  const double c0 = mode == kModeRk3b ? 3.0 : (mode == kModeRk3c || mode == kModeRk2b ? 1.0 : 0.0);
  const double cs_inv = mode == kModeRk3b ? 4.0 : (mode == kModeRk3c ? 3.0 : (mode == kModeRk2b ? 2.0 : 1.0));
  const double cs = mode == kModeRk3b ? 0.25 : (mode == kModeRk3c ? 1.0 / 3 : (mode == kModeRk2b ? 0.5 : 1.0));
#pragma unroll
  for (int k = 0; k < Q; ++k) {
    double r[5];
#pragma unroll
    for (int c = 0; c < 5; ++c) r[c] = own[(k * ND + 10 + c) * CPB];
#pragma unroll
    for (int m = 0; m < Q; ++m) {
#pragma unroll
      for (int c = 0; c < 5; ++c) {
        r[c] = fma(a0[m], s0[((m * FQ + k) * ND + c) * CPB], r[c]);
        r[c] = fma(a1[m], s1[((m * Q + k) * ND + c) * CPB], r[c]);
      }
    }
    double det;
    if (kAffine) {
      det = g.sA[15 * CPB + lc];
    } else {
      const double z = tab.xi[k];
      double r0[3], r1[3], nr[3];
#pragma unroll
      for (int x = 0; x < 3; ++x) { r0[x] = fma(z, J.B0[x], J.A0[x]); r1[x] = fma(z, J.B1[x], J.A1[x]); }
      Cross(r1, J.R2, nr);
      det = r0[0] * nr[0] + r0[1] * nr[1] + r0[2] * nr[2];
    }
    const double inv_det = RcpC(det);
#pragma unroll
    for (int c = 0; c < 5; ++c) {
      const double res = r[c] * inv_det;
      const double w = fma(res, dt, ld.uo[k][c]);
      // (w cw + u0 c0) cs with cs = 1, 1/4, 1/3, 1/2 (rk.hpp:16-49); the last two operations make the quotient by 3
      // correctly rounded in all but rare half-way cases and change nothing for the powers of two
      const double sum = fma(ld.u0[k][c], c0, w * cw);
      const double qv = sum * cs;
      const double val = fma(fma(-cs_inv, qv, sum), cs, qv);
      out[k][c] = mode == kModeResidual ? res : val;
      p.u_out[row0 + (k * 5 + c) * kTile] = out[k][c];
    }
  }
  if (p.peer.on && g.cell >= p.peer.first_cell) {   // a second copy of a send cell's rows, packed for the copy engines
    const int t0 = p.peer.tgt_offset[g.cell - p.peer.first_cell], t1 = p.peer.tgt_offset[g.cell - p.peer.first_cell + 1];
    for (int tt = t0; tt < t1; ++tt) {
      double *to = p.peer.send + (size_t)n0 * 5 * p.peer.n_send + p.peer.tgt_item[tt];
#pragma unroll
      for (int k = 0; k < Q; ++k) {
#pragma unroll
        for (int c = 0; c < 5; ++c) to[(size_t)(k * 5 + c) * p.peer.n_send] = out[k][c];
      }
    }
  }
}

#ifdef __CUDACC__
template <int Q, int kRiemann, int CPB, bool kAffine>
__global__ void __maxnreg__((LineShape<Q, CPB>::kMaxRegs))
LineStageKernel(const __grid_constant__ StageArgs p, const __grid_constant__ LineTables tab) {
  extern __shared__ double smem[];
  const int tid = threadIdx.x, bid = blockIdx.x;
  if (p.peer.on && p.block_begin + bid >= p.peer.first_block) {   // a block with inter cells: the neighbours' delivery
    if (tid < p.peer.n_neighbours &&
        !SpinUntil(p.peer.flags + tid, p.peer.counts[kPeerPlain] + p.peer.counts[kPeerStageBase] + (unsigned long long)p.peer.wait_add))
      *p.peer.error = 1;
    __syncthreads();
  }
  LineRegs<Q> t;
  LineLoad<Q, CPB>(p, tab, tid, bid, smem, &t);
  LinePrologue<Q, CPB, kAffine>(p, tab, tid, bid, smem);
  AsyncCopyWait();
  __syncthreads();
  LinePhaseA<Q, CPB, kAffine>(p, tab, tid, bid, smem, &t);
  LinePhaseF<Q, kRiemann, CPB, kAffine>(p, tab, tid, bid, smem);
  LineUpdateLoads<Q, CPB> ld;
  LineLoadUpdate<Q, CPB>(p, tab, tid, bid, smem, &ld);    // in flight across the barrier
  __syncthreads();
  LinePhaseN<Q, CPB, kAffine>(p, tab, tid, bid, smem, ld);
}
#else
template <int Q, int kRiemann, int CPB, bool kAffine>
void RunLineStageBlockOnHost(const StageArgs &p, const LineTables &tab, int bid, double *smem) {
  constexpr int T = LineShape<Q, CPB>::kThreads;
  static LineRegs<Q> regs[T];
  for (int t = 0; t < T; ++t) {
    LineLoad<Q, CPB>(p, tab, t, bid, smem, &regs[t]);
    LinePrologue<Q, CPB, kAffine>(p, tab, t, bid, smem);
  }
  static LineUpdateLoads<Q, CPB> ld[T];
  for (int t = 0; t < T; ++t) {
    LinePhaseA<Q, CPB, kAffine>(p, tab, t, bid, smem, &regs[t]);
    LinePhaseF<Q, kRiemann, CPB, kAffine>(p, tab, t, bid, smem);
    LineLoadUpdate<Q, CPB>(p, tab, t, bid, smem, &ld[t]);
  }
  for (int t = 0; t < T; ++t) LinePhaseN<Q, CPB, kAffine>(p, tab, t, bid, smem, ld[t]);
}
#endif

}  // namespace mcfd
#endif  // MINICFD_B200_STAGE_LINES_CUH_
