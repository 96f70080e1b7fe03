// stage_nodes.cuh — the inviscid stage kernel of the lumped Lobatto schemes at p = 2, "one warp per node", persistent:
// what mcfd_set_mesh picks for meshes of parallelepipeds (MCFD_STAGE_KERNEL=lines|nodes overrides).
// Included at the end of kernels.cuh, after stage_lines.cuh whose device functions (HllcFlux, ConsToPrimFlat,
// ComputeLineJacobian, ContravariantFlux, RcpC ...) and tables (LineTables) it shares.
//
// Same mathematics, same HBM layout, same results as LineStageKernel; a different mapping onto the SM:
//   * one CTA per SM walks over the tiles of the launch (tile = 32 cells; blockIdx.x, + gridDim.x, ...);
//   * warp n of its 27 warps owns node n = (i, j, k) of every cell of the tile, lanes are the 32 cells: every node
//     index, table entry and shared-memory offset is warp uniform, every global access a whole 256-byte row, and
//     "does my node lie on a face" is a warp-uniform branch.  27 warps per SM at 72 registers instead of 10 at 166;
//   * the tile's own state (135 rows = 34 560 contiguous bytes), geometry (21 rows), neighbour ids and codes arrive by
//     BULK ASYNC COPIES (cp.async.bulk ... mbarrier::complete_tx, one elected thread) issued one tile ahead; the
//     neighbours' states at a thread's own face points are gathered by that thread with cp.async (LDGSTS) one tile
//     ahead as well, into slots nobody else touches;
//   * per tile: A (primitive state, metric, the three contravariant fluxes -> shared memory) | barrier |
//     B (flux divergence from the three node lines through the node) | barrier | C (the node's 0..3 face points:
//     branch-free HLLC; the residual, the state and 1 / det J wait meanwhile in the thread's own 15 rows of the flux
//     array, which nobody else reads any more, because 72 registers do not hold them beside a Riemann problem) ->
//     D (1 / det J, RK update, store; a send cell's row also packed for the copy engines).  Nothing is shared after B,
//     so the warps drift apart until the next tile's first barrier;
//   * on a mesh of parallelepipeds (kAffine) the warp of the centre node, which has no face point, computes the
//     cell's metric — columns of J*, their lengths, 1 / det J — once per tile, one tile ahead, into a double-buffered
//     block guarded by an mbarrier; the other warps read it instead of deriving it per node and per face point.
// Measured (profiles/r2_kernel_history.md): 141 us per stage on the benchmark box (affine), 166 us on a general mesh.
// Reference members restated: as in stage_lines.cuh.
#ifndef MINICFD_B200_STAGE_NODES_CUH_
#define MINICFD_B200_STAGE_NODES_CUH_

namespace mcfd {

struct NodeShape {
  static constexpr int Q = 3, N = 27, F = 135, FQ = 9, NP = 54;
  static constexpr int kWarps = N, kThreads = kWarps * 32;
  static constexpr int kProducerWarp = 13;                   // node (1, 1, 1): no face point of its own
  // shared memory, in doubles
  static constexpr int kOffF = 0;                            // F^ [N][15][32]
  static constexpr int kOffN = kOffF + N * 15 * kTile;       // neighbour states [NP][5][32]
  static constexpr int kOffU = kOffN + NP * 5 * kTile;       // own state [F][32]
  static constexpr int kOffG = kOffU + F * kTile;            // geometry [2][21][32]
  static constexpr int kOffA = kOffG + 2 * kGeomRows * kTile;   // affine meshes: per-cell metric [2][16][32]
  static constexpr int kAffineRows = 16;                     // columns of J* (3 x 3), 1 / length and length of each (2 x 3), 1 / det J
  static constexpr int kDoubles = kOffA + 2 * kAffineRows * kTile;
  // then, in bytes from kDoubles * 8
  static constexpr int kNbrBytes = 6 * kTile * 4, kCodeBytes = 6 * kTile;
  static constexpr int kOffNbr = 0;                          // int [2][6][32]
  static constexpr int kOffCode = kOffNbr + 2 * kNbrBytes;   // uchar [2][6][32]
  static constexpr int kOffTab = kOffCode + 2 * kCodeBytes;  // nbr_node [6][8][FQ]
  static constexpr int kOffBar = (kOffTab + 6 * 8 * FQ + 15) / 16 * 16;   // four mbarriers: tile data [2], affine metric [2]
  static constexpr size_t kSharedBytes = (size_t)kDoubles * 8 + kOffBar + 32;
  static constexpr int kBulkBytes = F * kTile * 8 + kGeomRows * kTile * 8 + kNbrBytes + kCodeBytes;
};

#ifdef __CUDACC__
__device__ __forceinline__ unsigned SmemAddr(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void MbarInit(void *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(SmemAddr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void MbarExpect(void *bar, int bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(SmemAddr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void MbarWait(void *bar, int parity) {
  const unsigned addr = SmemAddr(bar);
  for (long long spins = 0;; ++spins) {
    unsigned done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(addr), "r"(parity) : "memory");
    if (done) return;
    // A copy that never lands: fail instead of hanging the device for ever.  The bound is minutes, not seconds: a tile
    // with inter cells is armed only after the neighbours' delivery has arrived, and the producer thread may rightly wait
    // for a late neighbour for up to kPeerSpinLimit polls (about ten seconds) before it gives up and reports it.
    if (spins > (1ll << 32)) __trap();
  }
}
__device__ __forceinline__ void MbarArrive(void *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(SmemAddr(bar)) : "memory");
}
// global -> shared, `bytes` (a multiple of 16) in one piece; completion is counted on the mbarrier
__device__ __forceinline__ void BulkCopy(void *dst_shared, const void *src, int bytes, void *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(SmemAddr(dst_shared)), "l"(src), "r"(bytes), "r"(SmemAddr(bar)) : "memory");
}
#endif

// what a thread knows about itself and the tile
struct NodeThread {
  int lane, n, i, j, k, tile, cell, buf;
  bool active;
  double *sF, *sN, *sU, *sG, *sA;
  int *sNbr;
  unsigned char *sCode, *sNbrNode;
  MCFD_HD NodeThread(const StageArgs &p, int tid, int tile_, int buf_, double *smem) {
    using S = NodeShape;
    lane = tid & 31; n = tid >> 5;
    i = n / S::FQ; j = (n / S::Q) % S::Q; k = n % S::Q;
    tile = tile_; buf = buf_;
    cell = tile * kTile + lane;
    active = cell < p.n_local;
    sF = smem + S::kOffF; sN = smem + S::kOffN; sU = smem + S::kOffU;
    sG = smem + S::kOffG + buf * kGeomRows * kTile;
    sA = smem + S::kOffA + buf * S::kAffineRows * kTile;
    unsigned char *bytes = reinterpret_cast<unsigned char *>(smem + S::kDoubles);
    sNbr = reinterpret_cast<int *>(bytes + S::kOffNbr + buf * S::kNbrBytes);
    sCode = bytes + S::kOffCode + buf * S::kCodeBytes;
    sNbrNode = bytes + S::kOffTab;
  }
};

// the faces node (i, j, k) lies on: direction dd -> local face id or -1
MCFD_HD int NodeFace(int i, int j, int k, int dd) {
  constexpr int Q = NodeShape::Q;
  if (dd == 0) return i == 0 ? 4 : (i == Q - 1 ? 2 : -1);
  if (dd == 1) return j == 0 ? 1 : (j == Q - 1 ? 3 : -1);
  return k == 0 ? 0 : (k == Q - 1 ? 5 : -1);
}

// The neighbours' states at this thread's own face points -> its own slots of sN (one tile ahead of their use).
MCFD_HD void NodeGather(const StageArgs &p, const LineTables &tab, const NodeThread &g) {
  using S = NodeShape;
  if (!g.active) return;
#pragma unroll
  for (int dd = 0; dd < 3; ++dd) {
    const int f = NodeFace(g.i, g.j, g.k, dd);
    if (f < 0) continue;
    const int q = tab.node_q[g.n][dd];
    const int nb = g.sNbr[f * kTile + g.lane], code = g.sCode[f * kTile + g.lane];
    if (nb < 0) continue;
    const int n2node = g.sNbrNode[((code & 7) * 8 + ((code >> 3) & 7)) * S::FQ + q];
    double *dst = g.sN + ((f * S::FQ + q) * 5) * kTile + g.lane;
    if (p.peer.on && nb >= p.n_local) {   // a ghost cell whose state lies in the receive stage [row][item] (peer.cuh)
      const double *src = p.peer.recv + (size_t)(n2node * 5) * p.peer.n_recv + p.peer.recv_index[nb - p.n_local];
#pragma unroll
      for (int c = 0; c < 5; ++c) AsyncCopy8(dst + c * kTile, src + (size_t)c * p.peer.n_recv);
    } else {
      const double *src = p.u_in + ((size_t)(nb >> 5) * S::F + n2node * 5) * kTile + (nb & 31);
#pragma unroll
      for (int c = 0; c < 5; ++c) AsyncCopy8(dst + c * kTile, src + c * kTile);
    }
  }
}

// A mesh of parallelepipeds: J is constant in a cell.  One warp (lanes = the cells of the tile) leaves the three
// columns of J*, their lengths and reciprocal lengths and 1 / det J in shared memory, one tile ahead — the same
// operations on the same numbers as every node and face point would perform (rows of J = c1, c2, c3).
MCFD_HD void NodeAffineRows(const NodeThread &g) {
  const double *gm = g.sG + g.lane;
  double c[3][3], nrm[3][3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
#pragma unroll
    for (int x = 0; x < 3; ++x) c[r][x] = gm[(r * 3 + x) * kTile];
  }
  Cross(c[1], c[2], nrm[0]); Cross(c[2], c[0], nrm[1]); Cross(c[0], c[1], nrm[2]);
  double *a = g.sA + g.lane;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    const double n2 = nrm[d][0] * nrm[d][0] + nrm[d][1] * nrm[d][1] + nrm[d][2] * nrm[d][2];
    const double inv_area = RsqrtC(n2);
#pragma unroll
    for (int x = 0; x < 3; ++x) a[(d * 5 + x) * kTile] = nrm[d][x];
    a[(d * 5 + 3) * kTile] = inv_area;
    a[(d * 5 + 4) * kTile] = n2 * inv_area;
  }
  a[15 * kTile] = RcpC(c[0][0] * nrm[0][0] + c[0][1] * nrm[0][1] + c[0][2] * nrm[0][2]);
}

// what a thread carries from phase to phase
struct NodeRegs {
  double u[5];       // own state
  PrimA q;           // its primitive form
  double inv_det;
  double res[5];
};

// ---- phase A: primitive state, metric, F^_d = F(u) . (column d of J*) for d = xi, eta, zeta -> shared memory ----
template <bool kAffine>
MCFD_HD void NodePhaseA(const StageArgs &p, const LineTables &tab, const NodeThread &g, NodeRegs *t) {
  using S = NodeShape;
#pragma unroll
  for (int c = 0; c < 5; ++c) t->u[c] = g.sU[(g.n * 5 + c) * kTile + g.lane];
  bool ok;
  t->q = ConsToPrimFlat(t->u, &ok);
  if (!ok) t->q = ConsToPrimA(t->u);             // clipped or vacuum state
  double *dst = g.sF + g.n * 15 * kTile + g.lane;
  double s[3], f[5];
  if (kAffine) {   // the columns of J* and 1 / det J of the cell are waiting in shared memory (NodeAffineRows)
    const double *a = g.sA + g.lane;
    t->inv_det = a[15 * kTile];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
#pragma unroll
      for (int x = 0; x < 3; ++x) s[x] = a[(d * 5 + x) * kTile];
      ContravariantFlux(t->u, t->q, s, f);
#pragma unroll
      for (int c = 0; c < 5; ++c) dst[(d * 5 + c) * kTile] = f[c];
    }
    return;
  }
  const LineJacobian J = ComputeLineJacobian(g.sG + g.lane, kTile, tab.xi[g.i], tab.xi[g.j]);
  const double z = tab.xi[g.k];
  double r0[3], r1[3];
#pragma unroll
  for (int x = 0; x < 3; ++x) { r0[x] = fma(z, J.B0[x], J.A0[x]); r1[x] = fma(z, J.B1[x], J.A1[x]); }
  Cross(r1, J.R2, s);
  t->inv_det = RcpC(r0[0] * s[0] + r0[1] * s[1] + r0[2] * s[2]);   // fr/general.hpp:256-266; dg/lobatto.hpp:143-158
  ContravariantFlux(t->u, t->q, s, f);
#pragma unroll
  for (int c = 0; c < 5; ++c) dst[c * kTile] = f[c];
  Cross(J.R2, r0, s);
  ContravariantFlux(t->u, t->q, s, f);
#pragma unroll
  for (int c = 0; c < 5; ++c) dst[(5 + c) * kTile] = f[c];
  Cross(r0, r1, s);
  ContravariantFlux(t->u, t->q, s, f);
#pragma unroll
  for (int c = 0; c < 5; ++c) dst[(10 + c) * kTile] = f[c];
}

// ---- phase B: the flux divergence at the node from the three node lines through it (fr/general.hpp:280-286;
// dg/lobatto.hpp:161-170).  The node's own face-flux terms +-coef_F F^_d(own) (fr/general.hpp:305-313) ride on the
// diagonal of the differentiation matrix.
MCFD_HD void NodePhaseB(const LineTables &tab, const NodeThread &g, NodeRegs *t) {
  using S = NodeShape;
  constexpr int Q = S::Q;
  const double cF = tab.coef_F;
  const int ijk[3] = {g.i, g.j, g.k};
  const int stride[3] = {S::FQ, Q, 1};
#pragma unroll
  for (int c = 0; c < 5; ++c) t->res[c] = 0.0;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    const int me = ijk[d];
    const double *line = g.sF + ((g.n - me * stride[d]) * 15 + d * 5) * kTile + g.lane;   // node with index 0 along d
#pragma unroll
    for (int m = 0; m < Q; ++m) {
      double a = tab.a[me][m];
      if (m == me && me == 0) a -= cF;
      if (m == me && me == Q - 1) a += cF;
#pragma unroll
      for (int c = 0; c < 5; ++c) t->res[c] = fma(a, line[(m * stride[d] * 15 + c) * kTile], t->res[c]);
    }
  }
}

// One face point of the node, any kind: boundary fluxes, clipped states, and every point of the solvers that are
// not written branch-free (exact, AUSM).  fl: flux along the own outward normal, to be multiplied by *scale.
template <int kRiemann>
MCFD_HD void NodePointGeneric(const StageArgs &p, const LineTables &tab, const NodeThread &g, int f, int q,
    const double *nr, double ox, double oy, double oz, double area, double *fl, double *scale) {
  using S = NodeShape;
  const int nb = g.sNbr[f * kTile + g.lane], code = g.sCode[f * kTile + g.lane];
  const double sgn = (f == 2 || f == 3 || f == 5) ? 1.0 : -1.0;
  double uo[5];
  const double *own = p.u_in + ((size_t)g.tile * S::F + g.n * 5) * kTile + g.lane;
#pragma unroll
  for (int c = 0; c < 5; ++c) uo[c] = own[c * kTile];
  *scale = -tab.coef_f * area;     // the lifted term  -g' |n| f*  (FR) or  -|n| f* / w_end  (DG)
  if (nb >= 0) {
    // posed in the holder's frame like the reference (the exact solver's Newton iteration starts from
    // the left pressure): swap the states, flip the normal and negate when the neighbour holds the face
    const double *src = g.sN + ((f * S::FQ + q) * 5) * kTile + g.lane;
    const bool holder = (code & 64) != 0;
    const double hs = holder ? 1.0 : -1.0;
    double ul[5], ur[5];
#pragma unroll
    for (int c = 0; c < 5; ++c) { const double v = src[c * kTile]; ul[c] = holder ? uo[c] : v; ur[c] = holder ? v : uo[c]; }
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
      const double *gv = p.bvals + ((size_t)bf * S::FQ + q) * 5;
#pragma unroll
      for (int c = 0; c < 5; ++c) { ui.v[c] = uo[c]; given.v[c] = gv[c]; }
      const State5 bfl = BoundaryFluxOutOfLine<kRiemann>(kind, ui, given, ox, oy, oz);
#pragma unroll
      for (int c = 0; c < 5; ++c) fl[c] = bfl.v[c];
    }
  }
}
// the rare points of the branch-free HLLC path: one shared out-of-line copy.  Takes and returns everything in
// registers or shared memory (no stack): the lifted term scale * flux replaces the neighbour's state in the thread's
// own slot of sN.
MCFD_NOINLINE void NodePointRare(const StageArgs &p, const LineTables &tab, int tid, int tile, int buf, double *smem,
    int f, int q, double n0, double n1, double n2, double ox, double oy, double oz, double area) {
  const NodeThread g(p, tid, tile, buf, smem);
  const double nr[3] = {n0, n1, n2};
  double fl[5], scale;
  NodePointGeneric<kHllc>(p, tab, g, f, q, nr, ox, oy, oz, area, fl, &scale);
  double *slot = g.sN + ((f * NodeShape::FQ + q) * 5) * kTile + g.lane;
#pragma unroll
  for (int c = 0; c < 5; ++c) slot[c * kTile] = scale * fl[c];
}

// rows of the Jacobian at the node that the normal of its dd-faces needs (column dd of J* = the cross product of the
// other two rows; the same FMA nesting as ComputeLineJacobian), straight from the tile's geometry rows
MCFD_HD void NodeNormal(const double *g, double x, double e, double z, int dd, double *nr) {
  double ra[3], rb[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const double c4 = g[(9 + k) * kTile], c6 = g[(15 + k) * kTile], c7 = g[(18 + k) * kTile];
    const double b1 = fma(x, c7, g[(12 + k) * kTile]);
    const double r0 = fma(z, fma(e, c7, c6), fma(e, c4, g[(0 + k) * kTile]));
    const double r1 = fma(z, b1, fma(x, c4, g[(3 + k) * kTile]));
    const double r2 = fma(e, b1, fma(x, c6, g[(6 + k) * kTile]));
    ra[k] = dd == 0 ? r1 : (dd == 1 ? r2 : r0);
    rb[k] = dd == 0 ? r2 : (dd == 1 ? r0 : r1);
  }
  Cross(ra, rb, nr);
}

// ---- phase C: the node's face points (0..3).  Everything a point needs is the thread's own: its primitive state in
// registers, the neighbour's state in its slot of sN, the metric from the tile's geometry rows.  The lifted term
// goes straight into the residual in registers (fr/general.hpp:296-315; dg/lobatto.hpp:171-191).
template <int kRiemann, bool kAffine>
MCFD_HD void NodePhaseC(const StageArgs &p, const LineTables &tab, const NodeThread &g, double *smem, NodeRegs *t) {
  using S = NodeShape;
  if (!g.active) return;
#pragma unroll
  for (int dd = 0; dd < 3; ++dd) {
    const int f = NodeFace(g.i, g.j, g.k, dd);
    if (f < 0) continue;                       // warp uniform
    const int q = tab.node_q[g.n][dd];
    double nr[3], inv_area, area;
    const double sgn = (f == 2 || f == 3 || f == 5) ? 1.0 : -1.0;
    if (kAffine) {   // the cell's column dd of J*, its length and the reciprocal wait in shared memory
      const double *a = g.sA + dd * 5 * kTile + g.lane;
      nr[0] = a[0]; nr[1] = a[kTile]; nr[2] = a[2 * kTile];
      inv_area = a[3 * kTile]; area = a[4 * kTile];
    } else {
      NodeNormal(g.sG + g.lane, tab.xi[g.i], tab.xi[g.j], tab.xi[g.k], dd, nr);
      const double n2 = nr[0] * nr[0] + nr[1] * nr[1] + nr[2] * nr[2];
      inv_area = RsqrtC(n2);
      area = n2 * inv_area;
    }
    const double os_ = sgn * inv_area;
    const double ox = os_ * nr[0], oy = os_ * nr[1], oz = os_ * nr[2];   // own outward unit normal
    double fl[5], scale;
    if (kRiemann == kHllc) {
      const double *v = g.sN + ((f * S::FQ + q) * 5) * kTile + g.lane;
      double r[5];
#pragma unroll
      for (int c = 0; c < 5; ++c) r[c] = v[c * kTile];
      bool ok;
      const PrimA R = ConsToPrimFlat(r, &ok);
      HllcFlux(t->q, R, ox, oy, oz, fl);
      scale = -tab.coef_f * area;
      if (!(ok && g.sNbr[f * kTile + g.lane] >= 0 && t->q.a > 0)) {   // boundary face, clipped or vacuum state
        NodePointRare(p, tab, g.n * 32 + g.lane, g.tile, g.buf, smem, f, q, nr[0], nr[1], nr[2], ox, oy, oz, area);
#pragma unroll
        for (int c = 0; c < 5; ++c) fl[c] = v[c * kTile];
        scale = 1.0;
      }
    } else {
      NodePointGeneric<kRiemann>(p, tab, g, f, q, nr, ox, oy, oz, area, fl, &scale);
    }
    double *acc = g.sF + (g.n * 15 + 6) * kTile + g.lane;   // the residual waits in the thread's own rows (NodeStash)
#pragma unroll
    for (int c = 0; c < 5; ++c) acc[c * kTile] = fma(scale, fl[c], acc[c * kTile]);
  }
}

// ---- phase D: 1 / det J, RK update (rk.hpp:16-56), store; a send cell's row also goes packed to the send stage ----
// Between the second barrier and the next tile's phase A a thread's own 15 rows of sF are nobody's: its state and
// 1 / det J wait there while the face points need the registers, and the residual gathers the lifted terms there.
MCFD_HD void NodeStash(const NodeThread &g, const NodeRegs &t) {
  double *own = g.sF + g.n * 15 * kTile + g.lane;
#pragma unroll
  for (int c = 0; c < 5; ++c) own[c * kTile] = t.u[c];
  own[5 * kTile] = t.inv_det;
#pragma unroll
  for (int c = 0; c < 5; ++c) own[(6 + c) * kTile] = t.res[c];
}
MCFD_HD void NodePhaseD(const StageArgs &p, const NodeThread &g) {
  using S = NodeShape;
  if (!g.active) return;
  const double *own = g.sF + g.n * 15 * kTile + g.lane;
  double u[5];
#pragma unroll
  for (int c = 0; c < 5; ++c) u[c] = own[c * kTile];
  const double inv_det = own[5 * kTile];
  double r[5];
#pragma unroll
  for (int c = 0; c < 5; ++c) r[c] = own[(6 + c) * kTile];
  const size_t row0 = ((size_t)g.tile * S::F + g.n * 5) * kTile + g.lane;
  const double dt = p.dt_dev ? *p.dt_dev : p.dt;
  const int mode = p.mode;
  // kModeEuler: w;  kModeRk3b: (u0 3 + w) / 4;  kModeRk3c: (w 2 + u0) / 3;  kModeRk2b: (w + u0) / 2
  const double cw = mode == kModeRk3c ? 2.0 : 1.0;
  const double c0 = mode == kModeRk3b ? 3.0 : (mode == kModeRk3c || mode == kModeRk2b ? 1.0 : 0.0);
  const double cs_inv = mode == kModeRk3b ? 4.0 : (mode == kModeRk3c ? 3.0 : (mode == kModeRk2b ? 2.0 : 1.0));
  const double cs = mode == kModeRk3b ? 0.25 : (mode == kModeRk3c ? 1.0 / 3 : (mode == kModeRk2b ? 0.5 : 1.0));
  double u0[5], out[5];
#pragma unroll
  for (int c = 0; c < 5; ++c) u0[c] = mode >= kModeRk3b ? p.u0[row0 + c * kTile] : 0.0;
#pragma unroll
  for (int c = 0; c < 5; ++c) {
    const double res = r[c] * inv_det;
    const double w = fma(res, dt, u[c]);
    // (w cw + u0 c0) cs with cs = 1, 1/4, 1/3, 1/2 (rk.hpp:16-49); the last two operations make the quotient by 3
    // correctly rounded in all but rare half-way cases and change nothing for the powers of two
    const double sum = fma(u0[c], c0, w * cw);
    const double qv = sum * cs;
    const double val = fma(fma(-cs_inv, qv, sum), cs, qv);
    out[c] = mode == kModeResidual ? res : val;
    p.u_out[row0 + c * kTile] = out[c];
  }
  if (p.peer.on && g.cell >= p.peer.first_cell) {   // a second copy of a send cell's rows, packed for the copy engines
    const int t0 = p.peer.tgt_offset[g.cell - p.peer.first_cell], t1 = p.peer.tgt_offset[g.cell - p.peer.first_cell + 1];
    for (int tt = t0; tt < t1; ++tt) {
      double *to = p.peer.send + (size_t)(g.n * 5) * p.peer.n_send + p.peer.tgt_item[tt];
#pragma unroll
      for (int c = 0; c < 5; ++c) to[(size_t)c * p.peer.n_send] = out[c];
    }
  }
}

#ifdef __CUDACC__
// tile `tile` (absolute) -> buffer `buf`: wait for the neighbours' delivery if the tile has inter cells, arm the
// mbarrier, issue the four bulk copies.  One thread.
__device__ __forceinline__ void NodeProduce(const StageArgs &p, int tile, int buf, double *smem, bool *waited) {
  using S = NodeShape;
  unsigned char *bytes = reinterpret_cast<unsigned char *>(smem + S::kDoubles);
  void *bar = bytes + S::kOffBar + 8 * buf;
  if (p.peer.on && tile >= p.peer.first_block && !*waited) {   // ghost states come from the receive stage: delivered?
    const unsigned long long want = p.peer.counts[kPeerPlain] + p.peer.counts[kPeerStageBase] + (unsigned long long)p.peer.wait_add;
    for (int x = 0; x < p.peer.n_neighbours; ++x) if (!SpinUntil(p.peer.flags + x, want)) *p.peer.error = 1;
    *waited = true;
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the buffers' last generic reads are behind a barrier
  MbarExpect(bar, S::kBulkBytes);
  BulkCopy(smem + S::kOffU, p.u_in + (size_t)tile * S::F * kTile, S::F * kTile * 8, bar);
  BulkCopy(smem + S::kOffG + buf * kGeomRows * kTile, p.geom + (size_t)tile * kGeomRows * kTile, kGeomRows * kTile * 8, bar);
  BulkCopy(bytes + S::kOffNbr + buf * S::kNbrBytes, p.nbr + (size_t)tile * 6 * kTile, S::kNbrBytes, bar);
  BulkCopy(bytes + S::kOffCode + buf * S::kCodeBytes, p.code + (size_t)tile * 6 * kTile, S::kCodeBytes, bar);
}

template <int kRiemann, bool kAffine>
__global__ void __launch_bounds__(NodeShape::kThreads, 1)
NodeStageKernel(const __grid_constant__ StageArgs p, const __grid_constant__ LineTables tab, int n_tiles) {
  using S = NodeShape;
  extern __shared__ __align__(16) double smem[];
  const int tid = threadIdx.x;
  unsigned char *bytes = reinterpret_cast<unsigned char *>(smem + S::kDoubles);
  const bool producer = tid == S::kProducerWarp * 32;
  bool waited = false;
  for (int x = tid; x < 6 * 8 * S::FQ; x += S::kThreads) bytes[S::kOffTab + x] = p.line_bytes[x];
  if (producer) {
    MbarInit(bytes + S::kOffBar, 1);
    MbarInit(bytes + S::kOffBar + 8, 1);
    MbarInit(bytes + S::kOffBar + 16, 32);   // the affine metric of a tile: the 32 lanes of the warp that computes it
    MbarInit(bytes + S::kOffBar + 24, 32);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  int t = blockIdx.x;
  if (t >= n_tiles) return;
  if (producer) NodeProduce(p, p.block_begin + t, 0, smem, &waited);
  MbarWait(bytes + S::kOffBar, 0);
  {
    const NodeThread g(p, tid, p.block_begin + t, 0, smem);
    NodeGather(p, tab, g);
    asm volatile("cp.async.commit_group;" ::: "memory");
    if (kAffine && g.n == S::kProducerWarp) { NodeAffineRows(g); MbarArrive(bytes + S::kOffBar + 16); }
  }
  for (int it = 0; t < n_tiles; t += gridDim.x, ++it) {
    const int buf = it & 1;
    const bool more = t + (int)gridDim.x < n_tiles;
    const NodeThread g(p, tid, p.block_begin + t, buf, smem);
    MbarWait(bytes + S::kOffBar + 8 * buf, (it >> 1) & 1);
    if (kAffine) MbarWait(bytes + S::kOffBar + 16 + 8 * buf, (it >> 1) & 1);
    if (p.mode >= kModeRk3b && g.active) {   // the second stream of the update: bring it towards the SM early
      const double *u0 = p.u0 + ((size_t)g.tile * S::F + g.n * 5) * kTile + g.lane;
#pragma unroll
      for (int c = 0; c < 5; ++c) MCFD_PREFETCH_L2(u0 + c * kTile);
    }
    NodeRegs r;
    NodePhaseA<kAffine>(p, tab, g, &r);
    __syncthreads();                       // F^ complete; the own-state buffer has been read
    if (producer && more) NodeProduce(p, p.block_begin + t + gridDim.x, buf ^ 1, smem, &waited);
    NodePhaseB(tab, g, &r);
    __syncthreads();                       // F^ may be overwritten by the next tile
    NodeStash(g, r);
    if (kAffine && more && g.n == S::kProducerWarp) {   // this warp has no face point: the next tile's metric meanwhile.  Its
      // buffer was last read in the tile before this one, which every warp left before this tile's first barrier.
      MbarWait(bytes + S::kOffBar + 8 * (buf ^ 1), ((it + 1) >> 1) & 1);
      NodeAffineRows(NodeThread(p, tid, p.block_begin + t + gridDim.x, buf ^ 1, smem));
      MbarArrive(bytes + S::kOffBar + 16 + 8 * (buf ^ 1));
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");   // my gathers for this tile
    NodePhaseC<kRiemann, kAffine>(p, tab, g, smem, &r);
    if (more) {                            // my slots of sN are free: the next tile's neighbours
      MbarWait(bytes + S::kOffBar + 8 * (buf ^ 1), ((it + 1) >> 1) & 1);
      const NodeThread gn(p, tid, p.block_begin + t + gridDim.x, buf ^ 1, smem);
      NodeGather(p, tab, gn);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    NodePhaseD(p, g);
  }
}
#else
// the same tile by tile on the host (tests/host_emulation): every thread of a phase before the next phase
template <int kRiemann, bool kAffine>
void RunNodeStageTileOnHost(const StageArgs &p, const LineTables &tab, int tile, double *smem) {
  using S = NodeShape;
  unsigned char *bytes = reinterpret_cast<unsigned char *>(smem + S::kDoubles);
  for (int x = 0; x < 6 * 8 * S::FQ; ++x) bytes[S::kOffTab + x] = p.line_bytes[x];
  for (int x = 0; x < S::F * kTile; ++x) smem[S::kOffU + x] = p.u_in[(size_t)tile * S::F * kTile + x];
  for (int x = 0; x < kGeomRows * kTile; ++x) smem[S::kOffG + x] = p.geom[(size_t)tile * kGeomRows * kTile + x];
  for (int x = 0; x < 6 * kTile; ++x) {
    reinterpret_cast<int *>(bytes + S::kOffNbr)[x] = p.nbr[(size_t)tile * 6 * kTile + x];
    bytes[S::kOffCode + x] = p.code[(size_t)tile * 6 * kTile + x];
  }
  static NodeRegs regs[S::kThreads];
  if (kAffine) for (int lane = 0; lane < kTile; ++lane) NodeAffineRows(NodeThread(p, S::kProducerWarp * 32 + lane, tile, 0, smem));
  for (int t = 0; t < S::kThreads; ++t) {
    const NodeThread g(p, t, tile, 0, smem);
    NodeGather(p, tab, g);
    NodePhaseA<kAffine>(p, tab, g, &regs[t]);
  }
  for (int t = 0; t < S::kThreads; ++t) NodePhaseB(tab, NodeThread(p, t, tile, 0, smem), &regs[t]);
  for (int t = 0; t < S::kThreads; ++t) {
    const NodeThread g(p, t, tile, 0, smem);
    NodeStash(g, regs[t]);
    NodePhaseC<kRiemann, kAffine>(p, tab, g, smem, &regs[t]);
    NodePhaseD(p, g);
  }
}
#endif

}  // namespace mcfd
#endif  // MINICFD_B200_STAGE_NODES_CUH_
