/* TEST INFRASTRUCTURE — NOT PART OF THE PRODUCT PATH.
 *
 * minicfd_oracle.c: a plain-C, single-threaded restatement of the reference's algorithm for the hot
 * path (pvcStillinGradSchool/miniCFD @ e14f896): FR-Lobatto / DG-Lobatto Euler residual on
 * Hexahedron8 cells, Exact / HLLC / AUSM Riemann solvers, the inviscid boundary kinds, RK1/2/3 and the
 * CFL time step.  It is deliberately written the way the reference works — face loops that scatter to
 * holder and sharer, flux points matched to solution points by COORDINATES, the Riemann problem solved
 * in the rotated (normal, tangent, bitangent) frame, dense volume loops — and not the way the CUDA
 * kernels work, so that the two are independent.  Citations are paths under the reference's
 * include/mini/ unless stated otherwise.
 *
 * PINNED: tests/test_oracle_pinned.py checks this file against the reference's own code compiled from
 * /root/reference (oracle/_ref via ref_harness.cpp) and against the golden vectors under
 * tests/golden/ generated from it (tests/golden/make_golden.py), and the Riemann solvers against the
 * Toro known-answer states the reference's tests use (test/riemann/euler/exact.cpp:57-138).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this library.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define GAMMA 1.4
#define GAS_R 287.05
#define MAXQ 4
#define MAXN 64

/* ---- riemann/euler/types.hpp ----------------------------------------------------------------- */
typedef struct { double rho, u, v, w, p; } prim_t;

static prim_t cons_to_prim(const double *c) { /* types.hpp:289-301 with :194-202 */
  prim_t q = {c[0], c[1], c[2], c[3], c[4]};
  if (q.rho < 0) memset(&q, 0, sizeof q);
  if (q.p < 0) memset(&q, 0, sizeof q);
  if (q.rho != 0) {
    q.u /= q.rho; q.v /= q.rho; q.w /= q.rho;
    q.p -= q.rho * (0.5 * (q.u * q.u + q.v * q.v + q.w * q.w));
    q.p *= GAMMA - 1;
    if (q.rho < 0) memset(&q, 0, sizeof q);
    if (q.p < 0) memset(&q, 0, sizeof q);
  }
  return q;
}
static double sound_speed(double rho, double p) { /* types.hpp:252-254 */
  return (rho > 0 && p > 0) ? sqrt(GAMMA * p / rho) : 0;
}
static void prim_to_flux(const prim_t *q, double *f) { /* types.hpp:313-320 */
  double ke = 0.5 * (q->u * q->u + q->v * q->v + q->w * q->w);
  double c[5] = {q->rho, q->rho * q->u, q->rho * q->v, q->rho * q->w, q->p * (1 / (GAMMA - 1)) + q->rho * ke};
  for (int k = 0; k < 5; ++k) f[k] = c[k] * q->u;
  f[1] += q->p;
  f[4] += q->p * q->u;
}
static void flux_matrix(const double *c, double m[5][3]) { /* types.hpp:322-333 */
  prim_t q = cons_to_prim(c);
  double vel[3] = {q.u, q.v, q.w};
  double h0 = c[4] + q.p;
  for (int d = 0; d < 3; ++d) {
    m[0][d] = c[1 + d];
    m[1][d] = c[1] * vel[d];
    m[2][d] = c[2] * vel[d];
    m[3][d] = c[3] * vel[d];
    m[1 + d][d] += q.p;
    m[4][d] = h0 * vel[d];
  }
}

/* ---- riemann/euler/hllc.hpp -------------------------------------------------------------------- */
static double hllc_q(double pe, double pk) { /* hllc.hpp:69-77 */
  if (pe <= pk) return 1.0;
  return sqrt(1 + ((GAMMA + 1) / 2) * (pe / pk - 1) / GAMMA);
}
static void hllc_star(const prim_t *s, double wk, double ws, double *f) { /* hllc.hpp:94-109 */
  prim_to_flux(s, f);
  double ke = 0.5 * (s->u * s->u + s->v * s->v + s->w * s->w);
  double uk[5] = {s->rho, s->rho * s->u, s->rho * s->v, s->rho * s->w, s->p * (1 / (GAMMA - 1)) + s->rho * ke};
  double temp = s->rho * (wk - s->u) / (wk - ws);
  double us[5] = {temp, temp * ws, temp * s->v, temp * s->w, 0};
  us[4] = uk[4] / s->rho + (ws - s->u) * (ws + s->p / (s->rho * (wk - s->u)));
  us[4] *= temp;
  for (int k = 0; k < 5; ++k) f[k] += (us[k] - uk[k]) * wk;
}
static void hllc(const prim_t *l, const prim_t *r, double *f) { /* hllc.hpp:30-68 */
  double rho_avg = (l->rho + r->rho) / 2;
  double al = sound_speed(l->rho, l->p), ar = sound_speed(r->rho, r->p);
  double a_avg = (al + ar) / 2;
  double p_pvrs = (l->p + r->p - (r->u - l->u) * rho_avg * a_avg) / 2;
  double pe = p_pvrs > 0 ? p_pvrs : 0;
  double wl = l->u - al * hllc_q(pe, l->p);
  double wr = r->u + ar * hllc_q(pe, r->p);
  double ws = (r->p - l->p + l->rho * l->u * (wl - l->u) - r->rho * r->u * (wr - r->u)) /
              (l->rho * (wl - l->u) - r->rho * (wr - r->u));
  if (0.0 <= wl) prim_to_flux(l, f);
  else if (wr <= 0.0) prim_to_flux(r, f);
  else if (0.0 <= ws) hllc_star(l, wl, ws, f);
  else hllc_star(r, wr, ws, f);
}

/* ---- riemann/euler/ausm.hpp -------------------------------------------------------------------- */
static void ausm_signed(const prim_t *s, int negative, double *f) { /* ausm.hpp:52-71 */
  double a = sound_speed(s->rho, s->p);
  double h = a * a / (GAMMA - 1) + 0.5 * (s->u * s->u + s->v * s->v + s->w * s->w);
  double fl[5] = {1, s->u, s->v, s->w, h};
  double mach = s->u / a, p = s->p;
  double sm = (1 + (negative ? -mach : mach)) * 0.5;
  if (mach >= -1 && mach <= 1) {
    mach = negative ? -(sm * sm) : sm * sm;
    p *= sm;
  } else if (sm < 0) {
    mach = 0; p = 0;
  }
  for (int k = 0; k < 5; ++k) f[k] = fl[k] * (s->rho * a * mach);
  f[1] += p;
}
static void ausm(const prim_t *l, const prim_t *r, double *f) { /* ausm.hpp:30-33 */
  double a[5], b[5];
  ausm_signed(l, 0, a);
  ausm_signed(r, 1, b);
  for (int k = 0; k < 5; ++k) f[k] = a[k] + b[k];
}

/* ---- riemann/euler/exact.hpp ------------------------------------------------------------------- */
typedef struct { double rho_b, p_b, a_b, p_const; } speed_change_t;
static speed_change_t sc_make(const prim_t *b) { /* exact.hpp:155-160 */
  speed_change_t s = {b->rho, b->p, b->rho == 0 ? 0 : sound_speed(b->rho, b->p), b->p * ((GAMMA - 1) / 2)};
  return s;
}
static double sc_P(const speed_change_t *s, double pa) { return ((GAMMA + 1) / 2) * pa + s->p_const; }
static double sc_value(const speed_change_t *s, double pa) { /* exact.hpp:161-174 */
  if (pa >= s->p_b) return (pa - s->p_b) / sqrt(s->rho_b * sc_P(s, pa));
  double v = pow(pa / s->p_b, ((GAMMA - 1) / 2) / GAMMA) - 1;
  return v * ((2 / (GAMMA - 1)) * s->a_b);
}
static double sc_prime(const speed_change_t *s, double pa) { /* exact.hpp:175-191 */
  if (pa >= s->p_b) {
    double p = sc_P(s, pa);
    double v = ((GAMMA + 1) / 4) * (s->p_b - pa);
    return (1 + v / p) / sqrt(s->rho_b * p);
  }
  return pow(pa / s->p_b, -((GAMMA + 1) / 2) / GAMMA) / (s->rho_b * s->a_b);
}
static prim_t inside_expansion(double gri_1, double gri_2) { /* exact.hpp:255-266 */
  double a = ((GAMMA - 1) / (GAMMA + 1)) * gri_2, a2 = a * a;
  prim_t x = {0, 0, 0, 0, 0};
  x.rho = pow(a2 / gri_1 * (1 / GAMMA), 1 / (GAMMA - 1));
  x.u = a;
  x.p = a2 * x.rho * (1 / GAMMA);
  return x;
}
static prim_t near_shock(int field, const prim_t *b, prim_t *after) { /* exact.hpp:203-234 */
  double divisor = (after->u - b->u) * b->rho;
  double su = b->u + (divisor ? (after->p - b->p) / divisor : b->u);
  int is_after = field == 1 ? (su < 0) : (su > 0);
  if (is_after) {
    after->rho = b->rho * ((b->u - su) / (after->u - su));
    return *after;
  }
  return *b;
}
static prim_t near_expansion(int field, const prim_t *b, prim_t *after) { /* exact.hpp:236-292 */
  double a_before = sound_speed(b->rho, b->p);
  double gri_1 = b->p / pow(b->rho, GAMMA);
  double gri_2 = field == 1 ? b->u + (2 / (GAMMA - 1)) * a_before : b->u - (2 / (GAMMA - 1)) * a_before;
  double a_after = field == 1 ? a_before + ((GAMMA - 1) / 2) * (b->u - after->u)
                              : a_before - ((GAMMA - 1) / 2) * (b->u - after->u);
  int t_after = field == 1 ? (after->u - a_after < 0) : (after->u + a_after > 0);
  if (t_after) {
    after->rho = GAMMA * after->p / (a_after * a_after);
    return *after;
  }
  int t_before = field == 1 ? (b->u - a_before > 0) : (b->u + a_before < 0);
  if (t_before) return *b;
  return inside_expansion(gri_1, gri_2);
}
/* exact.hpp:62-138; *star_u as the solver member of :29 (fresh solver: 0) */
static prim_t exact_axis(const prim_t *l, const prim_t *r, double *star_u) {
  prim_t zero = {0, 0, 0, 0, 0};
  if (l->rho > 0) {
    if (!(r->rho > 0)) {
      double al = sound_speed(l->rho, l->p);
      if (l->u >= al) return *l;
      double us = l->u + al * (2 / (GAMMA - 1));
      if (us <= 0) return zero;
      return inside_expansion(l->p / pow(l->rho, GAMMA), us);
    }
  } else {
    if (r->rho > 0) {
      double ar = sound_speed(r->rho, r->p);
      if (r->u + ar <= 0) return *r;
      double us = r->u - ar * (2 / (GAMMA - 1));
      if (us >= 0) return zero;
      return inside_expansion(r->p / pow(r->rho, GAMMA), us);
    }
    return zero;
  }
  double du = r->u - l->u;
  speed_change_t cl = sc_make(l), cr = sc_make(r);
#define F_OF(x) (sc_value(&cl, (x)) + sc_value(&cr, (x)) + du)
  if (F_OF(0.0) < 0) {
    double x = l->p; /* FindRoot, exact.hpp:140-152 */
    while (F_OF(x) > 0) x *= 0.5;
    while (F_OF(x) < -1e-8) x -= F_OF(x) / (sc_prime(&cl, x) + sc_prime(&cr, x));
    prim_t star = {0, 0, 0, 0, 0};
    star.p = x;
    star.u = 0.5 * (r->u + sc_value(&cr, x) + l->u - sc_value(&cl, x));
    *star_u = star.u;
    if (0 < star.u) return star.p >= l->p ? near_shock(1, l, &star) : near_expansion(1, l, &star);
    return star.p >= r->p ? near_shock(3, r, &star) : near_expansion(3, r, &star);
  }
#undef F_OF
  double la = sound_speed(l->rho, l->p); /* PrimitiveNearVacuum, exact.hpp:293-320 */
  if (l->u > la) return *l;
  double ra = sound_speed(r->rho, r->p);
  if (r->u + ra < 0) return *r;
  double g1 = l->p / pow(l->rho, GAMMA), g2 = l->u + la * (2 / (GAMMA - 1));
  if (g2 >= 0) return inside_expansion(g1, g2);
  g1 = r->p / pow(r->rho, GAMMA);
  g2 = r->u - ra * (2 / (GAMMA - 1));
  if (g2 < 0) return inside_expansion(g1, g2);
  return zero;
}
static void exact(const prim_t *l, const prim_t *r, double *f) { /* exact.hpp:36-60 */
  double star_u = 0;
  prim_t s = exact_axis(l, r, &star_u);
  s.v = star_u > 0 ? l->v : r->v;
  s.w = star_u > 0 ? l->w : r->w;
  prim_to_flux(&s, f);
}

/* ---- riemann/rotated/euler.hpp ----------------------------------------------------------------- */
typedef struct { double a[3], b[3], c[3]; } frame_t; /* normal, tangent, bitangent */

static void to_normal(const frame_t *fr, double *m) { /* rotated/euler.hpp:117-125 */
  double x = m[0], y = m[1], z = m[2];
  m[0] = x * fr->a[0] + y * fr->a[1] + z * fr->a[2];
  m[1] = x * fr->b[0] + y * fr->b[1] + z * fr->b[2];
  m[2] = x * fr->c[0] + y * fr->c[1] + z * fr->c[2];
}
static void to_global(const frame_t *fr, double *m) { /* rotated/euler.hpp:126-134 */
  double a = m[0], b = m[1], c = m[2];
  m[0] = a * fr->a[0] + b * fr->b[0] + c * fr->c[0];
  m[1] = a * fr->a[1] + b * fr->b[1] + c * fr->c[1];
  m[2] = a * fr->a[2] + b * fr->b[2] + c * fr->c[2];
}
static void solve_1d(int riemann, const prim_t *l, const prim_t *r, double *f) {
  if (riemann == 0) exact(l, r, f);
  else if (riemann == 1) hllc(l, r, f);
  else ausm(l, r, f);
}
static void flux_upwind(int riemann, const frame_t *fr, const double *cl, const double *cr, double *f) {
  prim_t l = cons_to_prim(cl), r = cons_to_prim(cr); /* rotated/euler.hpp:153-159, 74-88 */
  to_normal(fr, &l.u);
  to_normal(fr, &r.u);
  solve_1d(riemann, &l, &r, f);
  to_global(fr, &f[1]);
}
static void prim_flux_global(const frame_t *fr, prim_t q, double *f) { /* rotated/euler.hpp:68-73 */
  to_normal(fr, &q.u);
  prim_to_flux(&q, f);
  to_global(fr, &f[1]);
}
/* kinds: 0 wall 1 supersonic outlet 2 supersonic inlet 3 subsonic inlet 4 subsonic outlet 5 smart */
static void boundary_flux(int riemann, int kind, const frame_t *fr, const double *ci, const double *given,
    double *f) {
  prim_t q;
  switch (kind) {
  case 0: /* rotated/euler.hpp:160-174 */
    q = cons_to_prim(ci);
    f[0] = 0; f[1] = q.p; f[2] = 0; f[3] = 0; f[4] = 0;
    to_global(fr, &f[1]);
    break;
  case 1: prim_flux_global(fr, cons_to_prim(ci), f); break;    /* :148-151 */
  case 2: prim_flux_global(fr, cons_to_prim(given), f); break; /* :144-147 */
  case 3: { /* :181-236 */
    q = cons_to_prim(ci);
    double c_int = sound_speed(q.rho, q.p);
    to_normal(fr, &q.u);
    double ri = -q.u - c_int * (2 / (GAMMA - 1));
    double uc = given[1], vc = given[2], wc = given[3], tt = given[4];
    double A = fr->a[0] * uc + fr->a[1] * vc + fr->a[2] * wc, A2 = A * A, gr = GAMMA * GAS_R;
    double qa = A2 + 2 / (GAMMA - 1), qb = ri, qc = ((GAMMA - 1) / 2) * (ri * ri) - A2 * gr * tt;
    double sd = sqrt(qb * qb - qa * qc);
    double x1 = (-qb + sd) / qa, x2 = (-qb - sd) / qa;
    double cb = x1 > x2 ? x1 : x2;
    double tb = cb * cb / gr;
    double mb = sqrt((tt / tb - 1) / ((GAMMA - 1) / 2));
    double pb = given[0] / pow(1 + ((GAMMA - 1) / 2) * mb * mb, GAMMA / (GAMMA - 1));
    double sp = mb * cb;
    prim_t b = {pb / GAS_R / tb, uc * sp, vc * sp, wc * sp, pb};
    prim_flux_global(fr, b, f);
    break;
  }
  case 4: { /* :260-280 */
    q = cons_to_prim(ci);
    double rt = q.p / q.rho;
    q.p = given[4];
    q.rho = q.p / rt;
    prim_flux_global(fr, q, f);
    break;
  }
  default: flux_upwind(riemann, fr, ci, given, f); /* :296-299 */
  }
}

/* ---- fields (same formulas as ref_harness.cpp::EvalField) -------------------------------------- */
static void eval_field(int kind, const double *prm, const double *x, double t, double *v) {
  if (kind == 0) {
    for (int k = 0; k < 5; ++k) v[k] = prm[k];
  } else if (kind == 1) {
    const double g = GAMMA, pi = 3.14159265358979323846;
    double dx = x[0] - prm[0] - prm[2] * t, dy = x[1] - prm[1] - prm[3] * t, beta = prm[5];
    double e = exp(0.5 * (1 - (dx * dx + dy * dy)));
    double du = -beta / (2 * pi) * e * dy, dv = beta / (2 * pi) * e * dx;
    double T = 1 - (g - 1) * beta * beta / (8 * g * pi * pi) * e * e;
    double rho = pow(T, 1 / (g - 1)), p = pow(rho, g);
    double u = prm[2] + du, vv = prm[3] + dv, w = prm[4];
    v[0] = rho; v[1] = rho * u; v[2] = rho * vv; v[3] = rho * w;
    v[4] = p / (g - 1) + 0.5 * rho * (u * u + vv * vv + w * w);
  } else {
    double a = prm[0] ? prm[0] : 1.0;
    double rho = 1.0 + 0.2 * sin(a * (x[0] + 2 * x[1] - x[2]) + 0.3 * t);
    double u = 0.5 + 0.3 * cos(a * (x[0] - x[1]) + 0.1);
    double vv = -0.2 + 0.25 * sin(a * (x[1] + x[2]) + 0.7);
    double w = 0.1 + 0.2 * cos(a * (x[2] - 2 * x[0]) - 0.4);
    double p = 1.0 + 0.3 * sin(a * (x[0] + x[1] + x[2]) + 1.1);
    v[0] = rho; v[1] = rho * u; v[2] = rho * vv; v[3] = rho * w;
    v[4] = p / 0.4 + 0.5 * rho * (u * u + vv * vv + w * w);
  }
}

/* ---- discretisation ------------------------------------------------------------------------------ */
static const int FACE_NODES[6][4] = { /* coordinate/hexahedron.hpp:378-389 */
  {0, 3, 2, 1}, {0, 1, 5, 4}, {1, 2, 6, 5}, {2, 3, 7, 6}, {0, 4, 7, 3}, {4, 5, 6, 7}};
static const double VERT[8][3] = {{-1, -1, -1}, {1, -1, -1}, {1, 1, -1}, {-1, 1, -1},
                                  {-1, -1, 1}, {1, -1, 1}, {1, 1, 1}, {-1, 1, 1}};
static const double QUAD[4][2] = {{-1, -1}, {1, -1}, {1, 1}, {-1, 1}};

static void lobatto(int q, double *x, double *w) { /* integrator/lobatto.hpp:38-57 */
  if (q == 3) { x[0] = -1; x[1] = 0; x[2] = 1; w[0] = 1.0 / 3; w[1] = 4.0 / 3; w[2] = 1.0 / 3; }
  else { x[0] = -1; x[1] = -sqrt(1.0 / 5); x[2] = sqrt(1.0 / 5); x[3] = 1;
         w[0] = 1.0 / 6; w[1] = 5.0 / 6; w[2] = 5.0 / 6; w[3] = 1.0 / 6; }
}
static double lagrange_d(int q, const double *x, int m, int i) { /* l_m'(x_i), basis/lagrange.hpp:130-133 */
  double sum = 0;
  for (int k = 0; k < q; ++k) {
    if (k == m) continue;
    double prod = 1.0 / (x[m] - x[k]);
    for (int l = 0; l < q; ++l) if (l != m && l != k) prod *= (x[i] - x[l]) / (x[m] - x[l]);
    sum += prod;
  }
  return sum;
}
static double lagrange_d2(int q, const double *x, int m, int i) { /* l_m''(x_i), basis/lagrange.hpp:130-133 with k = 2 */
  double sum = 0;
  for (int k = 0; k < q; ++k) {
    if (k == m) continue;
    for (int j = 0; j < q; ++j) {
      if (j == m || j == k) continue;
      double prod = 1.0 / ((x[m] - x[k]) * (x[m] - x[j]));
      for (int l = 0; l < q; ++l) if (l != m && l != k && l != j) prod *= (x[i] - x[l]) / (x[m] - x[l]);
      sum += prod;
    }
  }
  return sum;
}
static double g_prime_right_at_one(int degree) { /* basis/vincent.hpp:31-33, 75-91 */
  double cn = degree / (2.0 * degree + 1), cp = 1 - cn, prev = 0, curr = 0, next = 0;
  for (int k = 0; k <= degree; ++k) {
    if (k > 0) { prev = curr; curr = next; }
    next = (k + 1) * 1.0 + 1.0 * curr; /* P_k(1) = 1 */
  }
  return 0.5 * (curr + cp * prev + cn * next);
}

static double legendre_p(int k, double x) { /* std::legendre */
  double p0 = 1, p1 = x;
  if (k == 0) return p0;
  for (int n = 1; n < k; ++n) { double p2 = ((2 * n + 1) * x * p1 - n * p0) / (n + 1); p0 = p1; p1 = p2; }
  return p1;
}
static double vincent_right_derivative(int degree, double c_next, double x) { /* basis/vincent.hpp:75-91 */
  double cp = 1 - c_next, prev = 0, curr = 0, next = 0;
  for (int k = 0; k <= degree; ++k) {
    if (k > 0) { prev = curr; curr = next; }
    next = (k + 1) * legendre_p(k, x) + x * curr;
  }
  return 0.5 * (curr + cp * prev + c_next * next);
}

typedef struct oracle_s {
  int degree, q, n, scheme, riemann;   /* scheme: 0 fr::Lobatto, 1 dg::Lobatto, 2 fr::General,
                                          3 fr::Lobatto on ConvectionDiffusion<Euler, Direct<NavierStokes>> */
  double beta0, beta1, nu, prandtl;    /* direct.hpp:31-32, navier_stokes.hpp:54 (class statics there) */
  double d2[MAXQ][MAXQ];               /* d2[i][m] = l_m''(x_i) */
  double *jtgrad;   /* [cell][n][3][3][3] d(J^-T)/d(xi_a) = mat_after_grad_of_U_, hexahedron.hpp:233-236 */
  double *djac;     /* [cell][n][3][3][3] dJ/d(xi_a), coordinate/cell.hpp:100-119 */
  double *d2jac;    /* [cell][3] the one non-zero row of d2J/d(xi_a)d(xi_b), a != b (row 3-a-b): sum_v s_v P_v / 8 */
  int klocal;       /* Hexahedron<..., true>: a column holds det(J) u (polynomial/hexahedron.hpp:338-343) */
  double *dist;     /* [face][q*q] normal . HolderToSharer, fem.hpp:83-85 */
  /* scheme 9 = dg::General on the modal Projection polynomial with Gauss-Legendre rules */
  int kf;           /* doubles per cell in a column: 5 n (nodal) or 5 M (modal) */
  void *dgm;        /* dgm_t, see the end of this file */
  double *cell_nu;  /* scheme 6: [cell][5] viscosity of every component, constant inside a cell: what
                       EnergyBasedViscosity::GetPropertyOnCell returns (viscosity.hpp:68-71); null = uniform nu */
  int limited;      /* WithLimiter<.., weno::Lazy>(w0, eps): the limiter follows every SetSolutionColumn */
  double weno_w0, weno_eps;
  double c_next;                       /* fr::General only (fr/general.hpp:218-219) */
  int64_t n_cells;
  double x[MAXQ], w[MAXQ], d1[MAXQ][MAXQ]; /* d1[i][m] = l_m'(x_i) */
  double *xyz;      /* [cell][n][3] node coordinates */
  double *jstar;    /* [cell][n][3][3]  J*(x, d) */
  double *jinv;     /* [cell][n][3][3]  J^-1(x, d) */
  double *det;      /* [cell][n] */
  double *length;   /* [cell] */
  /* faces */
  int64_t n_faces;  /* interior first, then boundary */
  int64_t n_interior;
  int64_t *holder, *sharer;
  int *tag;
  int *h_node, *s_node;   /* [face][q*q] cell node of each flux point */
  int *h_dir, *s_dir;     /* [face] direction of the face on each side */
  double *h_scale, *s_scale; /* [face][q*q] +-|J*.col(d)| */
  frame_t *frame;         /* [face][q*q] */
  double *fweight;        /* [face][q*q] global face weight */
  double *fxyz;           /* [face][q*q][3] */
  int bc_kind[16], field_kind[16];
  double field_prm[16][8];
} oracle_t;

static void cross(const double *a, const double *b, double *c) {
  c[0] = a[1] * b[2] - a[2] * b[1]; c[1] = a[2] * b[0] - a[0] * b[2]; c[2] = a[0] * b[1] - a[1] * b[0];
}
static double norm3(const double *a) { return sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]); }

static int cmp_i64(const void *a, const void *b) {
  int64_t x = *(const int64_t *)a, y = *(const int64_t *)b;
  return x < y ? -1 : x > y;
}
typedef struct { int64_t key[4]; int64_t cell; int face; } side_t;
static int cmp_side(const void *a, const void *b) {
  const side_t *x = a, *y = b;
  for (int k = 0; k < 4; ++k) if (x->key[k] != y->key[k]) return x->key[k] < y->key[k] ? -1 : 1;
  return x->cell < y->cell ? -1 : x->cell > y->cell;
}

/* fill the per-face caches for one side: fr/lobatto.hpp:72-163 by coordinate matching
 * (polynomial/hexahedron.hpp:821-928) */
static void match_side(const oracle_t *o, int64_t cell, int lf, const double *fxyz, int *node, int *dir,
    double *scale) {
  int q = o->q, n = o->n, fq = q * q;
  int d = (lf == 2 || lf == 4) ? 0 : (lf == 1 || lf == 3) ? 1 : 2;
  int plus = (lf == 2 || lf == 3 || lf == 5);
  *dir = d;
  for (int p = 0; p < fq; ++p) {
    int found = -1;
    for (int c = 0; c < n && found < 0; ++c) {
      const double *y = &o->xyz[(cell * n + c) * 3];
      double e[3] = {y[0] - fxyz[p * 3], y[1] - fxyz[p * 3 + 1], y[2] - fxyz[p * 3 + 2]};
      int ijk[3] = {c / fq, (c / q) % q, c % q};
      if (ijk[d] != (plus ? q - 1 : 0)) continue;
      if (norm3(e) < 1e-8) found = c;
    }
    if (found < 0) abort();
    node[p] = found;
    const double *js = &o->jstar[(cell * n + found) * 9];
    double nrm[3] = {js[0 * 3 + d], js[1 * 3 + d], js[2 * 3 + d]};
    scale[p] = plus ? norm3(nrm) : -norm3(nrm);
  }
}

/* scheme 8 = fr::General (Vincent correction over the whole node line) on the Navier-Stokes Riemann type */
#define IS_NS(o) ((o)->scheme == 3 || (o)->scheme == 8 || (o)->scheme == 14)
/* scheme 14: EnergyBasedViscosity over the Navier-Stokes base (viscosity.hpp:76-139): both fluxes, the base's first */
#define HAS_CELL_NU(o) ((o)->scheme == 6 || (o)->scheme == 14)
#define IS_GENERAL(o) ((o)->scheme == 2 || (o)->scheme == 8)
#define IS_LUMPED_FR(o) ((o)->scheme == 0 || (o)->scheme == 3 || (o)->scheme == 6 || (o)->scheme == 14)

static void dgm_setup(struct oracle_s *o, const double *nodes, const int64_t *cn);
static void dgm_free(struct oracle_s *o);
static void dgm_residual(struct oracle_s *o, double t, const double *u, double *res);
static double dgm_min_dt(struct oracle_s *o, const double *u);
static void dgm_approximate(struct oracle_s *o, int field_kind, const double *prm, double *column);

void *oracle_create(int degree, int scheme, int riemann, int64_t n_nodes, const double *nodes,
    int64_t n_cells, const int64_t *cn, int64_t n_bfaces, const int64_t *bf, const int32_t *btag) {
  (void)n_nodes;
  oracle_t *o = calloc(1, sizeof *o);
  int q = degree + 1, n = q * q * q, fq = q * q;
  o->degree = degree; o->q = q; o->n = n; o->scheme = scheme; o->riemann = riemann; o->n_cells = n_cells;
  lobatto(q, o->x, o->w);
  for (int i = 0; i < q; ++i) for (int m = 0; m < q; ++m) o->d1[i][m] = lagrange_d(q, o->x, m, i);
  o->xyz = malloc(sizeof(double) * n_cells * n * 3);
  o->jstar = malloc(sizeof(double) * n_cells * n * 9);
  o->jinv = malloc(sizeof(double) * n_cells * n * 9);
  o->det = malloc(sizeof(double) * n_cells * n);
  o->length = malloc(sizeof(double) * n_cells);
  o->jtgrad = malloc(sizeof(double) * n_cells * n * 27);
  o->djac = malloc(sizeof(double) * n_cells * n * 27);
  o->d2jac = calloc(n_cells * 3, sizeof(double));
  for (int64_t c = 0; c < n_cells; ++c)
    for (int v = 0; v < 8; ++v) for (int k = 0; k < 3; ++k)
      o->d2jac[c * 3 + k] += VERT[v][0] * VERT[v][1] * VERT[v][2] / 8 * nodes[cn[c * 8 + v] * 3 + k];
  for (int i = 0; i < q; ++i) for (int m = 0; m < q; ++m) o->d2[i][m] = lagrange_d2(q, o->x, m, i);
  o->beta0 = 2.0; o->beta1 = 1.0 / 12; o->nu = 0.0; o->prandtl = 0.708;
  for (int64_t c = 0; c < n_cells; ++c) {
    double vol = 0;
    for (int a = 0; a < n; ++a) {
      int ijk[3] = {a / fq, (a / q) % q, a % q};
      double loc[3] = {o->x[ijk[0]], o->x[ijk[1]], o->x[ijk[2]]};
      double X[3] = {0, 0, 0}, J[3][3] = {{0}}, dJ[3][3][3] = {{{0}}}; /* dJ[a] = dJ/dxi_a, coordinate/cell.hpp:100-119 */
      for (int v = 0; v < 8; ++v) { /* coordinate/hexahedron.hpp:413-456 */
        const double *P = &nodes[cn[c * 8 + v] * 3];
        double fx = 1 + VERT[v][0] * loc[0], fy = 1 + VERT[v][1] * loc[1], fz = 1 + VERT[v][2] * loc[2];
        double shape = fx * fy * fz / 8;
        double g[3] = {VERT[v][0] * (fy * fz / 8), VERT[v][1] * (fz * fx / 8), VERT[v][2] * (fx * fy / 8)};
        double h01 = VERT[v][0] * VERT[v][1] * (fz / 8), h02 = VERT[v][0] * VERT[v][2] * (fy / 8);
        double h12 = VERT[v][1] * VERT[v][2] * (fx / 8); /* shape Hessian; the diagonal vanishes for Hexahedron8 */
        for (int k = 0; k < 3; ++k) {
          X[k] += shape * P[k];
          for (int d = 0; d < 3; ++d) J[d][k] += g[d] * P[k]; /* J rows = local directions */
          dJ[0][1][k] += h01 * P[k]; dJ[1][0][k] += h01 * P[k];
          dJ[0][2][k] += h02 * P[k]; dJ[2][0][k] += h02 * P[k];
          dJ[1][2][k] += h12 * P[k]; dJ[2][1][k] += h12 * P[k];
        }
      }
      memcpy(&o->xyz[(c * n + a) * 3], X, sizeof X);
      double c0[3], c1[3], c2[3];
      cross(J[1], J[2], c0); cross(J[2], J[0], c1); cross(J[0], J[1], c2);
      double det = J[0][0] * c0[0] + J[0][1] * c0[1] + J[0][2] * c0[2];
      double *inv = &o->jinv[(c * n + a) * 9], *js = &o->jstar[(c * n + a) * 9];
      for (int k = 0; k < 3; ++k) { /* inverse: columns are the cross products / det */
        inv[k * 3 + 0] = c0[k] / det; inv[k * 3 + 1] = c1[k] / det; inv[k * 3 + 2] = c2[k] / det;
      }
      for (int k = 0; k < 9; ++k) js[k] = det * inv[k]; /* polynomial/hexahedron.hpp:228-230 */
      o->det[c * n + a] = det;
      memcpy(&o->djac[(c * n + a) * 27], dJ, sizeof dJ);
      for (int e = 0; e < 3; ++e) { /* -(J^-1 dJ_e J^-1)^T, polynomial/hexahedron.hpp:233-236 */
        double t1[3][3], t2[3][3];
        for (int r = 0; r < 3; ++r) for (int s2 = 0; s2 < 3; ++s2) {
          t1[r][s2] = 0;
          for (int m = 0; m < 3; ++m) t1[r][s2] += inv[r * 3 + m] * dJ[e][m][s2];
        }
        for (int r = 0; r < 3; ++r) for (int s2 = 0; s2 < 3; ++s2) {
          t2[r][s2] = 0;
          for (int m = 0; m < 3; ++m) t2[r][s2] += t1[r][m] * inv[m * 3 + s2];
        }
        for (int r = 0; r < 3; ++r) for (int s2 = 0; s2 < 3; ++s2)
          o->jtgrad[((c * n + a) * 3 + e) * 9 + s2 * 3 + r] = -t2[r][s2];
      }
      vol += o->w[ijk[0]] * o->w[ijk[1]] * o->w[ijk[2]] * det; /* integrator/element.hpp:101-114 */
    }
    o->length[c] = cbrt(vol) * 0.5; /* mesh/part.hpp:230 */
  }
  /* faces from shared node sets (mesh/part.hpp:939-966, 1508-1531) */
  side_t *sides = malloc(sizeof(side_t) * n_cells * 6);
  for (int64_t c = 0; c < n_cells; ++c)
    for (int f = 0; f < 6; ++f) {
      side_t *s = &sides[c * 6 + f];
      for (int m = 0; m < 4; ++m) s->key[m] = cn[c * 8 + FACE_NODES[f][m]];
      qsort(s->key, 4, sizeof(int64_t), cmp_i64);
      s->cell = c; s->face = f;
    }
  qsort(sides, n_cells * 6, sizeof(side_t), cmp_side);
  int64_t max_faces = n_cells * 6;
  o->holder = malloc(sizeof(int64_t) * max_faces); o->sharer = malloc(sizeof(int64_t) * max_faces);
  o->tag = malloc(sizeof(int) * max_faces);
  int *hface = malloc(sizeof(int) * max_faces), *sface = malloc(sizeof(int) * max_faces);
  int64_t nf = 0;
  for (int64_t i = 0; i + 1 < n_cells * 6; ++i)
    if (memcmp(sides[i].key, sides[i + 1].key, sizeof sides[i].key) == 0) {
      o->holder[nf] = sides[i].cell; hface[nf] = sides[i].face;       /* holder = smaller id */
      o->sharer[nf] = sides[i + 1].cell; sface[nf] = sides[i + 1].face;
      o->tag[nf] = -1; nf++; i++;
    }
  o->n_interior = nf;
  for (int64_t b = 0; b < n_bfaces; ++b) {
    side_t k;
    for (int m = 0; m < 4; ++m) k.key[m] = bf[b * 4 + m];
    qsort(k.key, 4, sizeof(int64_t), cmp_i64);
    k.cell = -1;
    int64_t lo = 0, hi = n_cells * 6;
    while (lo < hi) { int64_t mid = (lo + hi) / 2; if (cmp_side(&sides[mid], &k) < 0) lo = mid + 1; else hi = mid; }
    if (lo >= n_cells * 6 || memcmp(sides[lo].key, k.key, sizeof k.key) != 0) abort();
    o->holder[nf] = sides[lo].cell; hface[nf] = sides[lo].face; o->sharer[nf] = -1; sface[nf] = -1;
    o->tag[nf] = btag[b]; nf++;
  }
  o->n_faces = nf;
  free(sides);
  o->h_node = malloc(sizeof(int) * nf * fq); o->s_node = malloc(sizeof(int) * nf * fq);
  o->h_dir = malloc(sizeof(int) * nf); o->s_dir = malloc(sizeof(int) * nf);
  o->h_scale = malloc(sizeof(double) * nf * fq); o->s_scale = malloc(sizeof(double) * nf * fq);
  o->frame = malloc(sizeof(frame_t) * nf * fq);
  o->fweight = malloc(sizeof(double) * nf * fq);
  o->fxyz = malloc(sizeof(double) * nf * fq * 3);
  o->dist = malloc(sizeof(double) * nf * fq);
  for (int64_t f = 0; f < nf; ++f) {
    /* face quad in the holder's outward order (coordinate/cell.hpp:220-258) */
    const double *P[4];
    for (int m = 0; m < 4; ++m) P[m] = &nodes[cn[o->holder[f] * 8 + FACE_NODES[hface[f]][m]] * 3];
    for (int ia = 0; ia < q; ++ia) for (int ib = 0; ib < q; ++ib) { /* integrator/quadrangle.hpp:66-79 */
      int p = ia * q + ib;
      double a = o->x[ia], b = o->x[ib], X[3] = {0, 0, 0}, Ja[3] = {0, 0, 0}, Jb[3] = {0, 0, 0};
      for (int m = 0; m < 4; ++m) { /* coordinate/quadrangle.hpp:77-105 */
        double fa = (1 + QUAD[m][0] * a), fb = (1 + QUAD[m][1] * b);
        for (int k = 0; k < 3; ++k) {
          X[k] += fa * fb / 4 * P[m][k];
          Ja[k] += QUAD[m][0] * (fb / 4) * P[m][k];
          Jb[k] += QUAD[m][1] * (fa / 4) * P[m][k];
        }
      }
      memcpy(&o->fxyz[(f * fq + p) * 3], X, sizeof X);
      frame_t *fr = &o->frame[f * fq + p]; /* coordinate/face.hpp:54-62 */
      double nn[3]; cross(Ja, Jb, nn);
      double len = norm3(nn), la = norm3(Ja);
      for (int k = 0; k < 3; ++k) { fr->a[k] = nn[k] / len; fr->b[k] = Ja[k] / la; }
      cross(fr->a, fr->b, fr->c);
      /* sqrt(det(J J^T)), integrator/element.hpp:105-108 */
      double aa = Ja[0] * Ja[0] + Ja[1] * Ja[1] + Ja[2] * Ja[2], bb = Jb[0] * Jb[0] + Jb[1] * Jb[1] + Jb[2] * Jb[2];
      double ab = Ja[0] * Jb[0] + Ja[1] * Jb[1] + Ja[2] * Jb[2];
      o->fweight[f * fq + p] = o->w[ia] * o->w[ib] * sqrt(aa * bb - ab * ab);
      /* HolderToSharer between the images of the local origins (mesh/part.hpp:118-133,
         coordinate/hexahedron.hpp:45, quadrangle.hpp:44); distance of fem.hpp:83-85 */
      double ch[3] = {0, 0, 0}, cs[3] = {0, 0, 0};
      for (int v = 0; v < 8; ++v) for (int k = 0; k < 3; ++k) ch[k] += nodes[cn[o->holder[f] * 8 + v] * 3 + k] / 8;
      if (o->sharer[f] >= 0) {
        for (int v = 0; v < 8; ++v) for (int k = 0; k < 3; ++k) cs[k] += nodes[cn[o->sharer[f] * 8 + v] * 3 + k] / 8;
        for (int k = 0; k < 3; ++k) cs[k] -= ch[k];
      } else {
        for (int m = 0; m < 4; ++m) for (int k = 0; k < 3; ++k) cs[k] += P[m][k] / 4;
        for (int k = 0; k < 3; ++k) cs[k] = (cs[k] - ch[k]) * 2;
      }
      o->dist[f * fq + p] = fr->a[0] * cs[0] + fr->a[1] * cs[1] + fr->a[2] * cs[2];
    }
    match_side(o, o->holder[f], hface[f], &o->fxyz[f * fq * 3], &o->h_node[f * fq], &o->h_dir[f], &o->h_scale[f * fq]);
    if (o->sharer[f] >= 0)
      match_side(o, o->sharer[f], sface[f], &o->fxyz[f * fq * 3], &o->s_node[f * fq], &o->s_dir[f], &o->s_scale[f * fq]);
  }
  free(hface); free(sface);
  o->kf = 5 * n;
  if (scheme == 9) dgm_setup(o, nodes, cn);
  return o;
}

void oracle_destroy(void *h) {
  oracle_t *o = h;
  dgm_free(o);
  free(o->djac); free(o->d2jac); free(o->cell_nu); free(o->jtgrad); free(o->dist); free(o->xyz); free(o->jstar); free(o->jinv); free(o->det); free(o->length); free(o->holder); free(o->sharer);
  free(o->tag); free(o->h_node); free(o->s_node); free(o->h_dir); free(o->s_dir); free(o->h_scale);
  free(o->s_scale); free(o->frame); free(o->fweight); free(o->fxyz); free(o);
}

void oracle_set_vincent(void *h, double c_next) { ((oracle_t *)h)->c_next = c_next; }
void oracle_set_cell_viscosity(void *h, const double *nu) { /* [n_cells][5], or null for the uniform value */
  oracle_t *o = h;
  free(o->cell_nu);
  o->cell_nu = NULL;
  if (nu) {
    o->cell_nu = malloc(sizeof(double) * o->n_cells * 5);
    memcpy(o->cell_nu, nu, sizeof(double) * o->n_cells * 5);
  }
}
void oracle_set_local(void *h, int klocal) { ((oracle_t *)h)->klocal = klocal; }
void oracle_node_determinants(void *h, double *det) {
  oracle_t *o = h;
  memcpy(det, o->det, sizeof(double) * o->n_cells * o->n);
}
void oracle_set_viscous(void *h, double beta0, double beta1, double nu, double prandtl) {
  oracle_t *o = h;
  o->beta0 = beta0; o->beta1 = beta1; o->nu = nu; o->prandtl = prandtl;
}

void oracle_set_bc(void *h, int tag, int bc_kind, int field_kind, const double *prm) {
  oracle_t *o = h;
  o->bc_kind[tag] = bc_kind; o->field_kind[tag] = field_kind;
  for (int k = 0; k < 8; ++k) o->field_prm[tag][k] = prm ? prm[k] : 0;
}

void oracle_approximate(void *h, int field_kind, const double *prm, double *column) { /* fem.hpp:147-154 */
  oracle_t *o = h;
  double p8[8] = {0};
  for (int k = 0; k < 8 && prm; ++k) p8[k] = prm[k];
  if (o->scheme == 9) { dgm_approximate(o, field_kind, p8, column); return; }
  for (int64_t i = 0; i < o->n_cells * o->n; ++i) eval_field(field_kind, p8, &o->xyz[i * 3], 0.0, &column[i * 5]);
}

/* FiniteElement::GetResidualColumn (fem.hpp:248-258) with fr/general.hpp:249-288, fr/lobatto.hpp:232-428
 * or dg/lobatto.hpp:143-302 */
/* ---- viscous part: ConvectionDiffusion<rotated::Euler<..>, Direct<NavierStokes<Gas>>> ---------------- */
/* gradients are [direction x,y,z][component], Hessians [XX,XY,XZ,YY,YZ,ZZ][component] (constant/index.hpp) */
static void basis_local_gradient(const oracle_t *o, int a, int b, double g[3]) { /* hexahedron.hpp:142-153 */
  int q = o->q, fq = q * q;
  int ia[3] = {a / fq, (a / q) % q, a % q}, ib[3] = {b / fq, (b / q) % q, b % q};
  g[0] = (ia[1] == ib[1] && ia[2] == ib[2]) ? o->d1[ia[0]][ib[0]] : 0;
  g[1] = (ia[0] == ib[0] && ia[2] == ib[2]) ? o->d1[ia[1]][ib[1]] : 0;
  g[2] = (ia[0] == ib[0] && ia[1] == ib[1]) ? o->d1[ia[2]][ib[2]] : 0;
}
static void basis_local_hessian(const oracle_t *o, int a, int b, double h[6]) { /* hexahedron.hpp:155-170 */
  int q = o->q, fq = q * q;
  int ia[3] = {a / fq, (a / q) % q, a % q}, ib[3] = {b / fq, (b / q) % q, b % q};
  double v[3] = {ia[0] == ib[0], ia[1] == ib[1], ia[2] == ib[2]}; /* l_b(x_a) on each line */
  double d[3] = {o->d1[ia[0]][ib[0]], o->d1[ia[1]][ib[1]], o->d1[ia[2]][ib[2]]};
  double dd[3] = {o->d2[ia[0]][ib[0]], o->d2[ia[1]][ib[1]], o->d2[ia[2]][ib[2]]};
  h[0] = dd[0] * v[1] * v[2]; h[1] = d[0] * d[1] * v[2]; h[2] = d[0] * v[1] * d[2];
  h[3] = v[0] * dd[1] * v[2]; h[4] = v[0] * d[1] * d[2]; h[5] = v[0] * v[1] * dd[2];
}
/* Hexahedron::GetGlobalValueGradient for kLocal == false (hexahedron.hpp:479-486, 221-226) */
static void global_gradient(const oracle_t *o, const double *u, int64_t c, int a, double g[3][5]) {
  int n = o->n;
  const double *inv = &o->jinv[(c * n + a) * 9];
  memset(g, 0, sizeof(double) * 15);
  for (int b = 0; b < n; ++b) {
    double lg[3], gg[3];
    basis_local_gradient(o, a, b, lg);
    if (lg[0] == 0 && lg[1] == 0 && lg[2] == 0) continue;
    for (int x = 0; x < 3; ++x) gg[x] = inv[x * 3] * lg[0] + inv[x * 3 + 1] * lg[1] + inv[x * 3 + 2] * lg[2];
    for (int x = 0; x < 3; ++x) for (int k = 0; k < 5; ++k) g[x][k] += gg[x] * u[(c * n + b) * 5 + k];
  }
}
/* Hexahedron::GetGlobalValueGradientHessian for kLocal == false (hexahedron.hpp:586-594, 496-524) */
static void global_gradient_hessian(const oracle_t *o, const double *u, int64_t c, int a, double g[3][5],
    double H[6][5]) {
  int n = o->n;
  const double *inv = &o->jinv[(c * n + a) * 9], *tg = &o->jtgrad[(c * n + a) * 27];
  double lg[3][5], lh[6][5];
  memset(lg, 0, sizeof lg); memset(lh, 0, sizeof lh);
  for (int b = 0; b < n; ++b) {
    double bg[3], bh[6];
    basis_local_gradient(o, a, b, bg); basis_local_hessian(o, a, b, bh);
    for (int k = 0; k < 5; ++k) {
      double ub = u[(c * n + b) * 5 + k];
      for (int x = 0; x < 3; ++x) lg[x][k] += bg[x] * ub;
      for (int x = 0; x < 6; ++x) lh[x][k] += bh[x] * ub;
    }
  }
  for (int x = 0; x < 3; ++x) for (int k = 0; k < 5; ++k)
    g[x][k] = inv[x * 3] * lg[0][k] + inv[x * 3 + 1] * lg[1][k] + inv[x * 3 + 2] * lg[2][k];
  static const int SYM[3][3] = {{0, 1, 2}, {1, 3, 4}, {2, 4, 5}};
  for (int k = 0; k < 5; ++k) {
    double S[3][3], G[3][3];
    for (int e = 0; e < 3; ++e) for (int j = 0; j < 3; ++j) {
      double v = 0;
      for (int b = 0; b < 3; ++b) v += lh[SYM[e][b]][k] * inv[j * 3 + b];   /* scalar_hess *= J^-T */
      for (int b = 0; b < 3; ++b) v += lg[b][k] * tg[e * 9 + b * 3 + j];    /* row(e) += grad^T d(J^-T)/dxi_e */
      S[e][j] = v;
    }
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j)
      G[i][j] = inv[i * 3] * S[0][j] + inv[i * 3 + 1] * S[1][j] + inv[i * 3 + 2] * S[2][j];
    H[0][k] = G[0][0]; H[1][k] = G[0][1]; H[2][k] = G[0][2]; H[3][k] = G[1][1]; H[4][k] = G[1][2]; H[5][k] = G[2][2];
  }
}
/* ---- kLocal == true (polynomial/hexahedron.hpp:171-219, 397-408, 526-562): U = det(J) u are the coefficients ---- */
typedef struct {
  double det, dg[3];       /* det J and its local gradient (coordinate/cell.hpp:157-167) */
  double mah[3][3];        /* mat_after_hess_of_U_ = J^-T / det */
  double mag[3][3][3];     /* mat_after_grad_of_U_[a] */
  double mbg[3];           /* mat_before_grad_of_U_ */
  double mbu[3][3];        /* mat_before_U_ */
} local_cache_t;
static void local_cache(const oracle_t *o, int64_t c, int a, local_cache_t *L) {
  int n = o->n;
  const double *inv = &o->jinv[(c * n + a) * 9], *tg = &o->jtgrad[(c * n + a) * 27], *dJ = &o->djac[(c * n + a) * 27];
  double det = o->det[c * n + a], det2 = det * det, det3 = det2 * det;
  double P[3][3][3], tr[3], dh[3][3];   /* P[e] = inv dJ_e */
  for (int e = 0; e < 3; ++e) {
    tr[e] = 0;
    for (int r = 0; r < 3; ++r) for (int s2 = 0; s2 < 3; ++s2) {
      double v = 0;
      for (int m = 0; m < 3; ++m) v += inv[r * 3 + m] * dJ[e * 9 + m * 3 + s2];
      P[e][r][s2] = v;
    }
    for (int r = 0; r < 3; ++r) tr[e] += P[e][r][r];
    L->dg[e] = det * tr[e];
  }
  for (int e = 0; e < 3; ++e) for (int f = 0; f < 3; ++f) { /* coordinate/cell.hpp:174-199 */
    double t2 = 0, tpp = 0;
    if (e != f) { int p = 3 - e - f; for (int x = 0; x < 3; ++x) t2 += inv[x * 3 + p] * o->d2jac[c * 3 + x]; }
    for (int r = 0; r < 3; ++r) for (int m = 0; m < 3; ++m) tpp += P[e][r][m] * P[f][m][r];
    dh[e][f] = (t2 + tr[e] * tr[f] - tpp) * det;
  }
  L->det = det;
  for (int b = 0; b < 3; ++b) for (int j = 0; j < 3; ++j) L->mah[b][j] = inv[j * 3 + b] / det;   /* inv_T(b, j) = inv(j, b) */
  for (int e = 0; e < 3; ++e) for (int b = 0; b < 3; ++b) for (int j = 0; j < 3; ++j)
    L->mag[e][b][j] = tg[e * 9 + b * 3 + j] / det + inv[j * 3 + b] * (-L->dg[e] / det2);
  for (int j = 0; j < 3; ++j) {
    double v = 0;
    for (int b = 0; b < 3; ++b) v += L->dg[b] * inv[j * 3 + b];
    L->mbg[j] = v / det2;
  }
  for (int e = 0; e < 3; ++e) for (int j = 0; j < 3; ++j) {
    double v = 0;
    for (int b = 0; b < 3; ++b) v += dh[e][b] * (inv[j * 3 + b] / det2);
    for (int b = 0; b < 3; ++b) v += L->dg[b] * (tg[e * 9 + b * 3 + j] / det2 + inv[j * 3 + b] * (-2 * L->dg[e] / det3));
    L->mbu[e][j] = v;
  }
}
/* GetGlobalValueGradient / GetGlobalValueGradientHessian for kLocal == true; U is the column, val = U / det */
static void local_gradient_hessian(const oracle_t *o, const double *U, int64_t c, int a, double *val, double g[3][5],
    double H[6][5]) {
  int n = o->n;
  const double *inv = &o->jinv[(c * n + a) * 9];
  local_cache_t L;
  local_cache(o, c, a, &L);
  double lg[3][5], lh[6][5];
  memset(lg, 0, sizeof lg); memset(lh, 0, sizeof lh);
  for (int b = 0; b < n; ++b) {
    double bg[3], bh[6];
    basis_local_gradient(o, a, b, bg); basis_local_hessian(o, a, b, bh);
    for (int k = 0; k < 5; ++k) {
      double ub = U[(c * n + b) * 5 + k];
      for (int x = 0; x < 3; ++x) lg[x][k] += bg[x] * ub;
      for (int x = 0; x < 6; ++x) lh[x][k] += bh[x] * ub;
    }
  }
  for (int k = 0; k < 5; ++k) val[k] = U[(c * n + a) * 5 + k] / L.det;
  for (int k = 0; k < 5; ++k) { /* hexahedron.hpp:397-408: J* (grad_local U - grad det u) / det^2 */
    double vg[3];
    for (int e = 0; e < 3; ++e) vg[e] = (lg[e][k] - L.dg[e] * val[k]) / (L.det * L.det);
    for (int x = 0; x < 3; ++x) g[x][k] = (L.det * inv[x * 3]) * vg[0] + (L.det * inv[x * 3 + 1]) * vg[1] + (L.det * inv[x * 3 + 2]) * vg[2];
  }
  if (!H) return;
  static const int SYM[3][3] = {{0, 1, 2}, {1, 3, 4}, {2, 4, 5}};
  for (int k = 0; k < 5; ++k) { /* hexahedron.hpp:526-562 */
    double S[3][3], G[3][3];
    for (int e = 0; e < 3; ++e) for (int j = 0; j < 3; ++j) {
      double v = 0;
      for (int b = 0; b < 3; ++b) v += lh[SYM[e][b]][k] * L.mah[b][j];
      for (int b = 0; b < 3; ++b) v += lg[b][k] * L.mag[e][b][j];
      v -= L.mbg[j] * lg[e][k];
      v -= L.mbu[e][j] * U[(c * n + a) * 5 + k];
      S[e][j] = v;
    }
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j)
      G[i][j] = ((L.det * inv[i * 3]) * S[0][j] + (L.det * inv[i * 3 + 1]) * S[1][j] + (L.det * inv[i * 3 + 2]) * S[2][j]) / L.det;
    H[0][k] = G[0][0]; H[1][k] = G[0][1]; H[2][k] = G[0][2]; H[3][k] = G[1][1]; H[4][k] = G[1][2]; H[5][k] = G[2][2];
  }
}
/* NavierStokes::ConservativeToPrimitive (navier_stokes.hpp:91-125); column 4 of gp is grad T */
static prim_t ns_primitive_gradient(const double *c, const double gc[3][5], double gp[3][5], double gpress[3]) {
  prim_t q = cons_to_prim(c);
  for (int x = 0; x < 3; ++x) {
    gp[x][0] = gc[x][0];
    gp[x][1] = (gc[x][1] - q.u * gc[x][0]) / q.rho;
    gp[x][2] = (gc[x][2] - q.v * gc[x][0]) / q.rho;
    gp[x][3] = (gc[x][3] - q.w * gc[x][0]) / q.rho;
    double v = q.u * gc[x][1] + c[1] * gp[x][1];
    v += q.v * gc[x][2] + c[2] * gp[x][2];
    v += q.w * gc[x][3] + c[3] * gp[x][3];
    v *= -0.5;
    v += gc[x][4];
    v *= (GAMMA - 1);
    gpress[x] = v;
    gp[x][4] = (v / q.rho - (q.p / (q.rho * q.rho)) * gc[x][0]) / GAS_R;
  }
  return q;
}
static void ns_stress(const double gp[3][5], double rho, double nu, double tau[6]) { /* :127-142, :70-74 */
  double div = gp[0][1] + gp[1][2] + gp[2][3], mu = nu * rho, zeta = -2.0 / 3 * mu;
  tau[0] = 2 * mu * gp[0][1] + zeta * div;
  tau[3] = 2 * mu * gp[1][2] + zeta * div;
  tau[5] = 2 * mu * gp[2][3] + zeta * div;
  tau[1] = mu * (gp[1][1] + gp[0][2]);
  tau[4] = mu * (gp[2][2] + gp[1][3]);
  tau[2] = mu * (gp[0][3] + gp[2][1]);
}
static double ns_kappa(const oracle_t *o, double rho) { /* :76-79; Cp = R gamma/(gamma-1), types.hpp:245 */
  return o->nu * rho * (GAS_R * (GAMMA / (GAMMA - 1))) / o->prandtl;
}
static void ns_minus_flux_matrix(const oracle_t *o, double F[5][3], const double *c, const double gc[3][5]) { /* :150-175 */
  double gp[3][5], gpr[3], tau[6];
  prim_t q = ns_primitive_gradient(c, gc, gp, gpr);
  ns_stress(gp, c[0], o->nu, tau);
  double kappa = ns_kappa(o, c[0]);
  static const int SYM[3][3] = {{0, 1, 2}, {1, 3, 4}, {2, 4, 5}};
  for (int x = 0; x < 3; ++x) {
    F[4][x] -= kappa * gp[x][4];
    F[1][x] -= tau[SYM[x][0]]; F[2][x] -= tau[SYM[x][1]]; F[3][x] -= tau[SYM[x][2]];
    F[4][x] -= tau[SYM[x][0]] * q.u + tau[SYM[x][1]] * q.v + tau[SYM[x][2]] * q.w;
  }
}
static void ns_minus_flux_core(const oracle_t *o, double *f, double rho, const double gp[3][5], const double *nrm,
    const double *uvw, double n_dot_grad_t) { /* :178-190 */
  double tau[6];
  ns_stress(gp, rho, o->nu, tau);
  static const int SYM[3][3] = {{0, 1, 2}, {1, 3, 4}, {2, 4, 5}};
  double work[3];
  for (int x = 0; x < 3; ++x) {
    f[1 + x] -= tau[SYM[x][0]] * nrm[0] + tau[SYM[x][1]] * nrm[1] + tau[SYM[x][2]] * nrm[2];
    work[x] = tau[SYM[x][0]] * uvw[0] + tau[SYM[x][1]] * uvw[1] + tau[SYM[x][2]] * uvw[2];
  }
  f[4] -= (work[0] * nrm[0] + work[1] * nrm[1] + work[2] * nrm[2]) + ns_kappa(o, rho) * n_dot_grad_t;
}
static void ns_minus_flux_normal(const oracle_t *o, double *f, const double *c, const double gc[3][5],
    const double *nrm) { /* :193-202 */
  double gp[3][5], gpr[3];
  prim_t q = ns_primitive_gradient(c, gc, gp, gpr);
  double uvw[3] = {q.u, q.v, q.w};
  ns_minus_flux_core(o, f, c[0], gp, nrm, uvw, nrm[0] * gp[0][4] + nrm[1] * gp[1][4] + nrm[2] * gp[2][4]);
}
static void ns_minus_flux_no_slip(const oracle_t *o, double *f, const double *wall, const double *c,
    const double gc[3][5], const double *nrm, double value_penalty) { /* :204-224 */
  double gp[3][5], gpr[3];
  prim_t q = ns_primitive_gradient(c, gc, gp, gpr);
  double pen[3] = {value_penalty * (wall[1] - q.u), value_penalty * (wall[2] - q.v), value_penalty * (wall[3] - q.w)};
  for (int x = 0; x < 3; ++x) for (int k = 0; k < 3; ++k) gp[x][1 + k] += nrm[x] * pen[k];
  ns_minus_flux_core(o, f, c[0], gp, nrm, &wall[1], wall[4]);
}
/* component-wise isotropic diffusion (diffusive/linear.hpp:66-92 with nu.x = nu.y = nu.z, or
   EnergyBasedViscosity::MinusViscousFlux, spatial/viscosity.hpp:114-129, with one nu per component) */
static void iso_cell_nu(const oracle_t *o, int64_t c, double *nu) {
  /* without a table: the uniform value of scheme 6; over the Navier-Stokes base (scheme 14) o->nu is the base's, the
     artificial viscosities start at zero */
  for (int k = 0; k < 5; ++k) nu[k] = o->cell_nu ? o->cell_nu[c * 5 + k] : (o->scheme == 14 ? 0.0 : o->nu);
}
static void iso_minus_flux_matrix(double F[5][3], const double *nu, const double gc[3][5]) {
  for (int k = 0; k < 5; ++k) for (int x = 0; x < 3; ++x) F[k][x] -= nu[k] * gc[x][k];
}
static void iso_minus_flux_normal(double *f, const double *nu, const double gc[3][5], const double *nrm) {
  for (int k = 0; k < 5; ++k) f[k] -= nu[k] * (nrm[0] * gc[0][k] + nrm[1] * gc[1][k] + nrm[2] * gc[2][k]);
}
/* Direct::GetCommonGradient with the Hessian penalty (direct.hpp:45-78, 85-90) */
static void ddg_common_gradient(const oracle_t *o, double dist, const double *nrm, const double *ul, const double *ur,
    const double gl[3][5], const double gr[3][5], const double hl[6][5], const double hr[6][5], double gc[3][5]) {
  static const int SYM[3][3] = {{0, 1, 2}, {1, 3, 4}, {2, 4, 5}};
  double n0[3], n1[3];
  for (int x = 0; x < 3; ++x) { n0[x] = nrm[x] * (o->beta0 / dist); n1[x] = nrm[x] * (o->beta1 * dist); }
  for (int x = 0; x < 3; ++x) for (int k = 0; k < 5; ++k) {
    double v = (gl[x][k] + gr[x][k]) / 2;
    v += n0[x] * (ur[k] - ul[k]);
    gc[x][k] = v;
  }
  for (int x = 0; x < 3; ++x) for (int y = 0; y < 3; ++y) for (int k = 0; k < 5; ++k)
    gc[x][k] += n1[y] * (hr[SYM[x][y]][k] - hl[SYM[x][y]][k]);
}

void oracle_residual(void *h, double t, const double *column, double *res) {
  oracle_t *o = h;
  if (o->scheme == 9) { dgm_residual(o, t, column, res); return; }
  int q = o->q, n = o->n, fq = q * q, K = 5;
  int64_t nc = o->n_cells;
  /* kLocal == true: the column holds U = det u; values are U / det (hexahedron.hpp:338-340), gradients and
     Hessians come from U's interpolant, the residual is dU/dt and is not divided by det (fr/general.hpp:251-253) */
  const double *U = column, *u = column;
  double *values = NULL;
  if (o->klocal) {
    values = malloc(sizeof(double) * nc * n * K);
    for (int64_t i = 0; i < nc * n; ++i) for (int k = 0; k < K; ++k) values[i * K + k] = U[i * K + k] / o->det[i];
    u = values;
  }
  double *fhat = malloc(sizeof(double) * nc * n * 15); /* flux_matrices_ cache, fr/general.hpp:213-215 */
  memset(res, 0, sizeof(double) * nc * n * K);
  for (int64_t c = 0; c < nc; ++c) {
    for (int a = 0; a < n; ++a) {
      double F[5][3];
      flux_matrix(&u[(c * n + a) * K], F);
      if (IS_NS(o) || HAS_CELL_NU(o)) { /* fem.hpp:397-404 */
        double gc[3][5], vdummy[5];
        if (o->klocal) local_gradient_hessian(o, U, c, a, vdummy, gc, NULL);
        else global_gradient(o, u, c, a, gc);
        if (IS_NS(o)) ns_minus_flux_matrix(o, F, &u[(c * n + a) * K], gc);
        if (HAS_CELL_NU(o)) {
          double nu[5];
          iso_cell_nu(o, c, nu);
          iso_minus_flux_matrix(F, nu, gc);
        }
      }
      const double *js = &o->jstar[(c * n + a) * 9];
      double *fh = &fhat[(c * n + a) * 15];
      for (int k = 0; k < 5; ++k) for (int d = 0; d < 3; ++d) /* hexahedron.hpp:607-611 */
        fh[k * 3 + d] = F[k][0] * js[0 * 3 + d] + F[k][1] * js[1 * 3 + d] + F[k][2] * js[2 * 3 + d];
    }
    for (int a = 0; a < n; ++a) { /* dense over (a, b) like the reference; gradients are tensor products */
      int ia[3] = {a / fq, (a / q) % q, a % q};
      for (int b = 0; b < n; ++b) {
        int ib[3] = {b / fq, (b / q) % q, b % q};
        /* grad of basis b at node a: d/dxi_d = l_b'(x_a) in direction d when the other two indices agree */
        double g[3];
        g[0] = (ia[1] == ib[1] && ia[2] == ib[2]) ? o->d1[ia[0]][ib[0]] : 0;
        g[1] = (ia[0] == ib[0] && ia[2] == ib[2]) ? o->d1[ia[1]][ib[1]] : 0;
        g[2] = (ia[0] == ib[0] && ia[1] == ib[1]) ? o->d1[ia[2]][ib[2]] : 0;
        if (g[0] == 0 && g[1] == 0 && g[2] == 0) continue;
        if (o->scheme != 1) { /* fr/general.hpp:280-286: res[a] -= Fhat_b . grad phi_b(xi_a) */
          const double *fh = &fhat[(c * n + b) * 15];
          for (int k = 0; k < 5; ++k)
            res[(c * n + a) * K + k] -= fh[k * 3] * g[0] + fh[k * 3 + 1] * g[1] + fh[k * 3 + 2] * g[2];
        } else { /* dg/lobatto.hpp:161-170: res[b] += w_a det_a F_a . (J_a^-1 grad_local phi_b(xi_a)) */
          double F[5][3];
          flux_matrix(&u[(c * n + a) * K], F);
          const double *inv = &o->jinv[(c * n + a) * 9];
          double wa = o->w[ia[0]] * o->w[ia[1]] * o->w[ia[2]] * o->det[c * n + a];
          double gg[3];
          for (int x = 0; x < 3; ++x) gg[x] = inv[x * 3] * g[0] + inv[x * 3 + 1] * g[1] + inv[x * 3 + 2] * g[2];
          for (int k = 0; k < 5; ++k)
            res[(c * n + b) * K + k] += wa * (F[k][0] * gg[0] + F[k][1] * gg[1] + F[k][2] * gg[2]);
        }
      }
    }
  }
  double gp = g_prime_right_at_one(o->degree);
  for (int64_t f = 0; f < o->n_faces; ++f) {
    int64_t ch = o->holder[f], cs = o->sharer[f];
    for (int p = 0; p < fq; ++p) {
      const frame_t *fr = &o->frame[f * fq + p];
      int ah = o->h_node[f * fq + p];
      const double *uh = &u[(ch * n + ah) * K];
      double fu[5];
      if (cs >= 0) {
        int as = o->s_node[f * fq + p];
        flux_upwind(o->riemann, fr, uh, &u[(cs * n + as) * K], fu);
        if (IS_NS(o) || HAS_CELL_NU(o)) { /* fr/general.hpp:316-350 */
          const double *us = &u[(cs * n + as) * K];
          double gh_[3][5], gs_[3][5], hh[6][5], hs[6][5], gcm[3][5], ucm[5];
          if (o->klocal) {
            double vd[5];
            local_gradient_hessian(o, U, ch, ah, vd, gh_, hh);
            local_gradient_hessian(o, U, cs, as, vd, gs_, hs);
          } else {
            global_gradient_hessian(o, u, ch, ah, gh_, hh);
            global_gradient_hessian(o, u, cs, as, gs_, hs);
          }
          ddg_common_gradient(o, o->dist[f * fq + p], fr->a, uh, us, gh_, gs_, hh, hs, gcm);
          for (int k = 0; k < 5; ++k) ucm[k] = (uh[k] + us[k]) / 2;
          if (IS_NS(o)) ns_minus_flux_normal(o, fu, ucm, gcm, fr->a);
          if (HAS_CELL_NU(o)) { /* property = (holder's + sharer's) / 2, fr/general.hpp:331-335 */
            double nh[5], nsh[5];
            iso_cell_nu(o, ch, nh); iso_cell_nu(o, cs, nsh);
            for (int k = 0; k < 5; ++k) { nh[k] += nsh[k]; nh[k] *= 0.5; }
            iso_minus_flux_normal(fu, nh, gcm, fr->a);
          }
        }
        if (IS_GENERAL(o)) { /* fr/general.hpp:296-315, 409-434: the whole node line is corrected */
          double sh = o->h_scale[f * fq + p], ss = o->s_scale[f * fq + p];
          const double *Fh = &fhat[(ch * n + ah) * 15], *Fs = &fhat[(cs * n + as) * 15];
          int strd[3] = {fq, q, 1};
          for (int side = 0; side < 2; ++side) {
            int64_t cell = side ? cs : ch;
            int a0 = side ? as : ah, d = side ? o->s_dir[f] : o->h_dir[f];
            double scale = side ? ss : sh;
            int id = (a0 / strd[d]) % q, base = a0 - id * strd[d];
            for (int i = 0; i < q; ++i) { /* fr/general.hpp:113-119, 147-153: left or right derivative */
              double gpr = scale > 0 ? vincent_right_derivative(o->degree, o->c_next, o->x[i])
                                     : -vincent_right_derivative(o->degree, o->c_next, -o->x[i]);
              for (int k = 0; k < 5; ++k) {
                double fc = side ? fu[k] * (-ss) - Fs[k * 3 + d] : fu[k] * sh - Fh[k * 3 + d];
                res[(cell * n + base + i * strd[d]) * K + k] -= fc * gpr;
              }
            }
          }
        } else if (IS_LUMPED_FR(o)) { /* fr/general.hpp:296-315, fr/lobatto.hpp:232-251 */
          double sh = o->h_scale[f * fq + p], ss = o->s_scale[f * fq + p];
          double gh = sh > 0 ? gp : -gp, gs = ss > 0 ? gp : -gp;
          const double *Fh = &fhat[(ch * n + ah) * 15], *Fs = &fhat[(cs * n + as) * 15];
          for (int k = 0; k < 5; ++k) {
            double f_holder = fu[k] * sh - Fh[k * 3 + o->h_dir[f]];
            double f_sharer = fu[k] * (-ss) - Fs[k * 3 + o->s_dir[f]];
            res[(ch * n + ah) * K + k] -= f_holder * gh;
            res[(cs * n + as) * K + k] -= f_sharer * gs;
          }
        } else { /* dg/lobatto.hpp:171-191 */
          double wf = o->fweight[f * fq + p];
          for (int k = 0; k < 5; ++k) {
            res[(ch * n + ah) * K + k] -= fu[k] * wf;
            res[(cs * n + as) * K + k] += fu[k] * wf;
          }
        }
      } else {
        int tag = o->tag[f], kind = o->bc_kind[tag];
        double given[5] = {0, 0, 0, 0, 0};
        if (kind >= 2) eval_field(o->field_kind[tag], o->field_prm[tag], &o->fxyz[(f * fq + p) * 3], t, given);
        if (kind == 6) { /* no-slip wall: fr/lobatto.hpp:301-319, fr/general.hpp:460-479; no term at all for
                            inviscid systems (fr/general.hpp:455-459) */
          /* fr::General does not override AddFluxOnNoSlipWalls (fem.hpp:382-383): nothing at all there either */
          if ((!IS_NS(o) && !HAS_CELL_NU(o)) || IS_GENERAL(o)) continue;
          double gh_[3][5];
          boundary_flux(o->riemann, 0, fr, uh, given, fu);
          if (o->klocal) { double vd[5]; local_gradient_hessian(o, U, ch, ah, vd, gh_, NULL); }
          else global_gradient(o, u, ch, ah, gh_);
          if (IS_NS(o)) ns_minus_flux_no_slip(o, fu, given, uh, gh_, fr->a, o->beta0 / o->dist[f * fq + p]);
          /* viscosity.hpp:132-139 over a diffusive base: MinusViscousFlux(flux, nu, ...) follows, which subtracts the
             base's normal flux of the unpenalised gradient once more before the artificial one */
          if (o->scheme == 14) ns_minus_flux_normal(o, fu, uh, gh_, fr->a);
          if (HAS_CELL_NU(o)) { /* linear.hpp:85-91: the plain normal flux of the cell's own gradient */
            double nu[5];
            iso_cell_nu(o, ch, nu);
            iso_minus_flux_normal(fu, nu, gh_, fr->a);
          }
        } else {
          boundary_flux(o->riemann, kind, fr, uh, given, fu);
        }
        if (IS_GENERAL(o)) { /* fr/general.hpp:435-640 */
          double sh = o->h_scale[f * fq + p];
          const double *Fh = &fhat[(ch * n + ah) * 15];
          int strd[3] = {fq, q, 1}, d = o->h_dir[f];
          int id = (ah / strd[d]) % q, base = ah - id * strd[d];
          for (int i = 0; i < q; ++i) {
            double gpr = sh > 0 ? vincent_right_derivative(o->degree, o->c_next, o->x[i])
                                : -vincent_right_derivative(o->degree, o->c_next, -o->x[i]);
            for (int k = 0; k < 5; ++k)
              res[(ch * n + base + i * strd[d]) * K + k] -= (fu[k] * sh - Fh[k * 3 + d]) * gpr;
          }
        } else if (IS_LUMPED_FR(o)) { /* fr/lobatto.hpp:278-428 */
          double sh = o->h_scale[f * fq + p], gh = sh > 0 ? gp : -gp;
          const double *Fh = &fhat[(ch * n + ah) * 15];
          for (int k = 0; k < 5; ++k) res[(ch * n + ah) * K + k] -= (fu[k] * sh - Fh[k * 3 + o->h_dir[f]]) * gh;
        } else { /* dg/lobatto.hpp:192-302 */
          double wf = o->fweight[f * fq + p];
          for (int k = 0; k < 5; ++k) res[(ch * n + ah) * K + k] -= fu[k] * wf;
        }
      }
    }
  }
  for (int64_t c = 0; c < nc; ++c)
    for (int a = 0; a < n; ++a) {
      int ia[3] = {a / fq, (a / q) % q, a % q};
      if (o->klocal) continue;
      double s = o->scheme != 1 ? 1.0 / o->det[c * n + a] /* fr/general.hpp:256-266 */
          : 1.0 / (o->w[ia[0]] * o->w[ia[1]] * o->w[ia[2]] * o->det[c * n + a]); /* dg/lobatto.hpp:143-158 */
      for (int k = 0; k < 5; ++k) res[(c * n + a) * K + k] *= s;
    }
  free(fhat);
}

void oracle_weno_lazy(void *h, double w0, double eps, double *u);
void oracle_weno_eigen(void *h, double w0, double eps, double *u);
void oracle_dgm_weno_lazy(void *h, double w0, double eps, double *u);
void oracle_set_limiter(void *h, int on, double w0, double eps) {
  oracle_t *o = h;
  o->limited = on; o->weno_w0 = w0; o->weno_eps = eps;
}
/* temporal/rk.hpp:16-49 with temporal/ode.hpp:121-133, on whole columns; SET() = System::SetSolutionColumn, which
   under WithLimiter is followed by the limiter (spatial/with_limiter.hpp:57-61) */
void oracle_rk_update(void *h, int order, double t, double dt, double *u) {
  oracle_t *o = h;
  int64_t m = o->n_cells * o->kf;
  double *r = malloc(sizeof(double) * m), *u0 = malloc(sizeof(double) * m), *un = malloc(sizeof(double) * m);
  memcpy(u0, u, sizeof(double) * m);
#define EULER_NEXT(tt, dst) do { oracle_residual(o, (tt), u, r); \
    for (int64_t i = 0; i < m; ++i) { r[i] *= dt; dst[i] = u[i] + r[i]; } } while (0)
#define LIMIT() do { if (o->limited == 1 && o->scheme == 9) oracle_dgm_weno_lazy(o, o->weno_w0, o->weno_eps, u); \
    else if (o->limited == 1) oracle_weno_lazy(o, o->weno_w0, o->weno_eps, u); \
    else if (o->limited == 2) oracle_weno_eigen(o, o->weno_w0, o->weno_eps, u); } while (0)
  if (order == 1) {
    EULER_NEXT(t, un); memcpy(u, un, sizeof(double) * m); LIMIT();
  } else if (order == 2) {
    EULER_NEXT(t, un); memcpy(u, un, sizeof(double) * m); LIMIT();
    EULER_NEXT(t + dt, un);
    for (int64_t i = 0; i < m; ++i) { un[i] += u0[i]; un[i] *= 0.5; u[i] = un[i]; }
    LIMIT();
  } else {
    EULER_NEXT(t, un); memcpy(u, un, sizeof(double) * m); LIMIT();
    EULER_NEXT(t + dt, un);
    for (int64_t i = 0; i < m; ++i) { double v = u0[i] * 3; v += un[i]; v /= 4; u[i] = v; }
    LIMIT();
    EULER_NEXT(t + dt / 2, un);
    for (int64_t i = 0; i < m; ++i) { double v = un[i] * 2; v += u0[i]; v /= 3; u[i] = v; }
    LIMIT();
  }
#undef LIMIT
#undef EULER_NEXT
  free(r); free(u0); free(un);
}

/* fem.hpp:26-37, 411-435; fr/lobatto.hpp:220-222; rotated/euler.hpp:340-345 */
double oracle_time_step(void *h, const double *u, double dt_guess, int rk_order) {
  static const double cfl[5][9] = {
    {1.000}, {1.000, 0.333}, {1.000, 0.333, 0.209, 0.130, 0.089, 0.066, 0.051, 0.040, 0.033},
    {1.256, 0.409, 0.209, 0.130, 0.089, 0.066, 0.051, 0.040, 0.033},
    {1.392, 0.464, 0.235, 0.145, 0.100, 0.073, 0.056, 0.045, 0.037}};
  oracle_t *o = h;
  double min_dt = 1e100;
  if (o->scheme == 9) {
    double dt9 = cfl[rk_order][o->degree] * dgm_min_dt(o, u);
    return dt_guess < dt9 ? dt_guess : dt9;
  }
  for (int64_t c = 0; c < o->n_cells; ++c) {
    double vmax = 1e-5;
    for (int a = 0; a < o->n; ++a) {
      double val[5];
      for (int k = 0; k < 5; ++k) val[k] = o->klocal ? u[(c * o->n + a) * 5 + k] / o->det[c * o->n + a] : u[(c * o->n + a) * 5 + k];
      prim_t q = cons_to_prim(val);
      double cs = sound_speed(q.rho, q.p), v = sqrt(q.u * q.u + q.v * q.v + q.w * q.w);
      double s = cs + v > cs - v ? cs + v : cs - v;
      if (s > vmax) vmax = s;
    }
    double d = o->length[c] / vmax;
    if (d < min_dt) min_dt = d;
  }
  double dt = cfl[rk_order][o->degree] * min_dt;
  if (dt_guess < dt) dt = dt_guess;
  if (IS_LUMPED_FR(o)) { dt *= 2.0; if (dt_guess < dt) dt = dt_guess; }
  return dt;
}

/* limiter/average.hpp:11-28 (kind 0) and limiter/majority.hpp:14-42 (kind 1) on a column */
void oracle_limiter(void *h, int kind, double *u) {
  oracle_t *o = h;
  int q = o->q, n = o->n, fq = q * q;
  for (int64_t c = 0; c < o->n_cells; ++c) {
    double value[5] = {0, 0, 0, 0, 0};
    if (kind == 0) { /* polynomial/hexahedron.hpp:626-632 */
      double volume = 0;
      for (int a = 0; a < n; ++a) {
        double w = o->w[a / fq] * o->w[(a / q) % q] * o->w[a % q] * o->det[c * n + a];
        volume += w;
        for (int k = 0; k < 5; ++k) value[k] += u[(c * n + a) * 5 + k] * w;
      }
      for (int k = 0; k < 5; ++k) value[k] /= volume;
    } else {
      for (int k = 0; k < 5; ++k)
        for (int a = 0; a < n; ++a) {
          int cnt = 0;
          for (int b = 0; b < n; ++b) cnt += u[(c * n + b) * 5 + k] == u[(c * n + a) * 5 + k];
          if (cnt * 2 > n) { value[k] = u[(c * n + a) * 5 + k]; break; }
        }
    }
    for (int a = 0; a < n; ++a) for (int k = 0; k < 5; ++k) u[(c * n + a) * 5 + k] = value[k];
  }
}

/* ---- weno::Lazy on Extrapolation<Hexahedron> -------------------------------------- */
#define MAXM 20
static int taylor_count(int degree) { return degree == 2 ? 10 : 20; } /* basis/taylor.hpp:27-31 */
static void taylor_values(int degree, const double *d, double *t) { /* basis/taylor.hpp:204-218 */
  double x = d[0], y = d[1], z = d[2];
  t[0] = 1; t[1] = x; t[2] = y; t[3] = z;
  t[4] = x * x; t[5] = x * y; t[6] = x * z; t[7] = y * y; t[8] = y * z; t[9] = z * z;
  if (degree == 3) {
    t[10] = x * t[4]; t[11] = x * t[5]; t[12] = x * t[6]; t[13] = x * t[7]; t[14] = x * t[8]; t[15] = x * t[9];
    t[16] = y * t[7]; t[17] = y * t[8]; t[18] = y * t[9]; t[19] = z * t[9];
  }
}
/* Taylor::GetPartialDerivatives for one component (basis/taylor.hpp:241-372): entry i is the partial
   derivative named by Taylor index i, entry 0 is zero */
static void taylor_partial_derivatives(int degree, const double *d, const double *c, double *r) {
  double x = d[0], y = d[1], z = d[2];
  int m = taylor_count(degree);
  for (int i = 0; i < m; ++i) r[i] = c[i];
  r[0] = 0;
  r[1] += c[4] * (2 * x) + c[5] * y + c[6] * z;
  r[2] += c[5] * x + c[7] * (2 * y) + c[8] * z;
  r[3] += c[6] * x + c[8] * y + c[9] * (2 * z);
  if (degree == 2) {
    r[4] += c[4]; r[7] += c[7]; r[9] += c[9];
    return;
  }
  double xx = x * x, xy = x * y, xz = x * z, yy = y * y, yz = y * z, zz = z * z;
  r[1] += c[10] * (3 * xx) + c[11] * (2 * xy) + c[12] * (2 * xz) + c[13] * yy + c[14] * yz + c[15] * zz;
  r[2] += c[11] * xx + c[14] * xz + c[13] * (2 * xy) + c[16] * (3 * yy) + c[17] * (2 * yz) + c[18] * zz;
  r[3] += c[12] * xx + c[14] * xy + c[15] * (2 * xz) + c[17] * yy + c[18] * (2 * yz) + c[19] * (3 * zz);
  r[4] += c[11] * y + c[12] * z + c[10] * x * 3; r[4] += r[4];
  r[5] += c[11] * x * 2 + c[13] * y * 2 + c[14] * z;
  r[6] += c[12] * x * 2 + c[15] * z * 2 + c[14] * y;
  r[7] += c[13] * x + c[17] * z + c[16] * y * 3; r[7] += r[7];
  r[8] += c[14] * x + c[17] * y * 2 + c[18] * z * 2;
  r[9] += c[15] * x + c[18] * y + c[19] * z * 3; r[9] += r[9];
  r[10] *= 6; r[11] += r[11]; r[12] += r[12]; r[13] += r[13]; r[15] += r[15];
  r[16] *= 6; r[17] += r[17]; r[18] += r[18]; r[19] *= 6;
}
typedef struct {
  double center[3];
  double s[MAXM][MAXM];   /* OrthoNormal::coeff(), lower triangular (basis/linear.hpp, integrator/function.hpp:138-172) */
} modal_basis_t;
static void cell_center(const oracle_t *o, int64_t c, double *ctr) { /* image of the local origin */
  /* mean of the 8 corner nodes = trilinear map at (0,0,0); the corners are nodes (i,j,k) in {0,q-1}^3 */
  int q = o->q, n = o->n, fq = q * q;
  for (int k = 0; k < 3; ++k) ctr[k] = 0;
  for (int a = 0; a < n; ++a) {
    int i = a / fq, j = (a / q) % q, kk = a % q;
    if ((i != 0 && i != q - 1) || (j != 0 && j != q - 1) || (kk != 0 && kk != q - 1)) continue;
    for (int k = 0; k < 3; ++k) ctr[k] += o->xyz[(c * n + a) * 3 + k] / 8;
  }
}
static double global_weight(const oracle_t *o, int64_t c, int a) { /* integrator/hexahedron.hpp: local weight x det */
  int q = o->q, fq = q * q;
  return o->w[a / fq] * o->w[(a / q) % q] * o->w[a % q] * o->det[c * o->n + a];
}
static void modal_basis_build(const oracle_t *o, int64_t c, modal_basis_t *b) {
  int n = o->n, m = taylor_count(o->degree);
  double A[MAXM][MAXM] = {{0}}, S[MAXM][MAXM] = {{0}};
  cell_center(o, c, b->center);
  for (int a = 0; a < n; ++a) {
    double d[3], t[MAXM], w = global_weight(o, c, a);
    for (int k = 0; k < 3; ++k) d[k] = o->xyz[(c * n + a) * 3 + k] - b->center[k];
    taylor_values(o->degree, d, t);
    for (int i = 0; i < m; ++i) for (int j = 0; j < m; ++j) A[i][j] += t[i] * t[j] * w;
  }
  for (int i = 0; i < m; ++i) S[i][i] = 1;
  S[0][0] = 1 / sqrt(A[0][0]);
  for (int i = 1; i < m; ++i) {
    for (int j = 0; j < i; ++j) {
      double temp = 0;
      for (int k = 0; k <= j; ++k) temp += S[j][k] * A[k][i];
      for (int l = 0; l <= j; ++l) S[i][l] -= temp * S[j][l];
    }
    double norm_sq = 0;
    for (int j = 0; j <= i; ++j) {
      double sum = 0, sij = S[i][j];
      for (int k = 0; k < j; ++k) sum += 2 * S[i][k] * A[k][j];
      norm_sq += sij * (sij * A[j][j] + sum);
    }
    double nrm = sqrt(norm_sq);
    for (int j = 0; j < m; ++j) S[i][j] /= nrm;
  }
  memcpy(b->s, S, sizeof S);
}
static void modal_basis_values(const oracle_t *o, const modal_basis_t *b, const double *x, double *phi) {
  int m = taylor_count(o->degree);
  double d[3] = {x[0] - b->center[0], x[1] - b->center[1], x[2] - b->center[2]}, t[MAXM];
  taylor_values(o->degree, d, t);
  for (int i = 0; i < m; ++i) {
    double v = 0;
    for (int j = 0; j <= i; ++j) v += b->s[i][j] * t[j];
    phi[i] = v;
  }
}
/* weno::GetSmoothness (limiter/weno.hpp:78-95, 37-75) of one component given on my orthonormal basis */
static double weno_smoothness(const oracle_t *o, int64_t c, const modal_basis_t *b, const double *coef) {
  int n = o->n, m = taylor_count(o->degree);
  double tc[MAXM], integral[MAXM] = {0};
  for (int j = 0; j < m; ++j) { /* GetCoeffOnTaylorBasis = coeff * basis.coeff() */
    double v = 0;
    for (int i = j; i < m; ++i) v += coef[i] * b->s[i][j];
    tc[j] = v;
  }
  for (int a = 0; a < n; ++a) {
    double d[3], pdv[MAXM], w = global_weight(o, c, a);
    for (int k = 0; k < 3; ++k) d[k] = o->xyz[(c * n + a) * 3 + k] - b->center[k];
    taylor_partial_derivatives(o->degree, d, tc, pdv);
    for (int i = 0; i < m; ++i) integral[i] += pdv[i] * pdv[i] * w;
  }
  double volume = 1 / (b->s[0][0] * b->s[0][0]);
  double total = (integral[1] + integral[2] + integral[3]) * (pow(volume, 1 * 2. / 3 - 1) / (tgamma(2) * tgamma(2)));
  static const int begin[4] = {1, 4, 10, 20};
  for (int p = 2; p <= o->degree; ++p) {
    double sm = 0;
    for (int i = begin[p - 1]; i < begin[p]; ++i) sm += integral[i];
    total += sm * (pow(volume, p * 2. / 3 - 1) / (tgamma(p + 1) * tgamma(p + 1)));
  }
  return total;
}
/* limiter::Reconstruct(part, weno::Lazy(w0, eps)) on a single part: every cell is an inner cell, one Jacobi
   sweep (limiter/reconstruct.hpp:21-54, limiter/weno.hpp:97-195, polynomial/extrapolation.hpp:60-90) */
void oracle_weno_lazy(void *h, double w0, double eps, double *u) {
  oracle_t *o = h;
  int n = o->n, m = taylor_count(o->degree);
  int64_t nc = o->n_cells;
  modal_basis_t *basis = malloc(sizeof(modal_basis_t) * nc);
  double *modal = malloc(sizeof(double) * nc * 5 * m), *fresh = malloc(sizeof(double) * nc * 5 * m);
  for (int64_t c = 0; c < nc; ++c) { /* Extrapolation::UpdateModalCoeff */
    modal_basis_build(o, c, &basis[c]);
    double *mc = &modal[c * 5 * m];
    memset(mc, 0, sizeof(double) * 5 * m);
    for (int a = 0; a < n; ++a) {
      double phi[MAXM], w = global_weight(o, c, a);
      modal_basis_values(o, &basis[c], &o->xyz[(c * n + a) * 3], phi);
      for (int k = 0; k < 5; ++k) for (int i = 0; i < m; ++i) mc[k * m + i] += u[(c * n + a) * 5 + k] * (phi[i] * w);
    }
  }
  /* adjacency in face order */
  int *adj_cnt = calloc(nc, sizeof(int));
  int64_t *adj = malloc(sizeof(int64_t) * nc * 6);
  for (int64_t f = 0; f < o->n_interior; ++f) {
    int64_t a = o->holder[f], b = o->sharer[f];
    adj[a * 6 + adj_cnt[a]++] = b;
    adj[b * 6 + adj_cnt[b]++] = a;
  }
  for (int64_t c = 0; c < nc; ++c) {
    const modal_basis_t *mb = &basis[c];
    const double *mine = &modal[c * 5 * m];
    int cnt = adj_cnt[c];
    double cand[7][5][MAXM];
    double s00 = mb->s[0][0];
    for (int j = 0; j < cnt; ++j) { /* Borrow: project the neighbour's modal polynomial onto my basis */
      int64_t d = adj[c * 6 + j];
      memset(cand[j], 0, sizeof cand[j]);
      for (int a = 0; a < n; ++a) {
        double phi[MAXM], phn[MAXM], w = global_weight(o, c, a);
        const double *x = &o->xyz[(c * n + a) * 3];
        modal_basis_values(o, mb, x, phi);
        modal_basis_values(o, &basis[d], x, phn);
        for (int k = 0; k < 5; ++k) {
          double v = 0;
          for (int i = 0; i < m; ++i) v += modal[(d * 5 + k) * m + i] * phn[i];
          for (int i = 0; i < m; ++i) cand[j][k][i] += v * phi[i] * w;
        }
      }
      for (int k = 0; k < 5; ++k) /* adj_proj += my_average - adj_proj.average() (projection.hpp:257-261) */
        cand[j][k][0] += (mine[k * m] * s00 - cand[j][k][0] * s00) / s00;
    }
    for (int k = 0; k < 5; ++k) for (int i = 0; i < m; ++i) cand[cnt][k][i] = mine[k * m + i];
    for (int k = 0; k < 5; ++k) {
      double wgt[7], sum = 0;
      for (int j = 0; j <= cnt; ++j) {
        double beta = weno_smoothness(o, c, mb, cand[j][k]) + eps;
        double w = j < cnt ? w0 : w0 * (-cnt) + 1.0;
        wgt[j] = w / (beta * beta);
        sum += wgt[j];
      }
      for (int j = 0; j <= cnt; ++j) wgt[j] /= sum;
      for (int i = 0; i < m; ++i) {
        double v = cand[cnt][k][i] * wgt[cnt];
        for (int j = 0; j < cnt; ++j) v += cand[j][k][i] * wgt[j];
        fresh[(c * 5 + k) * m + i] = v;
      }
    }
  }
  for (int64_t c = 0; c < nc; ++c) /* Extrapolation::SetCoeff(modal) -> UpdateNodalCoeff */
    for (int a = 0; a < n; ++a) {
      double phi[MAXM];
      modal_basis_values(o, &basis[c], &o->xyz[(c * n + a) * 3], phi);
      for (int k = 0; k < 5; ++k) {
        double v = 0;
        for (int i = 0; i < m; ++i) v += fresh[(c * 5 + k) * m + i] * phi[i];
        u[(c * n + a) * 5 + k] = v;
      }
    }
  free(basis); free(modal); free(fresh); free(adj_cnt); free(adj);
}

/* EigenMatrices (riemann/euler/eigen.hpp:27-56) of a conservative state in the frame (nu, mu, pi) */
static void eigen_matrices(const double *cons, const frame_t *fr, double L[5][5], double R[5][5]) {
  prim_t q = cons_to_prim(cons);
  const double *nu = fr->a, *mu = fr->b, *pi = fr->c;
  double ux = q.u, uy = q.v, uz = q.w;
  double u_nu = ux * nu[0] + uy * nu[1] + uz * nu[2];
  double u_mu = ux * mu[0] + uy * mu[1] + uz * mu[2];
  double u_pi = ux * pi[0] + uy * pi[1] + uz * pi[2];
  double a = sound_speed(q.rho, q.p);
  double ek = (ux * ux + uy * uy + uz * uz) / 2, e0 = a * a * (1 / (GAMMA - 1)), h0 = e0 + ek;
  double rc[5][5] = { /* columns of R */
    {1, ux - a * nu[0], uy - a * nu[1], uz - a * nu[2], h0 - a * u_nu},
    {1, ux, uy, uz, ek},
    {0, mu[0], mu[1], mu[2], u_mu},
    {0, pi[0], pi[1], pi[2], u_pi},
    {1, ux + a * nu[0], uy + a * nu[1], uz + a * nu[2], h0 + a * u_nu}};
  double b1 = 1 / e0, b2 = b1 * ek;
  double lc[5][5] = { /* columns of L */
    {(b2 + u_nu / a) / 2, 1 - b2, -u_mu, -u_pi, (b2 - u_nu / a) / 2},
    {-(b1 * ux + nu[0] / a) / 2, b1 * ux, mu[0], pi[0], -(b1 * ux - nu[0] / a) / 2},
    {-(b1 * uy + nu[1] / a) / 2, b1 * uy, mu[1], pi[1], -(b1 * uy - nu[1] / a) / 2},
    {-(b1 * uz + nu[2] / a) / 2, b1 * uz, mu[2], pi[2], -(b1 * uz - nu[2] / a) / 2},
    {b1 / 2, -b1, 0, 0, b1 / 2}};
  for (int i = 0; i < 5; ++i) for (int j = 0; j < 5; ++j) { R[i][j] = rc[j][i]; L[i][j] = lc[j][i]; }
}
void oracle_eigen_matrices(const double *cons, const double *frame9, double *L, double *R) {
  frame_t fr;
  memcpy(&fr, frame9, sizeof fr);
  eigen_matrices(cons, &fr, (double (*)[5])L, (double (*)[5])R);
}
/* limiter::Reconstruct(part, weno::Eigen(w0, eps)) on a single part (limiter/weno.hpp:197-349): the KXRCF-like
   detector picks the troubled cells, each of which is rebuilt face by face in characteristic variables */
void oracle_weno_eigen(void *h, double w0, double eps, double *u) {
  oracle_t *o = h;
  int n = o->n, m = taylor_count(o->degree), fq = o->q * o->q;
  int64_t nc = o->n_cells;
  modal_basis_t *basis = malloc(sizeof(modal_basis_t) * nc);
  double *modal = malloc(sizeof(double) * nc * 5 * m), *fresh = malloc(sizeof(double) * nc * 5 * m);
  double *volume = malloc(sizeof(double) * nc);
  char *troubled = calloc(nc, 1);
  for (int64_t c = 0; c < nc; ++c) {
    modal_basis_build(o, c, &basis[c]);
    double *mc = &modal[c * 5 * m];
    memset(mc, 0, sizeof(double) * 5 * m);
    volume[c] = 0;
    for (int a = 0; a < n; ++a) {
      double phi[MAXM], w = global_weight(o, c, a);
      volume[c] += w;
      modal_basis_values(o, &basis[c], &o->xyz[(c * n + a) * 3], phi);
      for (int k = 0; k < 5; ++k) for (int i = 0; i < m; ++i) mc[k * m + i] += u[(c * n + a) * 5 + k] * (phi[i] * w);
    }
  }
  int *adj_cnt = calloc(nc, sizeof(int));
  int64_t *adj = malloc(sizeof(int64_t) * nc * 6), *adj_face = malloc(sizeof(int64_t) * nc * 6);
  for (int64_t f = 0; f < o->n_interior; ++f) {
    int64_t a = o->holder[f], b = o->sharer[f];
    adj_face[a * 6 + adj_cnt[a]] = f; adj[a * 6 + adj_cnt[a]++] = b;
    adj_face[b * 6 + adj_cnt[b]] = f; adj[b * 6 + adj_cnt[b]++] = a;
  }
  for (int64_t c = 0; c < nc; ++c) { /* Eigen::IsNotSmooth (weno.hpp:230-258) */
    const modal_basis_t *mb = &basis[c];
    const int comps[2] = {0, 4};
    double max_avg[5], sum_diff[5] = {0}, mine[5], phi[MAXM];
    modal_basis_values(o, mb, mb->center, phi);
    for (int k = 0; k < 5; ++k) {
      max_avg[k] = modal[(c * 5 + k) * m] * mb->s[0][0];
      mine[k] = 0;
      for (int i = 0; i < m; ++i) mine[k] += modal[(c * 5 + k) * m + i] * phi[i];
    }
    for (int t = 0; t < 2; ++t) { int k = comps[t]; max_avg[k] = fmax(1e-9, fabs(max_avg[k])); }
    for (int j = 0; j < adj_cnt[c]; ++j) {
      int64_t d = adj[c * 6 + j];
      double phn[MAXM];
      modal_basis_values(o, &basis[d], mb->center, phn);
      for (int t = 0; t < 2; ++t) {
        int k = comps[t];
        double v = 0;
        for (int i = 0; i < m; ++i) v += modal[(d * 5 + k) * m + i] * phn[i];
        sum_diff[k] += fabs(mine[k] - v);
        max_avg[k] = fmax(max_avg[k], fabs(modal[(d * 5 + k) * m] * basis[d].s[0][0]));
      }
    }
    double divisor = pow(volume[c], (o->degree + 1.0) / 2.0 / 3) * adj_cnt[c];
    double reference = o->degree < 3 ? 1.0 : 3.0;
    for (int t = 0; t < 2; ++t) {
      int k = comps[t];
      if (sum_diff[k] / max_avg[k] / divisor > reference) troubled[c] = 1;
    }
  }
  for (int64_t c = 0; c < nc; ++c) {
    if (!troubled[c]) continue;
    const modal_basis_t *mb = &basis[c];
    const double *mine = &modal[c * 5 * m];
    int cnt = adj_cnt[c];
    double cand[7][5][MAXM], big_u[5], total[5][MAXM], total_volume = 0;
    double s00 = mb->s[0][0];
    for (int j = 0; j < cnt; ++j) { /* Borrow, as in weno::Lazy */
      int64_t d = adj[c * 6 + j];
      memset(cand[j], 0, sizeof cand[j]);
      for (int a = 0; a < n; ++a) {
        double phi[MAXM], phn[MAXM], w = global_weight(o, c, a);
        const double *x = &o->xyz[(c * n + a) * 3];
        modal_basis_values(o, mb, x, phi);
        modal_basis_values(o, &basis[d], x, phn);
        for (int k = 0; k < 5; ++k) {
          double v = 0;
          for (int i = 0; i < m; ++i) v += modal[(d * 5 + k) * m + i] * phn[i];
          for (int i = 0; i < m; ++i) cand[j][k][i] += v * phi[i] * w;
        }
      }
      for (int k = 0; k < 5; ++k) cand[j][k][0] += (mine[k * m] * s00 - cand[j][k][0] * s00) / s00;
    }
    for (int k = 0; k < 5; ++k) { for (int i = 0; i < m; ++i) cand[cnt][k][i] = mine[k * m + i]; big_u[k] = mine[k * m] * s00; }
    memset(total, 0, sizeof total);
    for (int jf = 0; jf < cnt; ++jf) { /* ReconstructOnFace */
      int64_t f = adj_face[c * 6 + jf];
      double L[5][5], R[5][5], rot[7][5][MAXM], blend[5][MAXM];
      eigen_matrices(big_u, &o->frame[f * fq], L, R);
      for (int j = 0; j <= cnt; ++j)
        for (int k = 0; k < 5; ++k) for (int i = 0; i < m; ++i) {
          double v = 0;
          for (int e = 0; e < 5; ++e) v += L[k][e] * cand[j][e][i];
          rot[j][k][i] = v;
        }
      for (int k = 0; k < 5; ++k) {
        double wgt[7], sum = 0;
        for (int j = 0; j <= cnt; ++j) {
          double beta = weno_smoothness(o, c, mb, rot[j][k]) + eps;
          double w = j < cnt ? w0 : w0 * (-cnt) + 1.0;
          wgt[j] = w / (beta * beta);
          sum += wgt[j];
        }
        for (int j = 0; j <= cnt; ++j) wgt[j] /= sum;
        for (int i = 0; i < m; ++i) {
          double v = rot[cnt][k][i] * wgt[cnt];
          for (int j = 0; j < cnt; ++j) v += rot[j][k][i] * wgt[j];
          blend[k][i] = v;
        }
      }
      double adj_volume = volume[adj[c * 6 + jf]];
      for (int k = 0; k < 5; ++k) for (int i = 0; i < m; ++i) {
        double v = 0;
        for (int e = 0; e < 5; ++e) v += R[k][e] * blend[e][i];
        total[k][i] += v * adj_volume;
      }
      total_volume += adj_volume;
    }
    for (int k = 0; k < 5; ++k) for (int i = 0; i < m; ++i) fresh[(c * 5 + k) * m + i] = total[k][i] / total_volume;
  }
  for (int64_t c = 0; c < nc; ++c) {
    if (!troubled[c]) continue;
    for (int a = 0; a < n; ++a) {
      double phi[MAXM];
      modal_basis_values(o, &basis[c], &o->xyz[(c * n + a) * 3], phi);
      for (int k = 0; k < 5; ++k) {
        double v = 0;
        for (int i = 0; i < m; ++i) v += fresh[(c * 5 + k) * m + i] * phi[i];
        u[(c * n + a) * 5 + k] = v;
      }
    }
  }
  free(basis); free(modal); free(fresh); free(volume); free(troubled); free(adj_cnt); free(adj); free(adj_face);
}

/* ---- EnergyBasedViscosity (spatial/viscosity.hpp) on a single part ------------------------------------------- */
/* Scalar viscous residual (nu = 1, every neighbour and boundary side zero) of one cell for nodal values v[n]:
   what GetCellResidual(nu = 1) - GetCellResidual(nu = 0) leaves of UpdateCellResidual (viscosity.hpp:251-289, 313-331)
   in component 0: AddFluxDivergence + AddFluxToHolder / AddFluxToSharer (fr/lobatto.hpp:252-277,
   fr/general.hpp:356-408) on the six faces, then 1/detJ (viscosity.hpp:320-328). */
static void energy_probe_residual(const oracle_t *o, int64_t c, const double *v, double *res) {
  int q = o->q, n = o->n, fq = q * q;
  double *u = calloc((size_t)o->n_cells * n * 5, sizeof(double));   /* the probe lives in all 5 components */
  for (int a = 0; a < n; ++a) for (int k = 0; k < 5; ++k) u[(c * n + a) * 5 + k] = v[a];
  double *fhat = malloc(sizeof(double) * n * 3);
  for (int a = 0; a < n; ++a) {
    double g[3][5];
    global_gradient(o, u, c, a, g);
    const double *js = &o->jstar[(c * n + a) * 9];
    for (int d = 0; d < 3; ++d) fhat[a * 3 + d] = -(g[0][0] * js[0 * 3 + d] + g[1][0] * js[1 * 3 + d] + g[2][0] * js[2 * 3 + d]);
  }
  for (int a = 0; a < n; ++a) {
    res[a] = 0;
    for (int b = 0; b < n; ++b) {
      double lg[3];
      basis_local_gradient(o, a, b, lg);
      res[a] -= fhat[b * 3] * lg[0] + fhat[b * 3 + 1] * lg[1] + fhat[b * 3 + 2] * lg[2];
    }
  }
  double gp = g_prime_right_at_one(o->degree);
  for (int64_t f = 0; f < o->n_faces; ++f) {
    int holder = o->holder[f] == c, sharer = o->sharer[f] == c;
    if (!holder && !sharer) continue;
    for (int p = 0; p < fq; ++p) {
      const frame_t *fr = &o->frame[f * fq + p];
      int node = holder ? o->h_node[f * fq + p] : o->s_node[f * fq + p];
      int dir = holder ? o->h_dir[f] : o->s_dir[f];
      double scale = holder ? o->h_scale[f * fq + p] : o->s_scale[f * fq + p];
      double g[3][5], H[6][5], zero3[3][5] = {{0}}, zero6[6][5] = {{0}}, uz[5] = {0}, uc[5], gc[3][5];
      global_gradient_hessian(o, u, c, node, g, H);
      for (int k = 0; k < 5; ++k) uc[k] = v[node];
      if (holder) ddg_common_gradient(o, o->dist[f * fq + p], fr->a, uc, uz, g, zero3, H, zero6, gc);
      else ddg_common_gradient(o, o->dist[f * fq + p], fr->a, uz, uc, zero3, g, zero6, H, gc);
      double f_up = -(fr->a[0] * gc[0][0] + fr->a[1] * gc[1][0] + fr->a[2] * gc[2][0]);   /* nu = 1 */
      double f_side = (holder ? f_up * scale : f_up * (-scale)) - fhat[node * 3 + dir];
      res[node] -= f_side * (scale > 0 ? gp : -gp);
    }
  }
  for (int a = 0; a < n; ++a) res[a] /= o->det[c * n + a];
  free(u); free(fhat);
}
/* EnergyBasedViscosity::BuildDampingMatrices (viscosity.hpp:299-384): d[cell][r][c] */
void oracle_damping_matrices(void *h, double *d) {
  oracle_t *o = h;
  int n = o->n;
  double *v = malloc(sizeof(double) * n), *r = malloc(sizeof(double) * n);
  for (int64_t c = 0; c < o->n_cells; ++c)
    for (int col = 0; col < n; ++col) {
      for (int a = 0; a < n; ++a) v[a] = a == col;
      energy_probe_residual(o, c, v, r);
      for (int a = 0; a < n; ++a) d[(c * n + a) * n + col] = r[a] * global_weight(o, c, a);
    }
  free(v); free(r);
}
/* EnergyBasedViscosity::UpdateProperties (viscosity.hpp:556-572): IntegrateJumpOnFaces (:446-475) and
   GetViscosityValues (:477-544).  nu[cell][5]; returns min_dt_ (viscosity.hpp:574-578). */
double oracle_update_viscosity(void *h, const double *u, const double *damping, double time_scale, double *nu) {
  oracle_t *o = h;
  int q = o->q, n = o->n, fq = q * q, P = o->degree;
  int64_t nc = o->n_cells;
  double *jump = calloc(nc * 5, sizeof(double)), *volume = calloc(nc, sizeof(double));
  for (int64_t c = 0; c < nc; ++c) for (int a = 0; a < n; ++a) volume[c] += global_weight(o, c, a);
  for (int64_t f = 0; f < o->n_interior; ++f) {
    int64_t ch = o->holder[f], cs = o->sharer[f];
    double area = 0;
    for (int p = 0; p < fq; ++p) area += o->fweight[f * fq + p];
    double hh = volume[ch] / area, hs = volume[cs] / area;   /* mesh/part.hpp:123-124 */
    for (int p = 0; p < fq; ++p)
      for (int k = 0; k < 5; ++k) {
        double j = u[(ch * n + o->h_node[f * fq + p]) * 5 + k] - u[(cs * n + o->s_node[f * fq + p]) * 5 + k];
        j *= j;
        jump[ch * 5 + k] += j * (o->fweight[f * fq + p] * hh);
        jump[cs * 5 + k] += j * (o->fweight[f * fq + p] * hs);
      }
  }
  static const double spatial_factor[7] = {1., 4., 10., 40., 28., 50., 90.};
  double min_dt = 1e100;
  for (int64_t c = 0; c < nc; ++c) {
    double max_speed = 1e-5, ref[5] = {1e-5, 1e-5, 1e-5, 1e-5, 1e-5};
    for (int a = 0; a < n; ++a) { /* rotated/euler.hpp:340-355 */
      const double *uc = &u[(c * n + a) * 5];
      prim_t pr = cons_to_prim(uc);
      double cs = sound_speed(pr.rho, pr.p), vel = sqrt(pr.u * pr.u + pr.v * pr.v + pr.w * pr.w);
      double lam = fmax(cs + vel, cs - vel);
      double sq_mass = uc[0] * uc[0], sq[5] = {sq_mass, sq_mass * lam * lam, sq_mass * lam * lam, sq_mass * lam * lam, uc[4] * uc[4]};
      max_speed = fmax(max_speed, lam);
      for (int k = 0; k < 5; ++k) ref[k] = fmax(ref[k], sq[k]);
    }
    double len = o->length[c];
    min_dt = fmin(min_dt, len / max_speed);
    double max_visc = max_speed * len / P, time_base = len / max_speed, scale = pow(len, (P + 1) + 3);
    for (int k = 0; k < 5; ++k) {
      if (jump[c * 5 + k] < ref[k] * scale) { nu[c * 5 + k] = 0; continue; }
      double rate = 0;
      for (int r = 0; r < n; ++r) {
        double du = 0;
        for (int a = 0; a < n; ++a) du += damping[(c * n + r) * n + a] * u[(c * n + a) * 5 + k];
        rate += u[(c * n + r) * 5 + k] * du;
      }
      rate = -rate;
      nu[c * 5 + k] = fmin(max_visc, fmax(0.0, jump[c * 5 + k] / (rate * (time_base * time_scale))));
      min_dt = fmin(min_dt, len / (max_speed + spatial_factor[P] * nu[c * 5 + k] / len));
    }
  }
  free(jump); free(volume);
  return min_dt;
}

/* one rotated Riemann problem (test hook) */
void oracle_riemann_flux(int riemann, int bc_kind, const double *frame9, const double *ul, const double *ur,
    double *f) {
  frame_t fr;
  memcpy(fr.a, frame9, sizeof fr.a); memcpy(fr.b, frame9 + 3, sizeof fr.b); memcpy(fr.c, frame9 + 6, sizeof fr.c);
  if (bc_kind < 0) flux_upwind(riemann, &fr, ul, ur, f);
  else boundary_flux(riemann, bc_kind, &fr, ul, ur, f);
}
/* 1-D solver on primitive states (rho, u, v, w, p), for the Toro known-answer tests */
void oracle_riemann_1d(int riemann, const double *pl, const double *pr, double *f) {
  prim_t l = {pl[0], pl[1], pl[2], pl[3], pl[4]}, r = {pr[0], pr[1], pr[2], pr[3], pr[4]};
  solve_1d(riemann, &l, &r, f);
}


/* ---- dg::General on Projection<Scalar, 3, P, 5> ------------------------------------------------- */
/* spatial/dg/general.hpp:58-89 (+ boundary loops :92-200), polynomial/projection.hpp:67-176, basis/linear.hpp:87-141,
   integrator/legendre.hpp:20-80, integrator/hexahedron.hpp:55-135, integrator/quadrangle.hpp:66-110.
   Column: [cell][mode][component] (expansion.hpp:83-90 on the column-major K x M Coeff). */
typedef struct {
  int m;                    /* modes per component */
  double gx[MAXQ], gw[MAXQ];
  double *xyz, *wgt;        /* [cell][n][3], [cell][n]: Gauss points and global weights */
  modal_basis_t *basis;     /* [cell] */
  double *fxyz, *fw;        /* [face][fq][3], [face][fq] */
  frame_t *frame;           /* [face][fq] */
} dgm_t;

static void legendre_rule(int q, double *x, double *w) { /* integrator/legendre.hpp:50-73 */
  if (q == 3) {
    x[0] = -sqrt(0.6); x[1] = 0; x[2] = sqrt(0.6);
    w[0] = 5.0 / 9.0; w[1] = 8.0 / 9.0; w[2] = 5.0 / 9.0;
  } else {
    double a = sqrt((3 + 2 * sqrt(1.2)) / 7), b = sqrt((3 - 2 * sqrt(1.2)) / 7);
    x[0] = -a; x[1] = -b; x[2] = b; x[3] = a;
    w[0] = w[3] = (18 - sqrt(30.0)) / 36; w[1] = w[2] = (18 + sqrt(30.0)) / 36;
  }
}
static void dgm_basis_build(const oracle_t *o, const dgm_t *g, int64_t c, const double *ctr, modal_basis_t *b) {
  int n = o->n, m = g->m;
  double A[MAXM][MAXM] = {{0}}, S[MAXM][MAXM] = {{0}};
  memcpy(b->center, ctr, sizeof b->center);
  for (int a = 0; a < n; ++a) {
    double d[3], t[MAXM], w = g->wgt[c * n + a];
    for (int k = 0; k < 3; ++k) d[k] = g->xyz[(c * n + a) * 3 + k] - ctr[k];
    taylor_values(o->degree, d, t);
    for (int i = 0; i < m; ++i) for (int j = 0; j < m; ++j) A[i][j] += t[i] * t[j] * w;
  }
  for (int i = 0; i < m; ++i) S[i][i] = 1;
  S[0][0] = 1 / sqrt(A[0][0]);
  for (int i = 1; i < m; ++i) {
    for (int j = 0; j < i; ++j) {
      double temp = 0;
      for (int k = 0; k <= j; ++k) temp += S[j][k] * A[k][i];
      for (int l = 0; l <= j; ++l) S[i][l] -= temp * S[j][l];
    }
    double norm_sq = 0;
    for (int j = 0; j <= i; ++j) {
      double sum = 0, sij = S[i][j];
      for (int k = 0; k < j; ++k) sum += 2 * S[i][k] * A[k][j];
      norm_sq += sij * (sij * A[j][j] + sum);
    }
    double nrm = sqrt(norm_sq);
    for (int j = 0; j < m; ++j) S[i][j] /= nrm;
  }
  memcpy(b->s, S, sizeof S);
}
static void dgm_setup(oracle_t *o, const double *nodes, const int64_t *cn) {
  dgm_t *g = calloc(1, sizeof *g);
  int q = o->q, n = o->n, fq = q * q;
  int64_t nc = o->n_cells;
  o->dgm = g;
  g->m = taylor_count(o->degree);
  o->kf = 5 * g->m;
  legendre_rule(q, g->gx, g->gw);
  g->xyz = malloc(sizeof(double) * nc * n * 3); g->wgt = malloc(sizeof(double) * nc * n);
  g->basis = malloc(sizeof(modal_basis_t) * nc);
  for (int64_t c = 0; c < nc; ++c) {
    double ctr[3] = {0, 0, 0};
    for (int v = 0; v < 8; ++v) for (int k = 0; k < 3; ++k) ctr[k] += nodes[cn[c * 8 + v] * 3 + k] / 8;
    for (int a = 0; a < n; ++a) {
      int ijk[3] = {a / fq, (a / q) % q, a % q};
      double loc[3] = {g->gx[ijk[0]], g->gx[ijk[1]], g->gx[ijk[2]]}, X[3] = {0, 0, 0}, J[3][3] = {{0}};
      for (int v = 0; v < 8; ++v) {
        const double *P = &nodes[cn[c * 8 + v] * 3];
        double fx = 1 + VERT[v][0] * loc[0], fy = 1 + VERT[v][1] * loc[1], fz = 1 + VERT[v][2] * loc[2];
        double shape = fx * fy * fz / 8;
        double gr[3] = {VERT[v][0] * (fy * fz / 8), VERT[v][1] * (fz * fx / 8), VERT[v][2] * (fx * fy / 8)};
        for (int k = 0; k < 3; ++k) { X[k] += shape * P[k]; for (int d = 0; d < 3; ++d) J[d][k] += gr[d] * P[k]; }
      }
      double c0[3];
      cross(J[1], J[2], c0);
      double det = J[0][0] * c0[0] + J[0][1] * c0[1] + J[0][2] * c0[2];
      memcpy(&g->xyz[(c * n + a) * 3], X, sizeof X);
      g->wgt[c * n + a] = g->gw[ijk[0]] * g->gw[ijk[1]] * g->gw[ijk[2]] * det;
    }
    dgm_basis_build(o, g, c, ctr, &g->basis[c]);
  }
  int64_t nf = o->n_faces;
  g->fxyz = malloc(sizeof(double) * nf * fq * 3); g->fw = malloc(sizeof(double) * nf * fq);
  g->frame = malloc(sizeof(frame_t) * nf * fq);
  for (int64_t f = 0; f < nf; ++f) {
    /* the holder's local face: the one whose four nodes are the face's (h_dir and the sign of h_scale name it) */
    int d = o->h_dir[f], plus = o->h_scale[f * fq] > 0;
    int lf = d == 0 ? (plus ? 2 : 4) : d == 1 ? (plus ? 3 : 1) : (plus ? 5 : 0);
    const double *P[4];
    for (int mm = 0; mm < 4; ++mm) P[mm] = &nodes[cn[o->holder[f] * 8 + FACE_NODES[lf][mm]] * 3];
    for (int ia = 0; ia < q; ++ia) for (int ib = 0; ib < q; ++ib) {
      int p = ia * q + ib;
      double a = g->gx[ia], b = g->gx[ib], X[3] = {0, 0, 0}, Ja[3] = {0, 0, 0}, Jb[3] = {0, 0, 0};
      for (int mm = 0; mm < 4; ++mm) {
        double fa = (1 + QUAD[mm][0] * a), fb = (1 + QUAD[mm][1] * b);
        for (int k = 0; k < 3; ++k) {
          X[k] += fa * fb / 4 * P[mm][k];
          Ja[k] += QUAD[mm][0] * (fb / 4) * P[mm][k];
          Jb[k] += QUAD[mm][1] * (fa / 4) * P[mm][k];
        }
      }
      memcpy(&g->fxyz[(f * fq + p) * 3], X, sizeof X);
      frame_t *fr = &g->frame[f * fq + p];
      double nn[3]; cross(Ja, Jb, nn);
      double len = norm3(nn), la = norm3(Ja);
      for (int k = 0; k < 3; ++k) { fr->a[k] = nn[k] / len; fr->b[k] = Ja[k] / la; }
      cross(fr->a, fr->b, fr->c);
      double aa = Ja[0] * Ja[0] + Ja[1] * Ja[1] + Ja[2] * Ja[2], bb = Jb[0] * Jb[0] + Jb[1] * Jb[1] + Jb[2] * Jb[2];
      double ab = Ja[0] * Jb[0] + Ja[1] * Jb[1] + Ja[2] * Jb[2];
      g->fw[f * fq + p] = g->gw[ia] * g->gw[ib] * sqrt(aa * bb - ab * ab);
    }
  }
}
static void dgm_free(oracle_t *o) {
  dgm_t *g = o->dgm;
  if (!g) return;
  free(g->xyz); free(g->wgt); free(g->basis); free(g->fxyz); free(g->fw); free(g->frame); free(g);
  o->dgm = NULL;
}
static void dgm_value(const oracle_t *o, const dgm_t *g, int64_t c, const double *u, const double *x, double *phi,
    double *val) { /* Projection::GlobalToValue (projection.hpp:99-104) */
  modal_basis_values(o, &g->basis[c], x, phi);
  for (int k = 0; k < 5; ++k) {
    double v = 0;
    for (int i = 0; i < g->m; ++i) v += u[(c * g->m + i) * 5 + k] * phi[i];
    val[k] = v;
  }
}
/* OrthoNormal::GetGradValue (basis/linear.hpp:131-135, basis/taylor.hpp:374-470): grad[mode][xyz] */
static void dgm_basis_gradients(const oracle_t *o, const modal_basis_t *b, const double *xg, double grad[MAXM][3]) {
  int m = taylor_count(o->degree);
  double x = xg[0] - b->center[0], y = xg[1] - b->center[1], z = xg[2] - b->center[2];
  double tg[MAXM][3] = {{0}};
  tg[1][0] = 1; tg[2][1] = 1; tg[3][2] = 1;
  tg[4][0] = 2 * x; tg[5][0] = y; tg[5][1] = x; tg[6][0] = z; tg[6][2] = x;
  tg[7][1] = 2 * y; tg[8][1] = z; tg[8][2] = y; tg[9][2] = 2 * z;
  if (m == 20) {
    double xx = x * x, xy = x * y, xz = x * z, yy = y * y, yz = y * z, zz = z * z;
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
  for (int i = 0; i < m; ++i) for (int d = 0; d < 3; ++d) {
    double v = 0;
    for (int j = 0; j <= i; ++j) v += b->s[i][j] * tg[j][d];
    grad[i][d] = v;
  }
}
static void dgm_residual(oracle_t *o, double t, const double *u, double *res) {
  dgm_t *g = o->dgm;
  int n = o->n, fq = o->q * o->q, m = g->m;
  int64_t nc = o->n_cells;
  memset(res, 0, sizeof(double) * nc * m * 5);
  for (int64_t c = 0; c < nc; ++c)
    for (int a = 0; a < n; ++a) { /* dg/general.hpp:58-69 */
      double phi[MAXM], val[5], F[5][3], grad[MAXM][3];
      const double *x = &g->xyz[(c * n + a) * 3];
      dgm_value(o, g, c, u, x, phi, val);
      flux_matrix(val, F);
      dgm_basis_gradients(o, &g->basis[c], x, grad);
      double w = g->wgt[c * n + a];
      for (int i = 0; i < m; ++i) for (int k = 0; k < 5; ++k)
        res[(c * m + i) * 5 + k] += (F[k][0] * w) * grad[i][0] + (F[k][1] * w) * grad[i][1] + (F[k][2] * w) * grad[i][2];
    }
  for (int64_t f = 0; f < o->n_faces; ++f) {
    int64_t ch = o->holder[f], cs = o->sharer[f];
    for (int p = 0; p < fq; ++p) {
      const frame_t *fr = &g->frame[f * fq + p];
      const double *x = &g->fxyz[(f * fq + p) * 3];
      double ph[MAXM], ps[MAXM], uh[5], us[5], fu[5];
      dgm_value(o, g, ch, u, x, ph, uh);
      if (cs >= 0) { /* dg/general.hpp:70-89 */
        dgm_value(o, g, cs, u, x, ps, us);
        flux_upwind(o->riemann, fr, uh, us, fu);
      } else { /* dg/general.hpp:92-200 */
        int tag = o->tag[f], kind = o->bc_kind[tag];
        double given[5] = {0, 0, 0, 0, 0};
        if (kind == 6) continue; /* not overridden: fem.hpp:382-383 */
        if (kind >= 2) eval_field(o->field_kind[tag], o->field_prm[tag], x, t, given);
        boundary_flux(o->riemann, kind, fr, uh, given, fu);
      }
      double w = g->fw[f * fq + p];
      for (int k = 0; k < 5; ++k) fu[k] *= -w;
      for (int i = 0; i < m; ++i) for (int k = 0; k < 5; ++k) res[(ch * m + i) * 5 + k] += fu[k] * ph[i];
      if (cs >= 0) for (int i = 0; i < m; ++i) for (int k = 0; k < 5; ++k) res[(cs * m + i) * 5 + k] += -fu[k] * ps[i];
    }
  }
}
static double dgm_min_dt(oracle_t *o, const double *u) { /* fem.hpp:417-432 with Cell::N = M quadrature points */
  dgm_t *g = o->dgm;
  double min_dt = 1e100;
  for (int64_t c = 0; c < o->n_cells; ++c) {
    double vmax = 1e-5;
    for (int a = 0; a < g->m; ++a) {
      double phi[MAXM], val[5];
      dgm_value(o, g, c, u, &g->xyz[(c * o->n + a) * 3], phi, val);
      prim_t q = cons_to_prim(val);
      double cs = sound_speed(q.rho, q.p), v = sqrt(q.u * q.u + q.v * q.v + q.w * q.w);
      double s = cs + v > cs - v ? cs + v : cs - v;
      if (s > vmax) vmax = s;
    }
    double d = o->length[c] / vmax;
    if (d < min_dt) min_dt = d;
  }
  return min_dt;
}
static void dgm_approximate(oracle_t *o, int field_kind, const double *prm, double *column) { /* projection.hpp:28-43 */
  dgm_t *g = o->dgm;
  int n = o->n, m = g->m;
  memset(column, 0, sizeof(double) * o->n_cells * m * 5);
  for (int64_t c = 0; c < o->n_cells; ++c)
    for (int a = 0; a < n; ++a) {
      double phi[MAXM], val[5];
      const double *x = &g->xyz[(c * n + a) * 3];
      eval_field(field_kind, prm, x, 0.0, val);
      modal_basis_values(o, &g->basis[c], x, phi);
      for (int i = 0; i < m; ++i) for (int k = 0; k < 5; ++k) column[(c * m + i) * 5 + k] += val[k] * phi[i] * g->wgt[c * n + a];
    }
}

/* limiter::Reconstruct(part, weno::Lazy(w0, eps)) on Projection cells (single part): as oracle_weno_lazy, but the
   modal coefficients are the state itself and every integral uses the cell's Gauss-Legendre rule */
static double dgm_smoothness(const oracle_t *o, const dgm_t *g, int64_t c, const double *coef) {
  int n = o->n, m = g->m;
  const modal_basis_t *b = &g->basis[c];
  double tc[MAXM], integral[MAXM] = {0};
  for (int j = 0; j < m; ++j) {
    double v = 0;
    for (int i = j; i < m; ++i) v += coef[i] * b->s[i][j];
    tc[j] = v;
  }
  for (int a = 0; a < n; ++a) {
    double d[3], pdv[MAXM], w = g->wgt[c * n + a];
    for (int k = 0; k < 3; ++k) d[k] = g->xyz[(c * n + a) * 3 + k] - b->center[k];
    taylor_partial_derivatives(o->degree, d, tc, pdv);
    for (int i = 0; i < m; ++i) integral[i] += pdv[i] * pdv[i] * w;
  }
  double volume = 1 / (b->s[0][0] * b->s[0][0]);
  double total = (integral[1] + integral[2] + integral[3]) * (pow(volume, 1 * 2. / 3 - 1) / (tgamma(2) * tgamma(2)));
  static const int begin[4] = {1, 4, 10, 20};
  for (int p = 2; p <= o->degree; ++p) {
    double sm = 0;
    for (int i = begin[p - 1]; i < begin[p]; ++i) sm += integral[i];
    total += sm * (pow(volume, p * 2. / 3 - 1) / (tgamma(p + 1) * tgamma(p + 1)));
  }
  return total;
}
void oracle_dgm_weno_lazy(void *h, double w0, double eps, double *u) {
  oracle_t *o = h;
  dgm_t *g = o->dgm;
  int n = o->n, m = g->m;
  int64_t nc = o->n_cells;
  double *fresh = malloc(sizeof(double) * nc * 5 * m);
  int *adj_cnt = calloc(nc, sizeof(int));
  int64_t *adj = malloc(sizeof(int64_t) * nc * 6);
  for (int64_t f = 0; f < o->n_interior; ++f) {
    int64_t a = o->holder[f], b = o->sharer[f];
    adj[a * 6 + adj_cnt[a]++] = b;
    adj[b * 6 + adj_cnt[b]++] = a;
  }
  for (int64_t c = 0; c < nc; ++c) {
    const modal_basis_t *mb = &g->basis[c];
    int cnt = adj_cnt[c];
    double cand[7][5][MAXM], mine[5][MAXM];
    double s00 = mb->s[0][0];
    for (int k = 0; k < 5; ++k) for (int i = 0; i < m; ++i) mine[k][i] = u[(c * m + i) * 5 + k];
    for (int j = 0; j < cnt; ++j) {
      int64_t d = adj[c * 6 + j];
      memset(cand[j], 0, sizeof cand[j]);
      for (int a = 0; a < n; ++a) {
        double phi[MAXM], phn[MAXM], w = g->wgt[c * n + a];
        const double *x = &g->xyz[(c * n + a) * 3];
        modal_basis_values(o, mb, x, phi);
        modal_basis_values(o, &g->basis[d], x, phn);
        for (int k = 0; k < 5; ++k) {
          double v = 0;
          for (int i = 0; i < m; ++i) v += u[(d * m + i) * 5 + k] * phn[i];
          for (int i = 0; i < m; ++i) cand[j][k][i] += v * phi[i] * w;
        }
      }
      for (int k = 0; k < 5; ++k) cand[j][k][0] += (mine[k][0] * s00 - cand[j][k][0] * s00) / s00;
    }
    for (int k = 0; k < 5; ++k) for (int i = 0; i < m; ++i) cand[cnt][k][i] = mine[k][i];
    for (int k = 0; k < 5; ++k) {
      double wgt[7], sum = 0;
      for (int j = 0; j <= cnt; ++j) {
        double beta = dgm_smoothness(o, g, c, cand[j][k]) + eps;
        double w = j < cnt ? w0 : w0 * (-cnt) + 1.0;
        wgt[j] = w / (beta * beta);
        sum += wgt[j];
      }
      for (int j = 0; j <= cnt; ++j) wgt[j] /= sum;
      for (int i = 0; i < m; ++i) {
        double v = cand[cnt][k][i] * wgt[cnt];
        for (int j = 0; j < cnt; ++j) v += cand[j][k][i] * wgt[j];
        fresh[(c * m + i) * 5 + k] = v;
      }
    }
  }
  memcpy(u, fresh, sizeof(double) * nc * 5 * m);
  free(fresh); free(adj_cnt); free(adj);
}
