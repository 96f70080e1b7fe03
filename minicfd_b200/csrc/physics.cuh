// physics.cuh — ideal-gas Euler fluxes, Riemann solvers and boundary fluxes, device side.
//
// Restates the arithmetic of the reference's include/mini/riemann/euler/{types,exact,hllc,ausm}.hpp
// and include/mini/riemann/rotated/euler.hpp for one flux point.  Written tangent-free: the
// reference rotates the momentum into the face frame (normal, tangent, bitangent) and back
// (rotated/euler.hpp:117-134); all three solvers treat the tangential velocity as passively
// advected, so the tangential momentum flux is a scalar times the tangential velocity *vector* of
// one side (or, for AUSM, of each side).  We therefore carry  v_t = v - (v.n) n  as a 3-vector and
// never need the tangents.  Identical in exact arithmetic, differs by rounding only.
#ifndef MINICFD_B200_PHYSICS_CUH_
#define MINICFD_B200_PHYSICS_CUH_

#ifdef __CUDACC__
#include <cuda_runtime.h>
#define MCFD_HD __device__ __forceinline__
#define MCFD_NOINLINE __device__ __noinline__
#define MCFD_CONSTANT __constant__
#else  // host build of the same source: tests/host_emulation only
#include <cmath>
#define MCFD_HD inline
#define MCFD_NOINLINE inline
#define MCFD_CONSTANT static
using std::fma; using std::fmax; using std::fmin; using std::pow; using std::sqrt;
#endif

namespace mcfd {

// IdealGas<double, 1.4> constants (types.hpp:206-247)
constexpr double kGamma = 1.4;
constexpr double kGasR = 287.05;
constexpr double kGm1 = kGamma - 1;                       // GammaMinusOne
constexpr double kOneOverGm1 = 1 / kGm1;                  // OneOverGammaMinusOne
constexpr double kGp1Over2 = (kGamma + 1) / 2;            // GammaPlusOneOverTwo
constexpr double kGp1Over4 = (kGamma + 1) / 4;            // GammaPlusOneOverFour
constexpr double kGm1Over2 = kGm1 / 2;                    // GammaMinusOneOverTwo
constexpr double kGm1Under2 = 2 / kGm1;                   // GammaMinusOneUnderTwo
constexpr double kOneOverGamma = 1 / kGamma;
constexpr double kGammaOverGm1 = kGamma / kGm1;

enum : int { kExact = 0, kHllc = 1, kAusm = 2 };

// Reciprocal, reciprocal square root, square root and quotient to ~1 ulp for NORMAL positive (Rcp:
// non-zero) arguments: the 20-bit MUFU seed plus two Newton steps in FP64.  The IEEE-exact `/` and
// sqrt() expand to 20-30 instructions each with special-case branches; a face point needs about a
// dozen of them, so they dominated the FP64 pipe.  The results differ from the correctly rounded
// ones by at most a unit in the last place, far inside the 1e-12 parity bound; zero, subnormal,
// infinite and NaN arguments are the caller's business (every call site guards or cannot see them).
#ifdef __CUDACC__
MCFD_HD double Rcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  r = fma(r, fma(-x, r, 1.0), r);
  r = fma(r, fma(-x, r, 1.0), r);
  return r;
}
MCFD_HD double Rsqrt(double x) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  double e = fma(-x * y, y, 1.0);
  y = fma(0.5 * y, e, y);
  e = fma(-x * y, y, 1.0);
  y = fma(0.5 * y, e, y);
  return y;
}
MCFD_HD double Sqrt(double x) {
  const double y = Rsqrt(x);
  const double s = x * y;
  return fma(fma(-s, s, x), 0.5 * y, s);
}
#else
MCFD_HD double Rcp(double x) { return 1.0 / x; }
MCFD_HD double Rsqrt(double x) { return 1.0 / sqrt(x); }
MCFD_HD double Sqrt(double x) { return sqrt(x); }
#endif
MCFD_HD double Div(double a, double b) { return a * Rcp(b); }

// pow() expands to several hundred instructions; one shared copy keeps the stage kernel's hot loop
// inside the instruction cache.
MCFD_NOINLINE double Pow(double x, double y) { return pow(x, y); }

struct Prim {
  double rho, u, v, w, p;  // velocity in the global frame
};

// IdealGas::ConservativeToPrimitive incl. SetZeroIfNegative (types.hpp:194-202, 289-301).
MCFD_HD Prim ConsToPrim(double r, double mx, double my, double mz, double e) {
  Prim q;
  if (r < 0 || e < 0) {  // the energy slot is tested before it becomes a pressure
    q.rho = q.u = q.v = q.w = q.p = 0;
    return q;
  }
  if (r != 0) {
    double inv = Rcp(r);
    q.rho = r;
    q.u = mx * inv;
    q.v = my * inv;
    q.w = mz * inv;
    double ke = 0.5 * (q.u * q.u + q.v * q.v + q.w * q.w);
    q.p = (e - r * ke) * kGm1;
    if (q.p < 0) {
      q.rho = q.u = q.v = q.w = q.p = 0;
    }
  } else {  // rho == 0: returned unchanged by the reference
    q.rho = 0; q.u = mx; q.v = my; q.w = mz; q.p = e;
  }
  return q;
}

// IdealGas::GetSpeedOfSound (types.hpp:252-254)
MCFD_HD double SoundSpeed(double rho, double p) {
  return (rho > 0 && p > 0) ? Sqrt(kGamma * Div(p, rho)) : 0.0;
}

// Result of a 1-D solver in the normal frame: mass, normal momentum and energy flux plus the
// factors multiplying the left / right tangential velocity vectors.
struct NormalFlux {
  double mass, mom_n, energy, c_left, c_right;
};

// Gas::PrimitiveToFlux (types.hpp:313-320) for a state with normal velocity un and kinetic energy ke.
MCFD_HD void StateFlux(double rho, double un, double p, double ke,
    double *mass, double *mom_n, double *energy) {
  double e = p * kOneOverGm1 + rho * ke;
  double m = rho * un;
  *mass = m;
  *mom_n = m * un + p;
  *energy = e * un + p * un;
}

// ---- HLLC (hllc.hpp:30-44, 54-109) -------------------------------------------------------------
MCFD_HD double HllcQ(double p_est, double p_k) {   // hllc.hpp:69-77
  if (p_est <= p_k) return 1.0;
  return Sqrt(1 + (kGp1Over2 / kGamma) * (Div(p_est, p_k) - 1));
}

MCFD_HD NormalFlux SolveHllc(double rl, double ul, double pl, double kel,
    double rr, double ur, double pr, double ker) {
  double rho_avg = (rl + rr) / 2;
  double al = SoundSpeed(rl, pl), ar = SoundSpeed(rr, pr);
  double a_avg = (al + ar) / 2;
  double p_pvrs = (pl + pr - (ur - ul) * rho_avg * a_avg) / 2;
  double p_est = fmax(0.0, p_pvrs);
  double wl = ul - al * HllcQ(p_est, pl);
  double wr = ur + ar * HllcQ(p_est, pr);
  double ws = Div(pr - pl + rl * ul * (wl - ul) - rr * ur * (wr - ur),
                  rl * (wl - ul) - rr * (wr - ur));
  // The four cases of hllc.hpp:34-43 without divergent control flow: pick the side K whose state
  // the flux starts from, then add the star-region jump S_K (U*_K - U_K) unless the t-axis lies
  // outside the fan on that side.
  const bool left = (0.0 <= wl) || (!(wr <= 0.0) && (0.0 <= ws));
  const bool star = !(0.0 <= wl) && !(wr <= 0.0);
  const double rho = left ? rl : rr, un = left ? ul : ur, p = left ? pl : pr, ke = left ? kel : ker;
  const double wk = left ? wl : wr;
  NormalFlux f;
  StateFlux(rho, un, p, ke, &f.mass, &f.mom_n, &f.energy);
  if (star) {   // uniform across a warp almost everywhere (subsonic normal velocity)
    const double energy = p * kOneOverGm1 + rho * ke;
    const double dk = wk - un;
    const double temp = rho * Div(dk, wk - ws);
    double e_star = Div(energy, rho) + (ws - un) * (ws + Div(p, rho * dk));
    e_star *= temp;
    f.mass += (temp - rho) * wk;
    f.mom_n += (temp * ws - rho * un) * wk;
    f.energy += (e_star - energy) * wk;
  }
  f.c_left = left ? f.mass : 0.0;
  f.c_right = left ? 0.0 : f.mass;
  return f;
}

// ---- AUSM (ausm.hpp:30-33, 52-71) --------------------------------------------------------------
template <bool kNegative>
MCFD_HD void AusmSigned(double rho, double un, double p_in, double ke,
    double *m_out, double *p_out, double *h_out) {
  double a = SoundSpeed(rho, p_in);
  double h = a * a * kOneOverGm1 + ke;
  double mach = un / a;  // a may be 0 for a vacuum state: keep the IEEE quotient here
  double p = p_in;
  double signed_mach = (1 + (kNegative ? -mach : mach)) * 0.5;
  if (mach >= -1 && mach <= 1) {
    mach = kNegative ? -(signed_mach * signed_mach) : signed_mach * signed_mach;
    p *= signed_mach;
  } else if (signed_mach < 0) {
    mach = 0.0;
    p = 0.0;
  }
  *m_out = rho * a * mach;
  *p_out = p;
  *h_out = h;
}

MCFD_HD NormalFlux SolveAusm(double rl, double ul, double pl, double kel,
    double rr, double ur, double pr, double ker) {
  double ml, ppl, hl, mr, ppr, hr;
  AusmSigned<false>(rl, ul, pl, kel, &ml, &ppl, &hl);
  AusmSigned<true>(rr, ur, pr, ker, &mr, &ppr, &hr);
  NormalFlux f;
  f.mass = ml + mr;
  f.mom_n = (ml * ul + ppl) + (mr * ur + ppr);
  f.energy = ml * hl + mr * hr;
  f.c_left = ml;
  f.c_right = mr;
  return f;
}

// ---- Exact (exact.hpp:36-46, 62-320) -----------------------------------------------------------
struct SpeedChange {  // exact.hpp:153-200
  double rho_b, p_b, a_b, p_const;
  MCFD_HD SpeedChange(double rho, double p)
      : rho_b(rho), p_b(p), a_b(rho == 0 ? 0 : SoundSpeed(rho, p)), p_const(p * kGm1Over2) {}
  MCFD_HD double P(double pa) const { return kGp1Over2 * pa + p_const; }
  MCFD_HD double Value(double pa) const {
    if (pa >= p_b) {
      return (pa - p_b) / sqrt(rho_b * P(pa));
    }
    constexpr double e = kGm1Over2 / kGamma;
    return (Pow(pa / p_b, e) - 1) * (kGm1Under2 * a_b);
  }
  MCFD_HD double Prime(double pa) const {
    if (pa >= p_b) {
      double pp = P(pa);
      double v = kGp1Over4 * (p_b - pa);
      return (1 + v / pp) / sqrt(rho_b * pp);
    }
    constexpr double e = -kGp1Over2 / kGamma;
    return Pow(pa / p_b, e) / (rho_b * a_b);
  }
};

struct Rup {  // (rho, u, p) on the t-axis
  double rho, u, p;
};

MCFD_HD Rup InsideExpansion(double gri_1, double gri_2) {  // exact.hpp:255-266
  constexpr double r = kGm1 / (kGamma + 1);
  double a = r * gri_2;
  double a2 = a * a;
  double rho = Pow(a2 / gri_1 * kOneOverGamma, kOneOverGm1);
  return Rup{rho, a, a2 * rho * kOneOverGamma};
}

template <int kField>
MCFD_HD Rup NearShock(double rb, double ub, double pb, double us, double ps) {
  // exact.hpp:203-234
  double divisor = (us - ub) * rb;
  double shock_u = ub + (divisor != 0 ? (ps - pb) / divisor : ub);
  bool after = (kField == 1) ? (shock_u < 0) : (shock_u > 0);
  if (after) {
    double rho = rb * ((ub - shock_u) / (us - shock_u));
    return Rup{rho, us, ps};
  }
  return Rup{rb, ub, pb};
}

template <int kField>
MCFD_HD Rup NearExpansion(double rb, double ub, double pb, double us, double ps) {
  // exact.hpp:236-292
  double a_before = SoundSpeed(rb, pb);
  double a_after = (kField == 1) ? a_before + kGm1Over2 * (ub - us) : a_before - kGm1Over2 * (ub - us);
  bool after = (kField == 1) ? (us - a_after < 0) : (us + a_after > 0);
  if (after) {
    return Rup{kGamma * ps / (a_after * a_after), us, ps};
  }
  bool before = (kField == 1) ? (ub - a_before > 0) : (ub + a_before < 0);
  if (before) {
    return Rup{rb, ub, pb};
  }
  double gri_1 = pb / Pow(rb, kGamma);
  double gri_2 = (kField == 1) ? ub + kGm1Under2 * a_before : ub - kGm1Under2 * a_before;
  return InsideExpansion(gri_1, gri_2);
}

// Returns the t-axis state; *star_u receives the contact speed used to upwind the passive
// (tangential) velocity.  The reference keeps star_u as mutable solver state that is only written
// in the ordinary (non-vacuum) case (exact.hpp:29, 120); a freshly constructed solver has 0, which
// is what the vacuum branches see here.
MCFD_HD Rup SolveExactAxis(double rl, double ul, double pl,
    double rr, double ur, double pr, double *star_u) {
  *star_u = 0.0;
  if (rl > 0) {
    if (!(rr > 0)) {  // only right vacuum (exact.hpp:68-83)
      double al = SoundSpeed(rl, pl);
      if (ul >= al) return Rup{rl, ul, pl};
      double us = ul + al * kGm1Under2;
      if (us <= 0) return Rup{0, 0, 0};
      return InsideExpansion(pl / Pow(rl, kGamma), us);
    }
  } else {
    if (rr > 0) {  // only left vacuum (exact.hpp:85-98)
      double ar = SoundSpeed(rr, pr);
      if (ur + ar <= 0) return Rup{rr, ur, pr};
      double us = ur - ar * kGm1Under2;
      if (us >= 0) return Rup{0, 0, 0};
      return InsideExpansion(pr / Pow(rr, kGamma), us);
    }
    return Rup{0, 0, 0};
  }
  double du = ur - ul;
  SpeedChange cl(rl, pl), cr(rr, pr);
  if (cl.Value(0.0) + cr.Value(0.0) + du < 0) {  // ordinary case (exact.hpp:115-133)
    // FindRoot (exact.hpp:140-152), eps = 1e-8
    double x = pl;
    double fx = cl.Value(x) + cr.Value(x) + du;
    while (fx > 0) {
      x *= 0.5;
      fx = cl.Value(x) + cr.Value(x) + du;
    }
    while (fx < -1e-8) {
      x -= fx / (cl.Prime(x) + cr.Prime(x));
      fx = cl.Value(x) + cr.Value(x) + du;
    }
    double ps = x;
    double us = 0.5 * (ur + cr.Value(ps) + ul - cl.Value(ps));
    *star_u = us;
    if (0 < us) {
      return (ps >= pl) ? NearShock<1>(rl, ul, pl, us, ps) : NearExpansion<1>(rl, ul, pl, us, ps);
    }
    return (ps >= pr) ? NearShock<3>(rr, ur, pr, us, ps) : NearExpansion<3>(rr, ur, pr, us, ps);
  }
  // vacuum between wave 1 and wave 3 (exact.hpp:293-320)
  double al = SoundSpeed(rl, pl);
  if (ul > al) return Rup{rl, ul, pl};
  double ar = SoundSpeed(rr, pr);
  if (ur + ar < 0) return Rup{rr, ur, pr};
  double gri_1 = pl / Pow(rl, kGamma);
  double gri_2 = ul + al * kGm1Under2;
  if (gri_2 >= 0) return InsideExpansion(gri_1, gri_2);
  gri_1 = pr / Pow(rr, kGamma);
  gri_2 = ur - ar * kGm1Under2;
  if (gri_2 < 0) return InsideExpansion(gri_1, gri_2);
  return Rup{0, 0, 0};
}

MCFD_HD NormalFlux SolveExact(double rl, double ul, double pl, double vt2l,
    double rr, double ur, double pr, double vt2r) {
  double star_u;
  Rup s = SolveExactAxis(rl, ul, pl, rr, ur, pr, &star_u);
  bool from_left = star_u > 0;  // SetPassiveScalars (exact.hpp:56-60)
  double vt2 = from_left ? vt2l : vt2r;
  double ke = 0.5 * (s.u * s.u + vt2);
  NormalFlux f;
  StateFlux(s.rho, s.u, s.p, ke, &f.mass, &f.mom_n, &f.energy);
  f.c_left = from_left ? f.mass : 0.0;
  f.c_right = from_left ? 0.0 : f.mass;
  return f;
}

// rotated::Euler::GetFluxUpwind (rotated/euler.hpp:153-159): conservative states in, global flux out,
// along the unit normal n (pointing left -> right).
template <int kRiemann>
MCFD_HD void FluxUpwind(const double *ul, const double *ur,
    double nx, double ny, double nz, double *flux) {
  Prim l = ConsToPrim(ul[0], ul[1], ul[2], ul[3], ul[4]);
  Prim r = ConsToPrim(ur[0], ur[1], ur[2], ur[3], ur[4]);
  double unl = l.u * nx + l.v * ny + l.w * nz;
  double unr = r.u * nx + r.v * ny + r.w * nz;
  double tlx = l.u - unl * nx, tly = l.v - unl * ny, tlz = l.w - unl * nz;
  double trx = r.u - unr * nx, try_ = r.v - unr * ny, trz = r.w - unr * nz;
  NormalFlux f;
  if (kRiemann == kExact) {
    f = SolveExact(l.rho, unl, l.p, tlx * tlx + tly * tly + tlz * tlz,
                   r.rho, unr, r.p, trx * trx + try_ * try_ + trz * trz);
  } else {
    double kel = 0.5 * (l.u * l.u + l.v * l.v + l.w * l.w);
    double ker = 0.5 * (r.u * r.u + r.v * r.v + r.w * r.w);
    f = (kRiemann == kHllc) ? SolveHllc(l.rho, unl, l.p, kel, r.rho, unr, r.p, ker)
                            : SolveAusm(l.rho, unl, l.p, kel, r.rho, unr, r.p, ker);
  }
  flux[0] = f.mass;
  flux[1] = f.mom_n * nx + f.c_left * tlx + f.c_right * trx;
  flux[2] = f.mom_n * ny + f.c_left * tly + f.c_right * try_;
  flux[3] = f.mom_n * nz + f.c_left * tlz + f.c_right * trz;
  flux[4] = f.energy;
}

// Flux of one primitive state along n: GlobalPrimitiveToGlobalFlux (rotated/euler.hpp:68-73).
MCFD_HD void PrimFluxAlong(const Prim &q, double nx, double ny, double nz,
    double *flux) {
  double un = q.u * nx + q.v * ny + q.w * nz;
  double ke = 0.5 * (q.u * q.u + q.v * q.v + q.w * q.w);
  double e = q.p * kOneOverGm1 + q.rho * ke;
  double m = q.rho * un;
  flux[0] = m;
  flux[1] = m * q.u + q.p * nx;
  flux[2] = m * q.v + q.p * ny;
  flux[3] = m * q.w + q.p * nz;
  flux[4] = e * un + q.p * un;
}

enum : int {
  kBcWall = 0, kBcSupOut = 1, kBcSupIn = 2, kBcSubIn = 3, kBcSubOut = 4, kBcSmart = 5, kBcNoSlip = 6
};

// Boundary flux along the outward unit normal (rotated/euler.hpp:144-174, 181-236, 260-280, 296-299).
// `given` is the pre-sampled value of the boundary function at this flux point.
template <int kRiemann>
MCFD_HD void BoundaryFlux(int kind, const double *ui, const double *given,
    double nx, double ny, double nz, double *flux) {
  switch (kind) {
  case kBcWall: {
    Prim q = ConsToPrim(ui[0], ui[1], ui[2], ui[3], ui[4]);
    flux[0] = 0; flux[1] = q.p * nx; flux[2] = q.p * ny; flux[3] = q.p * nz; flux[4] = 0;
    break;
  }
  case kBcSupOut: {
    Prim q = ConsToPrim(ui[0], ui[1], ui[2], ui[3], ui[4]);
    PrimFluxAlong(q, nx, ny, nz, flux);
    break;
  }
  case kBcSupIn: {
    Prim q = ConsToPrim(given[0], given[1], given[2], given[3], given[4]);
    PrimFluxAlong(q, nx, ny, nz, flux);
    break;
  }
  case kBcSubIn: {  // Carlson (2011) 2.7, rotated/euler.hpp:181-236
    Prim q = ConsToPrim(ui[0], ui[1], ui[2], ui[3], ui[4]);
    double c_int = SoundSpeed(q.rho, q.p);
    double un = q.u * nx + q.v * ny + q.w * nz;
    double riemann_int = -un - c_int * kGm1Under2;
    double u_cos = given[1], v_cos = given[2], w_cos = given[3], t_total = given[4];
    double a_dir = nx * u_cos + ny * v_cos + nz * w_cos;
    double a2 = a_dir * a_dir;
    const double gamma_r = kGamma * kGasR;
    double qa = a2 + kGm1Under2;
    double qb_half = riemann_int;
    double qc = kGm1Over2 * (riemann_int * riemann_int) - a2 * gamma_r * t_total;
    double sqrt_delta = sqrt(qb_half * qb_half - qa * qc);
    double c_b = fmax((-qb_half + sqrt_delta) / qa, (-qb_half - sqrt_delta) / qa);
    double t_b = c_b * c_b / gamma_r;
    double mach_b = sqrt((t_total / t_b - 1) / kGm1Over2);
    double p_b = given[0] / Pow(1 + kGm1Over2 * mach_b * mach_b, kGammaOverGm1);
    double speed = mach_b * c_b;
    Prim b;
    b.p = p_b;
    b.rho = p_b / kGasR / t_b;
    b.u = u_cos * speed; b.v = v_cos * speed; b.w = w_cos * speed;
    PrimFluxAlong(b, nx, ny, nz, flux);
    break;
  }
  case kBcSubOut: {  // Carlson (2011) 2.4, rotated/euler.hpp:260-280
    Prim q = ConsToPrim(ui[0], ui[1], ui[2], ui[3], ui[4]);
    double rt = q.p / q.rho;
    q.p = given[4];
    q.rho = q.p / rt;
    PrimFluxAlong(q, nx, ny, nz, flux);
    break;
  }
  case kBcSmart:  // early return at rotated/euler.hpp:299
    FluxUpwind<kRiemann>(ui, given, nx, ny, nz, flux);
    break;
  default:  // kBcNoSlip: no-op for inviscid systems (fem.hpp:382-383) -> caller skips
    flux[0] = flux[1] = flux[2] = flux[3] = flux[4] = 0;
  }
}

struct State5 {
  double v[5];
};

// Out-of-line boundary flux with by-value arguments: keeps the (rare) boundary code, incl. its own
// copy of the Riemann solver for the smart boundary, out of the interior-face instruction stream
// without forcing the caller's state arrays into local memory.
template <int kRiemann>
MCFD_NOINLINE State5 BoundaryFluxOutOfLine(int kind, State5 ui, State5 given, double nx, double ny,
    double nz) {
  State5 f;
  BoundaryFlux<kRiemann>(kind, ui.v, given.v, nx, ny, nz, f.v);
  return f;
}

// ---- diffusive part: NavierStokes<Gas> (riemann/diffusive/navier_stokes.hpp) ----------------------
// Gradients are g[direction x,y,z][component].  `cond` = nu * Cp / Pr, so that kappa = cond * rho
// (navier_stokes.hpp:70-79; Cp = R gamma / (gamma - 1), types.hpp:245).
struct Viscosity {
  double beta0, beta1;   // Direct::SetBetaValues (direct.hpp:80-83)
  double nu, cond;       // NavierStokes::SetProperty (navier_stokes.hpp:57-60)
};

// NavierStokes::ConservativeToPrimitive with gradients (navier_stokes.hpp:91-125): velocity
// gradient gv[direction][velocity component] and temperature gradient gt[direction].
MCFD_HD Prim NsPrimitiveGradient(const double *u, const double (*g)[5], double (*gv)[3], double *gt) {
  Prim q = ConsToPrim(u[0], u[1], u[2], u[3], u[4]);
  const double inv_rho = Rcp(q.rho);
#pragma unroll
  for (int x = 0; x < 3; ++x) {
    gv[x][0] = (g[x][1] - q.u * g[x][0]) * inv_rho;
    gv[x][1] = (g[x][2] - q.v * g[x][0]) * inv_rho;
    gv[x][2] = (g[x][3] - q.w * g[x][0]) * inv_rho;
    double gp = q.u * g[x][1] + u[1] * gv[x][0];
    gp += q.v * g[x][2] + u[2] * gv[x][1];
    gp += q.w * g[x][3] + u[3] * gv[x][2];
    gp = (g[x][4] - 0.5 * gp) * kGm1;
    gt[x] = (gp * inv_rho - (q.p * inv_rho * inv_rho) * g[x][0]) * (1.0 / kGasR);
  }
  return q;
}

// GetViscousStressTensor with Stokes' hypothesis (navier_stokes.hpp:127-142); tau = XX XY XZ YY YZ ZZ
MCFD_HD void NsStress(const double (*gv)[3], double mu, double *tau) {
  const double div = gv[0][0] + gv[1][1] + gv[2][2];
  const double zeta = (-2.0 / 3) * mu;
  tau[0] = 2 * mu * gv[0][0] + zeta * div;
  tau[3] = 2 * mu * gv[1][1] + zeta * div;
  tau[5] = 2 * mu * gv[2][2] + zeta * div;
  tau[1] = mu * (gv[1][0] + gv[0][1]);
  tau[4] = mu * (gv[2][1] + gv[1][2]);
  tau[2] = mu * (gv[0][2] + gv[2][0]);
}

MCFD_HD double SymTau(const double *tau, int x, int y) {
  return tau[x == y ? (x == 0 ? 0 : (x == 1 ? 3 : 5)) : (x + y == 1 ? 1 : (x + y == 2 ? 2 : 4))];
}

// MinusViscousFlux on a flux matrix fx[direction][component] (navier_stokes.hpp:150-175)
MCFD_HD void NsMinusFluxMatrix(const Viscosity &vis, const double *u, const double (*g)[5], double (*fx)[5]) {
  double gv[3][3], gt[3], tau[6];
  const Prim q = NsPrimitiveGradient(u, g, gv, gt);
  NsStress(gv, vis.nu * u[0], tau);
  const double kappa = vis.cond * u[0];
#pragma unroll
  for (int x = 0; x < 3; ++x) {
    const double t0 = SymTau(tau, x, 0), t1 = SymTau(tau, x, 1), t2 = SymTau(tau, x, 2);
    fx[x][1] -= t0; fx[x][2] -= t1; fx[x][3] -= t2;
    fx[x][4] -= kappa * gt[x] + (t0 * q.u + t1 * q.v + t2 * q.w);
  }
}

// _MinusViscousFlux (navier_stokes.hpp:178-190): normal flux with given velocity and n . grad T
MCFD_HD void NsMinusFluxCore(const Viscosity &vis, double rho, const double (*gv)[3], double nx, double ny,
    double nz, double vu, double vv, double vw, double n_dot_grad_t, double *flux) {
  double tau[6];
  NsStress(gv, vis.nu * rho, tau);
  double work[3];
#pragma unroll
  for (int x = 0; x < 3; ++x) {
    const double t0 = SymTau(tau, x, 0), t1 = SymTau(tau, x, 1), t2 = SymTau(tau, x, 2);
    flux[1 + x] -= t0 * nx + t1 * ny + t2 * nz;
    work[x] = t0 * vu + t1 * vv + t2 * vw;
  }
  flux[4] -= (work[0] * nx + work[1] * ny + work[2] * nz) + (vis.cond * rho) * n_dot_grad_t;
}

// MinusViscousFlux along a normal (navier_stokes.hpp:193-202)
MCFD_HD void NsMinusFluxNormal(const Viscosity &vis, const double *u, const double (*g)[5], double nx,
    double ny, double nz, double *flux) {
  double gv[3][3], gt[3];
  const Prim q = NsPrimitiveGradient(u, g, gv, gt);
  NsMinusFluxCore(vis, u[0], gv, nx, ny, nz, q.u, q.v, q.w, nx * gt[0] + ny * gt[1] + nz * gt[2], flux);
}

// MinusViscousFluxOnNoSlipWall (navier_stokes.hpp:204-224): `wall` = (unused, u, v, w, n . grad T)
MCFD_HD void NsMinusFluxNoSlip(const Viscosity &vis, const double *wall, const double *u, const double (*g)[5],
    double nx, double ny, double nz, double value_penalty, double *flux) {
  double gv[3][3], gt[3];
  const Prim q = NsPrimitiveGradient(u, g, gv, gt);
  const double pen[3] = {value_penalty * (wall[1] - q.u), value_penalty * (wall[2] - q.v),
                         value_penalty * (wall[3] - q.w)};
  const double n[3] = {nx, ny, nz};
#pragma unroll
  for (int x = 0; x < 3; ++x) {
#pragma unroll
    for (int k = 0; k < 3; ++k) gv[x][k] += n[x] * pen[k];
  }
  NsMinusFluxCore(vis, u[0], gv, nx, ny, nz, wall[1], wall[2], wall[3], wall[4], flux);
}

// ---- component-wise isotropic diffusion --------------------------------------------------------------
// diffusive::Isotropic (riemann/diffusive/linear.hpp:66-92, 103-121) with one nu for everything, or the flux that
// EnergyBasedViscosity adds with one nu per component and cell (spatial/viscosity.hpp:114-129): F_k -= nu_k grad u_k.
MCFD_HD void IsoMinusFluxMatrix(const double *nu, const double (*g)[5], double (*fx)[5]) {
#pragma unroll
  for (int x = 0; x < 3; ++x) {
#pragma unroll
    for (int k = 0; k < 5; ++k) fx[x][k] -= nu[k] * g[x][k];
  }
}
MCFD_HD void IsoMinusFluxNormal(const double *nu, const double (*g)[5], double nx, double ny, double nz, double *flux) {
#pragma unroll
  for (int k = 0; k < 5; ++k) flux[k] -= nu[k] * (nx * g[0][k] + ny * g[1][k] + nz * g[2][k]);
}

// ---- EigenMatrices (riemann/euler/eigen.hpp:27-56) -----------------------------------------------
// Left / right eigenvector matrices of the Euler flux Jacobian of the conservative state `u` along the
// frame (nu, mu, pi) = f[0..2], f[3..5], f[6..8].  Row k of L and row c of R are all a thread of the
// characteristic limiter needs.
struct EigenCommon {
  double ux, uy, uz, u_nu, u_mu, u_pi, a, ek, h0, b1, b2;
};
MCFD_HD EigenCommon EigenPrepare(const double *u, const double *f) {
  const Prim q = ConsToPrim(u[0], u[1], u[2], u[3], u[4]);
  EigenCommon e;
  e.ux = q.u; e.uy = q.v; e.uz = q.w;
  e.u_nu = q.u * f[0] + q.v * f[1] + q.w * f[2];
  e.u_mu = q.u * f[3] + q.v * f[4] + q.w * f[5];
  e.u_pi = q.u * f[6] + q.v * f[7] + q.w * f[8];
  e.a = SoundSpeed(q.rho, q.p);
  e.ek = (q.u * q.u + q.v * q.v + q.w * q.w) / 2;
  const double e0 = e.a * e.a * kOneOverGm1;
  e.h0 = e0 + e.ek;
  e.b1 = 1 / e0;
  e.b2 = e.b1 * e.ek;
  return e;
}
MCFD_HD void EigenRowL(const EigenCommon &e, const double *f, int k, double *row) {
  const double vel[3] = {e.ux, e.uy, e.uz};
  if (k == 0 || k == 4) {
    const double s = k == 0 ? 1.0 : -1.0;
    row[0] = (e.b2 + s * e.u_nu / e.a) / 2;
    for (int x = 0; x < 3; ++x) row[1 + x] = -(e.b1 * vel[x] + s * f[x] / e.a) / 2;
    row[4] = e.b1 / 2;
  } else if (k == 1) {
    row[0] = 1 - e.b2;
    for (int x = 0; x < 3; ++x) row[1 + x] = e.b1 * vel[x];
    row[4] = -e.b1;
  } else {
    const double *t = f + (k == 2 ? 3 : 6);
    row[0] = k == 2 ? -e.u_mu : -e.u_pi;
    for (int x = 0; x < 3; ++x) row[1 + x] = t[x];
    row[4] = 0;
  }
}
MCFD_HD void EigenRowR(const EigenCommon &e, const double *f, int c, double *row) {
  if (c == 0) {
    row[0] = 1; row[1] = 1; row[2] = 0; row[3] = 0; row[4] = 1;
  } else if (c == 4) {
    row[0] = e.h0 - e.a * e.u_nu; row[1] = e.ek; row[2] = e.u_mu; row[3] = e.u_pi; row[4] = e.h0 + e.a * e.u_nu;
  } else {
    const int x = c - 1;
    const double v = x == 0 ? e.ux : (x == 1 ? e.uy : e.uz);
    row[0] = v - e.a * f[x]; row[1] = v; row[2] = f[3 + x]; row[3] = f[6 + x]; row[4] = v + e.a * f[x];
  }
}

}  // namespace mcfd
#endif  // MINICFD_B200_PHYSICS_CUH_
