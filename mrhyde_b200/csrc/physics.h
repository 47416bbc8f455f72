// Point-level weak forms of the four physics modules on the hot path.  Each routine restates the integrand
// of the reference module at ONE quadrature point and returns F[v][0] (multiplies the test function phi_i)
// and F[v][1+c] (multiplies d phi_i/dx_c), weights included:
//   thermal           /root/reference/src/physics/thermal.cpp:71-165 (volume), :172-281 (boundary)
//   cdr               /root/reference/src/physics/cdr.cpp:63-142
//   linearelasticity  /root/reference/src/physics/linearelasticity.cpp:92-241, :244-670, :1217-1255 (stress)
//   navierstokes      /root/reference/src/physics/navierstokes.cpp:494-848 (3-D volume), :855-1009, :1054-1079 (tau)
//   generic Flux BC   /root/reference/src/interfaces/physicsInterface.cpp:1698-1766
// S is double (residual only) or PAD<P> (residual + point Jacobian).  Reference quirks are kept: the 3-D uz
// momentum block is written with the uy offsets (navierstokes.cpp:688); |u| stays squared when |u|^2<=1e-12.
#pragma once
#include "params.h"

namespace mrh {

template <class S>
struct PointIn {
  const KParams* kp;
  PointState<S> ps;
  double w;      // integration weight at this point (volume or side measure)
  double h;      // element size (volume: (sum w)^(1/d); side: (sum w_side)^(1/(d-1)))
  MRH_HDN S fn(int k) const { return eval_prog<S>(kp->progs[k], kp->code, kp->cpool, ps); }
  MRH_HD const S& U(int m) const { return ps.fields[m * FIELD_KINDS]; }
  MRH_HD const S& dU(int m, int c) const { return ps.fields[m * FIELD_KINDS + 1 + c]; }
  MRH_HD const S& Ut(int m) const { return ps.fields[m * FIELD_KINDS + 4]; }
};

template <int PHYS, int DIM> struct PhysTraits;
template <int DIM> struct PhysTraits<PHYS_THERMAL, DIM> { static constexpr int NVAR = 1; };
template <int DIM> struct PhysTraits<PHYS_CDR, DIM> { static constexpr int NVAR = 1; };
template <int DIM> struct PhysTraits<PHYS_LE, DIM> { static constexpr int NVAR = DIM; };
template <int DIM> struct PhysTraits<PHYS_NS, DIM> { static constexpr int NVAR = DIM + 1; };
template <int DIM> struct PhysTraits<PHYS_STOKES, DIM> { static constexpr int NVAR = DIM + 1; };
template <int DIM> struct PhysTraits<PHYS_HELMHOLTZ, DIM> { static constexpr int NVAR = 2; };
template <int DIM> struct PhysTraits<PHYS_BURGERS, DIM> { static constexpr int NVAR = 1; };
// modules whose volume terms need the element size h = (sum w)^(1/d) (stabilisation parameters)
template <int PHYS> struct NeedsElementSize { static constexpr bool value = PHYS == PHYS_NS || PHYS == PHYS_STOKES || PHYS == PHYS_BURGERS; };

template <int PHYS, int DIM, class S> struct Physics;

// ---------------------------------------------------------------------------------------------- thermal
template <int DIM, class S>
struct Physics<PHYS_THERMAL, DIM, S> {
  static constexpr int NVAR = 1;
  template <class IN> MRH_HD static void volume(const IN& in, S (&F)[NVAR][1 + DIM]) {
    const S source = in.fn(TH_SOURCE), diff = in.fn(TH_DIFF), cp = in.fn(TH_CP), rho = in.fn(TH_RHO);
    const double w = in.w;
    F[0][0] = (rho * cp * in.Ut(0) - source) * w;
#pragma unroll
    for (int c = 0; c < DIM; ++c) F[0][1 + c] = diff * in.dU(0, c) * w;
    if (in.kp->ph.have_advection) {
      S adv = in.fn(TH_BX) * in.dU(0, 0);
      if (DIM > 1) adv = adv + in.fn(TH_BY) * in.dU(0, 1);
      if (DIM > 2) adv = adv + in.fn(TH_BZ) * in.dU(0, 2);
      F[0][0] = F[0][0] + adv * w;
    }
  }
  template <class IN> MRH_HDN static void boundary(const IN& in, S (&F)[NVAR][1 + DIM]) {
    const int bc = in.kp->ph.bc[0];
    const double w = in.w;
    if (bc == BC_NEUMANN) {
      F[0][0] = F[0][0] + (-in.fn(SIDE_BCSRC)) * w;
    } else if (bc == BC_WEAK_DIRICHLET) {
      const S nsource = in.fn(SIDE_BCSRC), diff = in.fn(SIDE_DIFF);
      const double epen = 10.0, sf = in.kp->ph.formparam;
      const S dT = in.U(0) - nsource;
      S gn = in.dU(0, 0) * in.ps.nrm[0];
#pragma unroll
      for (int c = 1; c < DIM; ++c) gn = gn + in.dU(0, c) * in.ps.nrm[c];
      F[0][0] = F[0][0] + epen / in.h * diff * dT * w + (-diff * gn) * w;
#pragma unroll
      for (int c = 0; c < DIM; ++c) F[0][1 + c] = F[0][1 + c] + (-sf * diff * dT * w) * in.ps.nrm[c];
    }
  }
};

// -------------------------------------------------------------------------------------------------- cdr
template <int DIM, class S>
struct Physics<PHYS_CDR, DIM, S> {
  static constexpr int NVAR = 1;
  template <class IN> MRH_HD static void volume(const IN& in, S (&F)[NVAR][1 + DIM]) {
    const S source = in.fn(CDR_SOURCE), diff = in.fn(CDR_DIFF), cp = in.fn(CDR_CP), rho = in.fn(CDR_RHO), reax = in.fn(CDR_REACTION);
    const double w = in.w;
    S f = in.Ut(0);
#pragma unroll
    for (int c = 0; c < DIM; ++c) f = f + in.fn(CDR_XVEL + c) * in.dU(0, c);
    F[0][0] = (f + reax - source) * w;
    const S k = 1.0 / (rho * cp) * diff;
#pragma unroll
    for (int c = 0; c < DIM; ++c) F[0][1 + c] = k * in.dU(0, c) * w;
  }
  template <class IN> MRH_HDN static void boundary(const IN&, S (&)[NVAR][1 + DIM]) {}   // cdr.cpp:148-150 is empty
};

// ------------------------------------------------------------------------------------ linear elasticity
template <int DIM, class S>
struct Physics<PHYS_LE, DIM, S> {
  static constexpr int NVAR = DIM;
  template <class IN> MRH_HD static void stress(const IN& in, const S& lambda, const S& mu, S (&Sg)[DIM][DIM]) {
#pragma unroll
    for (int a = 0; a < DIM; ++a)
#pragma unroll
      for (int b = 0; b < DIM; ++b) {
        if (a == b) {
          S s = (2.0 * mu + lambda) * in.dU(a, a);
          if (DIM == 2) s = s + lambda * in.dU(1 - a, 1 - a);
          if (DIM == 3) { const int o1 = (a == 0) ? 1 : 0, o2 = (a == 2) ? 1 : 2; s = s + lambda * (in.dU(o1, o1) + in.dU(o2, o2)); }
          Sg[a][b] = s;
        } else {
          const int lo = a < b ? a : b, hi = a < b ? b : a;
          Sg[a][b] = mu * (in.dU(lo, hi) + in.dU(hi, lo));
        }
      }
  }
  template <class IN> MRH_HD static void volume(const IN& in, S (&F)[NVAR][1 + DIM]) {
    const S lambda = in.fn(LE_LAMBDA), mu = in.fn(LE_MU);
    S Sg[DIM][DIM];
    stress(in, lambda, mu, Sg);
#pragma unroll
    for (int a = 0; a < DIM; ++a) {
      F[a][0] = (-in.fn(LE_SRC0 + a)) * in.w;
#pragma unroll
      for (int c = 0; c < DIM; ++c) F[a][1 + c] = Sg[a][c] * in.w;
    }
  }
  template <class IN> MRH_HDN static void boundary(const IN& in, S (&F)[NVAR][1 + DIM]) {
    const int* bc = in.kp->ph.bc;
    bool alld = true;
#pragma unroll
    for (int c = 0; c < DIM; ++c) if (bc[c] != BC_DIRICHLET) alld = false;
    if (alld) return;
    bool anyweak = false;
#pragma unroll
    for (int c = 0; c < DIM; ++c) if (bc[c] == BC_WEAK_DIRICHLET) anyweak = true;
    S src[DIM];
#pragma unroll
    for (int c = 0; c < DIM; ++c) src[c] = (bc[c] == BC_NEUMANN || bc[c] == BC_WEAK_DIRICHLET) ? in.fn(SIDE_BCSRC + c) : S(0.0);
#pragma unroll
    for (int a = 0; a < DIM; ++a) if (bc[a] == BC_NEUMANN) F[a][0] = F[a][0] + (-src[a]) * in.w;
    if (!anyweak) return;
    const S lam = in.fn(SIDE_LAMBDA), mu = in.fn(SIDE_MU);
    S Sg[DIM][DIM];
    stress(in, lam, mu, Sg);
    const double* n = in.ps.nrm;
    const S penalty = in.kp->ph.le_penalty * (lam + 2.0 * mu) / in.h;
    S delta[DIM];
#pragma unroll
    for (int c = 0; c < DIM; ++c) delta[c] = in.U(c) - src[c];
#pragma unroll
    for (int a = 0; a < DIM; ++a) {
      if (bc[a] != BC_WEAK_DIRICHLET) continue;
      S bvec[DIM];
#pragma unroll
      for (int c = 0; c < DIM; ++c) {
        if (c == a) {
          S t = ((0 == a) ? (lam + 2.0 * mu) * delta[0] * n[0] : lam * delta[0] * n[0]);
#pragma unroll
          for (int m = 1; m < DIM; ++m) t = t + ((m == a) ? (lam + 2.0 * mu) * delta[m] * n[m] : lam * delta[m] * n[m]);
          bvec[c] = t;
        } else {
          const int lo = a < c ? a : c, hi = a < c ? c : a;
          bvec[c] = mu * delta[hi] * n[lo] + mu * delta[lo] * n[hi];
        }
      }
      S sn = -Sg[a][0] * n[0];
#pragma unroll
      for (int c = 1; c < DIM; ++c) sn = sn - Sg[a][c] * n[c];
      F[a][0] = F[a][0] + (sn + penalty * delta[a]) * in.w;
#pragma unroll
      for (int c = 0; c < DIM; ++c) F[a][1 + c] = F[a][1 + c] + (-in.kp->ph.formparam * in.w) * bvec[c];
    }
  }
};

// ---------------------------------------------------------------------------------------- Navier-Stokes
template <int DIM, class S>
struct Physics<PHYS_NS, DIM, S> {
  static constexpr int NVAR = DIM + 1;
  MRH_HD static int uvar(int a) { return a == 0 ? 0 : a + 1; }       // varlist = ux, pr, uy, uz
  MRH_HD static S tau(const S& visc, const S* vel, double h, double dt, int transient) {
    const double C1 = 4.0, C2 = 2.0, C3 = transient ? 2.0 : 0.0;
    S nvel = vel[0] * vel[0];
#pragma unroll
    for (int c = 1; c < DIM; ++c) nvel = nvel + vel[c] * vel[c];
    if (val(nvel) > 1E-12) nvel = f_sqrt(nvel);
    const S a = C1 * visc / h / h, b = C2 * nvel / h;
    const double c3 = C3 / dt;
    S t = a * a + b * b + c3 * c3;
    return 1.0 / f_sqrt(t);
  }
  template <class IN> MRH_HD static void volume(const IN& in, S (&F)[NVAR][1 + DIM]) {
    const double w = in.w;
    const int srcslot[3] = {NS_SRC_UX, NS_SRC_UY, NS_SRC_UZ};
    const S dens = in.fn(NS_DENSITY), visc = in.fn(NS_VISCOSITY);
#pragma unroll
    for (int v = 0; v < NVAR; ++v) for (int a = 0; a <= DIM; ++a) F[v][a] = S(0.0);
    S vel[DIM], src[DIM], strong[DIM];
#pragma unroll
    for (int c = 0; c < DIM; ++c) { vel[c] = in.U(uvar(c)); src[c] = in.fn(srcslot[c]); }
    const S pr = in.U(1);
    const bool need_tau = in.kp->ph.useSUPG || in.kp->ph.usePSPG;
    S tauv = S(0.0);
    if (need_tau) tauv = tau(visc, vel, in.h, in.kp->tp.dt, in.kp->tp.transient);
#pragma unroll
    for (int a = 0; a < DIM; ++a) {
      const int ua = uvar(a);
      const int rowvar = (DIM == 3 && a == 2) ? uvar(1) : ua;        // reference quirk, navierstokes.cpp:688
      S conv = vel[0] * in.dU(ua, 0);
#pragma unroll
      for (int c = 1; c < DIM; ++c) conv = conv + vel[c] * in.dU(ua, c);
#pragma unroll
      for (int c = 0; c < DIM; ++c) {
        S Fc = visc * in.dU(ua, c);
        if (c == a) Fc = Fc - pr;
        F[rowvar][1 + c] = F[rowvar][1 + c] + Fc * w;
      }
      F[rowvar][0] = F[rowvar][0] + (in.Ut(ua) + conv - src[a]) * (dens * w);
      strong[a] = dens * in.Ut(ua) + dens * conv + in.dU(1, a) - dens * src[a];
      if (in.kp->ph.useSUPG) {
#pragma unroll
        for (int c = 0; c < DIM; ++c) F[rowvar][1 + c] = F[rowvar][1 + c] + tauv * strong[a] * vel[c] * w;
      }
    }
    S divu = in.dU(uvar(0), 0);
#pragma unroll
    for (int c = 1; c < DIM; ++c) divu = divu + in.dU(uvar(c), c);
    F[1][0] = F[1][0] + divu * w;
    if (in.kp->ph.usePSPG) {
#pragma unroll
      for (int c = 0; c < DIM; ++c) F[1][1 + c] = F[1][1 + c] + strong[c] * (tauv * w / dens);
    }
  }
  template <class IN> MRH_HDN static void boundary(const IN& in, S (&F)[NVAR][1 + DIM]) {
#pragma unroll
    for (int a = 0; a < DIM; ++a) {
      const int ua = uvar(a);
      if (in.kp->ph.bc[ua] == BC_NEUMANN) F[ua][0] = F[ua][0] + (-in.fn(SIDE_BCSRC + ua)) * in.w;
    }
  }
};

// ----------------------------------------------------------------------------------------------- Stokes
// /root/reference/src/physics/stokes.cpp:69-440 (volume), :445-447 (boundary: empty).  Variables ux, pr, uy, uz like navierstokes.
// Quirks kept: the PSPG term is tau (grad p PLUS source) with tau = h / (2 visc) in 2-D (:256-257) and h^2 / (2 visc) in 1-D / 3-D
// (:402); LSIC adds tau div(u) (d phi/dx + d phi/dy [+ d phi/dz]) to the PRESSURE rows (:268-286).  Program slots as for navierstokes.
template <int DIM, class S>
struct Physics<PHYS_STOKES, DIM, S> {
  static constexpr int NVAR = DIM + 1;
  MRH_HD static int uvar(int a) { return a == 0 ? 0 : a + 1; }
  template <class IN> MRH_HD static void volume(const IN& in, S (&F)[NVAR][1 + DIM]) {
    const double w = in.w;
    const int srcslot[3] = {NS_SRC_UX, NS_SRC_UY, NS_SRC_UZ};
    const S visc = in.fn(NS_VISCOSITY);
#pragma unroll
    for (int v = 0; v < NVAR; ++v) for (int a = 0; a <= DIM; ++a) F[v][a] = S(0.0);
    S src[DIM];
    const S pr = in.U(1);
#pragma unroll
    for (int a = 0; a < DIM; ++a) {
      const int ua = uvar(a);
      src[a] = in.fn(srcslot[a]);
#pragma unroll
      for (int c = 0; c < DIM; ++c) {
        S Fc = visc * in.dU(ua, c);
        if (c == a) Fc = Fc - pr;
        F[ua][1 + c] = Fc * w;
      }
      F[ua][0] = (-src[a]) * w;
    }
    S divu = in.dU(uvar(0), 0);
#pragma unroll
    for (int c = 1; c < DIM; ++c) divu = divu + in.dU(uvar(c), c);
    F[1][0] = divu * w;
    if (in.kp->ph.usePSPG) {
      const S tau = (DIM == 2) ? in.h / (2.0 * visc) : in.h * in.h / (2.0 * visc);
#pragma unroll
      for (int c = 0; c < DIM; ++c) F[1][1 + c] = F[1][1 + c] + (in.dU(1, c) + src[c]) * (tau * w);
    }
    if (in.kp->ph.useLSIC) {
      const S Sl = (in.h * in.h / (2.0 * visc)) * divu * w;
#pragma unroll
      for (int c = 0; c < DIM; ++c) F[1][1 + c] = F[1][1 + c] + Sl;
    }
  }
  template <class IN> MRH_HDN static void boundary(const IN&, S (&)[NVAR][1 + DIM]) {}     // stokes.cpp:445-447 is empty
};

// -------------------------------------------------------------------------------------------- helmholtz
// /root/reference/src/physics/helmholtz.cpp:71-240 (volume, non-fractional branch), :245-420 (boundary: Robin condition
// alpha u + du/dn - source = 0 on sides where ureal is "Neumann").  Variables ureal, uimag on the same basis, so the reference's
// test functions vr and vi are the same number: row (ureal, i) collects the coefficients of vr and vi together.
template <int DIM, class S>
struct Physics<PHYS_HELMHOLTZ, DIM, S> {
  static constexpr int NVAR = 2;
  template <class IN> MRH_HD static void volume(const IN& in, S (&F)[NVAR][1 + DIM]) {
    const double w = in.w;
    const S o2r = in.fn(HZ_OMEGA2R), o2i = in.fn(HZ_OMEGA2I), sr = in.fn(HZ_SRC_R), si = in.fn(HZ_SRC_I);
    const S ur = in.U(0), ui = in.U(1);
    F[0][0] = (-o2r * (ur + ui) + o2i * (ui - ur) - (sr + si)) * w;
    F[1][0] = (-o2r * (ui - ur) - o2i * (ur + ui) - (si - sr)) * w;
#pragma unroll
    for (int c = 0; c < DIM; ++c) {
      const S c2r = in.fn(HZ_C2 + 2 * c), c2i = in.fn(HZ_C2 + 2 * c + 1);
      const S dur = in.dU(0, c), dui = in.dU(1, c);
      F[0][1 + c] = (c2r * (dur + dui) - c2i * (dui - dur)) * w;
      F[1][1 + c] = (c2r * (dui - dur) + c2i * (dur + dui)) * w;
    }
  }
  template <class IN> MRH_HDN static void boundary(const IN& in, S (&F)[NVAR][1 + DIM]) {
    if (in.kp->ph.bc[0] != BC_NEUMANN) return;
    const double w = in.w;
    const S ar = in.fn(HZ_SIDE_ALPHA_R), ai = in.fn(HZ_SIDE_ALPHA_I), sr = in.fn(HZ_SIDE_SRC_R), si = in.fn(HZ_SIDE_SRC_I);
    const S ur = in.U(0), ui = in.U(1);
    S durdn = S(0.0), duidn = S(0.0), c2durdn = S(0.0), c2duidn = S(0.0);
#pragma unroll
    for (int c = 0; c < DIM; ++c) {
      const S c2r = in.fn(HZ_C2 + 2 * c), c2i = in.fn(HZ_C2 + 2 * c + 1);
      const S dur = in.dU(0, c), dui = in.dU(1, c);
      const double n = in.ps.nrm[c];
      durdn = durdn + dur * n; duidn = duidn + dui * n;
      c2durdn = c2durdn + (c2r * dur - c2i * dui) * n;
      c2duidn = c2duidn + (c2r * dui + c2i * dur) * n;
    }
    F[0][0] = F[0][0] + (((ar * (ur + ui) - ai * (ui - ur)) + (durdn + duidn) - (sr + si)) - (c2durdn + c2duidn)) * w;
    F[1][0] = F[1][0] + (((ar * (ui - ur) + ai * (ur + ui)) + (duidn - durdn) - (si - sr)) - (c2duidn - c2durdn)) * w;
  }
};

// ---------------------------------------------------------------------------------------------- Burgers
// /root/reference/src/physics/burgers.cpp:50-164 (volume: 2-D with optional entropy viscosity and SUPG, 3-D plain), :169-205 (2-D
// Neumann flux 1/2 u^2 v.n).  Quirks kept: the entropy viscosity is clamped to the plain number 0.1 (derivatives dropped), |v| stays
// squared when |v|^2 <= 1e-12.
template <int DIM, class S>
struct Physics<PHYS_BURGERS, DIM, S> {
  static constexpr int NVAR = 1;
  template <class IN> MRH_HD static void volume(const IN& in, S (&F)[NVAR][1 + DIM]) {
    const double w = in.w;
    const S source = in.fn(BG_SRC), eps = in.fn(BG_DIFF);
    const S u = in.U(0), dudt = in.Ut(0);
    S v[DIM];
#pragma unroll
    for (int c = 0; c < DIM; ++c) v[c] = in.fn(BG_VEL + c);
    S evisc = S(0.0);
    if (DIM == 2 && in.kp->ph.useEvisc) {
      const S entres = u * (dudt + u * in.dU(0, 0) + u * in.dU(0, 1));
      S a = 1.0e-12 + entres;
      if (val(a) < 0.0) a = -a;
      evisc = in.fn(BG_C1) * in.h * in.h * a / in.fn(BG_C2);
    }
    if (val(evisc) > 0.1) evisc = S(0.1);
    const S usq = 0.5 * u * u;
    F[0][0] = (dudt - source) * w;
#pragma unroll
    for (int c = 0; c < DIM; ++c) F[0][1 + c] = ((eps + evisc) * in.dU(0, c) - v[c] * usq) * w;
    if (DIM == 2 && in.kp->ph.useSUPG) {
      S nvel = v[0] * v[0] + v[1] * v[1];
      if (val(nvel) > 1E-12) nvel = f_sqrt(nvel);
      const S tau = in.fn(BG_SUPG_C) / (in.fn(BG_SUPG_C1) / in.kp->tp.dt + in.fn(BG_SUPG_C2) * nvel / in.h);
      const S sres = tau * (dudt + v[0] * u * in.dU(0, 0) + v[1] * u * in.dU(0, 1) - source) * w;
      F[0][1] = F[0][1] + sres * u * v[0];
      F[0][2] = F[0][2] + sres * u * v[1];
    }
  }
  template <class IN> MRH_HDN static void boundary(const IN& in, S (&F)[NVAR][1 + DIM]) {
    if (DIM != 2 || in.kp->ph.bc[0] != BC_NEUMANN) return;
    const S u = in.U(0);
    F[0][0] = F[0][0] + 0.5 * u * u * in.w * (in.fn(BG_SIDE_VEL) * in.ps.nrm[0] + in.fn(BG_SIDE_VEL + 1) * in.ps.nrm[1]);
  }
};

}  // namespace mrh
