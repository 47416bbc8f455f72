// ORACLE (test infrastructure, NOT product code).
// The four physics modules of the hot path, restated on the oracle's Workset/FunctionManager:
//   thermal           /root/reference/src/physics/thermal.cpp:48-65 (functions), :71-165 (volume), :172-281 (boundary)
//   cdr               /root/reference/src/physics/cdr.cpp:33-57, :63-142 (volume); boundary is empty (:148-150)
//   linearelasticity  /root/reference/src/physics/linearelasticity.cpp:67-86, :92-241 (volume), :244-670 (boundary),
//                     :913-1270 (computeStress, isotropic branch only)
//   navierstokes      /root/reference/src/physics/navierstokes.cpp:64-76, :82-849 (volume), :855-1009 (boundary),
//                     :1054-1079 (computeTau)
// Reference quirks are kept on purpose: NS 3-D uz momentum lands in the uy rows (navierstokes.cpp:688);
// cdr evaluates but never uses "SUPG tau" (cdr.cpp:82); |u| stays squared when |u|^2 <= 1e-12 (:1068).
// Written dimension-generic (loops over components) with the reference's arithmetic order.
#pragma once
#include <map>
#include <memory>
#include "funcman.hpp"

namespace orc {

typedef std::map<std::string, std::string> StrMap;
inline std::string getOr(const StrMap& m, const std::string& k, const std::string& d) {
  auto it = m.find(k); return it == m.end() ? d : it->second;
}
inline bool getBool(const StrMap& m, const std::string& k, bool d) {
  auto it = m.find(k); if (it == m.end()) return d; return it->second == "true" || it->second == "1";
}
inline double getDbl(const StrMap& m, const std::string& k, double d) {
  auto it = m.find(k); return it == m.end() ? d : std::stod(it->second);
}

template <class EvalT>
struct PhysicsBase {
  std::string label;
  std::vector<std::string> myvars, mybasistypes;
  Workset<EvalT>* wkset = nullptr;
  FunctionManager<EvalT>* functionManager = nullptr;
  virtual ~PhysicsBase() {}
  virtual void defineFunctions(const StrMap& fs, FunctionManager<EvalT>* fm) = 0;
  virtual void volumeResidual() = 0;
  virtual void boundaryResidual() {}
  virtual void setWorkset(Workset<EvalT>* w) { wkset = w; }
  int varnum(const std::string& v) const {
    for (size_t i = 0; i < wkset->varlist.size(); ++i) if (wkset->varlist[i] == v) return (int)i;
    return -1;
  }
};

static const char* const XYZ[3] = {"x", "y", "z"};
inline std::string gradname(const std::string& v, int c) { return "grad(" + v + ")[" + XYZ[c] + "]"; }

// ================================================================================================
template <class EvalT>
struct thermal : PhysicsBase<EvalT> {
  using PhysicsBase<EvalT>::wkset; using PhysicsBase<EvalT>::functionManager;
  double formparam = 1.0; bool have_advection = false, have_nsvel = false;
  int e_num = 0, ux_num = -1, uy_num = -1, uz_num = -1;
  thermal(const StrMap& settings, int) {
    this->label = "thermal"; this->myvars = {"e"}; this->mybasistypes = {"HGRAD"};
    formparam = getDbl(settings, "form_param", 1.0);
    have_advection = getBool(settings, "include advection", false);
  }
  void defineFunctions(const StrMap& fs, FunctionManager<EvalT>* fm) override {
    functionManager = fm;
    fm->addFunction("thermal source", getOr(fs, "thermal source", "0.0"), "ip");
    fm->addFunction("thermal diffusion", getOr(fs, "thermal diffusion", "1.0"), "ip");
    fm->addFunction("specific heat", getOr(fs, "specific heat", "1.0"), "ip");
    fm->addFunction("density", getOr(fs, "density", "1.0"), "ip");
    fm->addFunction("bx", getOr(fs, "advection x", "0.0"), "ip");
    fm->addFunction("by", getOr(fs, "advection y", "0.0"), "ip");
    fm->addFunction("bz", getOr(fs, "advection z", "0.0"), "ip");
    fm->addFunction("thermal diffusion", getOr(fs, "thermal diffusion", "1.0"), "side ip");
    fm->addFunction("robin alpha", getOr(fs, "robin alpha", "0.0"), "side ip");
  }
  void setWorkset(Workset<EvalT>* w) override {
    wkset = w; e_num = this->varnum("e"); ux_num = this->varnum("ux"); uy_num = this->varnum("uy"); uz_num = this->varnum("uz");
    have_nsvel = ux_num >= 0;
  }
  void volumeResidual() override {
    const int dim = wkset->dim, nb = wkset->nb, nq = wkset->gd->nq;
    auto source = functionManager->evaluate("thermal source", "ip");
    auto diff = functionManager->evaluate("thermal diffusion", "ip");
    auto cp = functionManager->evaluate("specific heat", "ip");
    auto rho = functionManager->evaluate("density", "ip");
    Vista<EvalT> bv[3];
    if (have_advection) { const char* nm[3] = {"bx", "by", "bz"}; for (int c = 0; c < dim; ++c) bv[c] = functionManager->evaluate(nm[c], "ip"); }
    auto& dTdt = wkset->getSolutionField("e_t");
    Arr2<EvalT>* dT[3]; for (int c = 0; c < dim; ++c) dT[c] = &wkset->getSolutionField(gradname("e", c));
    Arr2<EvalT>* U[3] = {nullptr, nullptr, nullptr};
    if (have_nsvel) { const char* nm[3] = {"ux", "uy", "uz"}; for (int c = 0; c < dim; ++c) U[c] = &wkset->getSolutionField(nm[c]); }
    auto& res = wkset->res; const GroupData& g = *wkset->gd;
    for (int elem = 0; elem < wkset->numElem; ++elem)
      for (int dof = 0; dof < nb; ++dof) {
        const int r = wkset->off(e_num, dof);
        for (int pt = 0; pt < nq; ++pt) {
          const double w = wkset->wt(elem, pt);
          res(elem, r) += (rho(elem, pt) * cp(elem, pt) * dTdt(elem, pt) - source(elem, pt)) * w * g.B(elem, dof, pt);
          for (int c = 0; c < dim; ++c) res(elem, r) += diff(elem, pt) * (*dT[c])(elem, pt) * w * g.G(elem, dof, pt, c);
          if (have_nsvel) {
            EvalT adv = (*U[0])(elem, pt) * (*dT[0])(elem, pt);
            for (int c = 1; c < dim; ++c) adv = adv + (*U[c])(elem, pt) * (*dT[c])(elem, pt);
            res(elem, r) += adv * w * g.B(elem, dof, pt);
          }
          if (have_advection) {
            EvalT adv = bv[0](elem, pt) * (*dT[0])(elem, pt);
            for (int c = 1; c < dim; ++c) adv = adv + bv[c](elem, pt) * (*dT[c])(elem, pt);
            res(elem, r) += adv * w * g.B(elem, dof, pt);
          }
        }
      }
  }
  void boundaryResidual() override {
    const int dim = wkset->dim, nb = wkset->nb, nq = wkset->gd->nq;
    const std::string bctype = wkset->var_bcs[e_num][wkset->currentside];
    Vista<EvalT> nsource;
    if (bctype == "weak Dirichlet") nsource = functionManager->evaluate("Dirichlet e " + wkset->sidename, "side ip");
    else if (bctype == "Neumann") nsource = functionManager->evaluate("Neumann e " + wkset->sidename, "side ip");
    auto diff_side = functionManager->evaluate("thermal diffusion", "side ip");
    functionManager->evaluate("robin alpha", "side ip");
    const double sf = formparam;
    auto h = wkset->getSideElementSize();
    auto& res = wkset->res; const GroupData& g = *wkset->gd;
    if (bctype == "Neumann") {
      for (int elem = 0; elem < wkset->numElem; ++elem)
        for (int dof = 0; dof < nb; ++dof)
          for (int pt = 0; pt < nq; ++pt)
            res(elem, wkset->off(e_num, dof)) += -nsource(elem, pt) * wkset->wt(elem, pt) * g.B(elem, dof, pt);
    } else if (bctype == "weak Dirichlet") {
      auto& T = wkset->getSolutionField("e");
      Arr2<EvalT>* dT[3]; Arr2<double>* n[3];
      for (int c = 0; c < dim; ++c) { dT[c] = &wkset->getSolutionField(gradname("e", c)); n[c] = &wkset->getScalarField(std::string("n[") + XYZ[c] + "]"); }
      const double epen = 10.0;
      for (int elem = 0; elem < wkset->numElem; ++elem)
        for (int dof = 0; dof < nb; ++dof) {
          const int r = wkset->off(e_num, dof);
          for (int pt = 0; pt < nq; ++pt) {
            const double w = wkset->wt(elem, pt);
            res(elem, r) += epen / h[elem] * diff_side(elem, pt) * (T(elem, pt) - nsource(elem, pt)) * w * g.B(elem, dof, pt);
            EvalT gn = (*dT[0])(elem, pt) * (*n[0])(elem, pt);
            for (int c = 1; c < dim; ++c) gn = gn + (*dT[c])(elem, pt) * (*n[c])(elem, pt);
            res(elem, r) += -diff_side(elem, pt) * gn * w * g.B(elem, dof, pt);
            double bn = g.G(elem, dof, pt, 0) * (*n[0])(elem, pt);
            for (int c = 1; c < dim; ++c) bn = bn + g.G(elem, dof, pt, c) * (*n[c])(elem, pt);
            res(elem, r) += -sf * diff_side(elem, pt) * (T(elem, pt) - nsource(elem, pt)) * w * bn;
          }
        }
    }
  }
};

// ================================================================================================
template <class EvalT>
struct cdr : PhysicsBase<EvalT> {
  using PhysicsBase<EvalT>::wkset; using PhysicsBase<EvalT>::functionManager;
  int cnum = 0;
  cdr(const StrMap&, int) { this->label = "cdr"; this->myvars = {"c"}; this->mybasistypes = {"HGRAD"}; }
  void defineFunctions(const StrMap& fs, FunctionManager<EvalT>* fm) override {
    functionManager = fm;
    fm->addFunction("source", getOr(fs, "source", "0.0"), "ip");
    fm->addFunction("diffusion", getOr(fs, "diffusion", "1.0"), "ip");
    fm->addFunction("specific heat", getOr(fs, "specific heat", "1.0"), "ip");
    fm->addFunction("density", getOr(fs, "density", "1.0"), "ip");
    fm->addFunction("reaction", getOr(fs, "reaction", "1.0"), "ip");
    fm->addFunction("xvel", getOr(fs, "xvel", "1.0"), "ip");
    fm->addFunction("yvel", getOr(fs, "yvel", "1.0"), "ip");
    fm->addFunction("zvel", getOr(fs, "zvel", "1.0"), "ip");
    fm->addFunction("SUPG tau", getOr(fs, "SUPG tau", "0.0"), "ip");
    fm->addFunction("diffusion", getOr(fs, "diffusion", "1.0"), "side ip");
    fm->addFunction("robin alpha", getOr(fs, "robin alpha", "0.0"), "side ip");
  }
  void setWorkset(Workset<EvalT>* w) override { wkset = w; cnum = this->varnum("c"); }
  void volumeResidual() override {
    const int dim = wkset->dim, nb = wkset->nb, nq = wkset->gd->nq;
    auto source = functionManager->evaluate("source", "ip");
    auto diff = functionManager->evaluate("diffusion", "ip");
    auto cp = functionManager->evaluate("specific heat", "ip");
    auto rho = functionManager->evaluate("density", "ip");
    auto reax = functionManager->evaluate("reaction", "ip");
    Vista<EvalT> vel[3]; const char* vn[3] = {"xvel", "yvel", "zvel"};
    for (int c = 0; c < 3; ++c) vel[c] = functionManager->evaluate(vn[c], "ip");
    functionManager->evaluate("SUPG tau", "ip");
    auto& dC_dt = wkset->getSolutionField("c_t");
    Arr2<EvalT>* dC[3]; for (int c = 0; c < dim; ++c) dC[c] = &wkset->getSolutionField(gradname("c", c));
    auto& res = wkset->res; const GroupData& g = *wkset->gd;
    for (int elem = 0; elem < wkset->numElem; ++elem)
      for (int pt = 0; pt < nq; ++pt) {
        const double w = wkset->wt(elem, pt);
        EvalT f = dC_dt(elem, pt);
        for (int c = 0; c < dim; ++c) f = f + vel[c](elem, pt) * (*dC[c])(elem, pt);
        f = (f + reax(elem, pt) - source(elem, pt)) * w;
        EvalT F[3];
        for (int c = 0; c < dim; ++c) F[c] = 1.0 / (rho(elem, pt) * cp(elem, pt)) * diff(elem, pt) * (*dC[c])(elem, pt) * w;
        for (int dof = 0; dof < nb; ++dof) {
          EvalT s = f * g.B(elem, dof, pt);
          for (int c = 0; c < dim; ++c) s = s + F[c] * g.G(elem, dof, pt, c);
          res(elem, wkset->off(cnum, dof)) += s;
        }
      }
  }
};

// ================================================================================================
template <class EvalT>
struct linearelasticity : PhysicsBase<EvalT> {
  using PhysicsBase<EvalT>::wkset; using PhysicsBase<EvalT>::functionManager;
  int spaceDim; int dnum[3] = {0, 0, 0};
  double mp[5];
  std::vector<EvalT> stress;    // (elem, pt, i, j)
  linearelasticity(const StrMap& settings, int dimension) : spaceDim(dimension) {
    this->label = "linearelasticity";
    const char* nm[3] = {"dx", "dy", "dz"};
    for (int c = 0; c < spaceDim; ++c) { this->myvars.push_back(nm[c]); this->mybasistypes.push_back("HGRAD"); }
    mp[0] = getDbl(settings, "form_param", 1.0); mp[1] = getDbl(settings, "penalty", 10.0);
    mp[2] = getDbl(settings, "Biot alpha", 0.0); mp[3] = getDbl(settings, "T_ambient", 0.0); mp[4] = getDbl(settings, "alpha_T", 1.0e-6);
  }
  void defineFunctions(const StrMap& fs, FunctionManager<EvalT>* fm) override {
    functionManager = fm;
    fm->addFunction("lambda", getOr(fs, "lambda", "1.0"), "ip");
    fm->addFunction("mu", getOr(fs, "mu", "0.5"), "ip");
    fm->addFunction("source dx", getOr(fs, "source dx", "0.0"), "ip");
    fm->addFunction("source dy", getOr(fs, "source dy", "0.0"), "ip");
    fm->addFunction("source dz", getOr(fs, "source dz", "0.0"), "ip");
    fm->addFunction("lambda", getOr(fs, "lambda", "1.0"), "side ip");
    fm->addFunction("mu", getOr(fs, "mu", "0.5"), "side ip");
  }
  void setWorkset(Workset<EvalT>* w) override {
    wkset = w; const char* nm[3] = {"dx", "dy", "dz"};
    for (int c = 0; c < spaceDim; ++c) dnum[c] = this->varnum(nm[c]);
  }
  EvalT& S(int e, int k, int i, int j, int npt) { return stress[(((size_t)e * npt + k) * spaceDim + i) * spaceDim + j]; }
  void computeStress(const Vista<EvalT>& lambda, const Vista<EvalT>& mu) {
    const int dim = spaceDim, npt = wkset->gd->nq; const char* nm[3] = {"dx", "dy", "dz"};
    stress.assign((size_t)wkset->numElem * npt * dim * dim, EvalT(0.0));
    Arr2<EvalT>* dd[3][3];
    for (int a = 0; a < dim; ++a) for (int c = 0; c < dim; ++c) dd[a][c] = &wkset->getSolutionField(gradname(nm[a], c));
    for (int e = 0; e < wkset->numElem; ++e)
      for (int k = 0; k < npt; ++k)
        for (int a = 0; a < dim; ++a)
          for (int b = 0; b < dim; ++b) {
            if (a == b) {
              EvalT s = (2.0 * mu(e, k) + lambda(e, k)) * (*dd[a][a])(e, k);
              if (dim == 2) s = s + lambda(e, k) * (*dd[1 - a][1 - a])(e, k);
              if (dim == 3) { const int o1 = (a == 0) ? 1 : 0, o2 = (a == 2) ? 1 : 2; s = s + lambda(e, k) * ((*dd[o1][o1])(e, k) + (*dd[o2][o2])(e, k)); }
              S(e, k, a, b, npt) = s;
            } else {
              const int lo = std::min(a, b), hi = std::max(a, b);
              S(e, k, a, b, npt) = mu(e, k) * ((*dd[lo][hi])(e, k) + (*dd[hi][lo])(e, k));
            }
          }
  }
  void volumeResidual() override {
    const int dim = spaceDim, nb = wkset->nb, nq = wkset->gd->nq;
    const char* sn[3] = {"source dx", "source dy", "source dz"};
    Vista<EvalT> src[3];
    for (int c = 0; c < dim; ++c) src[c] = functionManager->evaluate(sn[c], "ip");
    auto lambda = functionManager->evaluate("lambda", "ip");
    auto mu = functionManager->evaluate("mu", "ip");
    computeStress(lambda, mu);
    auto& res = wkset->res; const GroupData& g = *wkset->gd;
    for (int a = 0; a < dim; ++a)
      for (int elem = 0; elem < wkset->numElem; ++elem)
        for (int dof = 0; dof < nb; ++dof) {
          const int r = wkset->off(dnum[a], dof);
          for (int pt = 0; pt < nq; ++pt) {
            EvalT s = S(elem, pt, a, 0, nq) * g.G(elem, dof, pt, 0);
            for (int c = 1; c < dim; ++c) s = s + S(elem, pt, a, c, nq) * g.G(elem, dof, pt, c);
            res(elem, r) += (s - src[a](elem, pt) * g.B(elem, dof, pt)) * wkset->wt(elem, pt);
          }
        }
  }
  void boundaryResidual() override {
    const int dim = spaceDim, nb = wkset->nb, nq = wkset->gd->nq; const char* nm[3] = {"dx", "dy", "dz"};
    std::string st[3] = {"Dirichlet", "Dirichlet", "Dirichlet"};
    for (int c = 0; c < dim; ++c) st[c] = wkset->var_bcs[dnum[c]][wkset->currentside];
    if (st[0] == "Dirichlet" && st[1] == "Dirichlet" && st[2] == "Dirichlet") return;
    Vista<EvalT> src[3];
    for (int c = 0; c < dim; ++c) {
      if (st[c] == "Neumann") src[c] = functionManager->evaluate(std::string("Neumann ") + nm[c] + " " + wkset->sidename, "side ip");
      else if (st[c] == "weak Dirichlet") src[c] = functionManager->evaluate(std::string("Dirichlet ") + nm[c] + " " + wkset->sidename, "side ip");
    }
    auto lam = functionManager->evaluate("lambda", "side ip");
    auto mu = functionManager->evaluate("mu", "side ip");
    auto h = wkset->getSideElementSize();
    computeStress(lam, mu);
    Arr2<double>* n[3]; for (int c = 0; c < dim; ++c) n[c] = &wkset->getScalarField(std::string("n[") + XYZ[c] + "]");
    Arr2<EvalT>* d[3]; for (int c = 0; c < dim; ++c) d[c] = &wkset->getSolutionField(nm[c]);
    auto& res = wkset->res; const GroupData& g = *wkset->gd;
    for (int a = 0; a < dim; ++a) {
      if (st[a] == "Neumann") {
        for (int e = 0; e < wkset->numElem; ++e)
          for (int k = 0; k < nq; ++k)
            for (int i = 0; i < nb; ++i)
              res(e, wkset->off(dnum[a], i)) += (-src[a](e, k) * g.B(e, i, k)) * wkset->wt(e, k);
      } else if (st[a] == "weak Dirichlet") {
        for (int e = 0; e < wkset->numElem; ++e)
          for (int k = 0; k < nq; ++k) {
            EvalT penalty = mp[1] * (lam(e, k) + 2.0 * mu(e, k)) / h[e];
            EvalT delta[3];
            for (int c = 0; c < dim; ++c) delta[c] = (*d[c])(e, k) - src[c](e, k);
            // b_c: row a of the elasticity tensor applied to delta (x) n   (linearelasticity.cpp:505-507, :569-571, :632-634)
            EvalT bvec[3];
            for (int c = 0; c < dim; ++c) {
              if (c == a) {
                EvalT t = EvalT(0.0); bool first = true;
                for (int m = 0; m < dim; ++m) {
                  EvalT term = (m == a) ? (lam(e, k) + 2.0 * mu(e, k)) * delta[m] * (*n[m])(e, k) : lam(e, k) * delta[m] * (*n[m])(e, k);
                  if (first) { t = term; first = false; } else t = t + term;
                }
                bvec[c] = t;
              } else {
                const int lo = std::min(a, c), hi = std::max(a, c);
                bvec[c] = mu(e, k) * delta[hi] * (*n[lo])(e, k) + mu(e, k) * delta[lo] * (*n[hi])(e, k);
              }
            }
            EvalT sn = -S(e, k, a, 0, nq) * (*n[0])(e, k);
            for (int c = 1; c < dim; ++c) sn = sn - S(e, k, a, c, nq) * (*n[c])(e, k);
            for (int i = 0; i < nb; ++i) {
              EvalT bg = bvec[0] * g.G(e, i, k, 0);
              for (int c = 1; c < dim; ++c) bg = bg + bvec[c] * g.G(e, i, k, c);
              res(e, wkset->off(dnum[a], i)) += (sn * g.B(e, i, k) + penalty * delta[a] * g.B(e, i, k) - mp[0] * bg) * wkset->wt(e, k);
            }
          }
      }
    }
  }
};

// ================================================================================================
template <class EvalT>
struct navierstokes : PhysicsBase<EvalT> {
  using PhysicsBase<EvalT>::wkset; using PhysicsBase<EvalT>::functionManager;
  int spaceDim; bool useSUPG = false, usePSPG = false;
  int unum[3] = {0, 0, 0}, pr_num = 0;
  navierstokes(const StrMap& settings, int dimension) : spaceDim(dimension) {
    this->label = "navierstokes";
    this->myvars.push_back("ux"); this->myvars.push_back("pr");
    if (spaceDim > 1) this->myvars.push_back("uy");
    if (spaceDim > 2) this->myvars.push_back("uz");
    for (size_t i = 0; i < this->myvars.size(); ++i) this->mybasistypes.push_back("HGRAD");
    useSUPG = getBool(settings, "useSUPG", false); usePSPG = getBool(settings, "usePSPG", false);
  }
  void defineFunctions(const StrMap& fs, FunctionManager<EvalT>* fm) override {
    functionManager = fm;
    fm->addFunction("source ux", getOr(fs, "source ux", "0.0"), "ip");
    fm->addFunction("source pr", getOr(fs, "source pr", "0.0"), "ip");
    fm->addFunction("source uy", getOr(fs, "source uy", "0.0"), "ip");
    fm->addFunction("source uz", getOr(fs, "source uz", "0.0"), "ip");
    fm->addFunction("density", getOr(fs, "density", "1.0"), "ip");
    fm->addFunction("viscosity", getOr(fs, "viscosity", "1.0"), "ip");
  }
  void setWorkset(Workset<EvalT>* w) override {
    wkset = w; const char* nm[3] = {"ux", "uy", "uz"};
    for (int c = 0; c < spaceDim; ++c) unum[c] = this->varnum(nm[c]);
    pr_num = this->varnum("pr");
  }
  EvalT computeTau(const EvalT& localdiff, const EvalT* vel, double h, int dim, double dt, bool isTransient) const {
    const double C1 = 4.0, C2 = 2.0, C3 = isTransient ? 2.0 : 0.0;
    EvalT nvel = vel[0] * vel[0];
    for (int c = 1; c < dim; ++c) nvel = nvel + vel[c] * vel[c];
    if (nvel > 1E-12) nvel = sqrt(nvel);
    EvalT tau = (C1 * localdiff / h / h) * (C1 * localdiff / h / h) + (C2 * nvel / h) * (C2 * nvel / h) + (C3 / dt) * (C3 / dt);
    tau = 1. / sqrt(tau);
    return tau;
  }
  void volumeResidual() override {
    const int dim = spaceDim, nb = wkset->nb, nq = wkset->gd->nq;
    const double dt = wkset->deltat; const bool isTransient = wkset->isTransient;
    const char* un[3] = {"ux", "uy", "uz"}; const char* sn[3] = {"source ux", "source uy", "source uz"};
    Vista<EvalT> src[3];
    src[0] = functionManager->evaluate("source ux", "ip");
    functionManager->evaluate("source pr", "ip");
    for (int c = 1; c < dim; ++c) src[c] = functionManager->evaluate(sn[c], "ip");
    auto dens = functionManager->evaluate("density", "ip");
    auto visc = functionManager->evaluate("viscosity", "ip");
    Arr2<EvalT>* u[3]; Arr2<EvalT>* ut[3]; Arr2<EvalT>* du[3][3]; Arr2<EvalT>* dp[3];
    for (int a = 0; a < dim; ++a) {
      u[a] = &wkset->getSolutionField(un[a]); ut[a] = &wkset->getSolutionField(std::string(un[a]) + "_t");
      for (int c = 0; c < dim; ++c) du[a][c] = &wkset->getSolutionField(gradname(un[a], c));
      dp[a] = &wkset->getSolutionField(gradname("pr", a));
    }
    auto& pr = wkset->getSolutionField("pr");
    auto h = wkset->getElementSize();
    auto& res = wkset->res; const GroupData& g = *wkset->gd;
    for (int a = 0; a < dim; ++a) {
      // reference quirk: in 3-D the uz block scatters with the uy offsets (navierstokes.cpp:688)
      const int rowvar = (dim == 3 && a == 2) ? unum[1] : unum[a];
      for (int elem = 0; elem < wkset->numElem; ++elem)
        for (int pt = 0; pt < nq; ++pt) {
          const double w = wkset->wt(elem, pt);
          EvalT Fc[3];
          for (int c = 0; c < dim; ++c) {
            Fc[c] = visc(elem, pt) * (*du[a][c])(elem, pt);
            if (c == a) Fc[c] = Fc[c] - pr(elem, pt);
            Fc[c] *= w;
          }
          EvalT F = (*ut[a])(elem, pt);
          for (int c = 0; c < dim; ++c) F = F + (*u[c])(elem, pt) * (*du[a][c])(elem, pt);
          F = F - src[a](elem, pt);
          F *= dens(elem, pt) * w;
          for (int dof = 0; dof < nb; ++dof) {
            EvalT s = Fc[0] * g.G(elem, dof, pt, 0);
            for (int c = 1; c < dim; ++c) s = s + Fc[c] * g.G(elem, dof, pt, c);
            res(elem, wkset->off(rowvar, dof)) += s + F * g.B(elem, dof, pt);
          }
        }
      if (useSUPG) {
        for (int elem = 0; elem < wkset->numElem; ++elem)
          for (int pt = 0; pt < nq; ++pt) {
            const double w = wkset->wt(elem, pt);
            EvalT vel[3]; for (int c = 0; c < dim; ++c) vel[c] = (*u[c])(elem, pt);
            EvalT tau = computeTau(visc(elem, pt), vel, h[elem], dim, dt, isTransient);
            EvalT conv = vel[0] * (*du[a][0])(elem, pt);
            for (int c = 1; c < dim; ++c) conv = conv + vel[c] * (*du[a][c])(elem, pt);
            EvalT stabres = dens(elem, pt) * (*ut[a])(elem, pt) + dens(elem, pt) * conv + (*dp[a])(elem, pt) - dens(elem, pt) * src[a](elem, pt);
            EvalT Sc[3]; for (int c = 0; c < dim; ++c) Sc[c] = tau * stabres * vel[c] * w;
            for (int dof = 0; dof < nb; ++dof) {
              EvalT s = Sc[0] * g.G(elem, dof, pt, 0);
              for (int c = 1; c < dim; ++c) s = s + Sc[c] * g.G(elem, dof, pt, c);
              res(elem, wkset->off(rowvar, dof)) += s;
            }
          }
      }
    }
    // continuity
    for (int elem = 0; elem < wkset->numElem; ++elem)
      for (int pt = 0; pt < nq; ++pt) {
        EvalT divu = (*du[0][0])(elem, pt);
        for (int c = 1; c < dim; ++c) divu = divu + (*du[c][c])(elem, pt);
        divu = divu * wkset->wt(elem, pt);
        for (int dof = 0; dof < nb; ++dof) res(elem, wkset->off(pr_num, dof)) += divu * g.B(elem, dof, pt);
      }
    if (usePSPG) {
      for (int elem = 0; elem < wkset->numElem; ++elem)
        for (int pt = 0; pt < nq; ++pt) {
          const double w = wkset->wt(elem, pt);
          EvalT vel[3]; for (int c = 0; c < dim; ++c) vel[c] = (*u[c])(elem, pt);
          EvalT tau = computeTau(visc(elem, pt), vel, h[elem], dim, dt, isTransient);
          EvalT Sc[3];
          for (int a = 0; a < dim; ++a) {
            EvalT conv = vel[0] * (*du[a][0])(elem, pt);
            for (int c = 1; c < dim; ++c) conv = conv + vel[c] * (*du[a][c])(elem, pt);
            Sc[a] = dens(elem, pt) * (*ut[a])(elem, pt) + dens(elem, pt) * conv + (*dp[a])(elem, pt) - dens(elem, pt) * src[a](elem, pt);
            Sc[a] *= tau * w / dens(elem, pt);
          }
          for (int dof = 0; dof < nb; ++dof) {
            EvalT s = Sc[0] * g.G(elem, dof, pt, 0);
            for (int c = 1; c < dim; ++c) s = s + Sc[c] * g.G(elem, dof, pt, c);
            res(elem, wkset->off(pr_num, dof)) += s;
          }
        }
    }
  }
  void boundaryResidual() override {
    const int dim = spaceDim, nb = wkset->nb, nq = wkset->gd->nq; const char* un[3] = {"ux", "uy", "uz"};
    auto& res = wkset->res; const GroupData& g = *wkset->gd;
    for (int a = 0; a < dim; ++a) {
      if (wkset->var_bcs[unum[a]][wkset->currentside] != "Neumann") continue;
      auto src = functionManager->evaluate(std::string("Neumann ") + un[a] + " " + wkset->sidename, "side ip");
      for (int e = 0; e < wkset->numElem; ++e)
        for (int k = 0; k < nq; ++k)
          for (int i = 0; i < nb; ++i)
            res(e, wkset->off(unum[a], i)) += (-src(e, k) * g.B(e, i, k)) * wkset->wt(e, k);
    }
  }
};

// ---------------------------------------------------------------------------------------------- stokes
// /root/reference/src/physics/stokes.cpp: ctor :20-46, defineFunctions :52-63, volumeResidual :69-440 (1-D :90-197, 2-D :199-288,
// 3-D :290-438), boundaryResidual :445-447 (empty).  Quirks kept: the PSPG term uses grad(p) PLUS the source, and its parameter is
// h/(2 visc) in 2-D but h^2/(2 visc) in 1-D and 3-D (:256-257, :402); LSIC adds tau div(u) (d phi/dx + d phi/dy [+ d phi/dz]) to the
// PRESSURE rows.
template <class EvalT>
struct stokes : PhysicsBase<EvalT> {
  using PhysicsBase<EvalT>::wkset; using PhysicsBase<EvalT>::functionManager;
  int spaceDim; bool useLSIC = false, usePSPG = false;
  int unum[3] = {0, 0, 0}, pr_num = 0;
  stokes(const StrMap& settings, int dimension) : spaceDim(dimension) {
    this->label = "stokes";
    this->myvars.push_back("ux"); this->myvars.push_back("pr");
    if (spaceDim > 1) this->myvars.push_back("uy");
    if (spaceDim > 2) this->myvars.push_back("uz");
    for (size_t i = 0; i < this->myvars.size(); ++i) this->mybasistypes.push_back("HGRAD");
    useLSIC = getBool(settings, "useLSIC", false); usePSPG = getBool(settings, "usePSPG", false);
  }
  void defineFunctions(const StrMap& fs, FunctionManager<EvalT>* fm) override {
    functionManager = fm;
    fm->addFunction("source ux", getOr(fs, "source ux", "0.0"), "ip");
    fm->addFunction("source pr", getOr(fs, "source pr", "0.0"), "ip");
    fm->addFunction("source uy", getOr(fs, "source uy", "0.0"), "ip");
    fm->addFunction("source uz", getOr(fs, "source uz", "0.0"), "ip");
    fm->addFunction("viscosity", getOr(fs, "viscosity", "1.0"), "ip");
  }
  void setWorkset(Workset<EvalT>* w) override {
    wkset = w; const char* nm[3] = {"ux", "uy", "uz"};
    for (int c = 0; c < spaceDim; ++c) unum[c] = this->varnum(nm[c]);
    pr_num = this->varnum("pr");
  }
  void volumeResidual() override {
    const int dim = spaceDim, nb = wkset->nb, nq = wkset->gd->nq;
    const char* un[3] = {"ux", "uy", "uz"}; const char* sn[3] = {"source ux", "source uy", "source uz"};
    Vista<EvalT> src[3];
    src[0] = functionManager->evaluate("source ux", "ip");
    functionManager->evaluate("source pr", "ip");
    for (int c = 1; c < dim; ++c) src[c] = functionManager->evaluate(sn[c], "ip");
    auto visc = functionManager->evaluate("viscosity", "ip");
    Arr2<EvalT>* du[3][3]; Arr2<EvalT>* dp[3];
    for (int a = 0; a < dim; ++a) {
      for (int c = 0; c < dim; ++c) du[a][c] = &wkset->getSolutionField(gradname(un[a], c));
      dp[a] = &wkset->getSolutionField(gradname("pr", a));
    }
    auto& pr = wkset->getSolutionField("pr");
    auto h = wkset->getElementSize();
    auto& res = wkset->res; const GroupData& g = *wkset->gd;
    for (int a = 0; a < dim; ++a)
      for (int elem = 0; elem < wkset->numElem; ++elem)
        for (int pt = 0; pt < nq; ++pt) {
          const double w = wkset->wt(elem, pt);
          EvalT Fc[3];
          for (int c = 0; c < dim; ++c) {
            Fc[c] = visc(elem, pt) * (*du[a][c])(elem, pt);
            if (c == a) Fc[c] = Fc[c] - pr(elem, pt);
            Fc[c] *= w;
          }
          EvalT gsrc = -src[a](elem, pt) * w;
          for (int dof = 0; dof < nb; ++dof) {
            EvalT s = Fc[0] * g.G(elem, dof, pt, 0);
            for (int c = 1; c < dim; ++c) s = s + Fc[c] * g.G(elem, dof, pt, c);
            res(elem, wkset->off(unum[a], dof)) += s + gsrc * g.B(elem, dof, pt);
          }
        }
    for (int elem = 0; elem < wkset->numElem; ++elem)
      for (int pt = 0; pt < nq; ++pt) {
        EvalT divu = (*du[0][0])(elem, pt);
        for (int c = 1; c < dim; ++c) divu = divu + (*du[c][c])(elem, pt);
        divu *= wkset->wt(elem, pt);
        for (int dof = 0; dof < nb; ++dof) res(elem, wkset->off(pr_num, dof)) += divu * g.B(elem, dof, pt);
      }
    if (usePSPG)
      for (int elem = 0; elem < wkset->numElem; ++elem)
        for (int pt = 0; pt < nq; ++pt) {
          const double alpha = 1.0;
          EvalT tau = (dim == 2) ? alpha * h[elem] / (2. * visc(elem, pt)) : alpha * h[elem] * h[elem] / (2. * visc(elem, pt));
          EvalT S[3];
          for (int c = 0; c < dim; ++c) { S[c] = (*dp[c])(elem, pt) + src[c](elem, pt); S[c] *= tau * wkset->wt(elem, pt); }
          for (int dof = 0; dof < nb; ++dof) {
            EvalT s = S[0] * g.G(elem, dof, pt, 0);
            for (int c = 1; c < dim; ++c) s = s + S[c] * g.G(elem, dof, pt, c);
            res(elem, wkset->off(pr_num, dof)) += s;
          }
        }
    if (useLSIC)
      for (int elem = 0; elem < wkset->numElem; ++elem)
        for (int pt = 0; pt < nq; ++pt) {
          const double alpha = 1.0;
          EvalT tau = alpha * h[elem] * h[elem] / (2. * visc(elem, pt));
          EvalT stabres = (*du[0][0])(elem, pt);
          for (int c = 1; c < dim; ++c) stabres = stabres + (*du[c][c])(elem, pt);
          EvalT S = tau * stabres * wkset->wt(elem, pt);
          for (int dof = 0; dof < nb; ++dof) {
            double gs = g.G(elem, dof, pt, 0);
            for (int c = 1; c < dim; ++c) gs += g.G(elem, dof, pt, c);
            res(elem, wkset->off(pr_num, dof)) += S * gs;
          }
        }
  }
  void boundaryResidual() override {}
};

// ================================================================================================
// helmholtz (frequency-domain wave equation, real / imaginary parts as two HGRAD variables), non-fractional branch:
// /root/reference/src/physics/helmholtz.cpp:17-66 (variables, functions), :71-240 (volumeResidual), :245-420 (boundaryResidual:
// Robin condition on sides where ureal is "Neumann").  Both variables use the same basis, so the reference's vr / vi (and their
// gradients) are the same numbers; its formulas are kept term by term.
template <class EvalT>
struct helmholtz : PhysicsBase<EvalT> {
  using PhysicsBase<EvalT>::wkset; using PhysicsBase<EvalT>::functionManager;
  int ur_num = 0, ui_num = 1;
  helmholtz(const StrMap& settings, int) {
    this->label = "helmholtz"; this->myvars = {"ureal", "uimag"}; this->mybasistypes = {"HGRAD", "HGRAD"};
    if (getOr(settings, "fractional", "false") == "true") throw std::runtime_error("oracle: helmholtz 'fractional' branch is not restated");
  }
  void defineFunctions(const StrMap& fs, FunctionManager<EvalT>* fm) override {
    functionManager = fm;
    for (const char* n : {"c2r_x", "c2i_x", "c2r_y", "c2i_y", "c2r_z", "c2i_z", "omega2r", "omega2i", "omegar", "omegai", "source_r", "source_i"})
      fm->addFunction(n, getOr(fs, n, "0.0"), "ip");
    for (const char* n : {"omega2r", "omega2i", "source_r_side", "source_i_side", "robin_alpha_r", "robin_alpha_i", "c2r_x", "c2i_x", "c2r_y", "c2i_y", "c2r_z", "c2i_z"})
      fm->addFunction(n, getOr(fs, n, "0.0"), "side ip");
  }
  void setWorkset(Workset<EvalT>* w) override { wkset = w; ur_num = this->varnum("ureal"); ui_num = this->varnum("uimag"); }
  void volumeResidual() override {
    const int dim = wkset->dim, nb = wkset->nb, nq = wkset->gd->nq;
    const char* XN[3] = {"x", "y", "z"};
    Vista<EvalT> c2r[3], c2i[3];
    for (int c = 0; c < 3; ++c) { c2r[c] = functionManager->evaluate(std::string("c2r_") + XN[c], "ip"); c2i[c] = functionManager->evaluate(std::string("c2i_") + XN[c], "ip"); }
    auto omega2r = functionManager->evaluate("omega2r", "ip"), omega2i = functionManager->evaluate("omega2i", "ip");
    auto source_r = functionManager->evaluate("source_r", "ip"), source_i = functionManager->evaluate("source_i", "ip");
    auto& Ur = wkset->getSolutionField("ureal"); auto& Ui = wkset->getSolutionField("uimag");
    Arr2<EvalT>* dUr[3]; Arr2<EvalT>* dUi[3];
    for (int c = 0; c < dim; ++c) { dUr[c] = &wkset->getSolutionField(gradname("ureal", c)); dUi[c] = &wkset->getSolutionField(gradname("uimag", c)); }
    auto& res = wkset->res; const GroupData& g = *wkset->gd;
    for (int e = 0; e < wkset->numElem; ++e)
      for (int k = 0; k < nq; ++k) {
        const EvalT ur = Ur(e, k), ui = Ui(e, k);
        const double w = wkset->wt(e, k);
        for (int i = 0; i < nb; ++i) {
          const double vr = g.B(e, i, k), vi = g.B(e, i, k);
          EvalT sr = -omega2r(e, k) * (ur * vr + ui * vi) + omega2i(e, k) * (ui * vr - ur * vi);
          EvalT si = -omega2r(e, k) * (ui * vr - ur * vi) - omega2i(e, k) * (ur * vr + ui * vi);
          for (int c = 0; c < dim; ++c) {
            const double dvr = g.G(e, i, k, c), dvi = g.G(e, i, k, c);
            const EvalT dur = (*dUr[c])(e, k), dui = (*dUi[c])(e, k);
            sr = sr + c2r[c](e, k) * (dur * dvr + dui * dvi) - c2i[c](e, k) * (dui * dvr - dur * dvi);
            si = si + c2r[c](e, k) * (dui * dvr - dur * dvi) + c2i[c](e, k) * (dur * dvr + dui * dvi);
          }
          sr = sr - (source_r(e, k) * vr + source_i(e, k) * vi);
          si = si - (source_i(e, k) * vr - source_r(e, k) * vi);
          res(e, wkset->off(ur_num, i)) += sr * w;
          res(e, wkset->off(ui_num, i)) += si * w;
        }
      }
  }
  void boundaryResidual() override {
    const int dim = wkset->dim, nb = wkset->nb, nq = wkset->gd->nq;
    const char* XN[3] = {"x", "y", "z"};
    Vista<EvalT> c2r[3], c2i[3];
    for (int c = 0; c < 3; ++c) { c2r[c] = functionManager->evaluate(std::string("c2r_") + XN[c], "side ip"); c2i[c] = functionManager->evaluate(std::string("c2i_") + XN[c], "side ip"); }
    auto alpha_r = functionManager->evaluate("robin_alpha_r", "side ip"), alpha_i = functionManager->evaluate("robin_alpha_i", "side ip");
    auto src_r = functionManager->evaluate("source_r_side", "side ip"), src_i = functionManager->evaluate("source_i_side", "side ip");
    if (wkset->var_bcs[ur_num][wkset->currentside] != "Neumann") return;       // Robin condition alpha u + du/dn - source = 0
    auto& Ur = wkset->getSolutionField("ureal"); auto& Ui = wkset->getSolutionField("uimag");
    Arr2<EvalT>* dUr[3]; Arr2<EvalT>* dUi[3]; Arr2<double>* n[3];
    for (int c = 0; c < dim; ++c) {
      dUr[c] = &wkset->getSolutionField(gradname("ureal", c)); dUi[c] = &wkset->getSolutionField(gradname("uimag", c));
      n[c] = &wkset->getScalarField(std::string("n[") + XYZ[c] + "]");
    }
    auto& res = wkset->res; const GroupData& g = *wkset->gd;
    for (int e = 0; e < wkset->numElem; ++e)
      for (int k = 0; k < nq; ++k) {
        const EvalT ur = Ur(e, k), ui = Ui(e, k);
        EvalT durdn = EvalT(0.0), duidn = EvalT(0.0), c2durdn = EvalT(0.0), c2duidn = EvalT(0.0);
        for (int c = 0; c < dim; ++c) {
          const EvalT dur = (*dUr[c])(e, k), dui = (*dUi[c])(e, k);
          const double nc = (*n[c])(e, k);
          durdn = durdn + dur * nc; duidn = duidn + dui * nc;
          c2durdn = c2durdn + (c2r[c](e, k) * dur - c2i[c](e, k) * dui) * nc;
          c2duidn = c2duidn + (c2r[c](e, k) * dui + c2i[c](e, k) * dur) * nc;
        }
        const double w = wkset->wt(e, k);
        for (int i = 0; i < nb; ++i) {
          const double vr = g.B(e, i, k), vi = g.B(e, i, k);
          res(e, wkset->off(ur_num, i)) += (((alpha_r(e, k) * (ur * vr + ui * vi) - alpha_i(e, k) * (ui * vr - ur * vi)) + (durdn * vr + duidn * vi)
                                            - (src_r(e, k) * vr + src_i(e, k) * vi)) - (c2durdn * vr + c2duidn * vi)) * w;
          res(e, wkset->off(ui_num, i)) += (((alpha_r(e, k) * (ui * vr - ur * vi) + alpha_i(e, k) * (ur * vr + ui * vi)) + (duidn * vr - durdn * vi)
                                            - (src_i(e, k) * vr - src_r(e, k) * vi)) - (c2duidn * vr - c2durdn * vi)) * w;
        }
      }
  }
};

// ================================================================================================
// Burgers: du/dt + div(1/2 v u^2 - eps grad u) = source.  /root/reference/src/physics/burgers.cpp:17-26 (settings), :31-45 (functions),
// :50-164 (volumeResidual: 2-D with optional entropy viscosity and SUPG, 3-D plain), :169-205 (boundaryResidual: 2-D Neumann flux).
// Quirks kept: the entropy viscosity is clamped to the PLAIN number 0.1 (its derivatives vanish once clamped); |v| is left squared when
// |v|^2 <= 1e-12; "C1", "C2", "supg C", "supg C1", "supg C2" are user functions without module defaults.
template <class EvalT>
struct burgers : PhysicsBase<EvalT> {
  using PhysicsBase<EvalT>::wkset; using PhysicsBase<EvalT>::functionManager;
  bool use_evisc = false, use_SUPG = false; int unum = 0;
  burgers(const StrMap& settings, int) {
    this->label = "Burgers"; this->myvars = {"u"}; this->mybasistypes = {"HGRAD"};
    use_evisc = getOr(settings, "use entropy viscosity", "false") == "true";
    use_SUPG = getOr(settings, "use SUPG", "false") == "true";
  }
  void defineFunctions(const StrMap& fs, FunctionManager<EvalT>* fm) override {
    functionManager = fm;
    fm->addFunction("Burgers source", getOr(fs, "Burgers source", "0.0"), "ip");
    fm->addFunction("diffusion", getOr(fs, "diffusion", "0.0"), "ip");
    for (const char* n : {"xvel", "yvel", "zvel"}) fm->addFunction(n, getOr(fs, n, "1.0"), "ip");
    for (const char* n : {"xvel", "yvel", "zvel"}) fm->addFunction(n, getOr(fs, n, "1.0"), "side ip");
  }
  void setWorkset(Workset<EvalT>* w) override { wkset = w; unum = this->varnum("u"); }
  void volumeResidual() override {
    const int dim = wkset->dim, nb = wkset->nb, nq = wkset->gd->nq;
    auto source = functionManager->evaluate("Burgers source", "ip");
    auto eps = functionManager->evaluate("diffusion", "ip");
    const char* vn[3] = {"xvel", "yvel", "zvel"};
    Vista<EvalT> v[3]; for (int c = 0; c < dim; ++c) v[c] = functionManager->evaluate(vn[c], "ip");
    Vista<EvalT> C1, C2, sC, sC1, sC2;
    const bool ev = use_evisc && dim == 2, supg = use_SUPG && dim == 2;          // the 3-D branch has neither (burgers.cpp:141-163)
    if (ev) { C1 = functionManager->evaluate("C1", "ip"); C2 = functionManager->evaluate("C2", "ip"); }
    if (supg) { sC = functionManager->evaluate("supg C", "ip"); sC1 = functionManager->evaluate("supg C1", "ip"); sC2 = functionManager->evaluate("supg C2", "ip"); }
    auto& U = wkset->getSolutionField("u"); auto& dUdt = wkset->getSolutionField("u_t");
    Arr2<EvalT>* dU[3]; for (int c = 0; c < dim; ++c) dU[c] = &wkset->getSolutionField(gradname("u", c));
    auto h = wkset->getElementSize(); const double dt = wkset->deltat;
    auto& res = wkset->res; const GroupData& g = *wkset->gd;
    for (int elem = 0; elem < wkset->numElem; ++elem)
      for (int pt = 0; pt < nq; ++pt) {
        const EvalT u = U(elem, pt), dudt = dUdt(elem, pt);
        const double w = wkset->wt(elem, pt);
        EvalT evisc = EvalT(0.0);
        if (ev) {
          EvalT entres = u * (dudt + u * (*dU[0])(elem, pt) + u * (*dU[1])(elem, pt));
          EvalT a = 1.0e-12 + entres; if (value_of(a) < 0.0) a = -a;
          evisc = C1(elem, pt) * h[elem] * h[elem] * a / C2(elem, pt);
        }
        if (value_of(evisc) > 0.1) evisc = EvalT(0.1);
        const EvalT usq = 0.5 * u * u;
        const EvalT f = (dudt - source(elem, pt)) * w;
        EvalT F[3];
        for (int c = 0; c < dim; ++c) F[c] = ((eps(elem, pt) + evisc) * (*dU[c])(elem, pt) - v[c](elem, pt) * usq) * w;
        if (supg) {
          EvalT nvel = v[0](elem, pt) * v[0](elem, pt) + v[1](elem, pt) * v[1](elem, pt);
          if (value_of(nvel) > 1E-12) nvel = sqrt(nvel);
          const EvalT tau = sC(elem, pt) / (sC1(elem, pt) / dt + sC2(elem, pt) * nvel / h[elem]);
          const EvalT sres = tau * (dudt + v[0](elem, pt) * u * (*dU[0])(elem, pt) + v[1](elem, pt) * u * (*dU[1])(elem, pt) - source(elem, pt)) * w;
          F[0] = F[0] + sres * u * v[0](elem, pt);
          F[1] = F[1] + sres * u * v[1](elem, pt);
        }
        for (int dof = 0; dof < nb; ++dof) {
          EvalT r = f * g.B(elem, dof, pt);
          for (int c = 0; c < dim; ++c) r = r + F[c] * g.G(elem, dof, pt, c);
          res(elem, wkset->off(unum, dof)) += r;
        }
      }
  }
  void boundaryResidual() override {
    const int dim = wkset->dim, nb = wkset->nb, nq = wkset->gd->nq;
    if (!(dim == 2 && wkset->var_bcs[unum][wkset->currentside] == "Neumann")) return;
    auto& U = wkset->getSolutionField("u");
    auto vx = functionManager->evaluate("xvel", "side ip"), vy = functionManager->evaluate("yvel", "side ip");
    auto& nx = wkset->getScalarField("n[x]"); auto& ny = wkset->getScalarField("n[y]");
    auto& res = wkset->res; const GroupData& g = *wkset->gd;
    for (int elem = 0; elem < wkset->numElem; ++elem)
      for (int dof = 0; dof < nb; ++dof)
        for (int pt = 0; pt < nq; ++pt)
          res(elem, wkset->off(unum, dof)) += 0.5 * U(elem, pt) * U(elem, pt) * wkset->wt(elem, pt) * g.B(elem, dof, pt) * (vx(elem, pt) * nx(elem, pt) + vy(elem, pt) * ny(elem, pt));
  }
};

template <class EvalT>
inline std::unique_ptr<PhysicsBase<EvalT>> importPhysics(const std::string& name, const StrMap& settings, int dim) {
  if (name == "thermal") return std::unique_ptr<PhysicsBase<EvalT>>(new thermal<EvalT>(settings, dim));
  if (name == "cdr") return std::unique_ptr<PhysicsBase<EvalT>>(new cdr<EvalT>(settings, dim));
  if (name == "linearelasticity") return std::unique_ptr<PhysicsBase<EvalT>>(new linearelasticity<EvalT>(settings, dim));
  if (name == "navierstokes") return std::unique_ptr<PhysicsBase<EvalT>>(new navierstokes<EvalT>(settings, dim));
  if (name == "burgers" || name == "Burgers") return std::unique_ptr<PhysicsBase<EvalT>>(new burgers<EvalT>(settings, dim));
  if (name == "helmholtz") return std::unique_ptr<PhysicsBase<EvalT>>(new helmholtz<EvalT>(settings, dim));
  if (name == "stokes" || name == "Stokes") return std::unique_ptr<PhysicsBase<EvalT>>(new stokes<EvalT>(settings, dim));
  throw std::runtime_error("oracle: physics module not restated: " + name);
}

}  // namespace orc
