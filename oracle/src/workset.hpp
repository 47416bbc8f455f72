// ORACLE (test infrastructure, NOT product code).
// Per-group scratch ("workset") restating /root/reference/src/tools/workset.cpp:
//   createSolutionFields :118-330 (field names  var, grad(var)[x|y|z], var_t),
//   resetResidual :425-471, computeSolnTransientSeeded :479-792, computeSolnSteadySeeded :800-859,
//   evaluateSolutionField :937-1062 (fixed summation order dof0 then dof1..), evaluateSideSolutionField :1069-1185,
//   getElementSize / getSideElementSize :2666-2697.
#pragma once
#include <cmath>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>
#include "fad.hpp"

namespace orc {

template <class T>
struct Arr2 {
  int n0 = 0, n1 = 0;
  std::vector<T> a;
  Arr2() {}
  Arr2(int m0, int m1) : n0(m0), n1(m1), a((size_t)m0 * m1) {}
  void resize(int m0, int m1) { n0 = m0; n1 = m1; a.assign((size_t)m0 * m1, T()); }
  T& operator()(int i, int j) { return a[(size_t)i * n1 + j]; }
  const T& operator()(int i, int j) const { return a[(size_t)i * n1 + j]; }
};

// stored physical data of one element batch (Group / BoundaryGroup in the reference)
struct GroupData {
  int numElem = 0, nb = 0, nq = 0, dim = 0;
  std::vector<int> elems;                       // element ids of the batch
  std::vector<double> basis;                    // (e, nb, nq)
  std::vector<double> basis_grad;               // (e, nb, nq, dim)
  std::vector<double> wts;                      // (e, nq)
  std::vector<double> ipx[3];                   // (e, nq) each
  std::vector<double> nrm[3], tng[3];           // side only
  double B(int e, int i, int q) const { return basis[((size_t)e * nb + i) * nq + q]; }
  double G(int e, int i, int q, int c) const { return basis_grad[(((size_t)e * nb + i) * nq + q) * dim + c]; }
};

template <class EvalT>
struct SolnField {
  std::string expr; int var = 0; int deriv = 0;   // deriv: 0 value, 1 grad, 2 time
  int comp = 0; bool updated = false; Arr2<EvalT> data;
};
struct ScalarFieldO { std::string expr; Arr2<double> data; };

template <class EvalT>
struct Workset {
  int dim = 3, nvar = 1, nb = 8, maxElem = 0, numElem = 0, nq = 0, nqs = 0;
  std::vector<std::string> varlist;
  std::vector<int> offsets;                      // (nvar, nb)
  std::vector<std::vector<std::string>> var_bcs; // [var][side]
  bool isTransient = false, isOnSide = false;
  double time = 0.0, deltat = 1.0;
  int current_stage = 0;
  std::vector<double> butcher_A, butcher_b, BDF_wts;   // A is (ns,ns)
  int nstages = 1;
  std::string sidename; int currentside = 0;

  std::vector<Arr2<EvalT>> sol_vals, sol_dot_vals;     // [var](e, nb)
  Arr2<EvalT> res;                                      // (e, nvar*nb)
  const GroupData* gd = nullptr;                        // current basis/wts (volume or side)
  std::vector<SolnField<EvalT>> soln_fields, side_soln_fields;
  std::vector<ScalarFieldO> scalar_fields, side_scalar_fields;

  int off(int var, int dof) const { return offsets[(size_t)var * nb + dof]; }

  void setup(int dim_, const std::vector<std::string>& vars, int nb_, const std::vector<int>& offs,
             int maxElem_, int nq_, int nqs_, bool transient) {
    dim = dim_; varlist = vars; nvar = (int)vars.size(); nb = nb_; offsets = offs;
    maxElem = maxElem_; nq = nq_; nqs = nqs_; isTransient = transient;
    sol_vals.assign(nvar, Arr2<EvalT>(maxElem, nb));
    sol_dot_vals.assign(nvar, Arr2<EvalT>(maxElem, nb));
    res.resize(maxElem, nvar * nb);
    static const char* xyz[3] = {"x", "y", "z"};
    for (int v = 0; v < nvar; ++v) {
      auto add = [&](std::vector<SolnField<EvalT>>& L, const std::string& e, int deriv, int comp, int npt) {
        SolnField<EvalT> f; f.expr = e; f.var = v; f.deriv = deriv; f.comp = comp; f.data.resize(maxElem, npt); L.push_back(f);
      };
      add(soln_fields, vars[v], 0, 0, nq);
      for (int c = 0; c < 3; ++c) add(soln_fields, "grad(" + vars[v] + ")[" + xyz[c] + "]", 1, c, nq);
      add(soln_fields, vars[v] + "_t", 2, 0, nq);
      add(side_soln_fields, vars[v], 0, 0, nqs);
      for (int c = 0; c < 3; ++c) add(side_soln_fields, "grad(" + vars[v] + ")[" + xyz[c] + "]", 1, c, nqs);
    }
    for (int c = 0; c < 3; ++c) { ScalarFieldO s; s.expr = xyz[c]; s.data.resize(maxElem, nq); scalar_fields.push_back(s); }
    for (int c = 0; c < 3; ++c) { ScalarFieldO s; s.expr = xyz[c]; s.data.resize(maxElem, nqs); side_scalar_fields.push_back(s); }
    for (int c = 0; c < 3; ++c) { ScalarFieldO s; s.expr = std::string("n[") + xyz[c] + "]"; s.data.resize(maxElem, nqs); side_scalar_fields.push_back(s); }
    for (int c = 0; c < 3; ++c) { ScalarFieldO s; s.expr = std::string("t[") + xyz[c] + "]"; s.data.resize(maxElem, nqs); side_scalar_fields.push_back(s); }
  }

  void resetResidual() { for (auto& r : res.a) r = EvalT(0.0); }

  void resetSolutionFields() {
    for (auto& f : soln_fields) f.updated = false;
    for (auto& f : side_soln_fields) f.updated = false;
  }

  // attach a group's stored data; copies ip coordinates / normals into the scalar fields
  void attach(const GroupData* g, bool onside) {
    gd = g; numElem = g->numElem; isOnSide = onside;
    auto& sf = onside ? side_scalar_fields : scalar_fields;
    const int npt = g->nq;
    for (int c = 0; c < dim; ++c)
      for (int e = 0; e < numElem; ++e) for (int q = 0; q < npt; ++q) sf[c].data(e, q) = g->ipx[c][(size_t)e * npt + q];
    if (onside)
      for (int c = 0; c < dim; ++c)
        for (int e = 0; e < numElem; ++e) for (int q = 0; q < npt; ++q) {
          sf[3 + c].data(e, q) = g->nrm[c][(size_t)e * npt + q];
          sf[6 + c].data(e, q) = g->tng[c].empty() ? 0.0 : g->tng[c][(size_t)e * npt + q];
        }
    resetSolutionFields();
  }

  // ---- seeding (u is the gathered (e,var,dof) block) -------------------------------------------
  void computeSolnSteadySeeded(const std::vector<double>& u, int seedwhat) {
    for (int v = 0; v < nvar; ++v)
      for (int e = 0; e < numElem; ++e)
        for (int k = 0; k < nb; ++k) {
          const double val = u[((size_t)e * nvar + v) * nb + k];
          if constexpr (ad_traits<EvalT>::lanes > 0) {
            if (seedwhat == 1) sol_vals[v](e, k) = EvalT(EvalT::size(), off(v, k), val);
            else sol_vals[v](e, k) = EvalT(val);
          } else sol_vals[v](e, k) = val;
        }
  }
  // u_prev (e,var,dof,nsteps), u_stage (e,var,dof,nstages)
  void computeSolnTransientSeeded(const std::vector<double>& u, const std::vector<double>& u_prev, int nsteps,
                                  const std::vector<double>& u_stage, int seedwhat) {
    const int stage = current_stage;
    const double alpha_u = butcher_A[(size_t)stage * nstages + stage] / butcher_b[stage];
    const double timewt = 1.0 / deltat / butcher_b[stage];
    const double alpha_t = BDF_wts[0] * timewt;
    for (int v = 0; v < nvar; ++v)
      for (int e = 0; e < numElem; ++e)
        for (int k = 0; k < nb; ++k) {
          const size_t id = ((size_t)e * nvar + v) * nb + k;
          EvalT stageval;
          if constexpr (ad_traits<EvalT>::lanes > 0) {
            if (seedwhat == 1) stageval = EvalT(EvalT::size(), off(v, k), u[id]); else stageval = EvalT(u[id]);
          } else stageval = u[id];
          double beta_u = (1.0 - alpha_u) * u_prev[id * nsteps + 0];
          for (int s = 0; s < stage; ++s)
            beta_u += butcher_A[(size_t)stage * nstages + s] / butcher_b[s] * (u_stage[id * nstages + s] - u_prev[id * nsteps + 0]);
          sol_vals[v](e, k) = alpha_u * stageval + beta_u;
          double beta_t = 0.0;
          for (size_t s = 1; s < BDF_wts.size(); ++s) beta_t += BDF_wts[s] * u_prev[id * nsteps + (s - 1)];
          beta_t *= timewt;
          sol_dot_vals[v](e, k) = alpha_t * stageval + beta_t;
        }
  }

  // ---- lazy field evaluation ---------------------------------------------------------------------
  void evaluateField(SolnField<EvalT>& f) {
    bool proceed = true;
    if (f.deriv == 2 && (!isTransient || isOnSide)) proceed = false;
    if (!proceed) return;
    const Arr2<EvalT>& sv = (f.deriv == 2) ? sol_dot_vals[f.var] : sol_vals[f.var];
    const int npt = gd->nq;
    for (int e = 0; e < numElem; ++e) {
      for (int q = 0; q < npt; ++q) {
        const double b = (f.deriv == 1) ? ((f.comp < dim) ? gd->G(e, 0, q, f.comp) : 0.0) : gd->B(e, 0, q);
        f.data(e, q) = sv(e, 0) * b;
      }
      for (int k = 1; k < nb; ++k)
        for (int q = 0; q < npt; ++q) {
          const double b = (f.deriv == 1) ? ((f.comp < dim) ? gd->G(e, k, q, f.comp) : 0.0) : gd->B(e, k, q);
          f.data(e, q) += sv(e, k) * b;
        }
    }
    f.updated = true;
  }
  int findField(const std::string& name, bool side) const {
    const auto& L = side ? side_soln_fields : soln_fields;
    for (size_t i = 0; i < L.size(); ++i) if (L[i].expr == name) return (int)i;
    return -1;
  }
  Arr2<EvalT>& getSolutionField(const std::string& name) {
    auto& L = isOnSide ? side_soln_fields : soln_fields;
    const int i = findField(name, isOnSide);
    if (i < 0) throw std::runtime_error("oracle workset: unknown solution field " + name);
    if (!L[i].updated) evaluateField(L[i]);
    return L[i].data;
  }
  Arr2<double>& getScalarField(const std::string& name) {
    auto& L = isOnSide ? side_scalar_fields : scalar_fields;
    for (auto& s : L) if (s.expr == name) return s.data;
    throw std::runtime_error("oracle workset: unknown scalar field " + name);
  }
  std::vector<double> getElementSize() const {
    std::vector<double> h(numElem);
    for (int e = 0; e < numElem; ++e) {
      double vol = 0.0;
      for (int q = 0; q < gd->nq; ++q) vol += gd->wts[(size_t)e * gd->nq + q];
      h[e] = std::pow(vol, 1.0 / (double)dim);
    }
    return h;
  }
  std::vector<double> getSideElementSize() const {
    std::vector<double> h(numElem);
    for (int e = 0; e < numElem; ++e) {
      double vol = 0.0;
      for (int q = 0; q < gd->nq; ++q) vol += gd->wts[(size_t)e * gd->nq + q];
      h[e] = std::pow(vol, 1.0 / ((double)dim - 1.0));
    }
    return h;
  }
  double wt(int e, int q) const { return gd->wts[(size_t)e * gd->nq + q]; }
};

}  // namespace orc
