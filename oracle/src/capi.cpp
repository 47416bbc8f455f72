// ORACLE (test infrastructure, NOT product code).  C entry points for ctypes (tests, smoke, bench cpu legs).
// Parity status: basis / quadrature / J^-T pinned by regression/discretization/HGRAD/mrhyde.gold; the
// end-to-end path pinned by the 6-digit L2 errors of regression/{thermal,cdr,le,navierstokes} golds
// (see tests/test_oracle_golden.py).  DOF numbering / CRS conventions of the absent Trilinos pieces: UNPINNED.
#include <cstring>
#include "assembly.hpp"

using namespace orc;

static thread_local std::string g_err;

extern "C" {

const char* orc_last_error() { return g_err.c_str(); }

void* orc_create(const char* text) {
  try { return new Oracle(text); } catch (std::exception& e) { g_err = e.what(); return nullptr; }
}
void orc_destroy(void* h) { delete (Oracle*)h; }

// sizes: [dim, nelem, nnode, nvert, nvar, nb, ndof_elem, nlocal, nnz, nq, nqs, nsides, type_AD, ngroups, nbgroups]
void orc_sizes(void* h, int64_t* out) {
  Oracle* o = (Oracle*)h;
  int64_t v[15] = {o->cfg.dim, o->mesh.nelem, o->mesh.nnode, o->mesh.nvert, (int64_t)o->varlist.size(), o->ref.nb,
                   o->dofs.ndof_elem, o->dofs.nlocal, (int64_t)o->crs.colind.size(), o->ref.nq, o->ref.nqs,
                   (int64_t)o->mesh.sides.size(), o->type_AD, (int64_t)o->groups.size(), (int64_t)o->bgroups.size()};
  memcpy(out, v, sizeof(v));
}
void orc_get_mesh(void* h, double* nodes, int* elem2node) {
  Oracle* o = (Oracle*)h;
  memcpy(nodes, o->mesh.nodes.data(), o->mesh.nodes.size() * sizeof(double));
  memcpy(elem2node, o->mesh.elem2node.data(), o->mesh.elem2node.size() * sizeof(int));
}
void orc_get_dofmap(void* h, int* lids, int* offsets, uint8_t* is_fixed, double* dofxyz, int* dofvar) {
  Oracle* o = (Oracle*)h;
  if (lids) memcpy(lids, o->dofs.lids.data(), o->dofs.lids.size() * sizeof(int));
  if (offsets) memcpy(offsets, o->dofs.offsets.data(), o->dofs.offsets.size() * sizeof(int));
  if (is_fixed) memcpy(is_fixed, o->isFixed.data(), o->isFixed.size());
  if (dofxyz) memcpy(dofxyz, o->dofs.dofxyz.data(), o->dofs.dofxyz.size() * sizeof(double));
  if (dofvar) memcpy(dofvar, o->dofs.dofvar.data(), o->dofs.dofvar.size() * sizeof(int));
}
void orc_get_crs(void* h, int64_t* rowptr, int* colind) {
  Oracle* o = (Oracle*)h;
  memcpy(rowptr, o->crs.rowptr.data(), o->crs.rowptr.size() * sizeof(int64_t));
  memcpy(colind, o->crs.colind.data(), o->crs.colind.size() * sizeof(int));
}
// replace the vertex coordinates (nnode, dim); topology, numbering and graph stay
void orc_set_nodes(void* h, const double* nodes) { ((Oracle*)h)->setNodes(nodes); }
// side set s: name into buf, returns element count; elems/local side optional
int orc_get_side(void* h, int s, char* name, int namelen, int* local_side, int* elems) {
  Oracle* o = (Oracle*)h; const auto& S = o->mesh.sides[s];
  if (name) { strncpy(name, S.name.c_str(), namelen - 1); name[namelen - 1] = 0; }
  if (local_side) *local_side = S.local_side;
  if (elems) memcpy(elems, S.elems.data(), S.elems.size() * sizeof(int));
  return (int)S.elems.size();
}
// var_bcs(var, side) as a code: 0 none, 1 Dirichlet, 2 weak Dirichlet, 3 Neumann, 4 Flux
int orc_get_bc(void* h, int var, int side) {
  const std::string& b = ((Oracle*)h)->var_bcs[var][side];
  if (b == "Dirichlet") return 1; if (b == "weak Dirichlet") return 2; if (b == "Neumann") return 3; if (b == "Flux") return 4;
  return 0;
}
// stored physical basis of volume group g (reference-test print order: basis(elem,dof,pt), basis_grad(elem,dof,pt,c))
void orc_get_group_basis(void* h, int g, double* basis, double* basis_grad, double* wts, double* ip) {
  Oracle* o = (Oracle*)h; const GroupData& G = o->groups[g];
  if (basis) memcpy(basis, G.basis.data(), G.basis.size() * sizeof(double));
  if (basis_grad) memcpy(basis_grad, G.basis_grad.data(), G.basis_grad.size() * sizeof(double));
  if (wts) memcpy(wts, G.wts.data(), G.wts.size() * sizeof(double));
  if (ip) for (int c = 0; c < G.dim; ++c) memcpy(ip + c * G.ipx[c].size(), G.ipx[c].data(), G.ipx[c].size() * sizeof(double));
}
void orc_get_bgroup(void* h, int g, int* side, int* nelem, double* wts, double* normals, double* basis, double* basis_grad) {
  Oracle* o = (Oracle*)h; const auto& B = o->bgroups[g];
  if (side) *side = B.side; if (nelem) *nelem = B.g.numElem;
  if (wts) memcpy(wts, B.g.wts.data(), B.g.wts.size() * sizeof(double));
  if (normals) for (int c = 0; c < B.g.dim; ++c) memcpy(normals + c * B.g.nrm[c].size(), B.g.nrm[c].data(), B.g.nrm[c].size() * sizeof(double));
  if (basis) memcpy(basis, B.g.basis.data(), B.g.basis.size() * sizeof(double));
  if (basis_grad) memcpy(basis_grad, B.g.basis_grad.data(), B.g.basis_grad.size() * sizeof(double));
}

void orc_set_time(void* h, int nstages, const double* A, const double* b, int nbdf, const double* bdf,
                  double dt, double time, int stage) {
  Oracle* o = (Oracle*)h;
  o->ti.nstages = nstages; o->ti.A.assign(A, A + nstages * nstages); o->ti.b.assign(b, b + nstages);
  o->ti.bdf.assign(bdf, bdf + nbdf); o->ti.nsteps = nbdf - 1; o->ti.dt = dt; o->ti.time = time; o->ti.stage = stage;
}

// res / jac are accumulated into (caller zeroes).  returns 0 or -1.
int orc_assemble(void* h, const double* u, const double* u_prev, const double* u_stage, int compute_jacobian,
                 double* res, double* jac, int nthreads) {
  try { ((Oracle*)h)->assemble(u, u_prev, u_stage, compute_jacobian != 0, res, jac, nthreads); return 0; }
  catch (std::exception& e) { g_err = e.what(); return -1; }
}
// L2-projection system of the strong Dirichlet data (setDirichlet): rhs (nlocal) and mass (nnz) are accumulated into
int orc_assemble_dirichlet(void* h, double time, double* rhs, double* mass) {
  try { ((Oracle*)h)->assembleDirichlet(time, rhs, mass); return 0; }
  catch (std::exception& e) { g_err = e.what(); return -1; }
}
void orc_last_timing(void* h, double* out4) { memcpy(out4, ((Oracle*)h)->last_seconds, 4 * sizeof(double)); }

// evaluate a user expression at the volume ips of group g (double arithmetic): out (numElem, nq)
int orc_eval_expression(void* h, const char* expr, int g, double time, double* out) {
  try {
    Oracle* o = (Oracle*)h; const GroupData& G = o->groups[g];
    Workset<double> wk; wk.setup(o->cfg.dim, o->varlist, o->ref.nb, o->dofs.offsets, G.numElem, o->ref.nq, o->ref.nqs, false);
    wk.time = time; wk.attach(&G, false);
    FunctionManager<double> fm(G.numElem, o->ref.nq, o->ref.nqs); fm.wkset = &wk;
    fm.parameters = o->cfg.parameters;
    for (auto& f : o->cfg.functions) fm.addFunction(f.first, f.second, "ip");
    fm.addFunction("__expr__", expr, "ip");
    std::vector<double> zero((size_t)G.numElem * o->dofs.ndof_elem, 0.0);
    wk.computeSolnSteadySeeded(zero, 0);
    auto v = fm.evaluate("__expr__", "ip");
    for (int e = 0; e < G.numElem; ++e) for (int q = 0; q < o->ref.nq; ++q) out[(size_t)e * o->ref.nq + q] = v(e, q);
    return 0;
  } catch (std::exception& e) { g_err = e.what(); return -1; }
}

// L2 error of variable `var` against a "true solution" expression: sqrt( sum_e sum_q w (u_h - u_true)^2 )
// (postprocess "compute errors", /root/reference/src/managers/postprocessManager.cpp computeError)
int orc_l2_error(void* h, const double* u, int var, const char* true_expr, double time, double* out) {
  try {
    Oracle* o = (Oracle*)h; double err = 0.0;
    const int nb = o->ref.nb, nq = o->ref.nq, N = o->dofs.ndof_elem;
    for (size_t gi = 0; gi < o->groups.size(); ++gi) {
      const GroupData& G = o->groups[gi];
      std::vector<double> tv((size_t)G.numElem * nq);
      if (orc_eval_expression(h, true_expr, (int)gi, time, tv.data()) != 0) return -1;
      for (int le = 0; le < G.numElem; ++le) {
        const int* l = &o->dofs.lids[(size_t)G.elems[le] * N];
        for (int q = 0; q < nq; ++q) {
          double uh = 0.0;
          for (int k = 0; k < nb; ++k) uh += u[l[o->dofs.offsets[var * nb + k]]] * G.B(le, k, q);
          const double d = uh - tv[(size_t)le * nq + q];
          err += d * d * G.wts[(size_t)le * nq + q];
        }
      }
    }
    *out = std::sqrt(err); return 0;
  } catch (std::exception& e) { g_err = e.what(); return -1; }
}

}  // extern "C"
